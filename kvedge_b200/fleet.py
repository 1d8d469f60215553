"""Sharding a fleet of installs across GPUs.

One install is one kilobyte-scale unit of work and installs never depend on each other
(the reference's only multiplicity is re-installing the chart with another nameOverride,
README.md:72), so the fleet is simply partitioned by install index: no data-path collective,
one process per GPU.  torch.distributed is used for two things only: agreeing on the slowest
rank's time and gathering a few bytes of per-rank results to rank 0.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple


def shard_range(n_installs: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced partition: ranks differ by at most one install."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    base, extra = divmod(n_installs, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def max_over_ranks(seconds: float, device=None) -> float:
    """The step time of the whole job = the slowest rank's (all_reduce MAX)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return seconds
    t = torch.tensor([seconds], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_digests(local: Sequence[bytes]) -> List[List[bytes]]:
    """Per-install digests of every rank, in rank order, on every rank (tiny payloads)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [list(local)]
    out: List[List[bytes]] = [None] * dist.get_world_size()  # type: ignore[list-item]
    dist.all_gather_object(out, list(local))
    return out


def render_fleet_sharded(chart_dir: str, releases: Sequence[Tuple[str, dict]], rank: int, world: int,
                         render=None) -> Tuple[Tuple[int, int], List[bytes]]:
    """Render this rank's shard of ``releases`` (``kvedge_b200.helm.Chart.template_fleet`` on the
    rank's GPU unless ``render`` is given); returns ((lo, hi), manifests)."""
    lo, hi = shard_range(len(releases), rank, world)
    mine = list(releases[lo:hi])
    if render is None:
        from . import helm
        render = helm.Chart(chart_dir).template_fleet
    return (lo, hi), (render(mine) if mine else [])
