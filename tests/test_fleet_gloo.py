"""N>1 host logic on CPU: two gloo ranks each take their shard of a fleet, render it, and the
union must equal the single-process render.  The GPU encode is not available here, so the
ranks plan with the C++ engine and fill the pending sites with the CPU checker -- what is under
test is the partition, the gather and the max-over-ranks timing, not the kernel."""
import hashlib
import os
import socket
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIXTURE = os.path.join(ROOT, "tests", "fixtures", "edgevm")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _releases(n):
    from kvedge_b200 import workload
    return [(f"edge-{i:03d}", {"set": [f"nameOverride=site-{i:03d}"],
                               "blobs": {"runtimeConfig": workload.config_toml(i)[: 500 + 37 * i]}}) for i in range(n)]


def _render_with_checker(chart_dir):
    from kvedge_b200 import helm
    from oracle import b64 as ob64
    chart = helm.Chart(chart_dir)

    def render(rels):
        out = []
        for name, kw in rels:
            r = chart.plan(name, **kw)
            for i, p in enumerate(r.pending):
                r.fill(i, ob64.encode(p))
            out.append(r.manifest())
        return out
    return render


def _worker(rank, world, port, n, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from kvedge_b200 import fleet
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        (lo, hi), manifests = fleet.render_fleet_sharded(FIXTURE, _releases(n), rank, world,
                                                         render=_render_with_checker(FIXTURE))
        digests = [hashlib.sha256(m).digest() for m in manifests]
        gathered = fleet.gather_digests(digests)
        slowest = fleet.max_over_ranks(0.25 * (rank + 1))
        q.put((rank, lo, hi, [d for part in gathered for d in part], slowest))
    finally:
        dist.destroy_process_group()


def test_shard_range_covers_everything_once():
    from kvedge_b200 import fleet
    for n in (0, 1, 7, 64, 1001):
        for world in (1, 2, 3, 8):
            spans = [fleet.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        fleet.shard_range(4, 2, 2)


def test_two_gloo_ranks_render_the_same_fleet_as_one_process():
    import torch.multiprocessing as mp
    n, world = 13, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=180) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    single = [hashlib.sha256(m).digest() for m in _render_with_checker(FIXTURE)(_releases(n))]
    assert (results[0][1], results[0][2], results[1][1], results[1][2]) == (0, 7, 7, 13)
    for rank, lo, hi, all_digests, slowest in results:
        assert all_digests == single               # every rank sees the whole fleet's digests
        assert slowest == pytest.approx(0.5)       # max over ranks
