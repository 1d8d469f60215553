#!/usr/bin/env python
"""The copy-only floor under bench.py's `e2e` when N ranks share one host: every rank moves one
step's bytes (675.8 MB up, 899.8 MB down, pinned host memory, two streams, no kernel) at the same
time.  Run under torchrun (`python -m torch.distributed.run --nproc-per-node N tools/pcie_floor_multi.py`)
or alone.  Rank 0 prints one JSON line: the slowest rank's time for both directions at once, from
ordinary pinned memory and with the upload source in write-combined pinned memory."""
import ctypes, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

H2D, D2H = 675_755_289, 899_819_696
h_in = torch.empty(H2D, dtype=torch.uint8).pin_memory()
h_out = torch.empty(D2H, dtype=torch.uint8).pin_memory()
d_in = torch.empty(H2D, dtype=torch.uint8, device="cuda")
d_out = torch.zeros(D2H, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
rt = ctypes.CDLL("libcudart.so.12")
wc = ctypes.c_void_p()
WRITE_COMBINED = 0x04
assert rt.cudaHostAlloc(ctypes.byref(wc), ctypes.c_size_t(H2D), WRITE_COMBINED) == 0
ctypes.memset(wc, 0x41, H2D)
rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]


def sync_all():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def timed(fn, reps=5):
    fn(); sync_all()
    best = 1e9
    for _ in range(reps):
        sync_all()
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        best = min(best, float(dt.item()))
    return best


def up():
    with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)


def up_wc():
    assert rt.cudaMemcpyAsync(d_in.data_ptr(), wc, H2D, 1, ctypes.c_void_p(s1.cuda_stream)) == 0


def down():
    with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)


res = {"n_ranks": world, "bytes_up": H2D, "bytes_down": D2H}
for name, fn in (("up_ms", up), ("up_write_combined_ms", up_wc), ("down_ms", down),
                 ("both_ms", lambda: (up(), down())), ("both_write_combined_up_ms", lambda: (up_wc(), down()))):
    res[name] = timed(fn) * 1e3
res["aggregate_GBps_both"] = world * (H2D + D2H) / res["both_ms"] / 1e6
res["installs_per_s_if_copies_only"] = world * 49152 / (res["both_ms"] * 1e-3)
if rank == 0:
    print(json.dumps(res))
if world > 1:
    dist.destroy_process_group()
