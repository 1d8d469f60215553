#!/usr/bin/env python
"""The floor under bench.py's `e2e`: move one step's bytes (675.8 MB host->device, 899.8 MB
device->host, pinned host memory) with plain async copies and no kernel at all -- each direction
alone, then both at once on two streams, whole and in 32 MB chunks like kv_b64enc_batch does."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

H2D, D2H = 675_755_289, 899_819_696
h_in = torch.empty(H2D, dtype=torch.uint8).pin_memory()
h_out = torch.empty(D2H, dtype=torch.uint8).pin_memory()
d_in = torch.empty(H2D, dtype=torch.uint8, device="cuda")
d_out = torch.zeros(D2H, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best


def h2d():
    with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)


def both():
    h2d(); d2h()


def both_chunked(chunk=32 << 20):
    n = max(H2D, D2H)
    for o in range(0, n, chunk):
        if o < H2D:
            with torch.cuda.stream(s1): d_in[o:o + chunk].copy_(h_in[o:o + chunk], non_blocking=True)
        if o < D2H:
            with torch.cuda.stream(s2): h_out[o:o + chunk].copy_(d_out[o:o + chunk], non_blocking=True)


t_h, t_d, t_b, t_c = timed(h2d), timed(d2h), timed(both), timed(both_chunked)
print(json.dumps({
    "h2d_alone_GBps": H2D / t_h / 1e9, "d2h_alone_GBps": D2H / t_d / 1e9,
    "both_ms": t_b * 1e3, "both_combined_GBps": (H2D + D2H) / t_b / 1e9,
    "both_chunked_32MB_ms": t_c * 1e3,
    "installs_per_s_if_copies_only": 49152 / t_b,
}))


# Does the host address' alignment matter?  (kv_b64enc_batch copies packed units: a chunk's
# host addresses are 16-byte aligned at best.)
def both_chunked_skewed(skew_in, skew_out, chunk=32 << 20):
    n = max(H2D, D2H) - 4096
    for o in range(0, n, chunk):
        if o < H2D - 4096:
            m = min(chunk, H2D - 4096 - o)
            with torch.cuda.stream(s1): d_in[o + skew_in:o + skew_in + m].copy_(h_in[o + skew_in:o + skew_in + m], non_blocking=True)
        if o < D2H - 4096:
            m = min(chunk, D2H - 4096 - o)
            with torch.cuda.stream(s2): h_out[o + skew_out:o + skew_out + m].copy_(d_out[o + skew_out:o + skew_out + m], non_blocking=True)


for sk in ((0, 0), (16, 16), (1, 16), (64, 64), (256, 256), (1, 0), (0, 16)):
    t = timed(lambda: both_chunked_skewed(*sk))
    print(json.dumps({"skew_in_out": sk, "both_chunked_ms": t * 1e3}))
