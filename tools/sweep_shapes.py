#!/usr/bin/env python
"""Encode and decode rates (device-resident, plan + kernels) for unit-size mixes other than the
bench fleet, to see that the kernels are not tuned to one shape only.  One JSON line per shape."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kvedge_b200 import _lib as _kvlib
_kvlib.use_tools_build()   # sweep shapes and KVEDGE_* knobs live in libkvedge_b200_tools.so only
import numpy as np, torch
from kvedge_b200 import b64
from oracle import b64 as ob64

rng = np.random.default_rng(0)
SHAPES = {
    "cloudinit_only_2403B": np.full(280_000, 2403),
    "config_only_~11KB": rng.integers(9_000, 13_500, 60_000),
    "secret_ceiling_1MiB": np.full(640, 1_048_576 - 2_000),
    "tiny_300B": rng.integers(1, 600, 1_500_000),
    "ragged_1B..64KB": (rng.pareto(1.2, 120_000) * 600 + 1).clip(1, 65_536).astype(np.int64),
    "bench_mix": None,
}


def run(name, lens):
    if lens is None:
        import bench
        blob, in_off = bench.build_fleet(bench.DEFAULT_INSTALLS, 0)
        lens = np.diff(in_off)
    else:
        lens = lens.astype(np.int64)
        in_off = np.zeros(len(lens) + 1, dtype=np.int64); np.cumsum(lens, out=in_off[1:])
        blob = rng.integers(0, 256, int(in_off[-1]), dtype=np.uint8)
    n = len(lens)
    out_off = b64.encode_offsets(in_off, 16)
    d_in = torch.from_numpy(blob).cuda()
    d_in_off, d_out_off = torch.from_numpy(in_off).cuda(), torch.from_numpy(out_off).cuda()
    d_out = torch.zeros(int(out_off[-1]), dtype=torch.uint8, device="cuda")
    scratch = torch.empty(b64.scratch_bytes(n, max(int(in_off[-1]), int(out_off[-1]))), dtype=torch.uint8, device="cuda")

    def timed(fn):
        for _ in range(3): fn()
        best = 1e9
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5): fn()
            e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 5)
        return best

    ms_enc = timed(lambda: b64.b64enc_device(d_in, d_in_off, d_out, d_out_off, scratch))
    b64.b64enc_plan_device(d_in_off, d_out_off, scratch, int(in_off[-1]))
    ms_enc_kernel = timed(lambda: b64.b64enc_run_device(d_in, d_in_off, d_out, scratch, int(in_off[-1])))
    # spot check against the oracle
    got = d_out.cpu().numpy()
    ok = True
    for i in (0, 1, n // 2, n - 1):
        enc = ob64.encode(blob[int(in_off[i]):int(in_off[i + 1])].tobytes())
        ok = ok and got[int(out_off[i]):int(out_off[i]) + len(enc)].tobytes() == enc
    # decode the dense encodings back
    enc_len = (lens + 2) // 3 * 4
    dense_off = np.zeros(n + 1, dtype=np.int64); np.cumsum(enc_len, out=dense_off[1:])
    d_dense_off = torch.from_numpy(dense_off).cuda()
    d_dense = torch.zeros(int(dense_off[-1]), dtype=torch.uint8, device="cuda")
    b64.b64enc_device(d_in, d_in_off, d_dense, d_dense_off, scratch)
    dec_off = b64.decode_offsets(dense_off, align=16)
    d_dec_off = torch.from_numpy(dec_off).cuda()
    d_dec = torch.zeros(int(dec_off[-1]) + 16, dtype=torch.uint8, device="cuda")
    d_len = torch.zeros(n, dtype=torch.int64, device="cuda"); d_st = torch.zeros(n, dtype=torch.int64, device="cuda")
    ms_dec = timed(lambda: b64.b64dec_device(d_dense, d_dense_off, d_dec, d_dec_off, d_len, d_st, scratch, total_in_bytes=int(dense_off[-1])))
    ok = ok and bool((d_st == -1).all()) and bool((d_len.cpu() == torch.from_numpy(lens)).all())
    back = d_dec.cpu().numpy()
    for i in (0, 1, n // 2, n - 1):
        ok = ok and np.array_equal(back[int(dec_off[i]):int(dec_off[i]) + int(lens[i])], blob[int(in_off[i]):int(in_off[i + 1])])
    algo = int(in_off[-1]) + int(enc_len.sum())
    print(json.dumps({"shape": name, "units": n, "input_MB": round(int(in_off[-1]) / 1e6, 1),
                      "enc_ms": round(ms_enc, 4), "enc_kernel_ms": round(ms_enc_kernel, 4), "enc_GBps": round(algo / ms_enc / 1e6), "enc_frac_of_6538": round(algo / ms_enc / 1e6 / 6538, 3),
                      "dec_ms": round(ms_dec, 4), "dec_GBps": round(algo / ms_dec / 1e6), "dec_frac_of_6538": round(algo / ms_dec / 1e6 / 6538, 3),
                      "exact": ok}), flush=True)
    del d_in, d_out, d_dense, d_dec, scratch
    torch.cuda.empty_cache()


only = sys.argv[1:] or list(SHAPES)
for k in only:
    run(k, SHAPES[k])
