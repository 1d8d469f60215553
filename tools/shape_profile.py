#!/usr/bin/env python
"""A short, fixed launch sequence on one unit-size shape, for ncu: three encodes, then three decodes
of the same fleet (plan kernels in between).  `python tools/shape_profile.py cloudinit|bench|tiny`.
  ncu --set full -k regex:'b64enc_groups|b64dec_stream' -s 2 -c 2 ...   captures encode #3 and decode #1
Prints one JSON line with event timings (never quote a number taken under ncu)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if os.environ.get("KVEDGE_ENC_VARIANT") or os.environ.get("KVEDGE_DEC_VARIANT"):
    from kvedge_b200 import _lib as _kvlib
    _kvlib.use_tools_build()   # sweep shapes exist in libkvedge_b200_tools.so only
if os.environ.get("KVEDGE_LIB"):   # an experimental build of the library, by path
    from kvedge_b200 import _lib as _kvlib2
    _kvlib2._path = os.environ["KVEDGE_LIB"]
import numpy as np, torch
from kvedge_b200 import b64
from oracle import b64 as ob64

shape = sys.argv[1] if len(sys.argv) > 1 else "cloudinit"
rng = np.random.default_rng(0)
if shape == "bench":
    import bench
    blob, in_off = bench.build_fleet(bench.DEFAULT_INSTALLS, 0)
else:
    lens = {"cloudinit": np.full(196608, 2006), "tiny": rng.integers(1, 600, 1_000_000),
            "cloudinit_key": np.full(196608, 2403)}[shape].astype(np.int64)
    in_off = np.zeros(len(lens) + 1, dtype=np.int64); np.cumsum(lens, out=in_off[1:])
    blob = rng.integers(0, 256, int(in_off[-1]), dtype=np.uint8)
n = len(in_off) - 1
out_off = b64.encode_offsets(in_off, 16)
dense_off = b64.encode_offsets(in_off, 1)
d_in = torch.from_numpy(blob).cuda()
d_in_off, d_out_off, d_dense_off = (torch.from_numpy(x).cuda() for x in (in_off, out_off, dense_off))
d_out = torch.zeros(int(out_off[-1]), dtype=torch.uint8, device="cuda")
d_dense = torch.zeros(int(dense_off[-1]), dtype=torch.uint8, device="cuda")
scratch = torch.empty(b64.scratch_bytes(n, max(int(in_off[-1]), int(out_off[-1]))), dtype=torch.uint8, device="cuda")
b64.b64enc_device(d_in, d_in_off, d_dense, d_dense_off, scratch)       # encode #1 (dense layout, decode's input)
dec_off = b64.decode_offsets(dense_off, align=16)
d_dec_off = torch.from_numpy(dec_off).cuda()
d_dec = torch.zeros(int(dec_off[-1]) + 16, dtype=torch.uint8, device="cuda")
d_len, d_st = torch.zeros(n, dtype=torch.int64, device="cuda"), torch.zeros(n, dtype=torch.int64, device="cuda")


def timed(fn, reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

REPS = int(os.environ.get("REPS", "0"))
if REPS:   # timing mode (not for ncu): warm up, then time REPS launches of each
    timed(lambda: b64.b64enc_device(d_in, d_in_off, d_out, d_out_off, scratch), 3)
ms_enc = timed(lambda: b64.b64enc_device(d_in, d_in_off, d_out, d_out_off, scratch), REPS or 2)   # encodes #2, #3
ms_dec = timed(lambda: b64.b64dec_device(d_dense, d_dense_off, d_dec, d_dec_off, d_len, d_st, scratch,
                                         total_in_bytes=int(dense_off[-1])), REPS or 3)
ok = bool((d_st == -1).all())
got = d_out.cpu().numpy()
for i in (0, n // 2, n - 1):
    enc = ob64.encode(blob[int(in_off[i]):int(in_off[i + 1])].tobytes())
    ok = ok and got[int(out_off[i]):int(out_off[i]) + len(enc)].tobytes() == enc
algo = int(in_off[-1]) + int(dense_off[-1])
print(json.dumps({"shape": shape, "units": n, "enc_ms": ms_enc, "enc_frac_6538": algo / ms_enc / 1e6 / 6538,
                  "dec_ms": ms_dec, "dec_frac_6538": algo / ms_dec / 1e6 / 6538, "exact": ok,
                  "enc_variant": os.environ.get("KVEDGE_ENC_VARIANT", ""), "dec_variant": os.environ.get("KVEDGE_DEC_VARIANT", "")}))
