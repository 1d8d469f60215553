#!/usr/bin/env python
"""Kernel timing of the decode path (plan + b64dec_stream + b64dec_slow) on the encoded bench fleet."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kvedge_b200 import _lib as _kvlib
_kvlib.use_tools_build()   # sweep shapes and KVEDGE_* knobs live in libkvedge_b200_tools.so only
import numpy as np, torch
import bench
from kvedge_b200 import b64

installs = int(os.environ.get("INSTALLS", bench.DEFAULT_INSTALLS))
blob, in_off = bench.build_fleet(installs, 0)
n = len(in_off) - 1
lens = np.diff(in_off)
if os.environ.get("ROUND12"):   # every encoding a multiple of 16 characters: all units start 16-byte aligned
    keep = lens // 12 * 12
    idx = np.repeat(in_off[:-1], keep) + (np.arange(int(keep.sum())) - np.repeat(np.cumsum(keep) - keep, keep))
    blob = blob[idx]
    lens = keep
    in_off = np.zeros(n + 1, dtype=np.int64); np.cumsum(lens, out=in_off[1:])
enc_len = (lens + 2) // 3 * 4
align = int(os.environ.get("ENC_ALIGN", "16"))
enc_off = np.zeros(n + 1, dtype=np.int64); np.cumsum((enc_len + align - 1) // align * align, out=enc_off[1:])
d_in = torch.from_numpy(blob).cuda()
d_in_off, d_enc_off = torch.from_numpy(in_off).cuda(), torch.from_numpy(enc_off).cuda()
d_enc = torch.zeros(int(enc_off[-1]), dtype=torch.uint8, device="cuda")
scratch = torch.empty(b64.scratch_bytes(n, int(enc_off[-1])), dtype=torch.uint8, device="cuda")
b64.b64enc_device(d_in, d_in_off, d_enc, d_enc_off, scratch)
# decode input: exact unit extents inside the padded layout
dec_in_off = np.stack([enc_off[:-1], enc_off[:-1] + enc_len], 1)
# the API wants n+1 monotone offsets; use padded starts and pass lengths via a compacted copy
starts = torch.from_numpy(enc_off[:-1]).cuda()
cmp_off = np.zeros(n + 1, dtype=np.int64); np.cumsum(enc_len, out=cmp_off[1:])
idx = torch.repeat_interleave(starts - torch.from_numpy(cmp_off[:-1]).cuda(), torch.from_numpy(enc_len).cuda()) + torch.arange(int(cmp_off[-1]), device="cuda")
d_cmp = d_enc[idx].contiguous()
d_cmp_off = torch.from_numpy(cmp_off).cuda()
out_off = b64.decode_offsets(cmp_off, align=16)
d_out_off = torch.from_numpy(out_off).cuda()
d_out = torch.zeros(int(out_off[-1]) + 16, dtype=torch.uint8, device="cuda")
d_len = torch.zeros(n, dtype=torch.int64, device="cuda"); d_st = torch.zeros(n, dtype=torch.int64, device="cuda")
b64.b64dec_device(d_cmp, d_cmp_off, d_out, d_out_off, d_len, d_st, scratch, total_in_bytes=int(cmp_off[-1]))
torch.cuda.synchronize()
ok = bool((d_st == -1).all()) and bool((d_len.cpu() == torch.from_numpy(lens)).all())
# spot check bytes
got = d_out.cpu().numpy()
for i in (0, 1, n // 3, n - 1):
    ok = ok and np.array_equal(got[int(out_off[i]):int(out_off[i]) + int(lens[i])], blob[int(in_off[i]):int(in_off[i + 1])])
best = 1e9
for rep in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        b64.b64dec_device(d_cmp, d_cmp_off, d_out, d_out_off, d_len, d_st, scratch, total_in_bytes=int(cmp_off[-1]))
    e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / 10)
algo = int(cmp_off[-1]) + int(lens.sum())
print(json.dumps({"decode_ms": best, "GBps": algo / best / 1e6, "frac_of_6538": algo / best / 1e6 / 6538, "exact": ok,
                  "variant": os.environ.get("KVEDGE_DEC_VARIANT", "0")}))
