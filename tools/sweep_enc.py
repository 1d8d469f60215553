#!/usr/bin/env python
"""Kernel-only timing of the encode kernel on the bench fleet for one KVEDGE_ENC_VARIANT
(run once per variant: the variant is latched at first launch).  Prints one JSON line."""
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
out_off = b64.encode_offsets(in_off, 16)
d_in = torch.from_numpy(blob).cuda()
d_in_off, d_out_off = torch.from_numpy(in_off).cuda(), torch.from_numpy(out_off).cuda()
d_out = torch.zeros(int(out_off[-1]), dtype=torch.uint8, device="cuda")
scratch = torch.empty(b64.scratch_bytes(n, len(blob)), dtype=torch.uint8, device="cuda")
b64.b64enc_device(d_in, d_in_off, d_out, d_out_off, scratch)
torch.cuda.synchronize()
best = 1e9
for rep in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        b64.b64enc_run_device(d_in, d_in_off, d_out, scratch, len(blob))
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / 20)
from oracle import b64 as ob64
chk = 3000
ref = ob64.encode_batch(blob, in_off[:chk + 1], out_off[:chk + 1])
got = d_out[:int(out_off[chk])].cpu().numpy()
lens = np.diff(in_off[:chk + 1])
ok = all(np.array_equal(got[int(out_off[i]):int(out_off[i]) + (int(lens[i]) + 2) // 3 * 4],
                        ref[int(out_off[i]):int(out_off[i]) + (int(lens[i]) + 2) // 3 * 4]) for i in range(chk))
algo = bench.algorithmic_bytes(in_off)
print(json.dumps({"variant": os.environ.get("KVEDGE_ENC_VARIANT", "0"), "kernel_ms": best,
                  "GBps": algo / best / 1e6, "frac_of_6538": algo / best / 1e6 / 6538, "byte_exact": bool(ok)}))
