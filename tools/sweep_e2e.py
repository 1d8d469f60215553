#!/usr/bin/env python
"""e2e (host buffers through kv_b64enc_batch) on the bench fleet for one KVEDGE_CHUNK_MB."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from kvedge_b200 import b64, _lib
blob, in_off = bench.build_fleet(bench.DEFAULT_INSTALLS, 0)
n = len(in_off) - 1
out_off = b64.encode_offsets(in_off, 16)
h_in = torch.from_numpy(blob).pin_memory(); h_out = torch.empty(int(out_off[-1]), dtype=torch.uint8).pin_memory()
L = _lib.lib()
for _ in range(2): _lib.check(L.kv_b64enc_batch(h_in.data_ptr(), in_off.ctypes.data, n, h_out.data_ptr(), out_off.ctypes.data))
t0 = time.perf_counter()
for _ in range(5): _lib.check(L.kv_b64enc_batch(h_in.data_ptr(), in_off.ctypes.data, n, h_out.data_ptr(), out_off.ctypes.data))
dt = (time.perf_counter() - t0) / 5
print(json.dumps({"chunk_mb": os.environ.get("KVEDGE_CHUNK_MB", "32"), "ms": dt * 1e3, "installs_per_s": bench.DEFAULT_INSTALLS / dt,
                  "pcie_GBps": (len(blob) + int(out_off[-1])) / dt / 1e9}))
