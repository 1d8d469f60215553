#!/usr/bin/env python
"""e2e (host buffers through kv_b64enc_batch) on the bench fleet for one chunk size / slot count:
CHUNK_MB=16 SLOTS=4 python tools/sweep_e2e.py (KVEDGE_TRACE=1 adds the per-chunk timeline)."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kvedge_b200 import _lib as _kvlib
_kvlib.use_tools_build()   # KVEDGE_TRACE / KVEDGE_E2E_DIAG live in libkvedge_b200_tools.so only
import numpy as np, torch
import bench
from kvedge_b200 import b64, _lib
blob, in_off = bench.build_fleet(bench.DEFAULT_INSTALLS, 0)
n = len(in_off) - 1
out_off = b64.encode_offsets(in_off, 16)
h_in = torch.from_numpy(blob).pin_memory(); h_out = torch.empty(int(out_off[-1]), dtype=torch.uint8).pin_memory()
L = _lib.lib()
_lib.check(L.kv_host_pipeline_config(int(os.environ.get("CHUNK_MB", "16")) << 20, int(os.environ.get("SLOTS", "4"))))
for _ in range(2): _lib.check(L.kv_b64enc_batch(h_in.data_ptr(), in_off.ctypes.data, n, h_out.data_ptr(), out_off.ctypes.data))
t0 = time.perf_counter()
for _ in range(5): _lib.check(L.kv_b64enc_batch(h_in.data_ptr(), in_off.ctypes.data, n, h_out.data_ptr(), out_off.ctypes.data))
dt = (time.perf_counter() - t0) / 5
print(json.dumps({"chunk_mb": os.environ.get("CHUNK_MB", "16"), "slots": os.environ.get("SLOTS", "4"), "ms": dt * 1e3, "installs_per_s": bench.DEFAULT_INSTALLS / dt,
                  "pcie_GBps": (len(blob) + int(out_off[-1])) / dt / 1e9}))
