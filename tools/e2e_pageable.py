import os, sys, time, json
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
import numpy as np, torch
import bench
from kvedge_b200 import b64, _lib
blob, in_off = bench.build_fleet(bench.DEFAULT_INSTALLS, 0)
n = len(in_off) - 1
out_off = b64.encode_offsets(in_off, 16)
h_out = np.empty(int(out_off[-1]), dtype=np.uint8)
L = _lib.lib()
for _ in range(2): _lib.check(L.kv_b64enc_batch(blob.ctypes.data, in_off.ctypes.data, n, h_out.ctypes.data, out_off.ctypes.data))
t0 = time.perf_counter()
for _ in range(5): _lib.check(L.kv_b64enc_batch(blob.ctypes.data, in_off.ctypes.data, n, h_out.ctypes.data, out_off.ctypes.data))
dt = (time.perf_counter() - t0) / 5
print(json.dumps({"pageable_ms": dt * 1e3, "installs_per_s": bench.DEFAULT_INSTALLS / dt, "GBps": (len(blob) + int(out_off[-1])) / dt / 1e9}))
