#!/usr/bin/env python
"""Where the time of a fleet-scale `helm template` goes: kv_fleet_template renders N installs of
the test chart (every install with its own config.toml through a value blob), all Secret payloads
in one GPU batch.  Prints installs/s for the C call alone and for the Python wrapper around it."""
import ctypes as C
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kvedge_b200 import _lib as _kvlib
_kvlib.use_tools_build()   # sweep shapes and KVEDGE_* knobs live in libkvedge_b200_tools.so only
import numpy as np
from kvedge_b200 import helm, workload, _lib
from kvedge_b200._lib import KvOverrides

N = int(os.environ.get("INSTALLS", "4096"))
# the reference's own chart (deployment/helm, materialised from tests/golden); CHART=fixture picks the test chart
if os.environ.get("CHART") == "fixture":
    chart_dir, cfg_key, key_key = os.path.join(ROOT, "tests", "fixtures", "edgevm"), "runtimeConfig", "sshKey"
else:
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import refchart
    chart_dir, cfg_key, key_key = refchart.materialise(), "azIotEdgeConfig", "publicSshKey"
chart = helm.Chart(chart_dir)
rel = [(f"edge-{i}", {"blobs": {cfg_key: workload.config_toml(i)},
                      "set": [f"nameOverride=site-{i:05d},{key_key}={workload.ssh_key(i).decode()}"]})
       for i in range(N)]

t0 = time.perf_counter()
outs = chart.template_fleet(rel)
t_first = time.perf_counter() - t0
t0 = time.perf_counter()
outs = chart.template_fleet(rel)
t_py = time.perf_counter() - t0

# the C call alone, arguments prepared outside the timed region
ovs = [helm._Overrides(**kw) for _, kw in rel]
arr = (KvOverrides * N)(*[o.struct for o in ovs])
names = (C.c_char_p * N)(*[name.encode() for name, _ in rel])
out = C.c_void_p()
off = np.zeros(N + 1, dtype=np.int64)
L = _lib.lib()
best = 1e9
for _ in range(int(os.environ.get("REPS", "5"))):
    t0 = time.perf_counter()
    rc = L.kv_fleet_template(chart._h, N, C.cast(names, C.c_void_p), b"default", C.cast(arr, C.c_void_p), C.byref(out), off.ctypes.data)
    dt = time.perf_counter() - t0
    assert rc == 0, rc
    L.kv_free(out)
    best = min(best, dt)
print(json.dumps({"installs": N, "manifest_bytes_per_install": int(off[-1]) // N,
                  "c_call_ms": best * 1e3, "c_call_installs_per_s": N / best,
                  "python_ms": t_py * 1e3, "python_installs_per_s": N / t_py, "first_call_ms": t_first * 1e3,
                  "host_threads": os.cpu_count()}))
