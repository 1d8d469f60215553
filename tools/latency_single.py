#!/usr/bin/env python
"""The reference's REAL size: one install.  Latency of one b64enc (2 KB cloud-init text,
11 KiB config.toml) and of one full `helm template` through the GPU path, next to the CPU
oracle.  Prints one JSON line; meant for profiles/, not for bench.py."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kvedge_b200 import b64, helm, workload
from oracle import b64 as ob64, helmref

def best(f, n=200):
    f(); f()
    t = []
    for _ in range(n):
        t0 = time.perf_counter(); f(); t.append(time.perf_counter() - t0)
    t.sort()
    return t[len(t) // 2] * 1e6

cfg = workload.config_toml(0)
ci = b"#cloud-config\n" + bytes(2392)
chart = os.path.join(ROOT, "tests", "fixtures", "edgevm")
c = helm.Chart(chart)
out = {
    "gpu_b64enc_2KB_us": best(lambda: b64.b64enc(ci)),
    "gpu_b64enc_11KB_us": best(lambda: b64.b64enc(cfg)),
    "cpu_oracle_b64enc_2KB_us": best(lambda: ob64.encode(ci)),
    "cpu_oracle_b64enc_11KB_us": best(lambda: ob64.encode(cfg)),
    "gpu_helm_template_fixture_us": best(lambda: c.template("r", blobs={"runtimeConfig": cfg}), 100),
    "cpu_oracle_helm_template_fixture_us": best(lambda: helmref.helm_template(chart, "r"), 30),
}
# fleet render through the product API: n installs planned on all host threads, one GPU batch
n = 2048
rels = [(f"edge-{i}", {"set": [f"nameOverride=site-{i},sshKey={workload.ssh_key(i % 64).decode()}"],
                       "blobs": {"runtimeConfig": workload.config_toml(i % 64)}}) for i in range(n)]
c.template_fleet(rels[:64])
t0 = time.perf_counter(); res = c.template_fleet(rels); dt = time.perf_counter() - t0
out["fleet_template_installs"] = n
out["fleet_template_installs_per_s"] = n / dt
out["fleet_template_bytes_out"] = sum(len(r) for r in res)
out["host_threads"] = os.cpu_count()
print(json.dumps(out))
