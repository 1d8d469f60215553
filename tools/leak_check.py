import os, sys, resource, time
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
import numpy as np, torch
from kvedge_b200 import helm, workload, b64
chart = helm.Chart(os.path.join(sys.path[0], "tests", "fixtures", "edgevm"))
rel = [(f"edge-{i}", {"blobs": {"runtimeConfig": workload.config_toml(i)}, "set": [f"sshKey={workload.ssh_key(i).decode()}"]}) for i in range(512)]
units = [np.random.default_rng(i).bytes(5000 + 37 * i) for i in range(300)]
def rss(): return resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1024
def gpu(): return torch.cuda.mem_get_info()[0] / 2**20
for it in range(6):
    for _ in range(40):
        chart.template_fleet(rel)
        b64.b64dec_many(b64.b64enc_many(units))
    print(f"round {it}: maxrss {rss():.0f} MiB, gpu free {gpu():.0f} MiB", flush=True)
