#!/usr/bin/env python
"""Top stall-sample instructions of a .ncu-rep (source page, SASS view): tools/ncu_hot.py rep [N]"""
import csv, subprocess, sys
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(raw.splitlines()))
h = r[1]; c = {n: i for i, n in enumerate(h)}
rows = r[2:]
tot = sum(int(x[c["# Samples"]]) for x in rows)
stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
print("total samples", tot)
agg = {k: sum(int(x[c[k]]) for x in rows) for k in stalls}
print({k: round(100 * v / tot, 1) for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
idx = sorted(range(len(rows)), key=lambda i: -int(rows[i][c["# Samples"]]))[:N]
for i in sorted(idx):
    x = rows[i]
    top = sorted(((int(x[c[k]]), k) for k in stalls), reverse=True)[:2]
    print(f"{i:5d} {int(x[c['# Samples']]):6d} {100*int(x[c['# Samples']])/tot:5.1f}% exec={x[c['Instructions Executed']]:>9s} {x[c['Source']].strip()[:70]:70s} {top}")
