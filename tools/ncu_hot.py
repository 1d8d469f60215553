#!/usr/bin/env python
"""Top stall-sample instructions of a .ncu-rep (source page): tools/ncu_hot.py rep [N] [kernel-substring] [view]
The page holds one block per kernel and view (SASS first, then source lines); view = 0 (SASS) or 1."""
import csv, subprocess, sys
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
want = sys.argv[3] if len(sys.argv) > 3 else ""
view = int(sys.argv[4]) if len(sys.argv) > 4 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(raw.splitlines()):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "rows": []}
        blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(row)
seen = {}
for b in blocks:
    k = seen.get(b["name"], 0)
    seen[b["name"]] = k + 1
    if want not in b["name"] or k != view:
        continue
    h = b["rows"][0]; c = {n: i for i, n in enumerate(h)}
    rows = [x for x in b["rows"][1:] if len(x) == len(h)]
    tot = sum(int(x[c["# Samples"]]) for x in rows)
    stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
    print("==", b["name"][:90], "| total samples", tot, "| instructions", sum(int(x[c["Instructions Executed"]]) for x in rows))
    agg = {k: sum(int(x[c[k]]) for x in rows) for k in stalls}
    print({k: round(100 * v / tot, 1) for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
    idx = sorted(range(len(rows)), key=lambda i: -int(rows[i][c["# Samples"]]))[:N]
    for i in sorted(idx):
        x = rows[i]
        top = sorted(((int(x[c[k]]), k) for k in stalls), reverse=True)[:2]
        print(f"{i:5d} {int(x[c['# Samples']]):6d} {100*int(x[c['# Samples']])/tot:5.1f}% exec={x[c['Instructions Executed']]:>9s} {x[c['Source']].strip()[:70]:70s} {top}")
