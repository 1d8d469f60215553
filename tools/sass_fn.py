#!/usr/bin/env python
"""Print (or histogram) the SASS of one kernel from a built .so:  tools/sass_fn.py <so> <substr> [--hist]"""
import re, subprocess, sys
so, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
blocks = re.split(r"\n\s*Function : ", txt)
for b in blocks[1:]:
    name = b.split("\n", 1)[0]
    if pat in name:
        lines = [l for l in b.split("\n") if re.match(r"\s+/\*[0-9a-f]{4}\*/", l)]
        if "--hist" in sys.argv:
            from collections import Counter
            c = Counter()
            for l in lines:
                m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", l)
                if m: c[m.group(1)] += 1
            print(name, len(lines)); print(sorted(c.items(), key=lambda kv: -kv[1]))
        else:
            print(name); print("\n".join(l.split("/*", 2)[0] + l.split("*/", 1)[1].split("/*")[0] for l in lines))
