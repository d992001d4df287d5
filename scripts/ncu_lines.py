#!/usr/bin/env python
"""Per-source-line instruction counts and stall samples of an .ncu-rep (ncu --set full --import-source on; -lineinfo).
    python scripts/ncu_lines.py rep [kernel-regex] [top] [samples]"""
import csv, re, subprocess, sys
from collections import defaultdict

rep = sys.argv[1]
pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
by = 1 if len(sys.argv) > 4 and sys.argv[4] == "samples" else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
sections, cur = [], None
i = 0
while i < len(rows):
    r = rows[i]
    if r and r[0] == "File Path":
        cur = {"file": r[1], "func": rows[i + 1][1], "hdr": rows[i + 2], "lines": []}
        sections.append(cur)
        i += 3
        continue
    if cur is not None and r:
        cur["lines"].append(r)
    i += 1
def num(x):
    try:
        return int(x)
    except ValueError:
        return 0


seen = set()
kernels = defaultdict(lambda: defaultdict(lambda: [0, 0, ""]))
for s in sections:
    if pat and not pat.search(s["func"]):
        continue
    key = (s["func"], s["file"])
    if key in seen:      # same kernel captured twice: first instance only
        continue
    seen.add(key)
    h = s["hdr"]
    ci, cs = h.index("Instructions Executed"), h.index("# Samples")
    fn = s["file"].split("/")[-1]
    for r in s["lines"]:
        if r[0]:      # a source-line row (aggregated over its SASS)
            ln = num(r[0])
            if ln <= 0:
                continue
            e = kernels[s["func"]][(fn, ln)]
            e[0] += num(r[ci]); e[1] += num(r[cs]); e[2] = r[1]
for func, per in kernels.items():
    tot_i = sum(v[0] for v in per.values()) or 1
    tot_s = sum(v[1] for v in per.values()) or 1
    print(f"\n## {func[:80]}  inst={tot_i} samples={tot_s}")
    for (fn, ln), v in sorted(per.items(), key=lambda kv: -kv[1][by])[:top]:
        print(f"{fn[:22]:22s}{ln:5d} inst {100 * v[0] / tot_i:5.1f}%  smp {100 * v[1] / tot_s:5.1f}%  {v[2].strip()[:100]}")
