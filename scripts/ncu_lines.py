#!/usr/bin/env python
"""Per-source-line roll-up of an ncu SASS source page: joins `ncu --page source --csv` (per-instruction samples and
executed counts) with `nvdisasm -g` line info of the same cubin, by instruction order.
usage: ncu_lines.py rep kernel_regex cubin mangled_substring [top]"""
import csv, re, subprocess, sys, collections
rep, kre, cubin, mangled = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kre}"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ia, ismp, isrc = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
data = []
for r in rows[2:]:
    if len(r) > 10 and r[0].startswith("0x"):
        data.append(r)
    elif data:
        break
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
lines, cur, infn = [], ("?", 0), False
for l in dis:
    if l.startswith(".text."):
        infn = mangled in l
        continue
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append(cur)
print("sass instrs: ncu", len(data), "nvdisasm", len(lines))
n = min(len(data), len(lines))
agg = collections.defaultdict(lambda: [0, 0])
for r, ln in zip(data[:n], lines[:n]):
    agg[ln][0] += int(r[ismp]); agg[ln][1] += int(r[ia])
ts = sum(v[0] for v in agg.values()) or 1; ti = sum(v[1] for v in agg.values()) or 1
print("total samples", ts, "total inst", ti)
for ln, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{ln[0]}:{ln[1]:<5d} samples {100*v[0]/ts:5.1f}%  inst {100*v[1]/ti:5.1f}%")
