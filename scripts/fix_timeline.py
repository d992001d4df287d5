"""Timeline of the per-image fix-ups inside one k_assign launch (library built with -DYS_FIXTIME, which makes thread 0
of every fix-up print its phase stamps).  Reads the prints of the LAST launch from stdin and summarises them.
    YS_NVCC_EXTRA=-DYS_FIXTIME python yolo_scratch_b200/build.py --force
    python scripts/stage_times.py --imgsz 1280 --fixed-gt 300 --reps 1 | python scripts/fix_timeline.py"""
import re
import sys

runs, cur = [], None
for line in sys.stdin:
    if line.startswith("START"):
        cur = {"t": int(line.split("=")[1]), "fix": []}
        runs.append(cur)
    elif line.startswith("FIX") and cur is not None:
        cur["fix"].append({k: int(v) for k, v in re.findall(r"(\w+)=(\d+)", line)})
runs = [r for r in runs if r["fix"]]
if not runs:
    sys.exit("no FIX lines (library not built with -DYS_FIXTIME?)")
r = runs[-1]
t0 = r["t"]
fx = sorted(r["fix"], key=lambda f: f["t0"])
print(f"{len(runs)} launches; last one: {len(fx)} fix-ups")
print(" img  y    n needy uniq  start_us   A   p1   Bp   l2  flag  p3  total_us  end_us")
for f in fx:
    tot = f["A"] + f["p1"] + f["Bp"] + f["l2"] + f["fl"] + f["p3"]
    print(f"{f['b']:4d} {f['y']:3d} {f['n']:4d} {f['needy']:5d} {f['uniq']:4d} {(f['t0'] - t0) / 1e3:9.1f} "
          f"{f['A'] / 1e3:4.1f} {f['p1'] / 1e3:4.1f} {f['Bp'] / 1e3:4.1f} {f['l2'] / 1e3:4.1f} {f['fl'] / 1e3:5.1f} {f['p3'] / 1e3:4.1f} "
          f"{tot / 1e3:8.1f} {(f['t0'] - t0 + tot) / 1e3:8.1f}")
