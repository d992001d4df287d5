"""Phase times of a sample of k_assign blocks (library built with -DYS_SCANTIME); reads the prints of the last launch.
    YS_NVCC_EXTRA=-DYS_SCANTIME python yolo_scratch_b200/build.py --force
    python scripts/stage_times.py --reps 1 | python scripts/scan_timeline.py"""
import re
import sys

rows = []
for line in sys.stdin:
    if line.startswith("BLK"):
        rows.append({k: int(v) for k, v in re.findall(r"(\w+)=(\d+)", line)})
if not rows:
    sys.exit("no BLK lines")
# launches are separated by a big gap in t0
rows.sort(key=lambda r: r["t0"])
cut = 0
for i in range(1, len(rows)):
    if rows[i]["t0"] - rows[i - 1]["t0"] > 200_000:
        cut = i
rows = rows[cut:]
t0 = rows[0]["t0"]
keys = ["sync", "tma", "dec", "scan", "res", "fence", "arr"]
print(f"{len(rows)} sampled blocks of the last launch")
print("  x   y  n  start_us " + " ".join(f"{k:>6s}" for k in keys) + "  total")
for r in rows:
    tot = sum(r[k] for k in keys)
    print(f"{r['x']:3d} {r['y']:3d} {r['n']:2d} {(r['t0'] - t0) / 1e3:9.1f} " + " ".join(f"{r[k] / 1e3:6.2f}" for k in keys) + f" {tot / 1e3:6.2f}")
print("mean (us):  " + " ".join(f"{k}={sum(r[k] for r in rows) / len(rows) / 1e3:.2f}" for k in keys)
      + f" total={sum(sum(r[k] for k in keys) for r in rows) / len(rows) / 1e3:.2f}")
