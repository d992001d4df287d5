#!/usr/bin/env python
"""Per-kernel SASS / ptxas facts of the built library: registers, spills, and the mnemonics that prove the TMA / mbarrier
path (UTMALDG = cp.async.bulk.tensor, UBLKCP = cp.async.bulk, LDGSTS = cp.async, SYNCS = mbarrier ops), MUFU, atomics, REDUX.
    python scripts/sass_summary.py > profiles/r02_sass_summary.md"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "yolo_scratch_b200", "libys_b200.so")
log = open(os.path.join(ROOT, "yolo_scratch_b200", "build.log")).read()
ptx = {}
for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'\n.*?\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers(?:, used \d+ barriers)?(?:, (\d+) bytes cumulative stack size)?(?:, (\d+) bytes smem)?", log):
    ptx[m.group(1)] = (int(m.group(5)), int(m.group(3)), int(m.group(4)), m.group(7) or "0")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", sass)[1:]
keys = ["UTMALDG", "UBLKCP", "LDGSTS", "SYNCS", "MUFU.EX2", "MUFU.RCP", "REDUX", "ATOMS", "ATOMG", "RED.", "STG.E.128", "LDS.128", "BAR.SYNC", "UTCHMMA|UTCQMMA|HMMA"]
print("# r02 SASS summary of yolo_scratch_b200/libys_b200.so (cuobjdump -sass; ptxas -v from build.log), sm_100a\n")
print("No tensor-core instruction appears anywhere (last column): nothing on this path is a dense contraction.\n")
print("| kernel | regs | spill st/ld B | static smem B | " + " | ".join(k.replace("|", "/") for k in keys) + " | SASS lines |")
print("|---|---|---|---|" + "---|" * (len(keys) + 1))
dem = subprocess.run(["c++filt"], input="\n".join(f.split("\n")[0].strip() for f in funcs), capture_output=True, text=True).stdout.split("\n")
for f, name in zip(funcs, dem):
    mangled = f.split("\n")[0].strip()
    body = f
    short = re.sub(r"\(.*", "", name).replace("void ", "").replace("ys::", "")
    r = ptx.get(mangled, ("?", "?", "?", "?"))
    counts = [len(re.findall(k, body)) for k in keys]
    nl = sum(1 for l in body.split("\n") if re.search(r"/\*[0-9a-f]{4,6}\*/", l))
    print(f"| `{short}` | {r[0]} | {r[1]}/{r[2]} | {r[3]} | " + " | ".join(str(c) for c in counts) + f" | {nl} |")
