"""Build recipe for libys_b200.so: nvcc, sm_100a only, in-tree (the .so travels to the GPU box).

    python -m yolo_scratch_b200.build [--force] [--verbose]
"""
from __future__ import annotations

import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libys_b200.so")
SOURCES = ["ys_api.cu", "ys_loss.cu", "ys_nms.cu", "ys_decode.cu", "ys_metrics.cu", "ys_bboxloss.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    # host side of the bit-exact helpers must not contract a*b+c either
    "-Xcompiler", "-ffp-contract=off",
    "-Xptxas", "-v",
]
OBJ_DIR = os.path.join(HERE, "build")


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _deps():
    return glob.glob(os.path.join(CSRC, "*")) + glob.glob(os.path.join(HERE, "..", "include", "*.h")) + [__file__]


def needs_build() -> bool:
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    return any(os.path.getmtime(p) > t for p in _deps())


def build(force: bool = False, verbose: bool = False) -> str:
    """One nvcc -c per source (in parallel), then one link.  Objects are reused when neither their source nor any
    header changed and the extra flags are the same."""
    if not force and not needs_build():
        return SO
    from concurrent.futures import ThreadPoolExecutor
    extra = os.environ.get("YS_NVCC_EXTRA", "").split()      # tuning experiments only (scripts/gpu_tune.sh)
    os.makedirs(OBJ_DIR, exist_ok=True)
    stamp = os.path.join(OBJ_DIR, "flags.txt")
    flags_now = " ".join(NVCC_FLAGS + extra)
    flags_same = os.path.exists(stamp) and open(stamp).read() == flags_now
    headers = [p for p in _deps() if not p.endswith(".cu")]
    newest_header = max(os.path.getmtime(p) for p in headers)

    def compile_one(src):
        obj = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
        path = os.path.join(CSRC, src)
        if (not force and flags_same and os.path.exists(obj) and os.path.getmtime(obj) > os.path.getmtime(path)
                and os.path.getmtime(obj) > newest_header):
            return obj, 0, f"[cached] {src}\n"
        cmd = [_nvcc()] + NVCC_FLAGS + extra + ["-c", "-o", obj, path]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        return obj, proc.returncode, " ".join(cmd) + "\n" + proc.stdout + proc.stderr

    with ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 1)) as ex:
        results = list(ex.map(compile_one, SOURCES))
    log = "".join(r[2] for r in results)
    rc = max(r[1] for r in results)
    if rc == 0:
        cmd = [_nvcc(), "--shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", SO] + [r[0] for r in results]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        log += " ".join(cmd) + "\n" + proc.stdout + proc.stderr
        rc = proc.returncode
    with open(os.path.join(HERE, "build.log"), "w") as f:
        f.write(log)
    if verbose or rc != 0:
        print(log)
    if rc != 0:
        raise RuntimeError(f"nvcc failed ({rc}); see {os.path.join(HERE, 'build.log')}")
    with open(stamp, "w") as f:
        f.write(flags_now)
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
