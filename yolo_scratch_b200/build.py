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
    "--shared", "-Xcompiler", "-fPIC",
    # host side of the bit-exact helpers must not contract a*b+c either
    "-Xcompiler", "-ffp-contract=off",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build() -> bool:
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = glob.glob(os.path.join(CSRC, "*")) + glob.glob(os.path.join(HERE, "..", "include", "*.h")) + [__file__]
    return any(os.path.getmtime(p) > t for p in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return SO
    extra = os.environ.get("YS_NVCC_EXTRA", "").split()      # tuning experiments only (scripts/gpu_tune.sh)
    cmd = [_nvcc()] + NVCC_FLAGS + extra + ["-o", SO] + [os.path.join(CSRC, s) for s in SOURCES]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    log = proc.stdout + proc.stderr
    with open(os.path.join(HERE, "build.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + log)
    if verbose or proc.returncode != 0:
        print(log)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed ({proc.returncode}); see {os.path.join(HERE, 'build.log')}")
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
