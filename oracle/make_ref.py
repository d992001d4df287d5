"""ORACLE (test infrastructure): recipe for oracle/_ref -- the UNMODIFIED reference package, copied from where it lies
under /root/reference into oracle/_ref/yolo_scratch_cuda/ so that it travels to the GPU box (oracle/_ref is git-ignored,
not gpurun-ignored; nothing of it enters the repository's history).  The reference is pure Python: there is nothing to
compile.  Run by __graft_entry__.build() whenever /root/reference is present; bench.py's `--impl reference` arm and
`cpu_baseline` leg import the copy (kind "reference") and fall back to oracle/ref_port.py (kind "port") without it.

    python oracle/make_ref.py
"""
from __future__ import annotations

import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/yolo_scratch_cuda"
DST = os.path.join(HERE, "_ref", "yolo_scratch_cuda")


def build(verbose: bool = False) -> str | None:
    """Copy *.py / *.yaml of the reference package.  Returns the destination, or None when the reference is absent."""
    if not os.path.isdir(SRC):
        return DST if os.path.isdir(DST) else None
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    n = 0
    for root, dirs, files in os.walk(SRC):
        dirs[:] = [d for d in dirs if d not in ("__pycache__", ".git")]
        for f in files:
            if f.endswith((".py", ".yaml", ".yml")):
                rel = os.path.relpath(os.path.join(root, f), SRC)
                os.makedirs(os.path.dirname(os.path.join(DST, rel)), exist_ok=True)
                shutil.copyfile(os.path.join(root, f), os.path.join(DST, rel))
                n += 1
    with open(os.path.join(HERE, "_ref", "SOURCE.txt"), "w") as fh:
        fh.write(f"unmodified copy of {SRC} ({n} files), made by oracle/make_ref.py\n")
    if verbose:
        print(f"oracle/_ref: {n} files from {SRC}")
    return DST


def path() -> str | None:
    return DST if os.path.isfile(os.path.join(DST, "losses", "yolo_v8_loss.py")) else None


if __name__ == "__main__":
    print(build(verbose=True))
