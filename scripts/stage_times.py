#!/usr/bin/env python
"""Per-kernel CUDA-event times of the fused loss path at the bench workload (B=64, 640 px).  GPU box only.

    python scripts/stage_times.py [--reps 20] [--B 64] [--imgsz 640] [--max-gt 20]
Prints one JSON line: full step, each stage alone (state left by a full call), and k_assign without its fix-up
(YS_STAGE_DIAG_NOFIXUP) so the scan and the forced-assignment tail can be told apart.
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_scratch_b200 import _lib, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--B", type=int, default=64)
    ap.add_argument("--imgsz", type=int, default=640)
    ap.add_argument("--max-gt", type=int, default=20)
    ap.add_argument("--fixed-gt", type=int, default=0)
    ap.add_argument("--nc", type=int, default=80)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    lib = _lib.lib()
    B, nc = args.B, args.nc
    feats = [f.to(dev) for f in synth.make_head_levels(B, args.imgsz, nc, 16, seed=1000)]
    if args.fixed_gt:
        batch = synth.make_gt_batch(B, nc, args.fixed_gt, seed=1001, fixed_count=args.fixed_gt, dense=True)
    else:
        batch = synth.make_gt_batch(B, nc, args.max_gt, seed=1001)
    batch = {k: v.to(dev) for k, v in batch.items()}
    grads = [torch.empty_like(f) for f in feats]
    n = len(feats)
    A = sum(f.shape[2] * f.shape[3] for f in feats)
    n_gt = int(batch["cls"].shape[0])
    xp = (C.c_void_p * n)(*[f.data_ptr() for f in feats])
    gp = (C.c_void_p * n)(*[g.data_ptr() for g in grads])
    hs = (C.c_int * n)(*[f.shape[2] for f in feats])
    ws_ = (C.c_int * n)(*[f.shape[3] for f in feats])
    st = (C.c_float * n)(*[float(s) for s in synth.LEVEL_STRIDES])
    bi = batch["batch_idx"].contiguous()
    cl = batch["cls"].contiguous().view(-1)
    bb = batch["bboxes"].contiguous()
    out = torch.zeros(4, device=dev)
    wbytes = lib.ys_detection_loss_workspace_bytes(B, A, n_gt, nc)
    wsb = torch.zeros(wbytes + 4096, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)

    def call(stages=0, fwd_only=False):
        rc = lib.ys_detection_loss(n, xp, None if fwd_only else gp, hs, ws_, st, B, nc, 16, bi.data_ptr(), cl.data_ptr(), bb.data_ptr(),
                                   n_gt, float(args.imgsz), float(args.imgsz), 7.5, 0.5, 1.5, 0.5, 4.0,
                                   out.data_ptr(), None, None, None, stages, wsb.data_ptr(), wsb.numel(),
                                   stream.cuda_stream)
        _lib.check(rc, "ys_detection_loss")

    def timed(stages, reps, fwd_only=False):
        call()
        for _ in range(3):
            call(stages, fwd_only)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            call(stages, fwd_only)
        e1.record(stream)
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) / reps * 1e3

    # settle clocks
    for _ in range(300):
        call()
    torch.cuda.synchronize(dev)
    res = {"B": B, "A": A, "n_gt": n_gt}
    res["full_us"] = timed(0, args.reps)
    res["pack_us"] = timed(1, args.reps)
    res["assign_us"] = timed(2, args.reps)
    res["assign_nofix_us"] = timed(2 | 256, args.reps)
    call()
    res["loss_us"] = timed(4, args.reps)
    res["loss_fwd_only_us"] = timed(4, args.reps, fwd_only=True)
    call()
    res["full2_us"] = timed(0, args.reps)
    res["img_per_s"] = B / (res["full2_us"] * 1e-6)
    res["loss_items"] = out.tolist()
    print(json.dumps(res))


if __name__ == "__main__":
    main()
