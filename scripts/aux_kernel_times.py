#!/usr/bin/env python
"""CUDA-event times of the exported kernels that are not on the bench line (VERDICT r01 "missing" #6): the dense
TaskAlignedAssigner entry point (default and the two non-default modes), BboxLoss, bbox_iou, pairwise IoU, the mAP
matching and cda_nms -- each with its algorithmic bytes and the fraction of the measured HBM peak.  GPU box only.
    python scripts/aux_kernel_times.py > profiles/r02_aux_kernel_times.json"""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from yolo_scratch_b200 import _lib, synth  # noqa: E402
from yolo_scratch_b200.losses import TaskAlignedAssigner, BboxLoss, bbox_iou  # noqa: E402
from yolo_scratch_b200.utils import metrics as M  # noqa: E402

dev = torch.device("cuda", 0)
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def line(name, ms, alg_bytes, note=""):
    gbs = alg_bytes / (ms * 1e-3) / 1e9
    return {"kernel": name, "ms": ms, "algorithmic_bytes": alg_bytes, "achieved_gbs": gbs, "frac_hbm_peak": gbs / peak, "note": note}


out = {"peak_gbs": peak, "lines": []}
B, A, nc, G = 64, 8400, 80, 20
inp = [t.to(dev) for t in synth.make_assigner_inputs(B, ((80, 80), (40, 40), (20, 20)), synth.LEVEL_STRIDES, nc, G, 5)]
Gm = inp[4].shape[1]
# reads scores (B,A,nc) only where gathered, boxes (B,A,4); writes labels, boxes, scores (B,A,nc), fg, idx
alg_assign = B * A * (16 + 4 + 16 + nc * 4 + 1 + 8)
for name, kw in (("ys_assign (reference default)", {}), ("ys_assign_ex use_topk=13", {"use_topk": True}),
                 ("ys_assign_ex ciou_assign", {"ciou_assign": True})):
    a = TaskAlignedAssigner(topk=13, num_classes=nc, alpha=0.5, beta=4.0, **kw)
    ms = timed(lambda: a.forward(*inp))
    out["lines"].append(line(name, ms, alg_assign, f"B={B}, A={A}, G={Gm}, nc={nc}; includes the shim's output allocations; "
                                                    f"writes the dense (B,A,nc) target_scores the fused loss never materialises"))
# BboxLoss on the assigner's outputs
a = TaskAlignedAssigner(topk=13, num_classes=nc, alpha=0.5, beta=4.0)
tl, tb, ts, fg, tgi = a.forward(*inp)
pd = torch.randn(B, A, 64, device=dev, requires_grad=True)
anc = inp[2] / 8.0
pb = (inp[1] / 8.0).detach().requires_grad_(True)
tss = ts.sum().clamp(min=1)
bl = BboxLoss(16)
def bbox_step():
    li, ld = bl(pd, pb, anc, tb / 8.0, ts, tss, fg)
    (li + ld).backward()
    pd.grad = None; pb.grad = None
ms = timed(bbox_step)
out["lines"].append(line("ys_bbox_loss fwd+bwd (BboxLoss)", ms, B * A * (64 * 4 * 2 + 16 * 3 + nc * 4 + 1),
                         "dense inputs incl. the (B,A,nc) target_scores; autograd glue included"))
n = B * A
b1, b2 = inp[1].reshape(-1, 4).contiguous(), tb.reshape(-1, 4).contiguous()
ms = timed(lambda: bbox_iou(b1, b2, xywh=False, CIoU=True))
out["lines"].append(line("ys_bbox_iou CIoU, paired", ms, n * 36, f"n={n}"))
q1, q2 = b1[:4096].contiguous(), b2[:8192].contiguous()
ms = timed(lambda: M.calculate_iou_batch(q1, q2))
out["lines"].append(line("ys_pairwise_iou kind 0 (4096 x 8192)", ms, 4096 * 8192 * 4, "output-write bound"))
# mAP matching: B=64 images, 300 predictions and 20 GT each
g = torch.Generator().manual_seed(3)
preds, tgts = [], []
for _ in range(64):
    xy = torch.rand(300, 2, generator=g) * 0.8
    wh = torch.rand(300, 2, generator=g) * 0.2 + 0.01
    preds.append({"boxes": torch.cat((xy, xy + wh), 1).to(dev), "scores": torch.rand(300, generator=g).to(dev), "labels": torch.zeros(300, dtype=torch.long, device=dev)})
    xy = torch.rand(20, 2, generator=g) * 0.8
    wh = torch.rand(20, 2, generator=g) * 0.2 + 0.01
    tgts.append({"boxes": torch.cat((xy, xy + wh), 1).to(dev), "labels": torch.zeros(20, dtype=torch.long, device=dev)})
ms = timed(lambda: M.evaluate_detections(preds, tgts, 0.25, 0.5), reps=5)
out["lines"].append(line("evaluate_detections (ys_match_detections + host AP)", ms, 64 * (300 * 20 + 20 * 16), "64 images x 300 predictions x 20 GT x 11 thresholds; dominated by the shim's host-side packing and AP arithmetic"))
one = {"boxes": preds[0]["boxes"] * 640, "scores": preds[0]["scores"], "angles": torch.rand(300, generator=g).to(dev)}
ms = timed(lambda: M.cda_nms(one, 10, 0.15), reps=10)
out["lines"].append(line("cda_nms (ys_cda_nms, one image, 300 detections, top 10)", ms, 300 * 20, "latency bound: one warp, fp64"))
print(json.dumps(out, indent=1))
