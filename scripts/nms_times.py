#!/usr/bin/env python
"""configs[3] NMS (B=256, 8400 candidates x 80 classes, conf 0.001 / IoU 0.7) timing on the GPU box.
    python scripts/nms_times.py [--B 256] [--reps 5] [--clustered]
Run under `ncu --metrics gpu__time_duration.sum` for the per-kernel split."""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_scratch_b200 import synth
from yolo_scratch_b200.train_yolo11_cuda import nms_batched

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, default=256)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--conf", type=float, default=0.001)
ap.add_argument("--iou", type=float, default=0.7)
ap.add_argument("--clustered", action="store_true")
a = ap.parse_args()
dev = torch.device("cuda", 0)
base = synth.make_nms_predictions(16, 8400, 80, 640, seed=2000, clustered=a.clustered)
pred = base.repeat(a.B // 16, 1, 1).to(dev)
pred[:, :, :2] += torch.arange(a.B, device=dev).view(a.B, 1, 1) * 0.03125
for _ in range(2):
    out = nms_batched(pred, a.conf, a.iou, 640)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.reps):
    out = nms_batched(pred, a.conf, a.iou, 640)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.reps
print(json.dumps({"B": a.B, "ms": ms, "img_per_s": a.B / ms * 1e3, "kept_per_image": out[4].float().mean().item()}))
