"""Inference decode alone (both layouts), for ncu captures:  python scripts/decode_only.py [reps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_scratch_b200 import synth  # noqa: E402
from yolo_scratch_b200.head_decode import detect_inference  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
feats = [f.cuda() for f in synth.make_head_levels(64, 640, 80, 16, seed=1000)]
for am in (False, True):
    for _ in range(3):
        detect_inference(feats, 80, anchor_major=am)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        detect_inference(feats, 80, anchor_major=am)
    e1.record()
    torch.cuda.synchronize()
    print("anchor_major" if am else "channels_first", e0.elapsed_time(e1) / reps, "ms")
