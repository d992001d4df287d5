#!/usr/bin/env python
"""The "existing GPU path" (SURVEY §8d): the reference's algorithm in stock PyTorch ops, on the B200 -- i.e. the oracle
port (oracle/ref_port.py, test infrastructure) run on CUDA tensors instead of CPU tensors.  Informational only: it is
neither shipped nor part of bench.py; the number goes into profiles/ next to the bench line.

    python scripts/torch_gpu_baseline.py [--B 64] [--steps 5]
"""
import argparse, json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_scratch_b200 import synth
from oracle import ref_port as orc

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, default=64)
ap.add_argument("--steps", type=int, default=5)
a = ap.parse_args()
dev = torch.device("cuda", 0)
feats = [f.to(dev) for f in synth.make_head_levels(a.B, 640, 80, 16, seed=1000)]
batch = {k: v.to(dev) for k, v in synth.make_gt_batch(a.B, 80, 20, seed=1001).items()}
torch.set_default_device(dev)      # the port creates its helper tensors with plain factory calls


def step():
    leaves = [f.detach().requires_grad_(True) for f in feats]
    total, items = orc.detection_loss(leaves, batch, 80, 16, synth.LEVEL_STRIDES)
    total.backward()
    return items


step(); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(a.steps):
    items = step()
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / a.steps
print(json.dumps({"what": "oracle port (stock torch ops) on the B200, assign + loss fwd+bwd", "B": a.B, "ms_per_step": dt * 1e3,
                  "images_per_s": a.B / dt, "loss_items": items.tolist()}))
