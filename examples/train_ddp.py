#!/usr/bin/env python
"""BASELINE.json configs[2] / SURVEY §8(f) f1: a YOLO11 full train step -- PyTorch backbone + neck + head, the
hand-written loss path (yolo_scratch_b200.losses.v8DetectionLoss), DDP data-parallel, one process per GPU.

    python examples/train_ddp.py --scale s --batch 64 --steps 20                              # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \\
        --master-port 29533 examples/train_ddp.py --scale s --batch 64 --steps 20             # N GPUs (weak scaling)

The step is the reference's (train_yolo11_cuda.py:41-63): forward, criterion(preds, batch), backward,
clip_grad_norm_(10), AdamW step -- with the loss normalised PER RANK exactly as the reference does per process
(losses/yolo_v8_loss.py:472, :499); DDP's gradient mean then equals the single-process gradient of the mean loss.
The only collectives are DDP's bucketed gradient all-reduce (NCCL over NVLink) and one 3-float all-reduce of the loss
items for logging.  Prints one JSON line on rank 0: images/s of the whole job (CUDA events, max over ranks) and the
share of the step spent in the loss path (timed around criterion + the loss part of backward with events).
Synthetic images and COCO-shaped GT (no dataset on the box); random-init weights.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "examples"))

from yolo11_torch import YOLO11  # noqa: E402
from yolo_scratch_b200 import synth  # noqa: E402
from yolo_scratch_b200.losses import v8DetectionLoss  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", default="s")
    ap.add_argument("--batch", type=int, default=64, help="images per GPU")
    ap.add_argument("--imgsz", type=int, default=640)
    ap.add_argument("--nc", type=int, default=80)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--channels-last", action="store_true")
    ap.add_argument("--bf16", action="store_true", help="autocast the backbone to bf16 (the loss path stays fp32)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    torch.manual_seed(0)                                     # same initial weights on every rank
    model = YOLO11(args.scale, 3, args.nc).to(dev).train()
    if args.channels_last:
        model = model.to(memory_format=torch.channels_last)
    criterion = v8DetectionLoss(model)
    net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
    opt = torch.optim.AdamW(net.parameters(), lr=1e-3, weight_decay=5e-4)

    g = torch.Generator().manual_seed(100 + rank)            # a different shard of synthetic data per rank
    img = torch.rand(args.batch, 3, args.imgsz, args.imgsz, generator=g).to(dev)
    if args.channels_last:
        img = img.contiguous(memory_format=torch.channels_last)
    batch = {k: v.to(dev) for k, v in synth.make_gt_batch(args.batch, args.nc, 20, seed=200 + rank).items()}

    loss_ms = []

    def step(timed):
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=args.bf16):
            preds = net(img)
        e = [torch.cuda.Event(enable_timing=True) for _ in range(2)] if timed else None
        if timed:
            e[0].record()
        loss, items = criterion(preds, batch)                # forward AND backward kernels of the loss path run here
        if timed:
            e[1].record()
        loss.backward()
        torch.nn.utils.clip_grad_norm_(net.parameters(), 10.0)
        opt.step()
        if timed:
            loss_ms.append(e)
        return items

    for _ in range(max(args.warmup, 3)):
        items = step(False)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(args.steps):
        items = step(True)
    t1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ms = torch.tensor([t0.elapsed_time(t1)], device=dev, dtype=torch.float64)
    li = items.detach().clone()
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(li, op=dist.ReduceOp.AVG)
    loss_share_ms = sum(a.elapsed_time(b) for a, b in loss_ms) / len(loss_ms)
    if rank == 0:
        step_ms = float(ms.item()) / args.steps
        print(json.dumps({
            "workload": f"configs[2]: YOLO11{args.scale} full train step, {args.batch} img/GPU, {args.imgsz}px, nc={args.nc}, "
                        f"DDP x{world}, {'bf16 autocast backbone' if args.bf16 else 'fp32'}",
            "n_gpus": world, "images_per_s": world * args.batch * args.steps / (float(ms.item()) * 1e-3),
            "ms_per_step": step_ms, "loss_path_ms_per_step": loss_share_ms, "loss_path_share": loss_share_ms / step_ms,
            "params": sum(p.numel() for p in model.parameters()), "loss_items_mean_over_ranks": li.tolist(),
            "finite": bool(torch.isfinite(li).all().item())}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
