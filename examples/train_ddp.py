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


def check_grads(args, net, model, criterion, img, batch, rank, world, dev):
    """The multi-GPU oracle of SURVEY §8e: the path is per-image independent and each rank normalises its own loss, so DDP's
    gradient mean must equal the mean over ranks of the per-shard reference gradients.  Per rank: (1) the product step --
    forward through DDP, CUDA criterion, backward: parameter gradients averaged by NCCL; (2) the oracle -- the same head
    tensors through oracle/ref_port.detection_loss on the CPU, its head gradients back-propagated through the SAME
    autograd graph of the unwrapped module (no all-reduce), then averaged over ranks by hand.  Prints one JSON line."""
    from oracle import ref_port as orc                     # checker only (an example script, not the product path)
    params = [p for p in model.parameters() if p.requires_grad]
    for p in params:
        p.grad = None
    criterion.export_assignment = True                     # keep the CUDA path's decoded boxes: the oracle's assigner is fed
    preds = net(img)                                       # with them (spliced oracle), so a 1-ulp decode
    loss, items = criterion(preds, batch)                  # difference cannot flip a near-tie between two GT of one anchor
    loss.backward()
    boxes_px = criterion.last_export["pred_bboxes_px"].cpu()
    g_ddp = [p.grad.detach().clone() for p in params]
    for p in params:
        p.grad = None
    preds2 = model(img)                                    # same weights, no DDP hooks
    leaves = [p.detach().cpu().requires_grad_(True) for p in preds2]
    total, ritems = orc.detection_loss(leaves, {k: v.cpu() for k, v in batch.items()}, args.nc, 16, synth.LEVEL_STRIDES,
                                       assign_boxes_px=boxes_px)
    total.backward()
    g_shard = torch.autograd.grad(preds2, params, grad_outputs=[l.grad.to(dev) for l in leaves], allow_unused=True)
    g_shard = [torch.zeros_like(p) if g is None else g for g, p in zip(g_shard, params)]
    if world > 1:
        for g in g_shard:
            dist.all_reduce(g, op=dist.ReduceOp.AVG)
    num = sum(float((a - b).double().pow(2).sum()) for a, b in zip(g_ddp, g_shard)) ** 0.5
    den = sum(float(b.double().pow(2).sum()) for b in g_shard) ** 0.5
    worst = max(float((a - b).abs().max()) / (float(b.abs().max()) or 1.0) for a, b in zip(g_ddp, g_shard))
    item_rel = float(((items.cpu() - ritems).abs() / ritems.abs().clamp_min(1e-12)).max())
    ok = num / max(den, 1e-30) <= 1e-4 and item_rel <= 1e-5
    res = torch.tensor([num / max(den, 1e-30), worst, item_rel, 1.0 if ok else 0.0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(res, op=dist.ReduceOp.MAX)
        okt = torch.tensor([1.0 if ok else 0.0], device=dev)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        ok = bool(okt.item() > 0)
    if rank == 0:
        print(json.dumps({"check": "DDP-averaged parameter gradients vs mean over ranks of per-shard oracle gradients "
                                   "(oracle/ref_port.detection_loss on each rank's head tensors, back-propagated through the "
                                   "same graph)", "n_gpus": world, "model": f"YOLO11{args.scale}", "batch_per_gpu": args.batch,
                          "imgsz": args.imgsz, "params": sum(p.numel() for p in params),
                          "rel_l2_diff_all_params": float(res[0]), "worst_single_param_rel_max_diff": float(res[1]),
                          "note": "the worst single parameter is one whose true gradient is zero (a conv bias in front of "
                                  "BatchNorm): rounding noise over rounding noise; the all-parameter L2 figure is the check",
                          "loss_items_rel_diff_max_over_ranks": float(res[2]), "ok": ok}), flush=True)
    assert ok, "DDP gradients differ from the per-shard oracle mean"


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
    ap.add_argument("--check-grads", action="store_true",
                    help="multi-GPU oracle (SURVEY §8e): one step; DDP's averaged parameter gradients must equal the mean "
                         "over ranks of the gradients obtained by back-propagating each shard's ORACLE head gradients")
    args = ap.parse_args()

    if args.check_grads:       # an exact comparison: no TF32 in the convolutions (its 10-bit mantissa is a 5e-4 noise floor)
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
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

    if args.check_grads:
        check_grads(args, net, model, criterion, img, batch, rank, world, dev)
        if world > 1:
            dist.destroy_process_group()
        return

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
