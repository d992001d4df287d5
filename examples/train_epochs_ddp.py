#!/usr/bin/env python
"""SURVEY §8(f) f1 end to end: the reference's epoch driver (train_yolo11_cuda.py:595-653 -- cosine LR, criterion.epoch,
train_one_epoch, checkpoint) run data-parallel through yolo_scratch_b200.ddp_train with the hand-written loss path.

    python examples/train_epochs_ddp.py --scale n --epochs 2 --samples 64 --batch 8 --out /tmp/run    # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 \\
        examples/train_epochs_ddp.py --scale n --epochs 2 --samples 64 --batch 8 --out /tmp/run         # 2 GPUs

A synthetic dataset stands in for the crater images (no data on the box): item i is a seeded random image with
COCO-shaped boxes; `collate` builds the reference's flat batch dict (`img`, `batch_idx`, `cls`, `bboxes`,
datasets/crater_dataset_cuda.py:289-346).  Rank 0 prints one JSON line: per-epoch train metrics (means over all
ranks), the checkpoint keys, and that the files reload into a plain (unwrapped) model with identical weights."""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "examples"))

from yolo11_torch import YOLO11  # noqa: E402
from yolo_scratch_b200 import ddp_train as T  # noqa: E402
from yolo_scratch_b200 import synth  # noqa: E402
from yolo_scratch_b200.losses import v8DetectionLoss  # noqa: E402


class SyntheticDetection(torch.utils.data.Dataset):
    def __init__(self, n, imgsz, nc):
        self.n, self.imgsz, self.nc = n, imgsz, nc

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        g = torch.Generator().manual_seed(7000 + i)
        gt = synth.make_gt_batch(1, self.nc, 12, seed=9000 + i, imgsz=self.imgsz, p_empty=0.1)
        return {"img": torch.rand(3, self.imgsz, self.imgsz, generator=g), "cls": gt["cls"], "bboxes": gt["bboxes"]}


def collate(items):
    idx = torch.cat([torch.full((it["cls"].shape[0],), i, dtype=torch.int64) for i, it in enumerate(items)])
    return {"img": torch.stack([it["img"] for it in items]), "batch_idx": idx,
            "cls": torch.cat([it["cls"] for it in items]), "bboxes": torch.cat([it["bboxes"] for it in items])}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", default="n")
    ap.add_argument("--imgsz", type=int, default=320)
    ap.add_argument("--nc", type=int, default=80)
    ap.add_argument("--samples", type=int, default=64)
    ap.add_argument("--batch", type=int, default=8, help="images per GPU")
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--out", default="/tmp/ys_b200_run")
    args = ap.parse_args()

    rank, world, _, device = T.init_distributed()
    torch.manual_seed(0)                                     # same initial weights on every rank (DDP broadcasts anyway)
    model = YOLO11(args.scale, 3, args.nc)
    criterion = v8DetectionLoss(model.to(device))
    net = T.wrap_model(model, device)
    opt = torch.optim.AdamW(net.parameters(), lr=args.lr, weight_decay=5e-4)
    loader, sampler = T.make_loader(SyntheticDetection(args.samples, args.imgsz, args.nc), args.batch, collate_fn=collate,
                                    shuffle=True, seed=1, drop_last=True)
    best_loss, best_map, history = float("inf"), 0.0, []
    for epoch in range(args.epochs):
        lr = T.cosine_lr_schedule(opt, epoch, args.epochs, lr_min=args.lr * 0.01, lr_max=args.lr, warmup_epochs=1)
        criterion.epoch = epoch                              # train_yolo11_cuda.py:603-604
        tm = T.train_one_epoch(net, loader, opt, criterion, device, epoch + 1, args.epochs, sampler=sampler)
        history.append({"epoch": epoch, "lr": lr, **tm})
        best_loss, best_map = T.save_checkpoint(args.out, epoch, net, opt, tm, {"loss": tm["loss"]}, best_loss, best_map)
    if rank == 0:
        ck = torch.load(os.path.join(args.out, "last.pt"), map_location="cpu", weights_only=False)
        plain = YOLO11(args.scale, 3, args.nc)
        plain.load_state_dict(ck["model_state_dict"])
        same = all(torch.equal(a.detach().cpu(), b.detach().cpu())
                   for a, b in zip(plain.state_dict().values(), T.unwrap(net).state_dict().values()))
        print(json.dumps({"workload": f"YOLO11{args.scale} {args.imgsz}px, {args.samples} synthetic samples, {args.batch} img/GPU, "
                                      f"{args.epochs} epochs, DDP x{world}", "n_gpus": world, "steps_per_epoch_per_rank": len(loader),
                          "history": history, "checkpoint_keys": list(ck.keys()), "files": sorted(os.listdir(args.out)),
                          "reloads_into_plain_model": bool(same),
                          "finite": all(all(v == v and abs(v) != float("inf") for v in h.values()) for h in history)}), flush=True)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
