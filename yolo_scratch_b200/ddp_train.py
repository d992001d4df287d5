"""data-parallel training wrapper around the reference's epoch loop.

The reference trains on one device (`train_yolo11_cuda.py:31-98` train_one_epoch, `:521-543` loaders,
`:595-653` epoch driver + checkpoint); it has no multi-GPU path.  This module is the pure addition §8f asks for:
one process per GPU, `DistributedSampler` shards of the dataset, DDP gradient all-reduce over NCCL (NVLink /
NVSwitch) overlapped with the backward pass by buckets, and a rank-0 checkpoint in the reference's own
`last.pt` / `best.pt` dict format (`:628-653`) so that the files load in the reference unchanged.

What differs from the reference loop, on purpose:
  * no per-step host sync: the reference reads four `.item()`s per step (`:64-67`); here the running sums stay on the
    device and are read ONCE per epoch (after one 5-float all-reduce), so the launch queue never drains;
  * the metrics returned are means over ALL ranks' steps (same keys: loss / box_loss / cls_loss / dfl_loss);
  * `sampler.set_epoch(epoch)` reshuffles the shards every epoch.
The loss semantics are the reference's per process: `loss.sum() * B_local` (`losses/yolo_v8_loss.py:499`), gradients
averaged over ranks by DDP.  Nothing here touches the oracle; the criterion is
`yolo_scratch_b200.losses.v8DetectionLoss` (CUDA only, no CPU fallback).
"""
from __future__ import annotations

import math
import os
from pathlib import Path
from typing import Dict, Optional, Tuple

import torch
import torch.distributed as dist

CHECKPOINT_KEYS = ("epoch", "model_state_dict", "optimizer_state_dict", "train_metrics", "val_metrics",
                   "best_loss", "best_mAP50")      # train_yolo11_cuda.py:628-636


def init_distributed(backend: Optional[str] = None) -> Tuple[int, int, int, torch.device]:
    """Join the job torchrun started (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*).  Returns (rank, world, local, device).
    Without those variables this is a single-process run and no process group is created."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available()
    device = torch.device("cuda", local) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        backend = backend or ("nccl" if use_cuda else "gloo")
        kw = {"device_id": device} if backend == "nccl" else {}
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, world, local, device


def is_main() -> bool:
    return not (dist.is_available() and dist.is_initialized()) or dist.get_rank() == 0


def make_loader(dataset, batch_size: int, collate_fn=None, shuffle: bool = True, seed: int = 0, num_workers: int = 0,
                drop_last: bool = False, pin_memory: Optional[bool] = None):
    """Per-rank loader over this rank's `DistributedSampler` shard (`batch_size` is PER RANK: weak scaling, as the
    reference's `--batch` is per process, `train_yolo11_cuda.py:521-543`).  Returns (loader, sampler); sampler is None
    in a single-process run."""
    from torch.utils.data import DataLoader
    from torch.utils.data.distributed import DistributedSampler
    sampler = None
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        sampler = DistributedSampler(dataset, num_replicas=dist.get_world_size(), rank=dist.get_rank(),
                                     shuffle=shuffle, seed=seed, drop_last=drop_last)
    if pin_memory is None:
        pin_memory = torch.cuda.is_available()
    loader = DataLoader(dataset, batch_size=batch_size, shuffle=(shuffle and sampler is None), sampler=sampler,
                        collate_fn=collate_fn, num_workers=num_workers, pin_memory=pin_memory, drop_last=drop_last,
                        persistent_workers=num_workers > 0)
    return loader, sampler


def wrap_model(model: torch.nn.Module, device: torch.device, bucket_cap_mb: int = 25):
    """DDP over the current process group.  Buckets of `bucket_cap_mb` are all-reduced as soon as their gradients are
    ready, so the NCCL traffic (45.6 MB fp32 for YOLO11s) hides under the rest of the backward pass; the gradients are
    views into the buckets (no extra copy)."""
    model = model.to(device)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return model
    from torch.nn.parallel import DistributedDataParallel as DDP
    ids = [device.index] if device.type == "cuda" else None
    return DDP(model, device_ids=ids, bucket_cap_mb=bucket_cap_mb, gradient_as_bucket_view=True,
               broadcast_buffers=True)


def unwrap(model: torch.nn.Module) -> torch.nn.Module:
    return model.module if hasattr(model, "module") and isinstance(model.module, torch.nn.Module) else model


def cosine_lr_schedule(optimizer, epoch, epochs, lr_min=1e-6, lr_max=1e-3, warmup_epochs=3):
    """Linear warm-up then half-cosine decay, written into every param group (`train_yolo11_cuda.py:440-451`)."""
    if epoch < warmup_epochs:
        lr = lr_min + (lr_max - lr_min) * (epoch / warmup_epochs)
    else:
        span = max(epochs - warmup_epochs, 1)
        lr = lr_min + (lr_max - lr_min) * 0.5 * (1.0 + math.cos(math.pi * (epoch - warmup_epochs) / span))
    for group in optimizer.param_groups:
        group["lr"] = lr
    return lr


def _to_device(batch: Dict, device: torch.device) -> Dict:
    return {k: (v.to(device, non_blocking=True) if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}


def train_one_epoch(model, dataloader, optimizer, criterion, device, epoch, epochs, sampler=None, max_norm: float = 10.0,
                    progress: bool = False) -> Dict[str, float]:
    """The reference's epoch loop (`train_yolo11_cuda.py:31-98`: zero_grad(set_to_none) -> forward -> criterion ->
    backward -> clip_grad_norm_(10.0) -> step), data-parallel.  Same positional signature; returns the same four keys,
    each the mean over every step of every rank."""
    model.train()
    if sampler is not None:
        sampler.set_epoch(epoch)
    it = dataloader
    if progress and is_main():
        from tqdm import tqdm
        it = tqdm(dataloader, desc=f"Epoch {epoch}/{epochs}")
    acc = torch.zeros(5, device=device, dtype=torch.float64)      # loss, box, cls, dfl, steps
    for batch in it:
        batch = _to_device(batch, device)
        optimizer.zero_grad(set_to_none=True)
        preds = model(batch["img"])
        loss, loss_items = criterion(preds, batch)
        loss.backward()                                            # DDP: bucketed all-reduce overlaps this
        torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=max_norm)
        optimizer.step()
        acc[0] += loss.detach()
        acc[1:4] += loss_items.detach()[:3]
        acc[4] += 1.0
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(acc, op=dist.ReduceOp.SUM)
    a = acc.tolist()                                               # the epoch's only host sync
    n = max(a[4], 1.0)
    return {"loss": a[0] / n, "box_loss": a[1] / n, "cls_loss": a[2] / n, "dfl_loss": a[3] / n}


def save_checkpoint(save_dir, epoch: int, model, optimizer, train_metrics: Dict, val_metrics: Dict, best_loss: float,
                    best_mAP50: float) -> Tuple[float, float]:
    """Rank 0 writes `last.pt`, and `best.pt` when mAP50 (if present, else the validation loss) improved -- the
    reference's dict and rule (`train_yolo11_cuda.py:628-653`).  The state dict is the UNWRAPPED module's, so its keys
    carry no `module.` prefix and load straight into the reference's model.  Every rank gets the updated
    (best_loss, best_mAP50) back: the rule is a pure function of its arguments, no broadcast is needed."""
    old_loss, old_map = best_loss, best_mAP50
    improved = False
    if "mAP50" in val_metrics:
        if val_metrics["mAP50"] > best_mAP50:
            best_mAP50, improved = val_metrics["mAP50"], True
    elif val_metrics.get("loss", float("inf")) < best_loss:
        best_loss, improved = val_metrics["loss"], True
    if is_main():
        save_dir = Path(save_dir)
        save_dir.mkdir(parents=True, exist_ok=True)
        # the reference writes last.pt with the best_* values from BEFORE this epoch's update (:628-639) and best.pt with
        # the updated one (:646-652)
        ckpt = {"epoch": epoch, "model_state_dict": unwrap(model).state_dict(),
                "optimizer_state_dict": optimizer.state_dict(), "train_metrics": train_metrics,
                "val_metrics": val_metrics, "best_loss": old_loss, "best_mAP50": old_map}
        torch.save(ckpt, save_dir / "last.pt")
        if improved:
            ckpt["best_loss"], ckpt["best_mAP50"] = best_loss, best_mAP50
            torch.save(ckpt, save_dir / "best.pt")
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()                                             # nobody races ahead of the files
    return best_loss, best_mAP50


def load_checkpoint(path, model, optimizer=None, map_location="cpu") -> Dict:
    """Load a `last.pt` / `best.pt` written by this module OR by the reference trainer into a (possibly DDP-wrapped)
    model; returns the checkpoint dict."""
    ckpt = torch.load(path, map_location=map_location, weights_only=False)
    unwrap(model).load_state_dict(ckpt["model_state_dict"])
    if optimizer is not None and "optimizer_state_dict" in ckpt:
        optimizer.load_state_dict(ckpt["optimizer_state_dict"])
    return ckpt
