"""Image sharding of the hot path across ranks: every step of assign / loss / NMS is per image, so
rank r simply takes images [r*B/W, (r+1)*B/W) of the batch and its rows of the flat GT list, with batch_idx
re-based.  No data-path collective exists; the only exchange is the 3-float loss_items all-reduce for logging
(and DDP's gradient all-reduce around the unchanged PyTorch model).  Each rank normalises by its OWN
target_scores_sum, exactly as one reference process per shard would (losses/yolo_v8_loss.py:472, :499)."""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch


def shard_bounds(batch_size: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous split; the first `batch_size % world` ranks get one extra image."""
    base, extra = divmod(batch_size, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_batch(feats: Sequence[torch.Tensor], batch: Dict[str, torch.Tensor], rank: int, world: int):
    """Returns (feats_r, batch_r) for this rank."""
    B = feats[0].shape[0]
    lo, hi = shard_bounds(B, rank, world)
    bi = batch["batch_idx"]
    sel = (bi >= lo) & (bi < hi)
    out = {k: v for k, v in batch.items() if k not in ("batch_idx", "cls", "bboxes")}
    out["batch_idx"] = bi[sel] - lo
    out["cls"] = batch["cls"][sel]
    out["bboxes"] = batch["bboxes"][sel]
    return [f[lo:hi] for f in feats], out


def allreduce_loss_items(items: torch.Tensor, n_local: int) -> torch.Tensor:
    """Image-weighted mean of the per-rank loss items, for logging only (3 floats)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return items
    buf = torch.cat((items.detach().double() * n_local, torch.tensor([float(n_local)], dtype=torch.float64,
                                                                      device=items.device)))
    dist.all_reduce(buf)
    return (buf[:3] / buf[3]).to(items.dtype)
