"""Seeded synthetic inputs for the detection-head hot path.

Everything is generated on the CPU with an explicit ``torch.Generator`` so the
same seed yields the same tensors in this container, on the GPU box and in the
golden-fixture script.  Nothing here touches the oracle or the CUDA library.

Distributions:
  * head tensors: 4*reg_max box-distribution channels ~ N(0,1)+1.0 (the Detect
    head's box bias init, reference models/yolo11_modules.py:271), nc class
    channels ~ N(0,1)*1.5-4.6.
  * GT, COCO-shaped: count U{1..max_gt} per image, class U{0..nc-1}, centre
    U(0,1)^2, w,h = exp(U(ln .02, ln .6)), xyxy clamped to [0,1] as the
    reference collate does (datasets/crater_dataset_cuda.py:319-322); 5 % of
    images get zero GT, 5 % of GTs are tiny (< 4 px at 640) so no anchor centre
    falls inside (assigner loop 1), 5 % are nested in the previous GT (loop 2).
  * NMS candidates: cx,cy ~ U(0,S), w,h ~ U(5,105), class scores
    sigmoid(N(-4,2)); per-image best-class scores are made unique (reference
    argsort is unstable).
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence, Tuple

import torch

LEVEL_STRIDES = (8, 16, 32)


def level_shapes(imgsz: int, strides: Sequence[int] = LEVEL_STRIDES) -> List[Tuple[int, int]]:
    return [(imgsz // s, imgsz // s) for s in strides]


def make_head_levels(batch: int, imgsz: int = 640, nc: int = 80, reg_max: int = 16,
                     strides: Sequence[int] = LEVEL_STRIDES, seed: int = 0,
                     shapes: Sequence[Tuple[int, int]] | None = None) -> List[torch.Tensor]:
    """Three NCHW level tensors (B, 4*reg_max+nc, H_l, W_l), fp32, CPU."""
    g = torch.Generator().manual_seed(seed)
    shapes = list(shapes) if shapes is not None else level_shapes(imgsz, strides)
    out = []
    for (h, w) in shapes:
        box = torch.randn(batch, 4 * reg_max, h, w, generator=g) + 1.0
        cls = torch.randn(batch, nc, h, w, generator=g) * 1.5 - 4.6
        out.append(torch.cat((box, cls), 1).contiguous())
    return out


def make_gt_batch(batch: int, nc: int = 80, max_gt: int = 20, seed: int = 1,
                  fixed_count: int | None = None, dense: bool = False,
                  p_empty: float = 0.05, p_tiny: float = 0.05, p_nested: float = 0.05,
                  imgsz: int = 640) -> Dict[str, torch.Tensor]:
    """Flat GT list in the reference batch-dict format
    (datasets/crater_dataset_cuda.py:336-341): batch_idx (N,) i64, cls (N,1) i64,
    bboxes (N,4) fp32 normalised xyxy."""
    g = torch.Generator().manual_seed(seed)
    bi: List[int] = []
    cl: List[int] = []
    bx: List[List[float]] = []
    for b in range(batch):
        if fixed_count is None:
            if torch.rand((), generator=g).item() < p_empty:
                continue
            n = int(torch.randint(1, max_gt + 1, (), generator=g).item())
        else:
            n = fixed_count
        prev = None
        for _ in range(n):
            u = torch.rand(6, generator=g).tolist()
            cx, cy = u[0], u[1]
            if dense:
                w = 0.01 + 0.15 * u[2]
                h = 0.01 + 0.15 * u[3]
            else:
                w = math.exp(math.log(0.02) + (math.log(0.6) - math.log(0.02)) * u[2])
                h = math.exp(math.log(0.02) + (math.log(0.6) - math.log(0.02)) * u[3])
            if u[4] < p_tiny:
                # smaller than 4 px and placed on a stride-8 grid corner: no anchor centre inside
                w = h = 3.0 / imgsz
                cx = (8.0 * round(cx * imgsz / 8.0)) / imgsz
                cy = (8.0 * round(cy * imgsz / 8.0)) / imgsz
            elif u[4] < p_tiny + p_nested and prev is not None:
                pcx, pcy, pw, ph = prev
                cx, cy, w, h = pcx, pcy, pw * (0.2 + 0.3 * u[5]), ph * (0.2 + 0.3 * u[5])
            prev = (cx, cy, w, h)
            x1 = min(max(cx - w / 2, 0.0), 1.0)
            y1 = min(max(cy - h / 2, 0.0), 1.0)
            x2 = min(max(cx + w / 2, 0.0), 1.0)
            y2 = min(max(cy + h / 2, 0.0), 1.0)
            bi.append(b)
            cl.append(int(torch.randint(0, nc, (), generator=g).item()))
            bx.append([x1, y1, x2, y2])
    if not bi:
        return {"batch_idx": torch.zeros((0,), dtype=torch.long),
                "cls": torch.zeros((0, 1), dtype=torch.long),
                "bboxes": torch.zeros((0, 4), dtype=torch.float32)}
    return {"batch_idx": torch.tensor(bi, dtype=torch.long),
            "cls": torch.tensor(cl, dtype=torch.long).unsqueeze(1),
            "bboxes": torch.tensor(bx, dtype=torch.float32)}


def make_assigner_inputs(batch: int, anchors_hw: Sequence[Tuple[int, int]] = ((80, 80), (40, 40), (20, 20)),
                         strides: Sequence[int] = LEVEL_STRIDES, nc: int = 80, max_gt: int = 20,
                         seed: int = 2, jitter: float = 24.0, **gt_kw):
    """Dense inputs of TaskAlignedAssigner.forward (reference losses/yolo_v8_loss.py:79):
    pd_scores (B,A,nc) in (0,1), pd_bboxes (B,A,4) px xyxy, anc_points (A,2) px,
    gt_labels (B,G) float, gt_bboxes (B,G,4) px, mask_gt (B,G) float 0/1."""
    g = torch.Generator().manual_seed(seed)
    pts, sts = [], []
    for (h, w), s in zip(anchors_hw, strides):
        sy, sx = torch.meshgrid(torch.arange(h, dtype=torch.float32) + 0.5,
                                torch.arange(w, dtype=torch.float32) + 0.5, indexing="ij")
        pts.append(torch.stack((sx, sy), -1).view(-1, 2) * s)
        sts.append(torch.full((h * w,), float(s)))
    anc = torch.cat(pts)
    st = torch.cat(sts)
    A = anc.shape[0]
    imgsz = float(anchors_hw[0][0] * strides[0])
    ltrb = torch.rand(batch, A, 4, generator=g) * jitter * st.view(1, A, 1) / 8.0
    pd_bboxes = torch.cat((anc.view(1, A, 2) - ltrb[..., :2], anc.view(1, A, 2) + ltrb[..., 2:]), -1)
    pd_scores = torch.sigmoid(torch.randn(batch, A, nc, generator=g) * 1.5 - 2.0)
    gt = make_gt_batch(batch, nc=nc, max_gt=max_gt, seed=seed + 1000, imgsz=int(imgsz), **gt_kw)
    counts = torch.bincount(gt["batch_idx"], minlength=batch) if gt["batch_idx"].numel() else torch.zeros(batch, dtype=torch.long)
    G = int(counts.max().item()) if counts.numel() else 0
    gt_labels = torch.zeros(batch, G)
    gt_bboxes = torch.zeros(batch, G, 4)
    mask_gt = torch.zeros(batch, G)
    for b in range(batch):
        m = gt["batch_idx"] == b
        n = int(m.sum())
        if n:
            gt_labels[b, :n] = gt["cls"][m, 0].float()
            gt_bboxes[b, :n] = gt["bboxes"][m] * imgsz
            mask_gt[b, :n] = 1.0
    return pd_scores, pd_bboxes, anc, gt_labels, gt_bboxes, mask_gt


def make_nms_predictions(batch: int, anchors: int = 8400, nc: int = 80, imgsz: int = 640,
                         seed: int = 3, clustered: bool = False, n_clusters: int = 50) -> torch.Tensor:
    """pred (B, A, 4+nc) in the layout decode_predictions_for_metrics documents
    (reference train_yolo11_cuda.py:289-291): xywh pixels + sigmoided class scores.
    Best-class scores are unique per image."""
    g = torch.Generator().manual_seed(seed)
    if clustered:
        centres = torch.rand(batch, n_clusters, 2, generator=g) * imgsz
        sizes = 20.0 + 80.0 * torch.rand(batch, n_clusters, 2, generator=g)
        which = torch.randint(0, n_clusters, (batch, anchors), generator=g)
        idx = which.unsqueeze(-1).expand(-1, -1, 2)
        cxy = centres.gather(1, idx) + torch.randn(batch, anchors, 2, generator=g) * 4.0
        wh = sizes.gather(1, idx) * (1.0 + 0.1 * torch.randn(batch, anchors, 2, generator=g)).clamp(0.5, 1.5)
    else:
        cxy = torch.rand(batch, anchors, 2, generator=g) * imgsz
        wh = 5.0 + 100.0 * torch.rand(batch, anchors, 2, generator=g)
    scores = torch.sigmoid(torch.randn(batch, anchors, nc, generator=g) * 2.0 - 4.0)
    pred = torch.cat((cxy, wh, scores), -1).contiguous()
    make_best_scores_unique(pred)
    return pred


def make_best_scores_unique(pred: torch.Tensor) -> None:
    """In place: nudge the per-anchor best class score until all best scores of
    an image differ (ties have implementation-defined order in the reference)."""
    B = pred.shape[0]
    for b in range(B):
        sc = pred[b, :, 4:]
        for _ in range(64):
            best, arg = sc.max(1)
            srt, order = torch.sort(best, stable=True)
            dup = torch.zeros_like(srt, dtype=torch.bool)
            dup[1:] = srt[1:] == srt[:-1]
            if not bool(dup.any()):
                break
            rows = order[dup]
            # multiply by a distinct factor slightly below 1: stays the row max only if
            # the runner-up is lower, which the next iteration re-checks
            k = torch.arange(1, rows.numel() + 1, dtype=torch.float32)
            sc[rows, arg[rows]] = best[rows] * (1.0 - 1e-6 * k) + 1e-9 * k
        else:
            raise RuntimeError("could not make scores unique")
