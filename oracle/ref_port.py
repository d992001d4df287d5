"""ORACLE (test infrastructure, not product code).

CPU restatement, in torch CPU ops, of the detection-head hot path of
Pratye/yolo-scratch (`yolo_scratch_cuda`).  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import this module; the product package never does.

Parity pin: this port is checked bit-for-bit (integer outputs) and to 1e-6
(floating outputs) against the *imported, unmodified reference* run in the build
container; the reference's outputs on seeded inputs are committed as fixtures
under `tests/golden/` by `tests/golden/make_golden.py`, and
`tests/test_oracle_golden.py` replays them wherever the repo travels.

All `file:line` citations are relative to /root/reference/yolo_scratch_cuda/.
The port restates the algorithm in the order of SURVEY.md §8.1; it is written
per stage (pair metrics -> forced columns -> per-anchor resolve -> targets), not
as a transcription of the reference's classes.
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence, Tuple

import torch
import torch.nn.functional as F

EPS_IOU = 1e-7      # losses/yolo_v8_loss.py:12  (eps default)
EPS_ASSIGN = 1e-9   # losses/yolo_v8_loss.py:70, :210


# ----------------------------------------------------------------------------
# a6  bbox_iou(xywh=False)                      losses/yolo_v8_loss.py:31-61
# ----------------------------------------------------------------------------
def pair_iou_xyxy(p: torch.Tensor, q: torch.Tensor, ciou: bool = False, eps: float = EPS_IOU) -> torch.Tensor:
    """IoU / CIoU of broadcastable xyxy boxes (..., 4) -> (..., 1).

    Heights (only) carry +eps (:33-34); union carries +eps (:41).  CIoU adds the
    centre-distance and aspect terms with alpha held constant (:46-56)."""
    px1, py1, px2, py2 = p.unbind(-1)
    qx1, qy1, qx2, qy2 = q.unbind(-1)
    pw, ph = px2 - px1, py2 - py1 + eps
    qw, qh = qx2 - qx1, qy2 - qy1 + eps
    ix = (torch.minimum(px2, qx2) - torch.maximum(px1, qx1)).clamp(min=0)
    iy = (torch.minimum(py2, qy2) - torch.maximum(py1, qy1)).clamp(min=0)
    inter = ix * iy
    union = pw * ph + qw * qh - inter + eps
    iou = inter / union
    if ciou:
        cw = torch.maximum(px2, qx2) - torch.minimum(px1, qx1)
        ch = torch.maximum(py2, qy2) - torch.minimum(py1, qy1)
        c2 = cw.pow(2) + ch.pow(2) + eps
        rho2 = ((qx1 + qx2 - px1 - px2).pow(2) + (qy1 + qy2 - py1 - py2).pow(2)) / 4
        v = (4 / math.pi ** 2) * ((qw / qh).atan() - (pw / ph).atan()).pow(2)
        with torch.no_grad():
            alpha = v / (v - iou + (1 + eps))
        iou = iou - (rho2 / c2 + v * alpha)
    return iou.unsqueeze(-1)


def iou_general(box1, box2, xywh=True, GIoU=False, DIoU=False, CIoU=False, eps=EPS_IOU):
    """Full-signature variant (xywh input, GIoU/DIoU) : losses/yolo_v8_loss.py:12-61."""
    if xywh:
        (x1, y1, w1, h1), (x2, y2, w2, h2) = box1.chunk(4, -1), box2.chunk(4, -1)
        a = torch.cat((x1 - w1 / 2, y1 - h1 / 2, x1 + w1 / 2, y1 + h1 / 2), -1)
        b = torch.cat((x2 - w2 / 2, y2 - h2 / 2, x2 + w2 / 2, y2 + h2 / 2), -1)
        ax1, ay1, ax2, ay2 = a.unbind(-1)
        bx1, by1, bx2, by2 = b.unbind(-1)
        w1, h1, w2, h2 = w1.squeeze(-1), h1.squeeze(-1), w2.squeeze(-1), h2.squeeze(-1)
    else:
        ax1, ay1, ax2, ay2 = box1.unbind(-1)
        bx1, by1, bx2, by2 = box2.unbind(-1)
        w1, h1 = ax2 - ax1, ay2 - ay1 + eps
        w2, h2 = bx2 - bx1, by2 - by1 + eps
    inter = (torch.minimum(ax2, bx2) - torch.maximum(ax1, bx1)).clamp(min=0) * \
            (torch.minimum(ay2, by2) - torch.maximum(ay1, by1)).clamp(min=0)
    union = w1 * h1 + w2 * h2 - inter + eps
    iou = inter / union
    if CIoU or DIoU or GIoU:
        cw = torch.maximum(ax2, bx2) - torch.minimum(ax1, bx1)
        ch = torch.maximum(ay2, by2) - torch.minimum(ay1, by1)
        if CIoU or DIoU:
            c2 = cw.pow(2) + ch.pow(2) + eps
            rho2 = ((bx1 + bx2 - ax1 - ax2).pow(2) + (by1 + by2 - ay1 - ay2).pow(2)) / 4
            if CIoU:
                v = (4 / math.pi ** 2) * ((w2 / h2).atan() - (w1 / h1).atan()).pow(2)
                with torch.no_grad():
                    alpha = v / (v - iou + (1 + eps))
                return (iou - (rho2 / c2 + v * alpha)).unsqueeze(-1)
            return (iou - rho2 / c2).unsqueeze(-1)
        c_area = cw * ch + eps
        return (iou - (c_area - union) / c_area).unsqueeze(-1)
    return iou.unsqueeze(-1)


# ----------------------------------------------------------------------------
# a7-a10  TaskAlignedAssigner.forward            losses/yolo_v8_loss.py:78-270
# ----------------------------------------------------------------------------
def _resolve(mp: torch.Tensor, ov: torch.Tensor):
    """Per-anchor conflict resolution (:226-244).  mp, ov: (B, A, G)."""
    cnt = mp.sum(-1)
    if cnt.max() > 1:
        multi = (cnt.unsqueeze(-1) > 1).expand_as(mp)
        best = ov.argmax(-1)
        onehot = torch.zeros_like(mp).scatter_(-1, best.unsqueeze(-1), 1)
        mp = torch.where(multi, onehot, mp).float()
        cnt = mp.sum(-1)
    return mp.argmax(-1), cnt, mp


@torch.no_grad()
def assign(pd_scores, pd_bboxes, anc_points, gt_labels, gt_bboxes, mask_gt,
           num_classes: int = 80, alpha: float = 0.5, beta: float = 4.0, eps: float = EPS_ASSIGN,
           ciou_assign: bool = False, topk: int = 0):
    """Returns (target_labels f32 (B,A), target_bboxes (B,A,4), target_scores (B,A,nc),
    fg_mask bool (B,A), target_gt_idx i64 (B,A)) exactly as the reference does,
    including the odd G == 0 return (:100-108).

    Non-default modes (north_star wording; the reference itself runs neither):
      ciou_assign  the pair metric is the reference's own bbox_iou(CIoU=True) (:46-56) clamped at 0 instead of plain IoU
                   (:199) -- pinned by tests/golden/assign_ciou_s98.npz, a composition of reference code;
      topk > 0     the positive mask keeps, per GT, only the `topk` anchors with the largest alignment metric among the
                   anchors inside it (metric > 0) -- what `topk` would mean if the reference read it (:70-76, :363).
                   PARITY UNPINNED BY THE REFERENCE: the semantics are this repository's (torch.topk over the anchor
                   dimension of align_metric * mask_pos); the fixture assign_topk_s99.npz only freezes them."""
    B, A = pd_scores.shape[:2]
    G = gt_bboxes.shape[1]
    if G == 0:
        z = torch.zeros_like(pd_scores[..., 0])
        return (torch.full_like(z, num_classes).long(), torch.zeros_like(pd_bboxes),
                torch.zeros_like(pd_scores), z, z.clone())

    # step 1: pair metrics (:196-208)
    ov = pair_iou_xyxy(pd_bboxes.unsqueeze(2), gt_bboxes.unsqueeze(1), ciou=ciou_assign).squeeze(-1).clamp_(0)
    cls_at_gt = pd_scores.gather(-1, gt_labels.long().unsqueeze(1).expand(-1, A, -1))
    am = cls_at_gt.pow(alpha) * ov.pow(beta)
    # step 2: anchor centre strictly inside the GT (:210-224), times GT validity (:192)
    ax = anc_points[:, 0].view(1, A, 1)
    ay = anc_points[:, 1].view(1, A, 1)
    gx1, gy1, gx2, gy2 = (gt_bboxes[..., i].unsqueeze(1) for i in range(4))
    inside = torch.stack((ax - gx1, ay - gy1, gx2 - ax, gy2 - ay), -1).amin(-1).gt(eps).float()
    mp = inside * mask_gt.unsqueeze(1)
    if topk > 0:
        metric = am * mp
        vals, idx = torch.topk(metric, min(topk, A), dim=1)
        mp = mp * torch.zeros_like(metric).scatter_(1, idx, (vals > 0).to(metric.dtype))

    # step 3: valid GT columns that own no anchor get their best-IoU anchor (:117-139;
    # the "inside" branch at :130-133 is dead: it re-tests the condition found false at :121)
    valid = mask_gt > 0
    lonely = valid & (mp.sum(1) == 0)
    for b, g in lonely.nonzero().tolist():
        mp[b, int(ov[b, :, g].argmax()), g] = 1.0

    # step 4
    tgi, fg, mp = _resolve(mp, ov)

    # step 5: sequential in g; later GTs see earlier edits (:146-162)
    for b in range(B):
        for g in range(G):
            if not valid[b, g]:
                continue
            owned = bool(((tgi[b] == g) & (fg[b] > 0)).any())
            if not owned:
                k = int(ov[b, :, g].argmax())
                mp[b, k, g] = 1.0
                tgi[b, k] = g
                fg[b, k] = 1

    # step 6: recompute from mp alone (:165)
    tgi, fg, mp = _resolve(mp, ov)

    # step 7 (:246-270)
    flat = tgi + torch.arange(B).view(B, 1) * G
    labels = gt_labels.flatten()[flat].clamp(0, num_classes)
    tboxes = gt_bboxes.reshape(-1, 4)[flat]
    onehot = torch.zeros(B, A, num_classes).scatter_(2, labels.long().unsqueeze(-1), 1)
    tscores = torch.where((fg > 0).unsqueeze(-1), onehot, torch.zeros(()))

    # step 8 (:173-178) -- reduces over GTs per anchor
    am = am * mp
    pos_am = am.amax(-1, keepdim=True)
    pos_ov = (ov * mp).amax(-1, keepdim=True)
    norm = (am * pos_ov / (pos_am + eps)).amax(-1)
    return labels, tboxes, tscores * norm.unsqueeze(-1), fg.bool(), tgi


# ----------------------------------------------------------------------------
# a4  make_anchors                               losses/yolo_v8_loss.py:541-552
# ----------------------------------------------------------------------------
def anchors_for(shapes: Sequence[Tuple[int, int]], strides: Sequence[float], offset: float = 0.5):
    pts, sts = [], []
    for (h, w), s in zip(shapes, strides):
        ys, xs = torch.meshgrid(torch.arange(h, dtype=torch.float32) + offset,
                                torch.arange(w, dtype=torch.float32) + offset, indexing="ij")
        pts.append(torch.stack((xs, ys), -1).reshape(-1, 2))
        sts.append(torch.full((h * w, 1), float(s)))
    return torch.cat(pts), torch.cat(sts)


# ----------------------------------------------------------------------------
# a3  preprocess                                 losses/yolo_v8_loss.py:501-527
# ----------------------------------------------------------------------------
def pack_targets(cls, bboxes, batch_idx, batch_size: int, imgsz_hw: torch.Tensor) -> torch.Tensor:
    """(B, G, 6) = [cls, x1, y1, x2, y2 (px), valid].  The scale is
    [imgsz[0], imgsz[1], imgsz[0], imgsz[1]] with imgsz = (H, W)*stride0 (:400, :515)."""
    if len(cls) == 0:
        return torch.zeros(batch_size, 0, 6)
    counts = [int((batch_idx == i).sum()) for i in range(batch_size)]
    out = torch.zeros(batch_size, max(counts), 6)
    scaled = bboxes * imgsz_hw.repeat(2)[:4]
    for i, n in enumerate(counts):
        if n:
            m = batch_idx == i
            out[i, :n, 0] = cls[m, 0]
            out[i, :n, 1:5] = scaled[m]
            out[i, :n, 5] = 1.0
    return out


# ----------------------------------------------------------------------------
# a5  bbox_decode + dist2bbox                    losses/yolo_v8_loss.py:529-538, :555-564
# ----------------------------------------------------------------------------
def decode_ltrb(pred_dist: torch.Tensor, reg_max: int) -> torch.Tensor:
    b, a, c = pred_dist.shape
    bins = torch.arange(reg_max, dtype=pred_dist.dtype).view(-1, 1)
    return pred_dist.view(b, a, 4, reg_max).softmax(3).matmul(bins).view(b, a, 4)


def ltrb_to_xyxy(ltrb: torch.Tensor, anchor_points: torch.Tensor) -> torch.Tensor:
    lt, rb = ltrb.chunk(2, -1)
    return torch.cat((anchor_points - lt, anchor_points + rb), -1)


# ----------------------------------------------------------------------------
# a12  BboxLoss.forward + _df_loss + bbox2dist    losses/yolo_v8_loss.py:280-330
# ----------------------------------------------------------------------------
def bbox_loss(pred_dist, pred_bboxes, anchor_points, target_bboxes, target_scores, target_scores_sum, fg_mask,
              reg_max: int):
    """Stand-alone BboxLoss.forward on dense tensors -> (loss_iou, loss_dfl); autograd attached to
    pred_dist / pred_bboxes.  Boxes in grid units."""
    fgb = fg_mask.bool()
    w = target_scores.sum(-1)[fgb].unsqueeze(-1)                                   # :297
    ciou = pair_iou_xyxy(pred_bboxes[fgb], target_bboxes[fgb], ciou=True)          # :298
    loss_iou = ((1.0 - ciou) * w).sum() / target_scores_sum                        # :299
    if reg_max <= 1:
        return loss_iou, torch.tensor(0.0)                                         # :307-308
    x1y1, x2y2 = target_bboxes.chunk(2, -1)
    hi_clamp = reg_max - 1 - 0.01
    tgt = torch.cat((anchor_points - x1y1, x2y2 - anchor_points), -1).clamp(0, hi_clamp)[fgb]   # :303, :327-330, :314
    lo = tgt.long()
    w_lo = (lo + 1) - tgt
    logits = pred_dist[fgb].view(-1, reg_max)
    ce_lo = F.cross_entropy(logits, lo.view(-1), reduction="none").view(lo.shape) * w_lo          # :320
    ce_hi = F.cross_entropy(logits, (lo + 1).view(-1), reduction="none").view(lo.shape) * (1 - w_lo)   # :321
    loss_dfl = ((ce_lo + ce_hi).mean(-1, keepdim=True) * w).sum() / target_scores_sum               # :304-305, :323
    return loss_iou, loss_dfl


# ----------------------------------------------------------------------------
# a1, a11, a12  v8DetectionLoss.__call__          losses/yolo_v8_loss.py:372-499
# ----------------------------------------------------------------------------
def detection_loss(feats: Sequence[torch.Tensor], batch: Dict[str, torch.Tensor], nc: int, reg_max: int,
                   strides: Sequence[float], hyp=(7.5, 0.5, 1.5), alpha: float = 0.5, beta: float = 4.0,
                   return_parts: bool = False, assign_boxes_px: torch.Tensor = None):
    """Returns (loss.sum()*B, loss_items.detach()) with autograd attached to `feats`.
    With return_parts also a dict of the intermediate assignment tensors.
    assign_boxes_px (B, A, 4): "spliced" mode of SURVEY §4.2 -- the assigner is fed these pixel boxes (the CUDA path's own
    decode) instead of this function's, so that a 1-ulp decode difference cannot flip a near-tie of two IoUs; the
    loss terms still use this function's own decode."""
    B = feats[0].shape[0]
    no = nc + 4 * reg_max
    flat = torch.cat([f.view(B, no, -1) for f in feats], 2)                       # :391
    pred_distri = flat[:, : 4 * reg_max].permute(0, 2, 1).contiguous()            # :396
    pred_scores = flat[:, 4 * reg_max:].permute(0, 2, 1).contiguous()             # :395
    imgsz = torch.tensor(feats[0].shape[2:], dtype=torch.float32) * float(strides[0])   # :400
    t = pack_targets(batch["cls"], batch["bboxes"], batch["batch_idx"], B, imgsz)       # :403
    gt_labels, gt_bboxes, mask_gt = t[..., 0], t[..., 1:5], (t[..., 5] > 0).float()     # :404-405
    anc, st = anchors_for([f.shape[2:] for f in feats], strides)                   # :408
    pred_bboxes = ltrb_to_xyxy(decode_ltrb(pred_distri, reg_max), anc.unsqueeze(0))  # :409
    boxes_for_assign = (pred_bboxes * st).detach() if assign_boxes_px is None else assign_boxes_px
    labels, tboxes, tscores, fg, tgi = assign(pred_scores.detach().sigmoid(),
                                              boxes_for_assign, anc * st,
                                              gt_labels, gt_bboxes, mask_gt, nc, alpha, beta)  # :415-422
    # :424-470 emergency fallback is unreachable: loop 1 guarantees a positive per valid GT.
    tss = max(tscores.sum(), 1)                                                    # :472
    loss = torch.zeros(3)
    loss[1] = F.binary_cross_entropy_with_logits(pred_scores, tscores, reduction="none").sum() / tss  # :475
    fgb = fg.bool() if fg.dtype != torch.bool else fg
    if fgb.sum():                                                                  # :478
        tb = tboxes / st                                                           # :479
        w = tscores.sum(-1)[fgb].unsqueeze(-1)                                     # :298
        ciou = pair_iou_xyxy(pred_bboxes[fgb], tb[fgb], ciou=True)                 # :299
        loss[0] = ((1.0 - ciou) * w).sum() / tss                                   # :300
        x1y1, x2y2 = tb.chunk(2, -1)
        tl_ = torch.cat((anc - x1y1, x2y2 - anc), -1).clamp_(0, reg_max - 1 - 0.01)  # :304, :327-330
        tgt = tl_[fgb].clamp_(0, reg_max - 1 - 0.01)                               # :315
        lo = tgt.long()
        hi = lo + 1
        w_lo = hi - tgt
        w_hi = 1 - w_lo
        logits = pred_distri[fgb].view(-1, reg_max)
        ce_lo = F.cross_entropy(logits, lo.view(-1), reduction="none").view(lo.shape) * w_lo   # :321
        ce_hi = F.cross_entropy(logits, hi.view(-1), reduction="none").view(lo.shape) * w_hi   # :322
        loss[2] = ((ce_lo + ce_hi).mean(-1, keepdim=True) * w).sum() / tss         # :305-306, :324
    loss[0] *= hyp[0]
    loss[1] *= hyp[1]
    loss[2] *= hyp[2]
    total = loss.sum() * B                                                         # :499
    if return_parts:
        parts = {"pred_bboxes_px": (pred_bboxes * st).detach(), "fg_mask": fgb, "target_gt_idx": tgi,
                 "target_labels": labels, "target_scores_sum": torch.as_tensor(tss).detach().clone(),
                 "norm": tscores.sum(-1)}
        return total, loss.detach(), parts
    return total, loss.detach()


# ----------------------------------------------------------------------------
# a15  calculate_iou_batch_simple                train_yolo11_cuda.py:402-437
# a17  utils.metrics.calculate_iou_batch         utils/metrics.py:49-81
# ----------------------------------------------------------------------------
def iou_matrix(boxes1: torch.Tensor, boxes2: torch.Tensor) -> torch.Tensor:
    """(N,4) x (M,4) -> (N,M): inter / (area1 + area2 - inter + 1e-6); no eps on sides."""
    a = boxes1.unsqueeze(1)
    b = boxes2.unsqueeze(0)
    iw = (torch.minimum(a[..., 2], b[..., 2]) - torch.maximum(a[..., 0], b[..., 0])).clamp(min=0)
    ih = (torch.minimum(a[..., 3], b[..., 3]) - torch.maximum(a[..., 1], b[..., 1])).clamp(min=0)
    inter = iw * ih
    area_a = (a[..., 2] - a[..., 0]) * (a[..., 3] - a[..., 1])
    area_b = (b[..., 2] - b[..., 0]) * (b[..., 3] - b[..., 1])
    return inter / (area_a + area_b - inter + 1e-6)


# ----------------------------------------------------------------------------
# a14  nms_simple                                train_yolo11_cuda.py:361-399
# ----------------------------------------------------------------------------
def nms_greedy(boxes: torch.Tensor, scores: torch.Tensor, iou_threshold: float) -> List[int]:
    """Greedy class-agnostic NMS; keep while IoU <= thr (:396).  Ties in score are
    undefined in the reference (unstable argsort); here: lower index first."""
    if len(boxes) == 0:
        return []
    order = torch.sort(scores, descending=True, stable=True).indices
    keep: List[int] = []
    while order.numel() > 0:
        head = int(order[0])
        keep.append(head)
        if order.numel() == 1:
            break
        rest = order[1:]
        ious = iou_matrix(boxes[head:head + 1], boxes[rest]).squeeze(0)
        order = rest[ious <= iou_threshold]
    return keep


def nms_class_aware(boxes, scores, labels, iou_threshold: float) -> List[int]:
    """Non-default mode: the reference's greedy NMS run once per class
    subset, keep lists unioned and ordered by (score desc, index asc)."""
    keep: List[int] = []
    for c in torch.unique(labels).tolist():
        idx = (labels == c).nonzero().squeeze(1)
        keep += [int(idx[k]) for k in nms_greedy(boxes[idx], scores[idx], iou_threshold)]
    keep.sort(key=lambda i: (-float(scores[i]), i))
    return keep


# ----------------------------------------------------------------------------
# a13  decode_predictions_for_metrics            train_yolo11_cuda.py:265-358
# ----------------------------------------------------------------------------
def decode_for_metrics(pred: torch.Tensor, img_size, conf_threshold: float, iou_threshold: float,
                       class_aware: bool = False, return_keep: bool = False):
    """pred: (B, A, 4+nc) xywh px + sigmoided scores -> list of dicts
    {boxes (K,4) normalised xyxy clamped, scores (K,), labels (K,) i64}."""
    if isinstance(pred, tuple) and len(pred) == 2:
        pred = pred[0]
    out = []
    for row in pred:
        best, cid = row[:, 4:].max(1)
        m = best > conf_threshold
        if m.sum() == 0:
            d = {"boxes": torch.zeros((0, 4)), "scores": torch.zeros((0,)),
                 "labels": torch.zeros((0,), dtype=torch.long)}
            if return_keep:
                d["keep"] = torch.zeros((0,), dtype=torch.long)
            out.append(d)
            continue
        xywh, sc, cl = row[:, :4][m], best[m], cid[m]
        x, y, w, h = xywh.unbind(1)
        xyxy = torch.stack((x - w / 2, y - h / 2, x + w / 2, y + h / 2), 1)
        keep = nms_class_aware(xyxy, sc, cl, iou_threshold) if class_aware else nms_greedy(xyxy, sc, iou_threshold)
        d = {"boxes": torch.clamp(xyxy[keep] / img_size, 0.0, 1.0), "scores": sc[keep], "labels": cl[keep]}
        if return_keep:
            d["keep"] = m.nonzero().squeeze(1)[keep]      # indices into the A anchors
        out.append(d)
    return out


# ----------------------------------------------------------------------------
# a16  DFL + Detect.inference + dist2bbox(xywh=True, dim=1)
#      models/yolo11_modules.py:180-192, :248-266, :303-312
# ----------------------------------------------------------------------------
def detect_inference(feats: Sequence[torch.Tensor], nc: int, reg_max: int, strides: Sequence[float]) -> torch.Tensor:
    """list of NCHW levels -> (B, 4+nc, A): [cx, cy, w, h]*stride, sigmoid(cls)."""
    B = feats[0].shape[0]
    no = nc + 4 * reg_max
    flat = torch.cat([f.view(B, no, -1) for f in feats], 2)
    box, cls = flat.split((4 * reg_max, nc), 1)
    anc, st = anchors_for([f.shape[2:] for f in feats], strides)
    anc, st = anc.t(), st.t()
    a = box.shape[2]
    w = torch.arange(reg_max, dtype=torch.float32).view(1, reg_max, 1, 1)
    dist = F.conv2d(box.view(B, 4, reg_max, a).transpose(2, 1).softmax(1), w).view(B, 4, a)
    lt, rb = dist.chunk(2, 1)
    x1y1 = anc.unsqueeze(0) - lt
    x2y2 = anc.unsqueeze(0) + rb
    dbox = torch.cat(((x1y1 + x2y2) / 2, x2y2 - x1y1), 1) * st
    return torch.cat((dbox, cls.sigmoid()), 1)
