"""Drop-in for the IoU helpers of the reference's utils/metrics.py (:19-81), on the B200 kernels."""
from __future__ import annotations

import torch

from .. import _lib
from .._lib import check, ptr, stream_ptr


def calculate_iou_batch(boxes1, boxes2):
    """(N,4) x (M,4) xyxy -> (N,M) IoU = inter / (area1 + area2 - inter + 1e-6)
    (reference utils/metrics.py:49-81)."""
    _lib.require_cuda(boxes1, "boxes1")
    _lib.require_cuda(boxes2, "boxes2")
    b1 = boxes1.detach().float().contiguous().view(-1, 4)
    b2 = boxes2.detach().float().contiguous().view(-1, 4)
    out = torch.empty(b1.shape[0], b2.shape[0], device=b1.device, dtype=torch.float32)
    with _lib.on_device(b1, b2):
        check(_lib.lib().ys_pairwise_iou(ptr(b1), b1.shape[0], ptr(b2), b2.shape[0], 0, ptr(out), stream_ptr(b1.device)),
              "ys_pairwise_iou")
    return out


def calculate_iou(box1, box2):
    """Single-pair IoU with the reference's `union <= 0 -> 0.0` guard (utils/metrics.py:19-46).
    Scalar helper, not on the hot path: evaluated with tensor ops on whatever device the boxes are on."""
    x1 = torch.max(box1[0], box2[0]); y1 = torch.max(box1[1], box2[1])
    x2 = torch.min(box1[2], box2[2]); y2 = torch.min(box1[3], box2[3])
    inter = torch.clamp(x2 - x1, min=0.0) * torch.clamp(y2 - y1, min=0.0)
    union = (box1[2] - box1[0]) * (box1[3] - box1[1]) + (box2[2] - box2[0]) * (box2[3] - box2[1]) - inter
    if union <= 0:
        return 0.0
    return inter / union


# ----------------------------------------------------------------------------------------------
# mAP scoring: the matching loops run on the device, the AP arithmetic on the host
# ----------------------------------------------------------------------------------------------
def calculate_ap(tp, fp, n_gt):
    """All-point-interpolated AP from the scores of the true and false positives (reference utils/metrics.py:277-323):
    detections sorted by score (stable, true positives listed first), precision = tp/(tp+fp+1e-6), recall = tp/n_gt,
    monotone precision envelope, sum of recall steps times precision."""
    import numpy as np
    if n_gt == 0:
        return 0.0
    tp = np.asarray(tp, dtype=np.float64).reshape(-1)
    fp = np.asarray(fp, dtype=np.float64).reshape(-1)
    if tp.size + fp.size == 0:
        return 0.0
    # the reference's `sorted(tp + fp, reverse=True)` is stable: equal scores keep the listed order, true positives first
    order = np.argsort(-np.concatenate([tp, fp]), kind="stable")
    flags = np.concatenate([np.ones(tp.size, dtype=np.int64), np.zeros(fp.size, dtype=np.int64)])[order]
    ctp, cfp = np.cumsum(flags), np.cumsum(1 - flags)
    prec = ctp / (ctp + cfp + 1e-6)
    rec = ctp / n_gt
    mrec = np.concatenate([[0.0], rec, [1.0]])
    mpre = np.concatenate([[0.0], prec, [0.0]])
    mpre = np.maximum.accumulate(mpre[::-1])[::-1]          # the precision envelope (:311-313), without the Python loop
    idx = np.where(mrec[1:] != mrec[:-1])[0]
    return np.sum((mrec[idx + 1] - mrec[idx]) * mpre[idx + 1])


def match_detections(predictions, targets, conf_threshold, iou_thresholds):
    """Device part of evaluate_detections: returns (tp_flags, scores, n_valid, n_gt) with tp_flags (B, P, T) uint8 on
    the host, scores (B, P) fp32 on the host (score-descending per image), n_valid (B,), n_gt total."""
    import ctypes as C
    import numpy as np
    B = len(predictions)
    dev = None
    for d in list(predictions) + list(targets):
        if d["boxes"].is_cuda:
            dev = d["boxes"].device
            break
    if dev is None:
        raise _lib.YsError("evaluate_detections: the B200 path needs CUDA tensors (there is no CPU fallback)")
    P = max(1, max(int(p["boxes"].shape[0]) for p in predictions))
    pb = torch.zeros(B, P, 4, device=dev)
    ps = torch.full((B, P), float("-inf"), device=dev)
    pc = torch.zeros(B, dtype=torch.int32, device=dev)
    pcounts = [int(p["boxes"].shape[0]) for p in predictions]
    if sum(pcounts):                                            # one scatter for the whole batch instead of three ops per image
        rows = torch.repeat_interleave(torch.arange(B), torch.tensor(pcounts)).to(dev)
        cols = torch.cat([torch.arange(n) for n in pcounts]).to(dev)
        pb[rows, cols] = torch.cat([p["boxes"].to(dev, torch.float32).view(-1, 4) for p in predictions], 0)
        ps[rows, cols] = torch.cat([p["scores"].to(dev, torch.float32).view(-1) for p in predictions], 0)
    pc = torch.tensor(pcounts, dtype=torch.int32, device=dev)
    ps, order = torch.sort(ps, dim=1, descending=True, stable=True)            # :170 (scores are unique in practice)
    pb = torch.gather(pb, 1, order.unsqueeze(-1).expand(-1, -1, 4)).contiguous()
    counts = [int(t["boxes"].shape[0]) for t in targets]
    off = torch.tensor([0] + list(np.cumsum(counts)), dtype=torch.int32, device=dev)
    gt = (torch.cat([t["boxes"].to(dev, torch.float32).view(-1, 4) for t in targets], 0)
          if sum(counts) else torch.zeros(1, 4, device=dev)).contiguous()
    T = len(iou_thresholds)
    thr = (C.c_double * T)(*[float(x) for x in iou_thresholds])
    tp = torch.zeros(B, P, T, dtype=torch.uint8, device=dev)
    nv = torch.zeros(B, dtype=torch.int32, device=dev)
    with _lib.on_device(pb, gt):
        check(_lib.lib().ys_match_detections(ptr(pb), ptr(ps.contiguous()), ptr(pc), B, P, ptr(gt), ptr(off), max(counts + [0]),
                                             float(conf_threshold), thr, T, ptr(tp), ptr(nv), stream_ptr(dev)),
              "ys_match_detections")
    return tp.cpu().numpy(), ps.cpu().numpy(), nv.cpu().numpy(), int(sum(counts))


def evaluate_detections(predictions, targets, conf_threshold=0.25, iou_threshold=0.5):
    """Same signature and result dict as the reference (utils/metrics.py:84-274): precision and recall at IoU 0.5,
    mAP50 = AP at the first threshold, mAP50-95 = mean AP over the threshold ladder."""
    import numpy as np
    if iou_threshold == 0.5:
        ladder = np.arange(0.5, 0.95 + 1e-6, 0.05)                                                   # :134
    else:
        ladder = np.arange(iou_threshold, min(1.0, iou_threshold + 0.45) + 1e-6, 0.05)              # :137
    thresholds = list(ladder) + [0.5]                      # the last column serves precision / recall (:252)
    if len(predictions) == 0:
        return {"precision": 0.0, "recall": 0.0, "mAP50": 0.0, "mAP50-95": 0.0 if len(ladder) == 0 else float(np.mean([0.0] * len(ladder)))}
    tp, scores, nv, n_gt = match_detections(predictions, targets, conf_threshold, thresholds)
    aps = []
    valid = np.arange(tp.shape[1])[None, :] < nv[:, None]                    # (B, P): detections above the confidence cut
    for t in range(len(ladder)):
        f = tp[:, :, t].astype(bool)
        # row-major masks: image by image, score-descending inside an image -- the order the reference lists them in
        aps.append(calculate_ap(scores[valid & f], scores[valid & ~f], n_gt))
    n_tp = int(tp[:, :, -1][valid].sum())
    n_det = int(nv.sum())
    precision = n_tp / n_det if n_det > 0 else 0.0
    recall = n_tp / n_gt if n_gt > 0 else 0.0
    return {"precision": precision, "recall": recall, "mAP50": aps[0] if aps else 0.0,
            "mAP50-95": np.mean(aps) if aps else 0.0}


def cda_nms(predictions, max_detections=10, dga_thresh=0.15):
    """Same signature and result dict as the reference (utils/metrics.py:484-620): greedy suppression with the
    Gaussian-angle distance of the inscribed ellipses, top `max_detections` kept; keys boxes / scores / angles
    (+ labels when given), in score-descending order."""
    boxes, scores = predictions["boxes"], predictions["scores"]
    angles = predictions.get("angles", None)
    labels = predictions.get("labels", None)
    if len(boxes) == 0:
        out = {"boxes": boxes, "scores": scores, "angles": angles if angles is not None else torch.zeros(0, device=boxes.device)}
        if labels is not None:
            out["labels"] = labels
        return out
    _lib.require_cuda(boxes, "predictions['boxes']")
    dev = boxes.device
    order = torch.argsort(scores, descending=True, stable=True)                     # :569 (scores unique in practice)
    sb = boxes.detach().float()[order].contiguous().view(1, -1, 4)
    sa = angles.detach().float()[order].contiguous().view(1, -1) if angles is not None else None
    n = sb.shape[1]
    cnt = torch.tensor([n], dtype=torch.int32, device=dev)
    keep = torch.zeros(1, max_detections, dtype=torch.int32, device=dev)
    kc = torch.zeros(1, dtype=torch.int32, device=dev)
    with _lib.on_device(sb, sa):
        check(_lib.lib().ys_cda_nms(ptr(sb), ptr(sa) if sa is not None else None, ptr(cnt), 1, n, int(max_detections),
                                    float(dga_thresh), ptr(keep), ptr(kc), stream_ptr(dev)), "ys_cda_nms")
    kept = order[keep[0, :int(kc.item())].long()]
    out = {"boxes": boxes[kept], "scores": scores[kept],
           "angles": angles[kept] if angles is not None else torch.zeros(len(kept), device=dev)}
    if labels is not None:
        out["labels"] = labels[kept]
    return out
