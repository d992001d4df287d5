"""Drop-in for the NMS / decode-for-metrics functions of the reference's train script
(train_yolo11_cuda.py:265-437): same names, arguments and return types, on the B200 kernels
(two optional, non-reference keyword arguments are added: `labels=` on nms_simple and `class_aware=` on
decode_predictions_for_metrics; with their defaults the behaviour is the reference's).

    decode_predictions_for_metrics(preds, img_size, conf_threshold, iou_threshold, device) -> list[dict]
    nms_simple(boxes, scores, iou_threshold) -> list[int]
    calculate_iou_batch_simple(boxes1, boxes2)

Ties in score: the reference's argsort is unstable; here the lower index goes first.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr

_ws = _lib.Workspace()


def nms_batched(pred, conf_threshold, iou_threshold, img_size, class_aware=False, layout=1):
    """Batched device-side NMS.  pred: (B, A, 4+nc) [layout 1] or (B, 4+nc, A) [layout 0].
    Returns device tensors (keep_anchor i32 (B,A), boxes (B,A,4), scores (B,A), labels i32 (B,A), count i32 (B));
    rows [0, count[b]) of image b are valid and score-descending.  No host synchronisation."""
    _lib.require_cuda(pred, "pred")
    p = pred.detach().float().contiguous()
    if layout == 1:
        B, A, R = p.shape
    else:
        B, R, A = p.shape
    nc = R - 4
    dev = p.device
    lib = _lib.lib()
    ka = torch.empty(B, A, device=dev, dtype=torch.int32)
    kb = torch.empty(B, A, 4, device=dev)
    ks = torch.empty(B, A, device=dev)
    kl = torch.empty(B, A, device=dev, dtype=torch.int32)
    kc = torch.empty(B, device=dev, dtype=torch.int32)
    ws = _ws.get(lib.ys_nms_workspace_bytes(B, A), dev)
    with _lib.on_device(p):
        check(lib.ys_nms_batched(ptr(p), layout, B, A, nc, float(conf_threshold), float(iou_threshold), float(img_size),
                                 int(bool(class_aware)), ptr(ka), ptr(kb), ptr(ks), ptr(kl), ptr(kc), ptr(ws), ws.numel(),
                                 stream_ptr(dev)), "ys_nms_batched")
    return ka, kb, ks, kl, kc


def decode_predictions_for_metrics(preds, img_size, conf_threshold, iou_threshold, device, class_aware=False):
    """Reference train_yolo11_cuda.py:265-358.  `preds` is the inference output (decoded, raw) or the decoded
    tensor (B, A, 4+nc): xywh pixels + sigmoided class scores.  One host sync (the per-image counts)."""
    pred = preds[0] if isinstance(preds, tuple) and len(preds) == 2 else preds
    ka, kb, ks, kl, kc = nms_batched(pred, conf_threshold, iou_threshold, img_size, class_aware=class_aware, layout=1)
    counts = kc.tolist()
    out = []
    for b, k in enumerate(counts):
        if k == 0:
            out.append({"boxes": torch.zeros((0, 4), device=device), "scores": torch.zeros((0,), device=device),
                        "labels": torch.zeros((0,), dtype=torch.long, device=device)})
        else:
            out.append({"boxes": kb[b, :k], "scores": ks[b, :k], "labels": kl[b, :k].long()})
    return out


def nms_simple(boxes, scores, iou_threshold, labels=None):
    """Greedy NMS, returns the kept indices as a Python list in score-descending order
    (reference train_yolo11_cuda.py:361-399).  `labels` (optional, non-reference) makes it class-aware."""
    if len(boxes) == 0:
        return []
    _lib.require_cuda(boxes, "boxes")
    b = boxes.detach().float().contiguous().view(-1, 4)
    s = scores.detach().float().contiguous().view(-1)
    l = labels.detach().to(torch.int32).contiguous().view(-1) if labels is not None else None
    N = b.shape[0]
    dev = b.device
    lib = _lib.lib()
    keep = torch.empty(N, device=dev, dtype=torch.int32)
    cnt = torch.empty(1, device=dev, dtype=torch.int32)
    ws = _ws.get(lib.ys_nms_workspace_bytes(1, N), dev)
    with _lib.on_device(b, s, l):
        check(lib.ys_nms_single(ptr(b), ptr(s), ptr(l), N, float(iou_threshold), ptr(keep), ptr(cnt), ptr(ws), ws.numel(),
                                stream_ptr(dev)), "ys_nms_single")
    return keep[: int(cnt.item())].tolist()


def calculate_iou_batch_simple(boxes1, boxes2):
    """Reference train_yolo11_cuda.py:402-437: (1,4)|(N,4) x (M,4) -> (N,M), squeezed when N == 1."""
    if boxes1.dim() == 1:
        boxes1 = boxes1.unsqueeze(0)
    _lib.require_cuda(boxes1, "boxes1")
    b1 = boxes1.detach().float().contiguous()
    b2 = boxes2.detach().float().contiguous().view(-1, 4)
    out = torch.empty(b1.shape[0], b2.shape[0], device=b1.device, dtype=torch.float32)
    with _lib.on_device(b1, b2):
        check(_lib.lib().ys_pairwise_iou(ptr(b1), b1.shape[0], ptr(b2), b2.shape[0], 0, ptr(out), stream_ptr(b1.device)),
              "ys_pairwise_iou")
    return out.squeeze() if out.shape[0] == 1 else out      # `iou.squeeze()` (:437): (1,M) -> (M,), (1,1) -> 0-dim
