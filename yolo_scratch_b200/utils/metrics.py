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
