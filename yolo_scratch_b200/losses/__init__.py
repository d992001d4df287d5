"""Loss functions of the detection-head hot path -- same re-exports as the reference's
losses/__init__.py:3-11."""
from .yolo_v8_loss import (
    v8DetectionLoss, BboxLoss, TaskAlignedAssigner,
    bbox_iou, bbox2dist, make_anchors, dist2bbox,
)

__all__ = [
    "v8DetectionLoss", "BboxLoss", "TaskAlignedAssigner",
    "bbox_iou", "bbox2dist", "make_anchors", "dist2bbox",
]
