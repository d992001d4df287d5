"""utils of the hot path.  The reference's utils/__init__.py:1-3 re-exports evaluate_detections and
calculate_iou; only the pairwise IoU used around NMS is in scope here."""
from .metrics import calculate_iou_batch, calculate_iou

__all__ = ["calculate_iou_batch", "calculate_iou"]
