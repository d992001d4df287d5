"""utils of the hot path.  The reference's utils/__init__.py:1-3 re-exports evaluate_detections and
calculate_iou; the pairwise IoU used around NMS and the mAP matching (row f4) run on the B200 kernels."""
from .metrics import calculate_iou_batch, calculate_iou, evaluate_detections, calculate_ap, match_detections, cda_nms

__all__ = ["calculate_iou_batch", "calculate_iou", "evaluate_detections", "calculate_ap", "match_detections", "cda_nms"]
