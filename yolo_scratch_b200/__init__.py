"""yolo_scratch_b200 -- B200-native (sm_100a) detection-head hot path of Pratye/yolo-scratch:
task-aligned assignment, DFL + CIoU + BCE loss with backward, dist2bbox decode and batched NMS,
behind the reference's own Python call signatures.  See DESIGN.md / INTEGRATION.md."""
__version__ = "0.1.0"
