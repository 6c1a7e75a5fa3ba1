"""Drop-in for utils/metrics.py:233-255 box_iou (N x M, xyxy) on the GPU."""
from ..engine import box_iou  # noqa: F401
