"""Drop-in for the hot-path pieces of utils/metrics.py on the GPU.

    box_iou            utils/metrics.py:233-255
    ConfusionMatrix    utils/metrics.py:113-160  (process_batch)
"""
import numpy as np
import torch

from .. import engine
from ..engine import box_iou  # noqa: F401


class ConfusionMatrix:
    """Same constructor, process_batch(detections [N,6] xyxy,conf,cls; labels [M,5] cls,xyxy) and .matrix
    ((nc+1) x (nc+1), predicted x true, last index = background).  The counts live on the device and
    process_batch never synchronises; .matrix copies them out.  process_images() takes a whole batch."""

    def __init__(self, nc, conf=0.25, iou_thres=0.45, device=None):
        self.nc = nc
        self.conf = conf
        self.iou_thres = iou_thres
        self._dev = torch.device(device) if device is not None else None
        self._m = None

    def _matrix(self, dev):
        if self._m is None:
            self._dev = dev
            self._m = torch.zeros((self.nc + 1, self.nc + 1), dtype=torch.int64, device=dev)
        return self._m

    def process_batch(self, detections, labels):
        off_d = torch.tensor([0, detections.shape[0]], dtype=torch.int32)
        off_l = torch.tensor([0, labels.shape[0]], dtype=torch.int32)
        self.process_images(detections, off_d, labels, off_l)

    def process_images(self, detections, det_off, labels, lab_off):
        """detections [K,6] and labels [G,5] of several images, row offsets [n_img+1] each."""
        if not detections.is_cuda:
            raise ValueError("detections must live on a CUDA device (ypp has no CPU path)")
        m = self._matrix(detections.device)
        engine.confusion_update(m, detections, det_off, labels.to(detections.device), lab_off, self.nc, self.conf, self.iou_thres)

    @property
    def matrix(self):
        if self._m is None:
            return np.zeros((self.nc + 1, self.nc + 1))
        return self._m.cpu().numpy().astype(np.float64)
