"""Drop-in for the NMS of the vendored Ultralytics helpers, utils/general.py.

    xywh2xyxy              utils/general.py:546-553
    non_max_suppression    utils/general.py:640-731
The reference loops over the images in Python and calls torchvision.ops.nms once per image; here the
whole batch is four kernel launches.  Same arguments and return convention: a list with one [k,6]
tensor x1,y1,x2,y2,conf,cls per image.
"""
import torch

from .. import engine


def xywh2xyxy(x):
    y = x.clone()
    y[:, 0] = x[:, 0] - x[:, 2] / 2
    y[:, 1] = x[:, 1] - x[:, 3] / 2
    y[:, 2] = x[:, 0] + x[:, 2] / 2
    y[:, 3] = x[:, 1] + x[:, 3] / 2
    return y


def non_max_suppression(prediction, conf_thres=0.25, iou_thres=0.45, classes=None, agnostic=False, multi_label=False,
                        labels=(), max_det=300, **modes):
    """prediction [B, n, 5+nc] on a CUDA device.  `labels` (autolabelling, :667-674) is not provided.
    The reference's 10 s time limit (:726-728) has no counterpart: nothing here loops over images."""
    if labels:
        raise NotImplementedError("non_max_suppression(labels=...) (autolabelling) is not provided")
    res = engine.nms_ultralytics(prediction, conf_thres, iou_thres, classes, agnostic, multi_label, max_det, **modes)
    counts = res.keep_count.tolist()                       # one host sync, as the per-image loop needs the sizes
    out, off = [], 0
    for k in counts:
        out.append(res.det_rows[off: off + k])
        off += k
    return out
