"""Drop-in for the reference's utils/yolov1_utils.py.

    decode_predictions      utils/yolov1_utils.py:12-66
    DecodeYoloV1            utils/yolov1_utils.py:69-86
    MeanAveragePrecision    utils/yolov1_utils.py:89-135
"""
import torch
from torch import nn

from .. import engine
from ._accum import _GrowBuf
from .yolo_utils import _pp, nms_v3, mean_average_precision  # noqa: F401


def decode_predictions(input, num_classes, num_boxes, input_size):
    return engine.decode_yolov1(input, num_classes, num_boxes, input_size)


class DecodeYoloV1(nn.Module):
    def __init__(self, num_classes, num_boxes, input_size, conf_threshold=0.5, **modes):
        super().__init__()
        self.num_classes = num_classes
        self.num_boxes = num_boxes
        self.input_size = input_size
        self.conf_threshold = conf_threshold
        self._modes = modes

    def forward(self, x):
        assert x.size(0) == 1
        decode_pred = decode_predictions(x, self.num_classes, self.num_boxes, self.input_size)
        return nms_v3(decode_pred[0], conf_threshold=self.conf_threshold, num_classes=self.num_classes, **self._modes)


class MeanAveragePrecision:
    """update_state decodes the batch and runs the batched NMS in one chain (the reference loops over the
    images and grows its two tensors with torch.cat, :105-131)."""

    def __init__(self, num_classes, num_boxes, input_size, conf_threshold, **modes):
        self.num_classes = num_classes
        self.num_boxes = num_boxes
        self.input_size = input_size
        self.conf_threshold = conf_threshold
        self._modes = modes
        self.reset_states()

    def reset_states(self):
        self._true = _GrowBuf()
        self._pred = _GrowBuf()
        self.img_idx = 0

    @property
    def all_true_boxes_variable(self):
        return self._true.view() if self.img_idx else 0

    @property
    def all_pred_boxes_variable(self):
        return self._pred.view() if self.img_idx else 0

    def update_state(self, y_true, y_pred):
        dev = y_pred.device
        rows = decode_predictions(y_pred, self.num_classes, self.num_boxes, self.input_size)
        res = _pp(self.num_classes, self.conf_threshold, 0.45, **self._modes).nms_rows(rows, img_idx_base=self.img_idx, row_width=7)
        gt_rows, gt_cnt = engine.target_rows(y_true.to(dev, non_blocking=True), self.input_size, self.img_idx)
        n_det, n_gt = torch.cat([res.det_total, gt_cnt]).tolist()
        self._pred.append(res.det_rows[:n_det])
        self._true.append(gt_rows[:n_gt])
        self.img_idx += int(y_true.size(0))

    def result(self):
        return mean_average_precision(self._true.view(), self._pred.view(), self.num_classes)
