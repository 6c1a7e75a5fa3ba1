"""Shared accumulator behind MeanAveragePrecision (v2 and v3 flavours)."""
import torch

from .. import engine


class _GrowBuf:
    """[n,7] rows on the device with amortised doubling (replaces the O(total^2) torch.cat
    growth of utils/yolov3_utils.py:139-144)."""

    def __init__(self):
        self.buf = None
        self.n = 0

    def append(self, rows):
        k = rows.shape[0]
        if k == 0 and self.buf is not None:
            return
        if self.buf is None:
            self.buf = torch.empty((max(1024, 2 * k), 7), dtype=torch.float32, device=rows.device)
        if self.n + k > self.buf.shape[0]:
            nb = torch.empty((max(2 * self.buf.shape[0], self.n + k), 7), dtype=torch.float32, device=rows.device)
            nb[: self.n] = self.buf[: self.n]
            self.buf = nb
        self.buf[self.n: self.n + k] = rows
        self.n += k

    def view(self):
        return self.buf[: self.n]


class MeanAveragePrecisionBase:
    """reset_states / update_state / result with the reference's public attributes
    (utils/yolov3_utils.py:93-149, utils/yolov2_utils.py:81-128)."""

    def __init__(self, num_classes, anchors, input_size, conf_threshold, kind, **modes):
        self.num_classes = num_classes
        self.anchors = anchors
        self.input_size = input_size
        self.conf_threshold = conf_threshold
        self._pp = engine.PostProcessor(num_classes, anchors, input_size, conf_threshold, kind=kind, **modes)
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

    def update_state_features(self, y_true, features, weights):
        """update_state fed by the INPUTS of the heads' 1x1 convolutions and their weights (the fused head,
        PostProcessor.from_features): the head tensors are never materialised."""
        self._update(y_true, list(features), list(weights))

    def _update(self, y_true, heads, weights=None):
        dev = heads[0].device
        if weights is None:
            res = self._pp(heads, img_idx_base=self.img_idx, row_width=7)
        else:
            res = self._pp.from_features(heads, weights, img_idx_base=self.img_idx, row_width=7)
        gt_rows, gt_cnt = engine.target_rows(y_true.to(dev, non_blocking=True), self.input_size, self.img_idx)
        n_det, n_gt = torch.cat([res.det_total, gt_cnt]).tolist()      # the one host sync per batch
        self._pred.append(res.det_rows[:n_det])
        self._true.append(gt_rows[:n_gt])
        self.img_idx += int(y_true.size(0))

    def result(self):
        from .yolo_utils import mean_average_precision
        return mean_average_precision(self._true.view(), self._pred.view(), self.num_classes)
