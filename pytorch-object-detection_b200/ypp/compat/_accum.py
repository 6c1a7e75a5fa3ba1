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


class LiveParams:
    """The reference reads num_classes / anchors / input_size / conf_threshold from the object on every call
    (utils/yolov3_utils.py:79,88,113-125), so assigning one of them after construction changes the next call.
    Same here: assigning any name in `_live` drops the processor, and the next call builds one from the current
    values.  (Mutating the anchors list in place is not seen; assign a new list.)"""
    _live = ("num_classes", "anchors", "scaled_anchors", "input_size", "conf_threshold")

    def __setattr__(self, name, value):
        if name in self._live:
            self.__dict__.pop("_pp", None)
        super().__setattr__(name, value)

    def __getattr__(self, name):
        if name == "_pp":
            pp = self._make_processor()
            self.__dict__["_pp"] = pp
            return pp
        parent = getattr(super(), "__getattr__", None)        # nn.Module resolves parameters / buffers / submodules here
        if parent is not None:
            return parent(name)
        raise AttributeError("%r object has no attribute %r" % (type(self).__name__, name))


class MeanAveragePrecisionBase(LiveParams):
    """reset_states / update_state / result with the reference's public attributes
    (utils/yolov3_utils.py:93-149, utils/yolov2_utils.py:81-128)."""

    def __init__(self, num_classes, anchors, input_size, conf_threshold, kind, sync_every=16, **modes):
        self.sync_every = sync_every
        self.num_classes = num_classes
        self.anchors = anchors
        self.input_size = input_size
        self.conf_threshold = conf_threshold
        self._kind, self._modes = kind, modes
        self.reset_states()

    def _make_processor(self):
        return engine.PostProcessor(self.num_classes, self.anchors, self.input_size, self.conf_threshold,
                                    kind=self._kind, **self._modes)

    def reset_states(self):
        # validation epochs repeat: the next epoch's accumulators start at the size this one ended with (an upper
        # bound known without reading the device), so they are allocated once instead of doubling their way up
        prev = self.__dict__.get("_pred"), self.__dict__.get("_true")
        hints = self.__dict__.get("_hints", (0, 0))
        self._hints = tuple(max(h, min(a.bound, a.buf.shape[0]) if a is not None and a.buf is not None else 0) for h, a in zip(hints, prev))
        self._true = None          # RowAccumulators, made on the first batch's device
        self._pred = None
        self.img_idx = 0

    @property
    def all_true_boxes_variable(self):
        return self._true.view() if self._true is not None else 0

    @property
    def all_pred_boxes_variable(self):
        return self._pred.view() if self._pred is not None else 0

    def update_state_features(self, y_true, features, weights):
        """update_state fed by the INPUTS of the heads' 1x1 convolutions and their weights (the fused head,
        PostProcessor.from_features): the head tensors are never materialised."""
        self._update(y_true, list(features), list(weights))

    def _update(self, y_true, heads, weights=None):
        dev = heads[0].device
        if self._pred is None:
            self._pred = engine.RowAccumulator(dev, self.sync_every, self._hints[0])
            self._true = engine.RowAccumulator(dev, 64 * self.sync_every, self._hints[1])
        # both appends happen on the device behind the accumulators' cursors: no host read per batch
        self._pp.accumulate(self._pred, heads, weights, img_idx_base=self.img_idx)
        engine.target_rows_append(self._true, y_true.to(dev, non_blocking=True), self.input_size, self.img_idx)
        self.img_idx += int(y_true.size(0))

    def result(self):
        from .yolo_utils import mean_average_precision
        if self._pred is None:
            raise RuntimeError("result() before any update_state()")
        true, pred = engine.views(self._true, self._pred)                  # one host read for both counts
        return mean_average_precision(true, pred, self.num_classes)
