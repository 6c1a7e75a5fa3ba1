"""Drop-in for the reference's utils/yolov2_utils.py (single-scale YOLOv2 head).

    decode_predictions      utils/yolov2_utils.py:12-58   (anchors already in grid units)
    DecodeYoloV2            utils/yolov2_utils.py:61-78
    MeanAveragePrecision    utils/yolov2_utils.py:81-128
"""
from torch import nn

from .. import engine
from ._accum import LiveParams, MeanAveragePrecisionBase
from .yolo_utils import nms_v3, mean_average_precision  # noqa: F401


def decode_predictions(input, num_classes, scaled_anchors, input_size):
    pp = engine.PostProcessor(num_classes, scaled_anchors, input_size, kind="v2")
    return pp.decode([input])


class DecodeYoloV2(LiveParams, nn.Module):
    def __init__(self, num_classes, scaled_anchors, input_size, conf_threshold=0.5, **modes):
        super().__init__()
        self.num_classes = num_classes
        self.scaled_anchors = scaled_anchors
        self.input_size = input_size
        self.conf_threshold = conf_threshold
        self._modes = modes

    def _make_processor(self):
        return engine.PostProcessor(self.num_classes, self.scaled_anchors, self.input_size, self.conf_threshold, kind="v2",
                                    **self._modes)

    def forward(self, x):
        assert x.size(0) == 1
        return self._pp([x], row_width=6).rows()


class MeanAveragePrecision(MeanAveragePrecisionBase):
    def __init__(self, num_classes, scaled_anchors, input_size, conf_threshold, **modes):
        super().__init__(num_classes, scaled_anchors, input_size, conf_threshold, kind="v2", **modes)
        self.scaled_anchors = scaled_anchors

    def _make_processor(self):
        return engine.PostProcessor(self.num_classes, self.scaled_anchors, self.input_size, self.conf_threshold,
                                    kind="v2", **self._modes)

    def update_state(self, y_true, y_pred):
        self._update(y_true, [y_pred])
