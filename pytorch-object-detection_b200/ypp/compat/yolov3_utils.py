"""Drop-in for the reference's utils/yolov3_utils.py (YOLOv3 / YOLOv4-tiny heads).

    decode_predictions      utils/yolov3_utils.py:12-59
    DecodeYoloV3            utils/yolov3_utils.py:62-90
    MeanAveragePrecision    utils/yolov3_utils.py:93-149
"""
from torch import nn

from .. import engine
from ._accum import LiveParams, MeanAveragePrecisionBase
from .yolo_utils import nms_v3, mean_average_precision  # noqa: F401  (re-exported like the reference)


def decode_predictions(input, num_classes, anchors, input_size):
    """[B, A*(5+nc), H, W] -> [B, A*H*W, 6] rows cx,cy,w,h,score,cls; anchors in pixels."""
    pp = engine.PostProcessor(num_classes, anchors, input_size, kind="v3", anchors_per_scale=len(anchors))
    return pp.decode([input])


class DecodeYoloV3(LiveParams, nn.Module):
    """Decode YOLOv3 predictions of ONE image to boxes [K,6] (decode + nms_v3)."""

    def __init__(self, num_classes, anchors, input_size, conf_threshold=0.5, **modes):
        super().__init__()
        self.num_classes = num_classes
        self.anchors = anchors
        self.input_size = input_size
        self.conf_threshold = conf_threshold
        self._modes = modes

    def _make_processor(self):
        return engine.PostProcessor(self.num_classes, self.anchors, self.input_size, self.conf_threshold, kind="v3",
                                    **self._modes)

    def forward(self, x):
        assert x[0].size(0) == 1
        return self._pp(list(x), row_width=6).rows()

    def forward_batch(self, x):
        """Batched generalisation: list of per-image [K_b,6] tensors."""
        return self._pp(list(x), row_width=6).per_image()

    def forward_features(self, features, weights):
        """forward() fed by the inputs of the heads' 1x1 convolutions (models/detector/yolov3.py:56,62,68) and those
        layers' weights instead of their outputs: convolution, decode and NMS without the head tensors."""
        assert features[0].size(0) == 1
        return self._pp.from_features(list(features), list(weights), row_width=6).rows()


class MeanAveragePrecision(MeanAveragePrecisionBase):
    def __init__(self, num_classes, anchors, input_size, conf_threshold, **modes):
        super().__init__(num_classes, anchors, input_size, conf_threshold, kind="v3", **modes)

    def update_state(self, y_true, y_preds):
        self._update(y_true, list(y_preds))
