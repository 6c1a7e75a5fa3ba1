"""ypp -- B200-native YOLO post-processing hot path (decode / NMS / mAP@.5 matching).

Drop-in modules mirroring the reference's helpers live under ypp.compat:
    ypp.compat.yolov3_utils   decode_predictions, DecodeYoloV3, MeanAveragePrecision
    ypp.compat.yolov2_utils   decode_predictions, DecodeYoloV2, MeanAveragePrecision
    ypp.compat.yolo_utils     nms_v3, mean_average_precision, metrics_per_class, get_target_boxes_for_map
    ypp.compat.metrics        box_iou, ConfusionMatrix
    ypp.compat.writers        pred_file_lines, write_pred_files, coco_results, write_coco_results
    ypp.compat.general        non_max_suppression (the vendored Ultralytics semantics), xywh2xyxy
    ypp.compat.yolov1_utils   decode_predictions, DecodeYoloV1, MeanAveragePrecision
The CUDA library (ypp/libypp.so, built from csrc/) is mandatory; there is no CPU fallback.
"""
from . import synth  # noqa: F401  (pure torch, no extension needed)

__all__ = ["synth", "PostProcessor", "PostResult", "target_rows", "box_iou", "map_match", "launch_count"]


def __getattr__(name):
    # the op layer is imported lazily so that `import ypp.synth` works on a box without the .so,
    # while any use of an op without the extension raises ImportError (no silent fallback)
    if name in ("PostProcessor", "PostResult", "target_rows", "box_iou", "map_match", "launch_count", "engine"):
        import importlib
        engine = importlib.import_module(__name__ + ".engine")
        return engine if name == "engine" else getattr(engine, name)
    raise AttributeError(name)
