"""Drop-in for the hot-path functions of the reference's utils/yolo_utils.py.

    nms_v3                    utils/yolo_utils.py:321-348
    mean_average_precision    utils/yolo_utils.py:351-456
    metrics_per_class         utils/yolo_utils.py:458-574
    get_target_boxes_for_map  utils/yolo_utils.py:682-710
Same names, argument meaning and return conventions; tensors must live on a CUDA device.
"""
import torch

from .. import engine

_pp_cache = {}


def _pp(nc, conf, iou, **modes):
    key = (nc, float(conf), float(iou), tuple(sorted(modes.items())))
    pp = _pp_cache.get(key)
    if pp is None:
        pp = engine.PostProcessor(nc, [[1.0, 1.0]], 1, conf, iou, kind="v2", anchors_per_scale=1, **modes)
        _pp_cache[key] = pp
    return pp


def nms_v3(boxes, conf_threshold=0.25, iou_threshold=0.45, num_classes=None, **modes):
    """boxes [N,6] = cx,cy,w,h,score,cls -> rows of `boxes` kept by per-class NMS, in NMS order
    (score descending, ties by lower index); the empty [0,6] tensor when nothing passes
    (utils/yolo_utils.py:336-337).  num_classes bounds the class ids (one host sync to find it
    when omitted -- the reference itself syncs at :334)."""
    if boxes.shape[0] == 0:
        return boxes
    if num_classes is None:
        num_classes = int(boxes[..., 5].max().item()) + 1
    res = _pp(max(int(num_classes), 1), conf_threshold, iou_threshold, **modes).nms_rows(boxes, row_width=6)
    return res.rows()


def get_target_boxes_for_map(target, input_size):
    """[B,T,5] -> dict b -> FloatTensor[G_b,6] = cx,cy,w,h,1,cls (rows with sum<=0 dropped)."""
    rows, cnt = engine.target_rows(target, input_size, 0)
    rows = rows[: int(cnt.item())]
    img = rows[:, 0].to(torch.int64)
    return {b: rows[img == b][:, 1:] for b in range(target.size(0))}


def mean_average_precision(true_boxes, pred_boxes, num_classes, iou_threshold=0.5):
    """[G,7], [K,7] rows img,cx,cy,w,h,score,cls -> 0-dim mAP tensor (on the device)."""
    m, _, _ = engine.map_match(true_boxes, pred_boxes, num_classes, iou_threshold, want_tp=False)
    return m[0]


def metrics_per_class(true_boxes, pred_boxes, num_classes, iou_threshold=0.5):
    """-> [num_classes, 4] = AP, TP, FP, FN."""
    _, pc, _ = engine.map_match(true_boxes, pred_boxes, num_classes, iou_threshold, want_tp=False)
    return pc
