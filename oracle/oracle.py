"""ctypes front-end of the CPU oracle (oracle/ypp_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(pytorch-object-detection_b200/ypp) never imports this module.

Function names mirror the reference symbols they restate:
  decode_predictions_v3  utils/yolov3_utils.py:12-59
  decode_predictions_v2  utils/yolov2_utils.py:12-58
  decode_multi_scale     utils/yolov3_utils.py:76-84 / :112-120 (scale loop + cat)
  nms_v3                 utils/yolo_utils.py:321-348 (+ torchvision batched_nms)
  get_target_rows        utils/yolo_utils.py:682-710 + utils/yolov3_utils.py:134-137
  mean_average_precision utils/yolo_utils.py:351-456, metrics_per_class :458-574
  box_iou                utils/metrics.py:233-255
  bbox_iou_cxcywh        utils/yolo_utils.py:75-123,143-144 (IoU branch)
Parity status: pinned against the reference's own outputs (tests/golden/).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "ypp_oracle.c")
_LIB = os.path.join(_HERE, "libypp_oracle.so")

OFFSET_NONE, OFFSET_TRICK, OFFSET_AUTO_CPU, OFFSET_AUTO_CUDA = 0, 1, 2, 3
UNION_SEPARATE, UNION_FMA = 0, 1


def build(force=False):
    """gcc the C restatement into oracle/libypp_oracle.so (idempotent)."""
    if (not force and os.path.exists(_LIB)
            and os.path.getmtime(_LIB) >= os.path.getmtime(_SRC)):
        return _LIB
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-shared",
           "-fPIC", "-o", _LIB, _SRC, "-lm"]
    subprocess.run(cmd, check=True)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB)
        fp = ctypes.POINTER(ctypes.c_float)
        ip = ctypes.POINTER(ctypes.c_int)
        up = ctypes.POINTER(ctypes.c_ubyte)
        L.ypo_decode_scale.argtypes = [fp, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                       ctypes.c_int, fp, fp, ctypes.c_float, ctypes.c_float, fp,
                                       ctypes.c_int64, ctypes.c_int64]
        L.ypo_decode_scale.restype = None
        L.ypo_effective_thr.argtypes = [ctypes.c_double, ctypes.c_int]
        L.ypo_effective_thr.restype = ctypes.c_float
        L.ypo_nms_xyxy.argtypes = [fp, fp, ctypes.c_int, ctypes.c_float, ctypes.c_int, ip]
        L.ypo_nms_xyxy.restype = ctypes.c_int
        L.ypo_batched_nms.argtypes = [fp, fp, fp, ctypes.c_int, ctypes.c_float, ctypes.c_int,
                                      ctypes.c_int, ip]
        L.ypo_batched_nms.restype = ctypes.c_int
        L.ypo_nms_v3.argtypes = [fp, ctypes.c_int, ctypes.c_float, ctypes.c_float, ctypes.c_int,
                                 ctypes.c_int, ip, ip]
        L.ypo_nms_v3.restype = ctypes.c_int
        L.ypo_nms_v3_batch.argtypes = [fp, ctypes.c_int, ctypes.c_int, ctypes.c_float,
                                       ctypes.c_float, ctypes.c_int, ctypes.c_int, ip, ip, ip]
        L.ypo_nms_v3_batch.restype = None
        L.ypo_bbox_iou_cxcywh_pairs.argtypes = [fp, fp, ctypes.c_int, fp]
        L.ypo_bbox_iou_cxcywh_pairs.restype = None
        L.ypo_box_iou.argtypes = [fp, ctypes.c_int, fp, ctypes.c_int, fp]
        L.ypo_box_iou.restype = None
        L.ypo_target_rows.argtypes = [fp, ctypes.c_int, ctypes.c_int, ctypes.c_float,
                                      ctypes.c_float, fp]
        L.ypo_target_rows.restype = ctypes.c_int
        L.ypo_mean_average_precision.argtypes = [fp, ctypes.c_int, fp, ctypes.c_int, ctypes.c_int,
                                                 ctypes.c_float, up, fp]
        L.ypo_mean_average_precision.restype = ctypes.c_float
        L.ypo_num_threads.restype = ctypes.c_int
        L.ypo_set_num_threads.argtypes = [ctypes.c_int]
        _lib = L
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _ip(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int))


def num_threads():
    return lib().ypo_num_threads()


def set_num_threads(n):
    lib().ypo_set_num_threads(int(n))


def effective_thr(thr, compare_in_double):
    return float(lib().ypo_effective_thr(float(thr), int(bool(compare_in_double))))


def scale_params_v3(anchors, input_size, layer_h, layer_w):
    """anchors in pixels -> (anchor_w, anchor_h) fp32 grid units, stride_w, stride_h.
    Mirrors utils/yolov3_utils.py:29-31,38,40: python-float division, one fp32 rounding."""
    stride_h = input_size / layer_h
    stride_w = input_size / layer_w
    aw = np.array([a[0] / stride_w for a in anchors], dtype=np.float32)
    ah = np.array([a[1] / stride_h for a in anchors], dtype=np.float32)
    return aw, ah, np.float32(stride_w), np.float32(stride_h)


def scale_params_v2(scaled_anchors, input_size, layer_h, layer_w):
    """utils/yolov2_utils.py:27-30,37,39: anchors already in grid units."""
    stride_h = input_size / layer_h
    stride_w = input_size / layer_w
    aw = np.array([a[0] for a in scaled_anchors], dtype=np.float32)
    ah = np.array([a[1] for a in scaled_anchors], dtype=np.float32)
    return aw, ah, np.float32(stride_w), np.float32(stride_h)


def _decode_into(head, nc, aw, ah, sw, sh, out, box_offset):
    head = _f32(head)
    B, CH, H, W = head.shape
    A = len(aw)
    assert CH == A * (5 + nc), (CH, A, nc)
    lib().ypo_decode_scale(_fp(head), B, A, nc, H, W, _fp(aw), _fp(ah), float(sw), float(sh),
                           _fp(out), out.shape[1] * 6, box_offset)


def decode_predictions_v3(head, num_classes, anchors, input_size):
    head = _f32(head)
    B, _, H, W = head.shape
    aw, ah, sw, sh = scale_params_v3(anchors, input_size, H, W)
    out = np.empty((B, len(anchors) * H * W, 6), dtype=np.float32)
    _decode_into(head, num_classes, aw, ah, sw, sh, out, 0)
    return out


def decode_predictions_v2(head, num_classes, scaled_anchors, input_size):
    head = _f32(head)
    B, _, H, W = head.shape
    aw, ah, sw, sh = scale_params_v2(scaled_anchors, input_size, H, W)
    out = np.empty((B, len(scaled_anchors) * H * W, 6), dtype=np.float32)
    _decode_into(head, num_classes, aw, ah, sw, sh, out, 0)
    return out


def decode_multi_scale(heads, num_classes, anchors, input_size, anchors_per_scale=3):
    """Scale loop + torch.cat(dim=1) of DecodeYoloV3.forward / update_state."""
    heads = [_f32(h) for h in heads]
    B = heads[0].shape[0]
    n_tot = sum(anchors_per_scale * h.shape[2] * h.shape[3] for h in heads)
    out = np.empty((B, n_tot, 6), dtype=np.float32)
    off = 0
    for i, h in enumerate(heads):
        an = anchors[anchors_per_scale * i: anchors_per_scale * (i + 1)]
        aw, ah, sw, sh = scale_params_v3(an, input_size, h.shape[2], h.shape[3])
        _decode_into(h, num_classes, aw, ah, sw, sh, out, off)
        off += anchors_per_scale * h.shape[2] * h.shape[3]
    return out


def nms_v3_indices(rows, conf_threshold=0.25, iou_threshold=0.45,
                   offset_mode=OFFSET_AUTO_CPU, union_mode=UNION_SEPARATE,
                   compare_in_double=None):
    """Kept box indices (into rows [N,6]) in NMS order, plus candidate count."""
    rows = _f32(rows)
    N = rows.shape[0]
    if compare_in_double is None:
        compare_in_double = (union_mode == UNION_SEPARATE)
    thr = effective_thr(iou_threshold, compare_in_double)
    keep = np.empty(max(N, 1), dtype=np.int32)
    ncand = ctypes.c_int(0)
    k = lib().ypo_nms_v3(_fp(rows), N, float(np.float32(conf_threshold)), thr, offset_mode,
                         union_mode, _ip(keep), ctypes.byref(ncand))
    return keep[:k].copy(), int(ncand.value)


def nms_v3(rows, conf_threshold=0.25, iou_threshold=0.45, **kw):
    rows = _f32(rows)
    keep, _ = nms_v3_indices(rows, conf_threshold, iou_threshold, **kw)
    return rows[keep]


def nms_v3_batch(rows, conf_threshold=0.25, iou_threshold=0.45,
                 offset_mode=OFFSET_AUTO_CPU, union_mode=UNION_SEPARATE, compare_in_double=None):
    """rows [B,N,6] -> keep_src [B,N] int32, keep_count [B], cand_count [B] (OpenMP over images)."""
    rows = _f32(rows)
    B, N, _ = rows.shape
    if compare_in_double is None:
        compare_in_double = (union_mode == UNION_SEPARATE)
    thr = effective_thr(iou_threshold, compare_in_double)
    keep = np.full((B, max(N, 1)), -1, dtype=np.int32)
    kc = np.zeros(B, dtype=np.int32)
    cc = np.zeros(B, dtype=np.int32)
    lib().ypo_nms_v3_batch(_fp(rows), B, N, float(np.float32(conf_threshold)), thr, offset_mode,
                           union_mode, _ip(keep), _ip(kc), _ip(cc))
    return keep, kc, cc


def batched_nms(boxes, scores, idxs, iou_threshold, offset_mode=OFFSET_TRICK,
                union_mode=UNION_SEPARATE, compare_in_double=None):
    boxes, scores, idxs = _f32(boxes), _f32(scores), _f32(idxs)
    n = boxes.shape[0]
    if compare_in_double is None:
        compare_in_double = (union_mode == UNION_SEPARATE)
    thr = effective_thr(iou_threshold, compare_in_double)
    keep = np.empty(max(n, 1), dtype=np.int32)
    mode = offset_mode
    if mode == OFFSET_AUTO_CPU:
        mode = OFFSET_NONE if boxes.size > 4000 else OFFSET_TRICK
    if mode == OFFSET_AUTO_CUDA:
        mode = OFFSET_NONE if boxes.size > 100000 else OFFSET_TRICK
    k = lib().ypo_batched_nms(_fp(boxes), _fp(scores), _fp(idxs), n, thr, mode, union_mode, _ip(keep))
    return keep[:k].copy()


def get_target_rows(annot, input_size, img_base=0):
    """[B,T,5] padded annots -> [G,7] rows img,cx,cy,w,h,1,cls (image-major order)."""
    annot = _f32(annot)
    B, T, _ = annot.shape
    out = np.empty((max(B * T, 1), 7), dtype=np.float32)
    g = lib().ypo_target_rows(_fp(annot), B, T, float(input_size), float(img_base), _fp(out))
    return out[:g].copy()


def mean_average_precision(true_boxes, pred_boxes, num_classes, iou_threshold=0.5,
                           return_details=False):
    gt = _f32(true_boxes).reshape(-1, 7)
    det = _f32(pred_boxes).reshape(-1, 7)
    tp = np.empty(max(det.shape[0], 1), dtype=np.uint8)
    pc = np.zeros((num_classes, 4), dtype=np.float32)
    m = lib().ypo_mean_average_precision(_fp(gt), gt.shape[0], _fp(det), det.shape[0],
                                         int(num_classes), float(np.float32(iou_threshold)),
                                         tp.ctypes.data_as(ctypes.POINTER(ctypes.c_ubyte)), _fp(pc))
    if return_details:
        return np.float32(m), tp[:det.shape[0]].copy(), pc
    return np.float32(m)


def metrics_per_class(true_boxes, pred_boxes, num_classes, iou_threshold=0.5):
    _, _, pc = mean_average_precision(true_boxes, pred_boxes, num_classes, iou_threshold, True)
    return pc


def box_iou(box1, box2):
    a, b = _f32(box1), _f32(box2)
    out = np.empty((a.shape[0], b.shape[0]), dtype=np.float32)
    lib().ypo_box_iou(_fp(a), a.shape[0], _fp(b), b.shape[0], _fp(out))
    return out


def bbox_iou_cxcywh(a, b):
    a, b = _f32(a).reshape(-1, 4), _f32(b).reshape(-1, 4)
    out = np.empty(a.shape[0], dtype=np.float32)
    lib().ypo_bbox_iou_cxcywh_pairs(_fp(a), _fp(b), a.shape[0], _fp(out))
    return out


# ---- detection writers (numpy; byte/integer work) -----------------------------------------
WRITER_PRED_FILE, WRITER_COCO = 0, 1


def detection_ints(rows, mode=WRITER_PRED_FILE, scale_xy=None, img_idx_base=0):
    """Integer box columns of the reference's detection writers, rows [K,6|7] -> int64 [K,4].

    make_pred_file_yolov3.py:73-86: xmin = int(cx - (w / 2)), ..., xmax = int(cx + (w / 2)) on
    0-dim fp32 tensors.  yolo2coco_pred_file_yolov3.py:83-100: cx,cy,w,h are first multiplied by
    the Python floats width/input_size, height/input_size (fp32 tensor * Python float = fp32
    product with the scalar rounded to fp32), bbox = int(cx - w/2), int(cy - h/2), int(w), int(h).
    int() truncates toward zero."""
    r = _f32(rows)
    o = r.shape[1] - 6
    cx, cy, w, h = r[:, o], r[:, o + 1], r[:, o + 2], r[:, o + 3]
    two = np.float32(2)
    if mode == WRITER_COCO:
        s = np.asarray(scale_xy, dtype=np.float64).astype(np.float32).reshape(-1, 2)
        img = (r[:, 0] - np.float32(img_idx_base)).astype(np.int64) if o else np.zeros(r.shape[0], np.int64)
        sx, sy = s[img, 0], s[img, 1]
        cx, cy, w, h = cx * sx, cy * sy, w * sx, h * sy
        cols = [cx - w / two, cy - h / two, w, h]
    else:
        cols = [cx - w / two, cy - h / two, cx + w / two, cy + h / two]
    return np.stack([np.trunc(c).astype(np.int64) for c in cols], 1)


def pred_file_lines(boxes, class_name_list):
    """f'{class_name} {confidence_score} {xmin} {ymin} {xmax} {ymax}\\n' per box (make_pred_file_yolov3.py:72-86)."""
    r = _f32(boxes)
    v = detection_ints(r, WRITER_PRED_FILE)
    o = r.shape[1] - 6
    return ["%s %s %d %d %d %d\n" % (class_name_list[int(r[k, o + 5])], format(float(r[k, o + 4])),
                                     v[k, 0], v[k, 1], v[k, 2], v[k, 3]) for k in range(r.shape[0])]


# ---------------------------------------------------------------- alternate semantics
def nms_xyxy(boxes, scores, iou_threshold, union_mode=UNION_SEPARATE, compare_in_double=None):
    """torchvision.ops.nms on xyxy boxes: indices kept, score descending (stable)."""
    boxes, scores = _f32(boxes), _f32(scores)
    n = boxes.shape[0]
    if compare_in_double is None:
        compare_in_double = (union_mode == UNION_SEPARATE)
    keep = np.empty(max(n, 1), dtype=np.int32)
    k = lib().ypo_nms_xyxy(_fp(boxes), _fp(scores), n, effective_thr(iou_threshold, compare_in_double), union_mode, _ip(keep))
    return keep[:k].copy()


def non_max_suppression(prediction, conf_thres=0.25, iou_thres=0.45, classes=None, agnostic=False,
                        multi_label=False, max_det=300, max_nms=30000, max_wh=4096,
                        union_mode=UNION_SEPARATE, compare_in_double=None, return_index=False):
    """The vendored Ultralytics NMS, utils/general.py:640-731 (labels=() and merge=False as there).

    prediction [B,n,5+nc] -> list of [k,6] xyxy,conf,cls per image.  With return_index also the row
    number of every kept detection: box index, or box * nc + j with multi_label.  Where the
    reference takes the max_nms best through an unstable argsort, ties keep their order here."""
    p = _f32(prediction)
    nc = p.shape[2] - 5
    multi_label = bool(multi_label) and nc > 1                                   # :659
    out, idx_out = [], []
    two = np.float32(2)
    for x in p:
        sel = np.nonzero(x[:, 4] > np.float32(conf_thres))[0]                    # :645,665
        x = x[sel].copy()
        if x.shape[0]:
            x[:, 5:] *= x[:, 4:5]                                                # :669
        box = np.stack([x[:, 0] - x[:, 2] / two, x[:, 1] - x[:, 3] / two,
                        x[:, 0] + x[:, 2] / two, x[:, 1] + x[:, 3] / two], 1) if x.shape[0] else np.zeros((0, 4), np.float32)
        if multi_label:                                                          # :675-677
            i, j = np.nonzero(x[:, 5:] > np.float32(conf_thres))
            det = np.concatenate([box[i], x[i, j + 5, None], j[:, None].astype(np.float32)], 1)
            rows = sel[i] * nc + j
        else:                                                                    # :678-680
            if x.shape[0]:
                cls = np.empty(x.shape[0], np.int64)
                conf = np.empty(x.shape[0], np.float32)
                for r in range(x.shape[0]):          # first maximum, NaN maximal (torch.max)
                    v = x[r, 5:]
                    nan = np.isnan(v)
                    cls[r] = int(np.argmax(nan)) if nan.any() else int(np.argmax(v))
                    conf[r] = v[cls[r]]
            else:
                cls, conf = np.zeros(0, np.int64), np.zeros(0, np.float32)
            m = conf > np.float32(conf_thres)
            det = np.concatenate([box, conf[:, None], cls[:, None].astype(np.float32)], 1)[m]
            rows = sel[m]
        if classes is not None:                                                  # :683-684
            m = np.isin(det[:, 5], np.asarray(classes, dtype=np.float32))
            det, rows = det[m], rows[m]
        n = det.shape[0]
        if n > max_nms > 0:                                                      # :694-695
            o = np.argsort(-det[:, 4], kind="stable")[:max_nms]
            det, rows = det[o], rows[o]
        c = det[:, 5:6] * np.float32(0 if agnostic else max_wh)                  # :698
        keep = nms_xyxy(det[:, :4] + c, det[:, 4], iou_thres, union_mode, compare_in_double) if n else np.zeros(0, np.int32)
        if max_det > 0:
            keep = keep[:max_det]                                                # :701-702
        out.append(det[keep].astype(np.float32).reshape(-1, 6))
        idx_out.append(rows[keep].astype(np.int64))
    return (out, idx_out) if return_index else out


def decode_predictions_v1(inp, num_classes, num_boxes, input_size, S=7):
    """YOLO v1 decode, utils/yolov1_utils.py:12-66: [B, S*S*(5*num_boxes+nc)] -> [B, S*S, 6]."""
    x = _f32(inp).reshape(-1, S, S, num_boxes * 5 + num_classes)
    pred = (np.float32(1) / (np.float32(1) + np.exp(-x, dtype=np.float32))).astype(np.float32)    # torch.sigmoid
    B = pred.shape[0]
    stride = np.float32(input_size / S)
    conf = np.stack([pred[..., num_classes + 5 * k] for k in range(num_boxes)])          # [nb,B,S,S]
    nan = np.isnan(conf)
    best = np.where(nan.any(0), np.argmax(nan, 0), np.argmax(np.where(nan, -np.inf, conf), 0))
    pbox = np.zeros((B, S, S, 4), np.float32)
    pconf = np.zeros((B, S, S), np.float32)
    for k in range(num_boxes):
        hot = (best == k).astype(np.float32)
        pbox = pbox + hot[..., None] * pred[..., num_classes + 1 + 5 * k: num_classes + 5 + 5 * k]
        pconf = pconf + hot * pred[..., num_classes + 5 * k]
    gx = np.arange(S, dtype=np.float32)[None, None, :]
    gy = np.arange(S, dtype=np.float32)[None, :, None]
    out = np.empty((B, S, S, 6), np.float32)
    out[..., 0] = (pbox[..., 0] + gx) * stride
    out[..., 1] = (pbox[..., 1] + gy) * stride
    out[..., 2] = (pbox[..., 2] * np.float32(S)) * stride
    out[..., 3] = (pbox[..., 3] * np.float32(S)) * stride
    out[..., 4] = pconf
    cls_p = pred[..., :num_classes]
    cn = np.isnan(cls_p)
    out[..., 5] = np.where(cn.any(-1), np.argmax(cn, -1), np.argmax(np.where(cn, -np.inf, cls_p), -1)).astype(np.float32)
    return out.reshape(B, S * S, 6)


def confusion_process_batch(matrix, detections, labels, nc, conf=0.25, iou_thres=0.45):
    """ConfusionMatrix.process_batch, utils/metrics.py:122-160, accumulated into matrix [(nc+1),(nc+1)].

    Restated without the two argsort/unique rounds (:140-145): every detection keeps the ground truth of
    highest IoU above the threshold, then every ground truth keeps the detection of highest IoU among
    those; equal IoUs (where the reference's unstable argsort decides) go to the lower index."""
    det = _f32(detections).reshape(-1, 6)
    lab = _f32(labels).reshape(-1, 5)
    det = det[det[:, 4] > np.float32(conf)]                                      # :134
    gt_cls = lab[:, 0].astype(np.int32)
    det_cls = det[:, 5].astype(np.int32)
    M, N = lab.shape[0], det.shape[0]
    iou = box_iou(lab[:, 1:], det[:, :4]) if M and N else np.zeros((M, N), np.float32)     # :137
    ok = iou > np.float32(iou_thres)
    best_g = np.full(N, -1)
    for d in range(N):
        g = np.nonzero(ok[:, d])[0]
        if g.size:
            best_g[d] = g[np.argmax(iou[g, d])]
    matched_det = np.zeros(N, bool)
    for g in range(M):
        d = np.nonzero(best_g == g)[0]
        if d.size:
            dm = d[np.argmax(iou[g, d])]
            matrix[det_cls[dm], gt_cls[g]] += 1                                  # :151-152
            matched_det[dm] = True
        else:
            matrix[nc, gt_cls[g]] += 1                                           # :153-154
    if ok.any():                                                                 # :156 `if n:`
        for d in range(N):
            if not matched_det[d]:
                matrix[det_cls[d], nc] += 1                                      # :157-159
    return matrix


def head_conv1x1(feature, weight):
    """The head's last layer, nn.Conv2d(Cin, A*(5+nc), 1, 1, bias=False) (models/detector/yolov3.py:56,62,68):
    feature [B, Cin, H, W], weight [AC, Cin(,1,1)] -> (head [B, AC, H, W] in float64, bound), where
    bound = sum_k |x_k| |w_k| per output element: the tensor core cuts both factors to 10 mantissa bits
    (each loses up to 2^-10 relative), so |tf32 head - exact head| <= 2^-9 * bound * (1 + eps)."""
    x = np.asarray(feature, dtype=np.float64)
    w = np.asarray(weight, dtype=np.float64).reshape(np.asarray(weight).shape[0], -1)
    head = np.einsum("bkhw,ck->bchw", x, w)
    bound = np.einsum("bkhw,ck->bchw", np.abs(x), np.abs(w))
    return head, bound


# ---------------------------------------------------------------- utils/metrics.py: the AP side (compute_ap, ap_per_class, bbox_iou)
def best_class_scores(prediction, conf_thres=0.25):
    """The scoring lines of the vendored NMS alone, utils/general.py:645,669,679-680, for ONE image [n,5+nc]:
    obj > conf -> conf_j = cls_j * obj -> (conf, j) = first maximum over j -> conf > conf_thres.
    Returns (box numbers, conf, j): what YPP_SCORE_OBJ_TIMES_CLS must compute before NMS."""
    x = _f32(prediction)
    t = np.float32(conf_thres)
    idx, conf, cls = [], [], []
    for r in range(x.shape[0]):
        if not (x[r, 4] > t):
            continue
        best, bj = None, 0
        for j in range(x.shape[1] - 5):
            v = np.float32(x[r, 5 + j] * x[r, 4])
            if best is None or v > best:
                best, bj = v, j
        if best is not None and best > t:
            idx.append(r); conf.append(best); cls.append(bj)
    return np.array(idx, np.int64), np.array(conf, np.float32), np.array(cls, np.int64)


def compute_ap(recall, precision):
    """utils/metrics.py:86-111, float64, method 'interp': sentinels (0,1) / (1,0), right-to-left running maximum of
    the precision, trapezoid of the envelope sampled at 101 recall points.  Plain loops, no numpy reductions
    besides np.interp (whose definition is numpy's)."""
    mrec = [0.0] + [float(v) for v in recall] + [1.0]
    mpre = [1.0] + [float(v) for v in precision] + [0.0]
    for i in range(len(mpre) - 2, -1, -1):               # np.flip(np.maximum.accumulate(np.flip(mpre))); NaN propagates
        if mpre[i] == mpre[i] and not (mpre[i] >= mpre[i + 1]):
            mpre[i] = mpre[i + 1]
    mrec, mpre = np.array(mrec, np.float64), np.array(mpre, np.float64)
    x = np.linspace(0, 1, 101)
    y = np.interp(x, mrec, mpre)
    ap = float(np.add.reduce(np.diff(x) * (y[1:] + y[:-1]) / 2.0))      # np.trapz(y, x)
    return ap, mpre, mrec


def ap_per_class(tp, conf, pred_cls, target_cls):
    """utils/metrics.py:21-83 without the plots: -> (p, r, ap, f1, unique_classes.astype(int32))."""
    tp = np.asarray(tp)
    conf = np.asarray(conf)
    pred_cls = np.asarray(pred_cls)
    target_cls = np.asarray(target_cls)
    order = np.argsort(-conf)                                                    # :37 (numpy's default sort, as there)
    tp, conf, pred_cls = tp[order], conf[order], pred_cls[order]
    classes = np.unique(target_cls)
    nc = classes.shape[0]
    px = np.linspace(0, 1, 1000)
    ap = np.zeros((nc, tp.shape[1]))
    p = np.zeros((nc, 1000))
    r = np.zeros((nc, 1000))
    for ci in range(nc):
        c = classes[ci]
        sel = pred_cls == c
        n_l = int((target_cls == c).sum())
        n_p = int(sel.sum())
        if n_p == 0 or n_l == 0:
            continue
        fpc = (1 - tp[sel]).cumsum(0)
        tpc = tp[sel].cumsum(0)
        recall = tpc / (n_l + 1e-16)
        r[ci] = np.interp(-px, -conf[sel], recall[:, 0], left=0)
        precision = tpc / (tpc + fpc)
        p[ci] = np.interp(-px, -conf[sel], precision[:, 0], left=1)
        for j in range(tp.shape[1]):
            ap[ci, j] = compute_ap(recall[:, j], precision[:, j])[0]
    f1 = 2 * p * r / (p + r + 1e-16)
    i = f1.mean(0).argmax()
    return p[:, i], r[:, i], ap, f1[:, i], classes.astype("int32")


def bbox_iou_metrics(box1, box2, xywh=True, GIoU=False, DIoU=False, CIoU=False, eps=1e-7):
    """utils/metrics.py:192-230 (NOT utils/yolo_utils.py's bbox_iou: no abs, eps 1e-7, xywh flag): box1 [1|n,4]
    against box2 [n,4] -> [n,1], one fp32 rounding per torch operation, in the reference's operand order."""
    f = np.float32
    a, b = _f32(box1).reshape(-1, 4), _f32(box2).reshape(-1, 4)
    n = max(a.shape[0], b.shape[0])
    e = f(eps)
    out = np.empty((n, 1), np.float32)
    with np.errstate(all="ignore"):
        for k in range(n):
            p, q = a[k if a.shape[0] > 1 else 0], b[k if b.shape[0] > 1 else 0]
            if xywh:
                w1, h1, w2, h2 = p[2], p[3], q[2], q[3]
                ax1, ax2, ay1, ay2 = f(p[0] - f(w1 / f(2))), f(p[0] + f(w1 / f(2))), f(p[1] - f(h1 / f(2))), f(p[1] + f(h1 / f(2)))
                bx1, bx2, by1, by2 = f(q[0] - f(w2 / f(2))), f(q[0] + f(w2 / f(2))), f(q[1] - f(h2 / f(2))), f(q[1] + f(h2 / f(2)))
            else:
                ax1, ay1, ax2, ay2 = p
                bx1, by1, bx2, by2 = q
                w1, h1 = f(ax2 - ax1), f(f(ay2 - ay1) + e)
                w2, h2 = f(bx2 - bx1), f(f(by2 - by1) + e)
            iw = max(f(min(ax2, bx2) - max(ax1, bx1)), f(0))
            ih = max(f(min(ay2, by2) - max(ay1, by1)), f(0))
            inter = f(iw * ih)
            union = f(f(f(f(w1 * h1) + f(w2 * h2)) - inter) + e)
            iou = f(inter / union)
            res = iou
            if CIoU or DIoU or GIoU:
                cw = f(max(ax2, bx2) - min(ax1, bx1))
                ch = f(max(ay2, by2) - min(ay1, by1))
                if CIoU or DIoU:
                    c2 = f(f(f(cw * cw) + f(ch * ch)) + e)
                    dx = f(f(f(bx1 + bx2) - ax1) - ax2)
                    dy = f(f(f(by1 + by2) - ay1) - ay2)
                    rho2 = f(f(f(dx * dx) + f(dy * dy)) / f(4))
                    if CIoU:
                        import math
                        da = f(f(np.arctan(f(w2 / h2))) - f(np.arctan(f(w1 / h1))))
                        v = f(f(4 / math.pi ** 2) * f(da * da))
                        alpha = f(v / f(f(v - iou) + f(1 + eps)))
                        res = f(iou - f(f(rho2 / c2) + f(v * alpha)))
                    else:
                        res = f(iou - f(rho2 / c2))
                else:
                    c_area = f(f(cw * ch) + e)
                    res = f(iou - f(f(c_area - union) / c_area))
            out[k, 0] = res
    return out
