"""ctypes binding of libypp.so (C ABI declared in include/ypp.h).

There is no CPU fallback: if the CUDA library is missing the import of the op layer fails
loudly.  `build()` compiles it in-tree with nvcc for sm_100a.
"""
import ctypes
import os
import subprocess

_PKG = os.path.dirname(os.path.abspath(__file__))
_CSRC = os.path.join(os.path.dirname(_PKG), "csrc")
LIB_PATH = os.environ.get("YPP_LIB") or os.path.join(_PKG, "libypp.so")   # YPP_LIB: tuning builds only
SOURCES = ["api.cu", "decode.cu", "head.cu", "organise.cu", "nms.cu", "metrics.cu"]

YPP_MAX_SCALES = 4
YPP_STATIC_WS_BYTES = 4 << 20
YPP_MAX_ANCHORS = 8

SCORE_OBJ, SCORE_OBJ_TIMES_CLS = 0, 1
OFFSET_NONE, OFFSET_TRICK, OFFSET_AUTO_CPU, OFFSET_AUTO_CUDA = 0, 1, 2, 3
UNION_SEPARATE, UNION_FMA = 0, 1

STAGE_DECODE, STAGE_SORT, STAGE_SUPPRESS, STAGE_ROWS = 1, 2, 4, 8
STAGES = {"decode": STAGE_DECODE, "sort": STAGE_SORT, "suppress": STAGE_SUPPRESS, "rows": STAGE_ROWS,
          "nms": STAGE_SUPPRESS | STAGE_ROWS}

OFFSET_MODES = {"none": OFFSET_NONE, "trick": OFFSET_TRICK, "auto_cpu": OFFSET_AUTO_CPU,
                "auto_cuda": OFFSET_AUTO_CUDA}
UNION_MODES = {"separate": UNION_SEPARATE, "fma": UNION_FMA}
SCORE_MODES = {"obj": SCORE_OBJ, "obj_times_cls": SCORE_OBJ_TIMES_CLS}


class YppScale(ctypes.Structure):
    _fields_ = [("head", ctypes.c_void_p), ("A", ctypes.c_int), ("H", ctypes.c_int), ("W", ctypes.c_int),
                ("anchor_w", ctypes.c_float * YPP_MAX_ANCHORS), ("anchor_h", ctypes.c_float * YPP_MAX_ANCHORS),
                ("stride_w", ctypes.c_float), ("stride_h", ctypes.c_float)]


class YppCfg(ctypes.Structure):
    _fields_ = [("nc", ctypes.c_int), ("B", ctypes.c_int), ("n_scales", ctypes.c_int),
                ("conf_thr", ctypes.c_float), ("iou_thr", ctypes.c_float), ("score_mode", ctypes.c_int),
                ("offset_mode", ctypes.c_int), ("union_mode", ctypes.c_int)]


class YppUltraCfg(ctypes.Structure):
    _fields_ = [("nc", ctypes.c_int), ("B", ctypes.c_int), ("conf_thr", ctypes.c_float), ("iou_thr", ctypes.c_float),
                ("multi_label", ctypes.c_int), ("agnostic", ctypes.c_int), ("max_det", ctypes.c_int),
                ("max_nms", ctypes.c_int), ("max_wh", ctypes.c_float), ("class_allow", ctypes.c_void_p),
                ("union_mode", ctypes.c_int)]


class YppProbe(ctypes.Structure):
    _fields_ = [("host_word", ctypes.c_void_p), ("seq", ctypes.c_uint32)]


class YppHeadSrc(ctypes.Structure):
    _fields_ = [("feature", ctypes.c_void_p), ("weight", ctypes.c_void_p), ("cin", ctypes.c_int),
                ("head_out", ctypes.c_void_p)]


WRITER_MODES = {"pred_file": 0, "coco": 1}

EXPORTS = ["ypp_workspace_bytes", "ypp_effective_iou_thr", "ypp_decode_dense", "ypp_decode_compact",
           "ypp_rows_compact", "ypp_sort_segments", "ypp_nms", "ypp_postprocess", "ypp_nms_rows",
           "ypp_target_rows", "ypp_box_iou", "ypp_bbox_iou", "ypp_map_workspace_bytes", "ypp_map_match", "ypp_map_flags", "ypp_map_ap_from_flags", "ypp_map_mean",
           "ypp_detection_ints", "ypp_ultra_rows", "ypp_nms_ultralytics", "ypp_decode_yolov1",
           "ypp_confusion_update", "ypp_head_decode_compact", "ypp_head_postprocess", "ypp_nms_append", "ypp_postprocess_append", "ypp_run_stages", "ypp_postprocess_static", "ypp_peek_async", "ypp_pipeline_create", "ypp_pipeline_destroy", "ypp_pipeline_next_slot", "ypp_pipeline_stream",
           "ypp_pipeline_submit", "ypp_pipeline_query", "ypp_pipeline_wait_ticket", "ypp_pipeline_join", "ypp_target_rows_append", "ypp_last_error", "ypp_version", "ypp_launch_count", "ypp_adaptive_launch"]


def build(force=False, verbose=False, defines=(), out=None):
    """nvcc the kernels into ypp/libypp.so (sm_100a only, -lineinfo for ncu source pages)."""
    if out is not None:
        cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
               "-Xcompiler", "-fPIC", "-shared", "-o", out] + ["-D" + d for d in defines] + \
              [os.path.join(_CSRC, s) for s in SOURCES]
        subprocess.run(cmd, check=True)
        return out
    srcs = [os.path.join(_CSRC, s) for s in SOURCES]
    deps = srcs + [os.path.join(_CSRC, h) for h in ("ypp_common.cuh", "ypp_kernels.cuh", "decode_common.cuh")] + \
        [os.path.join(os.path.dirname(os.path.dirname(_PKG)), "include", "ypp.h")]
    if not force and os.path.exists(LIB_PATH):
        t = os.path.getmtime(LIB_PATH)
        if all(os.path.getmtime(d) <= t for d in deps if os.path.exists(d)):
            return LIB_PATH
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
           "-Xcompiler", "-fPIC", "-shared", "-o", LIB_PATH] + srcs
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    subprocess.run(cmd, check=True)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "ypp: %s is missing. The CUDA extension is the product; there is no CPU fallback. "
            "Build it with `python -c \"import __graft_entry__ as g; g.build()\"`." % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    vp, ci, cf, i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_int64
    cfgp, scp, prp = ctypes.POINTER(YppCfg), ctypes.POINTER(YppScale), ctypes.POINTER(YppProbe)
    L.ypp_workspace_bytes.argtypes = [ci, ci, ci]
    L.ypp_workspace_bytes.restype = ctypes.c_size_t
    L.ypp_effective_iou_thr.argtypes = [ctypes.c_double, ci]
    L.ypp_effective_iou_thr.restype = cf
    L.ypp_decode_dense.argtypes = [cfgp, scp, vp, vp]
    L.ypp_decode_compact.argtypes = [cfgp, scp, vp, ctypes.c_size_t, vp]
    L.ypp_rows_compact.argtypes = [cfgp, vp, ci, vp, ctypes.c_size_t, vp]
    L.ypp_sort_segments.argtypes = [cfgp, ci, vp, ctypes.c_size_t, vp]
    L.ypp_nms.argtypes = [cfgp, ci, vp, ctypes.c_size_t, vp, vp, vp, vp, ci, i64, vp, cf, vp]
    L.ypp_postprocess.argtypes = [cfgp, scp, vp, ctypes.c_size_t, vp, vp, vp, vp, ci, i64, vp, cf, vp]
    hsp = ctypes.POINTER(YppHeadSrc)
    L.ypp_head_decode_compact.argtypes = [cfgp, scp, hsp, vp, ctypes.c_size_t, vp]
    L.ypp_head_postprocess.argtypes = [cfgp, scp, hsp, vp, ctypes.c_size_t, vp, vp, vp, vp, ci, i64, vp, cf, vp]
    L.ypp_nms_rows.argtypes = [cfgp, vp, ci, vp, ctypes.c_size_t, vp, vp, vp, vp, ci, i64, vp, cf, vp]
    L.ypp_target_rows.argtypes = [vp, ci, ci, cf, cf, vp, vp, vp]
    L.ypp_target_rows_append.argtypes = [vp, ci, ci, cf, cf, vp, i64, vp, vp, prp]
    L.ypp_postprocess_append.argtypes = [cfgp, scp, vp, ctypes.c_size_t, vp, vp, vp, ci, i64, vp, vp, cf, vp, prp]
    L.ypp_nms_append.argtypes = [cfgp, ci, vp, ctypes.c_size_t, vp, vp, vp, ci, i64, vp, vp, cf, vp, prp]
    L.ypp_run_stages.argtypes = [cfgp, scp, ci, vp, ctypes.c_size_t, ci, vp, vp, vp, vp, ci, i64, vp, vp, vp, cf, vp, prp]
    L.ypp_postprocess_static.argtypes = [cfgp, scp, vp, vp, vp, vp, ci, i64, vp, cf, vp]
    L.ypp_postprocess_static.restype = ci
    L.ypp_pipeline_create.argtypes = [ci, ctypes.POINTER(vp)]
    L.ypp_pipeline_destroy.argtypes = [vp]
    L.ypp_pipeline_next_slot.argtypes = [vp]
    L.ypp_pipeline_stream.argtypes = [vp, ci]
    L.ypp_pipeline_stream.restype = vp
    L.ypp_pipeline_submit.argtypes = [vp, cfgp, scp, vp, ctypes.c_size_t, vp, vp, vp, ci, i64, vp, vp, prp, cf, vp, ctypes.POINTER(i64)]
    L.ypp_pipeline_query.argtypes = [vp, i64]
    L.ypp_pipeline_wait_ticket.argtypes = [vp, i64]
    L.ypp_pipeline_join.argtypes = [vp, vp]
    for name in ("ypp_pipeline_create", "ypp_pipeline_destroy", "ypp_pipeline_next_slot", "ypp_pipeline_submit", "ypp_pipeline_query",
                 "ypp_pipeline_wait_ticket", "ypp_pipeline_join"):
        getattr(L, name).restype = ci
    L.ypp_peek_async.argtypes = [vp, vp, ctypes.c_size_t, vp]
    L.ypp_peek_async.restype = ci
    L.ypp_box_iou.argtypes = [vp, ci, vp, ci, vp, vp]
    L.ypp_bbox_iou.argtypes = [vp, ci, vp, ci, ci, ci, ctypes.c_double, vp, vp]
    L.ypp_detection_ints.argtypes = [vp, i64, ci, ci, vp, ci, cf, vp, vp]
    L.ypp_ultra_rows.argtypes = [ci, ci, ci]
    L.ypp_ultra_rows.restype = i64
    L.ypp_nms_ultralytics.argtypes = [ctypes.POINTER(YppUltraCfg), vp, ci, vp, ctypes.c_size_t, vp, vp, vp, vp, ci, i64, vp, cf, vp]
    L.ypp_decode_yolov1.argtypes = [vp, ci, ci, ci, ci, cf, vp, vp]
    L.ypp_confusion_update.argtypes = [vp, vp, vp, vp, ci, i64, ci, cf, cf, vp, ctypes.c_size_t, vp, vp]
    L.ypp_map_workspace_bytes.argtypes = [i64, i64, ci]
    L.ypp_map_workspace_bytes.restype = ctypes.c_size_t
    L.ypp_map_match.argtypes = [vp, i64, vp, i64, ci, cf, vp, ctypes.c_size_t, vp, vp, vp, vp]
    L.ypp_map_flags.argtypes = [vp, i64, vp, i64, ci, cf, vp, ctypes.c_size_t, vp, vp, vp, vp]
    L.ypp_map_ap_from_flags.argtypes = [vp, i64, i64, vp, vp, ci, ci, ci, vp, ctypes.c_size_t, vp, vp]
    L.ypp_map_mean.argtypes = [vp, ci, vp, vp]
    L.ypp_last_error.restype = ctypes.c_char_p
    L.ypp_version.restype = ci
    L.ypp_launch_count.argtypes = [ci]
    L.ypp_launch_count.restype = i64
    L.ypp_adaptive_launch.argtypes = [ci]
    for name in ("ypp_decode_dense", "ypp_decode_compact", "ypp_rows_compact", "ypp_sort_segments", "ypp_nms",
                 "ypp_postprocess", "ypp_nms_rows", "ypp_target_rows", "ypp_box_iou", "ypp_bbox_iou", "ypp_map_match",
                 "ypp_detection_ints", "ypp_nms_ultralytics", "ypp_decode_yolov1", "ypp_confusion_update",
                 "ypp_head_decode_compact", "ypp_head_postprocess", "ypp_run_stages", "ypp_nms_append",
                 "ypp_postprocess_append", "ypp_target_rows_append", "ypp_bbox_iou", "ypp_map_flags",
                 "ypp_map_ap_from_flags", "ypp_map_mean"):
        getattr(L, name).restype = ci
    _lib = L
    return L


class YppError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise YppError("libypp error %d: %s" % (rc, lib().ypp_last_error().decode()))
