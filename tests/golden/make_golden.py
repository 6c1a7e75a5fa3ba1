"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference and torchvision 0.26.0 CPU):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

Every fixture stores the reference's own outputs; inputs are stored verbatim when small
and as (generator seed, crc32 of the bytes) when large, so tests can regenerate and verify
them.  The script also checks the CPU oracle (oracle/) against every fixture as it writes
it, which is what pins the oracle (it prints a line per case and exits non-zero on mismatch).
"""
import os
import sys
import types
import zlib

sys.dont_write_bytecode = True
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("YPP_REFERENCE", "/root/reference")
sys.path.insert(0, REF)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "pytorch-object-detection_b200"))

import numpy as np
import torch

# utils/metrics.py imports matplotlib at module level (absent here): empty stubs
for name in ("matplotlib", "matplotlib.pyplot"):
    if name not in sys.modules:
        sys.modules[name] = types.ModuleType(name)
sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]

import utils.yolo_utils as ref_yu                     # noqa: E402
from utils.yolov3_utils import decode_predictions as ref_decode_v3      # noqa: E402
from utils.yolov3_utils import DecodeYoloV3, MeanAveragePrecision as RefMAPv3   # noqa: E402
from utils.yolov2_utils import decode_predictions as ref_decode_v2      # noqa: E402
from utils.yolov2_utils import DecodeYoloV2, MeanAveragePrecision as RefMAPv2   # noqa: E402
from utils.metrics import box_iou as ref_box_iou      # noqa: E402
import utils.metrics as ref_metrics                   # noqa: E402

from oracle import oracle as O                        # noqa: E402
from ypp import synth                                 # noqa: E402

torch.manual_seed(0)
FAIL = []


def crc(t):
    return zlib.crc32(np.ascontiguousarray(t).tobytes()) & 0xFFFFFFFF


def check(name, ok, msg=""):
    print(("ok   " if ok else "FAIL ") + name + (" " + msg if msg else ""))
    if not ok:
        FAIL.append(name)


class _Capture:
    """Wraps the reference's batched_nms binding to record candidate/keep indices."""

    def __init__(self):
        self.keep = None
        self.orig = ref_yu.batched_nms

    def __enter__(self):
        def wrapped(b, s, i, t):
            k = self.orig(b, s, i, t)
            self.keep = k.clone()
            return k
        ref_yu.batched_nms = wrapped
        return self

    def __exit__(self, *a):
        ref_yu.batched_nms = self.orig


def ref_nms_indices(rows, conf=0.25, iou=0.45):
    """Run the reference nms_v3 on rows [N,6]; return (kept box indices in reference order, rows)."""
    cand = torch.where(rows[..., 4] > conf)[0]
    with _Capture() as cap:
        out = ref_yu.nms_v3(rows, conf_threshold=conf, iou_threshold=iou)
    if cand.numel() == 0:
        return np.zeros(0, np.int32), out.numpy()
    keep = cand[cap.keep]
    return keep.numpy().astype(np.int32), out.numpy()


def canon(keep, rows):
    """(score desc, index asc): the vanilla path's final sort is not stable."""
    k = np.asarray(keep, dtype=np.int64)
    order = np.lexsort((k, -rows[k, 4].astype(np.float64)))
    return k[order].astype(np.int32)


def cpu_modes(n_cand):
    """What torchvision 0.26.0 picks on CPU for n candidates."""
    return (O.OFFSET_NONE if 4 * n_cand > 4000 else O.OFFSET_TRICK), O.UNION_SEPARATE


# ------------------------------------------------------------------ decode + nms on heads

def case_heads(name, kind, nc, grids, A, batch, input_size, conf, mu_obj, g, store_rows_stride=1):
    heads = synth.make_heads(batch, nc, grids, A=A, g=g, mu_obj=mu_obj)
    if kind == "v2":
        anchors = synth.YOLOV2_SCALED_ANCHORS
        rows = ref_decode_v2(heads[0], nc, anchors, input_size)
        orows = O.decode_predictions_v2(heads[0].numpy(), nc, anchors, input_size)
    else:
        anchors = synth.YOLOV3_ANCHORS[: A * len(grids)] if A * len(grids) == 9 else synth.YOLOV3_ANCHORS[-A * len(grids):]
        rows = torch.cat([ref_decode_v3(h, nc, anchors[A * i: A * i + A], input_size)
                          for i, h in enumerate(heads)], dim=1)
        orows = O.decode_multi_scale([h.numpy() for h in heads], nc, anchors, input_size, A)
    rows_np = rows.numpy()
    rel = np.abs(orows[..., :5] - rows_np[..., :5]) / np.maximum(np.abs(rows_np[..., :5]), 1e-30)
    cls_mis = int((orows[..., 5] != rows_np[..., 5]).sum())
    check(name + ":decode", rel.max() <= 1e-5 and cls_mis <= 2,
          "max rel %.2e cls mismatches %d of %d" % (rel.max(), cls_mis, rows_np.shape[0] * rows_np.shape[1]))
    keeps, counts, cands = [], [], []
    all_ok = True
    for b in range(batch):
        k, out = ref_nms_indices(rows[b], conf)
        kc = canon(k, rows_np[b])
        ncand = int((rows_np[b, :, 4] > conf).sum())
        om, um = cpu_modes(ncand)
        ok_, onc = O.nms_v3_indices(rows_np[b], conf, 0.45, offset_mode=om, union_mode=um)
        if om == O.OFFSET_TRICK:
            same = np.array_equal(ok_, k)       # trick path: order is well defined
        else:
            same = np.array_equal(ok_, kc)
        all_ok &= bool(same) and onc == ncand
        keeps.append(kc if om == O.OFFSET_NONE else k)
        counts.append(len(k)); cands.append(ncand)
    check(name + ":nms(stage-wise on reference rows)", all_ok,
          "cand/img %.0f kept/img %.0f" % (np.mean(cands), np.mean(counts)))
    maxk = max(max(counts), 1)
    keep_arr = np.full((batch, maxk), -1, np.int32)
    for b, k in enumerate(keeps):
        keep_arr[b, :len(k)] = k
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"),
        kind=kind, nc=nc, grids=np.array(grids), A=A, batch=batch, input_size=input_size,
        conf=np.float32(conf), mu_obj=np.float32(mu_obj), g=g,
        head_crc=np.array([crc(h.numpy()) for h in heads], dtype=np.uint64),
        rows_stride=store_rows_stride, rows=rows_np[:, ::store_rows_stride].copy(),
        rows_crc=np.uint64(crc(rows_np)),
        keep=keep_arr, keep_count=np.array(counts, np.int32), cand_count=np.array(cands, np.int32))


# ------------------------------------------------------------------ hand-made NMS cases

def nms_case(rows, conf=0.25, iou=0.45):
    rows = torch.tensor(np.asarray(rows, dtype=np.float32)).reshape(-1, 6)
    k, out = ref_nms_indices(rows, conf, iou)
    return rows.numpy(), k, out


def case_adversarial():
    cases = {}
    # the commented 5-box example of utils/yolov3_utils.py:162-167
    cases["ref_comment_5box"] = ([[100, 100, 50, 50, .8, 1], [100, 100, 50, 50, .7, 1], [60, 60, 50, 50, .6, 1],
                                  [100, 100, 50, 50, .8, 2], [100, 100, 50, 50, .7, 2]], .25, .45)
    # all-equal scores, identical boxes: survivor = lowest index per class
    cases["identical_tied"] = ([[50, 50, 20, 20, .5, 0]] * 6 + [[50, 50, 20, 20, .5, 3]] * 3, .25, .45)
    # all-equal scores, chain of half-overlapping boxes
    cases["tied_chain"] = ([[10 + 6 * i, 10, 20, 20, .5, 0] for i in range(40)], .25, .45)
    # IoU exactly representable at threshold .5: boxes [0,0,2,1] vs [0,0,1,1]... inter 1, union 2
    cases["iou_eq_thr_half"] = ([[1, .5, 2, 1, .9, 0], [.5, .5, 1, 1, .8, 0]], .25, .5)
    # IoU == float(0.45)? use 9/20: boxes a=[0,0,20,1] area 20, b=[0,0,9,1] inside a -> iou 9/20 = 0.45 (double-rounded)
    cases["iou_9_20_thr_045"] = ([[10, .5, 20, 1, .9, 0], [4.5, .5, 9, 1, .8, 0]], .25, .45)
    # thr .3 rounds UP in fp32: iou == float(.3)?  3/10: a=[0,0,10,1], b=[0,0,3,1]
    cases["iou_3_10_thr_03"] = ([[5, .5, 10, 1, .9, 0], [1.5, .5, 3, 1, .8, 0]], .25, .3)
    # zero-area boxes (0/0 = NaN -> kept), and zero-area inside a real box
    cases["zero_area"] = ([[5, 5, 0, 0, .9, 0], [5, 5, 0, 0, .8, 0], [5, 5, 4, 4, .7, 0], [5, 5, 0, 4, .6, 0]], .25, .45)
    # negative w/h (inverted) rows fed straight to nms_v3
    cases["inverted"] = ([[5, 5, -4, -4, .9, 0], [5, 5, 4, 4, .8, 0], [5, 5, -4, 4, .7, 0]], .25, .45)
    # nothing passes the confidence threshold
    cases["empty"] = ([[5, 5, 4, 4, .2, 0], [6, 6, 4, 4, .25, 1]], .25, .45)
    # score exactly at conf (strict >): dropped
    cases["score_eq_conf"] = ([[5, 5, 4, 4, .25, 0], [50, 50, 4, 4, .2500001, 0]], .25, .45)
    # coordinate trick cross-class overlap: class-1 box at negative coords is shifted onto the
    # class-0 box near max_coord
    cases["trick_cross_class"] = ([[95, 95, 10, 10, .9, 0], [-6, -6, 10, 10, .8, 1], [30, 30, 10, 10, .7, 0]], .25, .45)
    # same but with a third class two bands away
    cases["trick_cross_class2"] = ([[95, 95, 10, 10, .9, 0], [-107, -107, 10, 10, .8, 2], [-6, -6, 10, 10, .85, 1],
                                    [-6.5, -6.5, 10, 10, .6, 1]], .25, .45)
    # single class, many overlapping
    rng = np.random.RandomState(7)
    r = np.concatenate([rng.uniform(0, 100, (300, 2)), rng.uniform(5, 60, (300, 2)),
                        rng.uniform(.2, 1, (300, 1)), np.zeros((300, 1))], 1)
    cases["single_class_300"] = (r, .25, .45)
    # quantised scores (many ties) x 5 classes
    r = np.concatenate([rng.uniform(0, 200, (600, 2)), rng.uniform(5, 80, (600, 2)),
                        np.round(rng.uniform(.2, 1, (600, 1)), 1), rng.randint(0, 5, (600, 1))], 1)
    cases["quantised_ties_600"] = (r, .25, .45)
    # heuristic flip of torchvision on CPU: 1000 candidates (trick) vs 1001 (vanilla)
    base = np.concatenate([rng.uniform(-30, 446, (1001, 2)), rng.uniform(5, 300, (1001, 2)),
                           rng.uniform(.3, 1, (1001, 1)), rng.randint(0, 20, (1001, 1))], 1)
    cases["flip_1000"] = (base[:1000], .25, .45)
    cases["flip_1001"] = (base, .25, .45)
    # large-coordinate boxes: rounding under the trick offsets
    r = np.concatenate([rng.uniform(0, 416, (800, 2)), np.exp(rng.normal(0, 1.2, (800, 2))) * 100,
                        rng.uniform(.3, 1, (800, 1)), rng.randint(0, 20, (800, 1))], 1)
    cases["trick_big_boxes_800"] = (r, .25, .45)
    out = {}
    for name, (rows, conf, iou) in cases.items():
        rows_np, k, _ = nms_case(rows, conf, iou)
        ncand = int((rows_np[:, 4] > np.float32(conf)).sum())
        om, um = cpu_modes(ncand)
        ok_, onc = O.nms_v3_indices(rows_np, conf, iou, offset_mode=om, union_mode=um)
        ref = k if om == O.OFFSET_TRICK else canon(k, rows_np)
        check("adv:" + name, np.array_equal(ok_, ref) and onc == ncand,
              "n=%d cand=%d keep=%d mode=%s" % (len(rows_np), ncand, len(k), "trick" if om else "vanilla"))
        out[name + "__rows"] = rows_np
        out[name + "__keep"] = ref
        out[name + "__conf_iou"] = np.array([conf, iou], np.float64)
        out[name + "__mode"] = np.array([om, um], np.int32)
    np.savez_compressed(os.path.join(HERE, "nms_adversarial.npz"), **out)


# ------------------------------------------------------------------ decode edge cases

def case_decode_edges():
    nc, A, H = 4, 3, 5
    gen = torch.Generator("cpu").manual_seed(99)
    t = torch.randn(2, A, 5 + nc, H, H, generator=gen)
    t[0, 0, 5:, 0, 0] = 20.0                # saturated class logits: all sigmoid == 1 -> argmax 0
    t[0, 1, 5:, 1, 1] = torch.tensor([17., 30., 18., 40.])   # saturation collisions
    t[0, 2, 5:, 2, 2] = -200.0              # sigmoid underflow -> all 0 -> argmax 0
    t[1, 0, 2:4, 0, 0] = 90.0               # exp overflow -> inf w/h
    t[1, 1, 2:4, 1, 1] = -110.0             # exp underflow -> 0 w/h
    t[1, 2, 4, 3, 3] = 30.0                 # score saturates to 1
    t[1, 2, 4, 3, 4] = -120.0               # score underflows to 0
    head = t.reshape(2, A * (5 + nc), H, H).contiguous()
    anchors = synth.YOLOV3_ANCHORS[3:6]
    rows = ref_decode_v3(head, nc, anchors, 160).numpy()
    orows = O.decode_predictions_v3(head.numpy(), nc, anchors, 160)
    with np.errstate(invalid="ignore", divide="ignore"):
        fin = np.isfinite(rows[..., :5])
        rel = np.where(fin, np.abs(orows[..., :5] - rows[..., :5]) / np.maximum(np.abs(rows[..., :5]), 1e-30), 0)
    same_inf = np.array_equal(np.isinf(rows), np.isinf(orows))
    check("decode_edges", rel.max() <= 1e-5 and same_inf and np.array_equal(rows[..., 5], orows[..., 5]),
          "max rel %.2e" % rel.max())
    np.savez_compressed(os.path.join(HERE, "decode_edges.npz"), head=head.numpy(), rows=rows, nc=nc, A=A,
                        anchors=np.array(anchors, np.float64), input_size=160)


# ------------------------------------------------------------------ wrappers (DecodeYoloV2/V3)

def case_wrappers():
    out = {}
    heads = synth.make_heads(1, 20, [52, 26, 13], g=5, mu_obj=-3.0)
    dec = DecodeYoloV3(20, synth.YOLOV3_ANCHORS, 416, conf_threshold=0.25)
    r3 = dec(heads).numpy()
    out["v3_rows"] = r3
    out["v3_head_crc"] = np.array([crc(h.numpy()) for h in heads], np.uint64)
    h2 = synth.make_heads(1, 20, [13], A=5, g=6, mu_obj=-1.0)
    dec2 = DecodeYoloV2(20, synth.YOLOV2_SCALED_ANCHORS, 416, conf_threshold=0.25)
    r2 = dec2(h2[0]).numpy()
    out["v2_rows"] = r2
    out["v2_head_crc"] = np.array([crc(h2[0].numpy())], np.uint64)
    orows = O.decode_multi_scale([h.numpy() for h in heads], 20, synth.YOLOV3_ANCHORS, 416)[0]
    ok3 = O.nms_v3(orows, 0.25)
    check("wrap:DecodeYoloV3", ok3.shape == r3.shape and np.allclose(ok3, r3, rtol=1e-5, atol=0), "K=%d" % len(r3))
    o2 = O.nms_v3(O.decode_predictions_v2(h2[0].numpy(), 20, synth.YOLOV2_SCALED_ANCHORS, 416)[0], 0.25)
    check("wrap:DecodeYoloV2", o2.shape == r2.shape and np.allclose(o2, r2, rtol=1e-5, atol=0), "K=%d" % len(r2))
    np.savez_compressed(os.path.join(HERE, "wrappers.npz"), **out)


# ------------------------------------------------------------------ mAP

def case_map():
    out = {}
    nc = 20
    # accumulator run: 2 batches of 8 images, sparse-ish heads, through the reference class
    m = RefMAPv3(nc, synth.YOLOV3_ANCHORS, 416, 0.25)
    m.reset_states()
    annots = []
    for g in range(2):
        heads = synth.make_heads(8, nc, [52, 26, 13], g=40 + g, mu_obj=-4.0)
        an = synth.make_annots(8, nc, g=40 + g)
        annots.append(an.numpy())
        m.update_state(an, heads)
    res = m.result()
    gt_rows = m.all_true_boxes_variable.numpy()
    det_rows = m.all_pred_boxes_variable.numpy()
    pc = ref_yu.metrics_per_class(m.all_true_boxes_variable, m.all_pred_boxes_variable, nc).numpy()
    out["acc_annots"] = np.stack(annots)
    out["acc_gt_rows"] = gt_rows
    out["acc_det_rows"] = det_rows
    out["acc_map"] = np.float32(res.item())
    out["acc_per_class"] = pc
    ogt = np.concatenate([O.get_target_rows(a, 416, 8 * i) for i, a in enumerate(annots)])
    check("map:get_target_rows", np.array_equal(ogt, gt_rows), "G=%d" % len(gt_rows))
    om, otp, opc = O.mean_average_precision(gt_rows, det_rows, nc, 0.5, True)
    check("map:accumulator", abs(float(om) - float(res)) <= 1e-6 and np.allclose(opc, pc, atol=1e-6),
          "mAP ref %.6f oracle %.6f dets %d" % (float(res), float(om), len(det_rows)))

    # synthetic matching stress: detections jittered around GT so TP/FP/duplicates all occur
    rng = np.random.RandomState(11)
    n_img, nc2 = 48, 6
    gts, dets = [], []
    for i in range(n_img):
        ng = rng.randint(0, 6)
        g = np.concatenate([np.full((ng, 1), i), rng.uniform(60, 350, (ng, 2)), rng.uniform(20, 120, (ng, 2)),
                            np.ones((ng, 1)), rng.randint(0, nc2 - 1, (ng, 1))], 1)   # last class never has GT
        gts.append(g)
        nd = rng.randint(0, 14)
        d = []
        for _ in range(nd):
            if ng and rng.rand() < .7:
                src = g[rng.randint(ng)]
                box = src[1:5] * rng.uniform(.8, 1.25, 4) + np.concatenate([rng.normal(0, 6, 2), [0, 0]])
                cls = src[6] if rng.rand() < .85 else rng.randint(0, nc2)
            else:
                box = np.concatenate([rng.uniform(0, 416, 2), rng.uniform(10, 150, 2)])
                cls = rng.randint(0, nc2)
            d.append(np.concatenate([[i], box, [rng.uniform(.25, 1)], [cls]]))
        dets.append(np.array(d).reshape(-1, 7))
    gt = torch.tensor(np.concatenate(gts), dtype=torch.float32)
    det = torch.tensor(np.concatenate(dets), dtype=torch.float32)
    r = ref_yu.mean_average_precision(gt, det, nc2)
    pc = ref_yu.metrics_per_class(gt, det, nc2).numpy()
    om, otp, opc = O.mean_average_precision(gt.numpy(), det.numpy(), nc2, 0.5, True)
    check("map:synthetic", abs(float(om) - float(r)) <= 1e-6 and np.allclose(opc, pc, atol=1e-6),
          "mAP ref %.6f oracle %.6f G=%d K=%d TP=%d" % (float(r), float(om), len(gt), len(det), int(pc[:, 1].sum())))
    out["syn_gt"] = gt.numpy(); out["syn_det"] = det.numpy(); out["syn_nc"] = nc2
    out["syn_map"] = np.float32(r.item()); out["syn_per_class"] = pc
    # per-detection TP flags are not exposed by the reference; the oracle's flags are stored and
    # their per-class sums are pinned by metrics_per_class above
    out["syn_tp_oracle"] = otp
    # no detections at all / no GT at all
    r0 = ref_yu.mean_average_precision(gt, torch.zeros((0, 7)), nc2)
    o0 = O.mean_average_precision(gt.numpy(), np.zeros((0, 7), np.float32), nc2)
    check("map:no_dets", float(r0) == float(o0) == 0.0)
    np.savez_compressed(os.path.join(HERE, "map_cases.npz"), **out)


# ------------------------------------------------------------------ IoU helpers

def case_iou():
    rng = np.random.RandomState(3)
    a = np.concatenate([rng.uniform(0, 300, (37, 2)), rng.uniform(0, 300, (37, 2))], 1).astype(np.float32)
    a[:, 2:] += a[:, :2]
    b = np.concatenate([rng.uniform(0, 300, (53, 2)), rng.uniform(0, 300, (53, 2))], 1).astype(np.float32)
    b[:, 2:] += b[:, :2]
    a[3] = b[5]                     # identical pair -> iou 1
    b[7, 2:] = b[7, :2]             # zero-area column
    with np.errstate(invalid="ignore"):
        ref = ref_box_iou(torch.tensor(a), torch.tensor(b)).numpy()
        ora = O.box_iou(a, b)
    check("iou:box_iou", np.array_equal(ref, ora, equal_nan=True))
    c1 = np.concatenate([rng.uniform(0, 416, (200, 2)), rng.uniform(1, 200, (200, 2))], 1).astype(np.float32)
    c2 = np.concatenate([rng.uniform(0, 416, (200, 2)), rng.uniform(1, 200, (200, 2))], 1).astype(np.float32)
    c2[:20] = c1[:20]
    refc = ref_yu.bbox_iou(torch.tensor(c1), torch.tensor(c2)).numpy()[:, 0]
    orac = O.bbox_iou_cxcywh(c1, c2)
    check("iou:bbox_iou_cxcywh", np.array_equal(refc, orac))
    np.savez_compressed(os.path.join(HERE, "iou_cases.npz"), a=a, b=b, box_iou=ref, c1=c1, c2=c2, bbox_iou=refc)


# ------------------------------------------------------------------ detection writers

def case_writers():
    """The reference's writer loops are inline in CLI scripts that need a checkpoint and a dataset
    (make_pred_file_yolov3.py:60-86, yolo2coco_pred_file_yolov3.py:66-100), so they cannot be called;
    the loop BODIES below are those lines verbatim in meaning, run on real torch 0-dim tensors."""
    rng = np.random.RandomState(21)
    n = 400
    boxes = np.concatenate([rng.uniform(-20, 436, (n, 2)), rng.uniform(0.2, 300, (n, 2)),
                            rng.uniform(0.25, 1, (n, 1)), rng.randint(0, 20, (n, 1))], 1).astype(np.float32)
    boxes[0, :4] = [10.5, 10.5, 21.0, 21.0]          # xmin = 0 exactly
    boxes[1, :4] = [3.0, 4.0, 10.0, 9.5]             # negative corners truncate toward zero
    boxes[2, :4] = [100.25, 50.75, 0.5, 0.5]
    names = ["cls%02d" % i for i in range(20)]
    cats = {nm: 100 + 3 * i for i, nm in enumerate(names)}
    t = torch.tensor(boxes)
    # make_pred_file_yolov3.py:72-86
    lines, ints = [], []
    for bbox in t:
        class_name = names[int(bbox[-1])]
        confidence_score = bbox[4]
        cx, cy, w, h = bbox[0], bbox[1], bbox[2], bbox[3]
        xmin = int((cx - (w / 2)))
        ymin = int((cy - (h / 2)))
        xmax = int((cx + (w / 2)))
        ymax = int((cy + (h / 2)))
        lines.append(f'{class_name} {confidence_score} {xmin} {ymin} {xmax} {ymax}\n')
        ints.append([xmin, ymin, xmax, ymax])
    ints = np.array(ints, dtype=np.int64)
    check("writers:pred_file ints", np.array_equal(O.detection_ints(boxes, O.WRITER_PRED_FILE), ints))
    check("writers:pred_file lines", O.pred_file_lines(boxes, names) == lines)
    # yolo2coco_pred_file_yolov3.py:83-100, three images of different original sizes
    input_size = 416
    sizes = [(500, 375), (1280, 720), (333, 500)]
    img_of = rng.randint(0, 3, n)
    coco = []
    for k, bbox in enumerate(t):
        width, height = sizes[img_of[k]]
        cx = bbox[0] * (width / input_size)
        cy = bbox[1] * (height / input_size)
        w = bbox[2] * (width / input_size)
        h = bbox[3] * (height / input_size)
        xmin = int((cx - (w / 2)))
        ymin = int((cy - (h / 2)))
        w = int(w)
        h = int(h)
        coco.append([xmin, ymin, w, h])
    coco = np.array(coco, dtype=np.int64)
    rows7 = np.concatenate([(img_of + 7).astype(np.float32)[:, None], boxes], 1)
    scale = [[w_ / input_size, h_ / input_size] for w_, h_ in sizes]
    check("writers:coco ints", np.array_equal(O.detection_ints(rows7, O.WRITER_COCO, scale, img_idx_base=7), coco))
    np.savez_compressed(os.path.join(HERE, "writers.npz"), boxes=boxes, pred_ints=ints, lines=np.array(lines),
                        img_of=img_of, sizes=np.array(sizes), input_size=input_size, coco_ints=coco,
                        names=np.array(names), cat_ids=np.array([cats[nm] for nm in names]))


ULTRA_VARIANTS = [dict(), dict(multi_label=True), dict(agnostic=True), dict(classes=[1, 3]),
                  dict(multi_label=True, max_det=9), dict(conf_thres=0.6, iou_thres=0.3),
                  dict(multi_label=True, agnostic=True, classes=[0, 2, 4]), dict(max_det=5000),
                  dict(multi_label=True, max_det=5000, conf_thres=0.1)]


def ultra_pred(seed, B, n, nc, far=False):
    """[B,n,5+nc] rows like the ones a YOLOv5 Detect layer emits; `far` adds boxes whose corners leave
    [0, max_wh): classes then share coordinates after the + cls * 4096 shift."""
    rng = np.random.RandomState(seed)
    p = np.zeros((B, n, 5 + nc), np.float32)
    cxy = rng.uniform(0, 640, (B, n // 4, 2))
    p[..., 0:2] = np.repeat(cxy, 4, axis=1)[:, :n] + rng.normal(0, 6, (B, n, 2))      # crowds of ~4 around a centre
    p[..., 2:4] = rng.uniform(8, 160, (B, n, 2))
    p[..., 4] = rng.uniform(0, 1, (B, n))
    p[..., 5:] = rng.uniform(0, 1, (B, n, nc)) ** 2
    p[:, ::7, 4] = np.float32(0.8125)                    # tied objectness
    p[:, ::7, 5:] = np.float32(0.5)                      # ... and tied class scores: first maximum, (box, j) order
    if far:
        k = n // 10
        p[:, :k, 0] = rng.uniform(4000, 4200, (B, k))    # straddle the first class band's end
        p[:, k:2 * k, 0] = rng.uniform(-80, 10, (B, k))
        p[:, :2 * k, 2] = rng.uniform(100, 400, (B, 2 * k))
    return p


def case_ultralytics():
    """utils/general.py:640-731 run as is (CPU tensors -> torchvision's CPU nms)."""
    from utils.general import non_max_suppression as ref_nms
    store = {}
    for tag, far in (("near", False), ("far", True)):
        p = ultra_pred(31 if not far else 32, 3, 600, 5, far)
        store[tag + "_pred"] = p
        for vi, kw in enumerate(ULTRA_VARIANTS):
            ref = ref_nms(torch.from_numpy(p.copy()), **kw)
            got = O.non_max_suppression(p, **kw)
            ok = all(np.array_equal(r.numpy(), g) for r, g in zip(ref, got))
            check("ultralytics:%s:%d %s" % (tag, vi, kw), ok, "kept %s" % [len(r) for r in ref])
            store["%s_%d_rows" % (tag, vi)] = np.concatenate([r.numpy() for r in ref], 0)
            store["%s_%d_counts" % (tag, vi)] = np.array([len(r) for r in ref])
    # single class: multi_label is switched off (:659)
    p1 = ultra_pred(33, 2, 300, 1)
    ref = ref_nms(torch.from_numpy(p1.copy()), multi_label=True)
    check("ultralytics:nc1", all(np.array_equal(r.numpy(), g) for r, g in zip(ref, O.non_max_suppression(p1, multi_label=True))))
    store["nc1_pred"] = p1
    store["nc1_rows"] = np.concatenate([r.numpy() for r in ref], 0)
    store["nc1_counts"] = np.array([len(r) for r in ref])
    np.savez_compressed(os.path.join(HERE, "ultralytics.npz"), **store)


def case_yolov1():
    """utils/yolov1_utils.py:12-66 on random logits (values differ from a libm sigmoid by an ulp: 1e-5 relative)."""
    from utils.yolov1_utils import decode_predictions as ref_v1
    rng = np.random.RandomState(41)
    store = {}
    for tag, nc, nb in (("voc_b2", 20, 2), ("nc3_b3", 3, 3)):
        x = rng.normal(0, 2.5, (5, 7 * 7 * (5 * nb + nc))).astype(np.float32)
        x[0, :40] = 30.0                                  # saturated sigmoids tie: first index wins
        r = ref_v1(torch.from_numpy(x), nc, nb, 448).numpy()
        g = O.decode_predictions_v1(x, nc, nb, 448)
        check("yolov1:%s" % tag, np.array_equal(r[..., 5], g[..., 5]) and np.allclose(r, g, rtol=1e-5, atol=1e-6))
        store[tag + "_in"] = x
        store[tag + "_out"] = r
    np.savez_compressed(os.path.join(HERE, "yolov1.npz"), **store)


def confusion_inputs(seed, n_img, nc):
    rng = np.random.RandomState(seed)
    dets, labs = [], []
    for it in range(n_img):
        M, N = rng.randint(0, 12), rng.randint(0, 40)
        lab = np.zeros((M, 5), np.float32)
        xy, wh = rng.uniform(0, 300, (M, 2)), rng.uniform(20, 120, (M, 2))
        lab[:, 0] = rng.randint(0, nc, M); lab[:, 1:3] = xy; lab[:, 3:5] = xy + wh
        det = np.zeros((N, 6), np.float32)
        for k in range(N):
            if M and rng.rand() < 0.7:
                det[k, :4] = lab[rng.randint(0, M), 1:] + rng.normal(0, 12, 4)
            else:
                c = rng.uniform(0, 300, 2)
                det[k, :2] = c; det[k, 2:4] = c + rng.uniform(20, 120, 2)
        det[:, 4] = rng.uniform(0, 1, N); det[:, 5] = rng.randint(0, nc, N)
        if it == 3:
            det = det[:0]
        if it == 5:
            det[:, :4] += 2000.0                          # no overlap at all: unmatched detections are NOT counted (:156)
        dets.append(det); labs.append(lab)
    return dets, labs


def case_confusion():
    """ConfusionMatrix.process_batch, utils/metrics.py:122-160, image by image."""
    from utils.metrics import ConfusionMatrix
    nc = 6
    dets, labs = confusion_inputs(51, 40, nc)
    cm = ConfusionMatrix(nc, conf=0.3, iou_thres=0.4)
    m = np.zeros((nc + 1, nc + 1))
    for d, l in zip(dets, labs):
        cm.process_batch(torch.from_numpy(d), torch.from_numpy(l))
        O.confusion_process_batch(m, d, l, nc, 0.3, 0.4)
    check("confusion", np.array_equal(cm.matrix, m), "sum %d" % m.sum())
    np.savez_compressed(os.path.join(HERE, "confusion.npz"), det=np.concatenate(dets), lab=np.concatenate(labs),
                        det_off=np.cumsum([0] + [len(d) for d in dets]), lab_off=np.cumsum([0] + [len(l) for l in labs]),
                        matrix=cm.matrix, nc=nc, conf=0.3, iou_thres=0.4)


# ------------------------------------------------------------------ utils/metrics.py: AP side + bbox_iou

def case_metrics_ap():
    """ap_per_class (:21-83), compute_ap (:86-111) and bbox_iou (:192-230) of utils/metrics.py run as they are."""
    rng = np.random.RandomState(61)
    out = {}
    for tag, n, T, ncp, nct, m in (("t1", 900, 1, 7, 6, 260), ("t10", 700, 10, 5, 5, 300), ("few", 12, 1, 3, 4, 9)):
        conf = rng.uniform(0, 1, n)
        conf[::11] = 0.5                                  # tied confidences
        tp = rng.uniform(0, 1, (n, T)) < np.linspace(0.7, 0.2, T)[None]
        pred_cls = rng.randint(0, ncp, n).astype(np.float64)
        target_cls = rng.randint(0, nct, m).astype(np.float64)
        ref = ref_metrics.ap_per_class(tp, conf, pred_cls, target_cls, names={})
        got = O.ap_per_class(tp, conf, pred_cls, target_cls)
        check("metrics:ap_per_class:" + tag, all(np.array_equal(a, b) for a, b in zip(ref, got)), "mAP %.4f" % ref[2].mean())
        for k, v in zip(("tp", "conf", "pred_cls", "target_cls"), (tp, conf, pred_cls, target_cls)):
            out["%s_%s" % (tag, k)] = v
        for k, v in zip(("p", "r", "ap", "f1", "classes"), ref):
            out["%s_%s" % (tag, k)] = v
    rec = np.sort(rng.uniform(0, 1, 64))
    pre = rng.uniform(0, 1, 64)
    ap, mpre, mrec = ref_metrics.compute_ap(rec, pre)
    oap, ompre, omrec = O.compute_ap(rec, pre)
    check("metrics:compute_ap", ap == oap and np.array_equal(mpre, ompre) and np.array_equal(mrec, omrec), "ap %.6f" % ap)
    out.update(ca_recall=rec, ca_precision=pre, ca_ap=np.float64(ap), ca_mpre=mpre, ca_mrec=mrec)
    # bbox_iou: xywh and xyxy, n-to-n and 1-to-n, the four returns
    b1 = np.concatenate([rng.uniform(0, 416, (240, 2)), rng.uniform(1, 200, (240, 2))], 1).astype(np.float32)
    b2 = np.concatenate([rng.uniform(0, 416, (240, 2)), rng.uniform(1, 200, (240, 2))], 1).astype(np.float32)
    b2[:12] = b1[:12]                                    # identical pairs
    b2[12:24, :2] = b1[12:24, :2] + 1000.0               # disjoint pairs
    b1[24, 2:] = 0.0                                     # a zero-area box
    out["bi_box1"], out["bi_box2"] = b1, b2
    for xywh in (True, False):
        A, B = b1.copy(), b2.copy()
        if not xywh:
            A[:, 2:] += A[:, :2]
            B[:, 2:] += B[:, :2]
        for mode, kw in (("iou", {}), ("giou", {"GIoU": True}), ("diou", {"DIoU": True}), ("ciou", {"CIoU": True})):
            for bc, a_in in (("nn", A), ("1n", A[:1])):
                with np.errstate(all="ignore"):
                    ref = ref_metrics.bbox_iou(torch.tensor(a_in), torch.tensor(B), xywh=xywh, **kw).numpy()
                    got = O.bbox_iou_metrics(a_in, B, xywh=xywh, **kw)
                ok = np.array_equal(ref, got, equal_nan=True) if mode != "ciou" else np.allclose(ref, got, rtol=0, atol=1e-6, equal_nan=True)
                check("metrics:bbox_iou:%s:%s:%s" % ("xywh" if xywh else "xyxy", mode, bc), ok)
                out["bi_%s_%s_%s" % ("xywh" if xywh else "xyxy", mode, bc)] = ref
    np.savez_compressed(os.path.join(HERE, "metrics_ap.npz"), **out)


# ------------------------------------------------------------------ objectness x class scores (YPP_SCORE_OBJ_TIMES_CLS)

def decoded_prediction(heads, nc, anchors, input_size, A=3):
    """[B, n, 5+nc] rows cx,cy,w,h,obj,cls_0.. the way a Detect layer would emit them: the boxes and the objectness
    are the reference decode's own (utils/yolov3_utils.py:12-59), the class columns torch.sigmoid of the class
    logits in the same box order (:45-46,51: view -> permute -> sigmoid)."""
    rows = torch.cat([ref_decode_v3(h, nc, anchors[A * i: A * i + A], input_size) for i, h in enumerate(heads)], dim=1)
    cls = []
    for h in heads:
        B, _, H, W = h.shape
        t = h.view(B, A, 5 + nc, H, W).permute(0, 1, 3, 4, 2)[..., 5:]
        cls.append(torch.sigmoid(t).reshape(B, A * H * W, nc))
    return torch.cat([rows[..., :5], torch.cat(cls, dim=1)], dim=2).contiguous()


def case_objcls():
    """The scoring of the vendored NMS (utils/general.py:645,669,679-680) on decoded heads: run the reference function
    with iou_thres = 1 (nothing is ever suppressed: IoU <= 1) and no max_det cap, so its output IS the scored
    candidate list; pins oracle.best_class_scores and tests/util.ultra_best_class (the torch-op restatement the GPU
    test runs on the device)."""
    from utils.general import non_max_suppression as ref_nms
    sys.path.insert(0, os.path.dirname(HERE))
    import util as T
    nc, conf = 20, 0.25
    heads = synth.make_heads(2, nc, [52, 26, 13], g=70, mu_obj=-1.5)
    # class logits that tie after the sigmoid (saturation) and products that collide after the rounding
    heads[0][0, 5:25, 0, 0] = 20.0
    heads[0][0, 4, 0, 0] = 2.0
    # 208 cells whose two best class probabilities sit an ulp or two apart, the larger at the HIGHER index: where their
    # products with the objectness round to the same float, the first maximum of the products is the lower index
    gen = torch.Generator("cpu").manual_seed(71)
    v = torch.rand(8, 26, generator=gen) * 3 - 1
    delta = 2.0 ** -23 * (1 + torch.exp(v)) * (0.3 + 0.9 * torch.rand(8, 26, generator=gen))
    heads[1][1, 5:25, :8, :] = -3.0
    heads[1][1, 5 + 7, :8, :] = v
    heads[1][1, 5 + 2, :8, :] = v - delta
    heads[1][1, 4, :8, :] = torch.rand(8, 26, generator=gen) * 2
    pred = decoded_prediction(heads, nc, synth.YOLOV3_ANCHORS, 416)
    ref = ref_nms(pred.clone(), conf_thres=conf, iou_thres=1.0, max_det=1000000)
    store = {"g": 70, "mu_obj": np.float32(-1.5), "nc": nc, "conf": np.float32(conf),
             "head_crc": np.array([crc(h.numpy()) for h in heads], np.uint64)}
    ok_o, ok_t = True, True
    for b in range(2):
        idx, sc, cl = O.best_class_scores(pred[b].numpy(), conf)
        r = ref[b].numpy()
        order = np.lexsort((idx, -sc.astype(np.float64)))         # the reference sorts by conf (stable on CPU)
        ok_o &= len(r) == len(idx) and np.array_equal(r[:, 4], sc[order]) and np.array_equal(r[:, 5], cl[order].astype(np.float32))
        ti, ts, tc = T.ultra_best_class(pred[b], conf)
        ok_t &= np.array_equal(ti.numpy(), idx) and np.array_equal(ts.numpy(), sc) and np.array_equal(tc.numpy(), cl)
        store["idx_%d" % b], store["conf_%d" % b], store["cls_%d" % b] = idx, sc, cl
    check("objcls:oracle.best_class_scores == reference scoring", ok_o, "cand %s" % [len(store["idx_%d" % b]) for b in range(2)])
    check("objcls:tests/util.ultra_best_class == oracle", ok_t)
    am = pred[1][store["idx_1"]][:, 5:].argmax(1).numpy()
    n_coll = int((am != store["cls_1"]).sum())
    check("objcls:fixture holds colliding products", n_coll >= 10, "%d boxes whose class is not argmax(sigmoid(cls))" % n_coll)
    store["n_collisions"] = n_coll
    np.savez_compressed(os.path.join(HERE, "objcls.npz"), **store)


# ------------------------------------------------------------------ heads of the random-init eval-mode reference model

def case_model_tie():
    """SURVEY 8(d), third input regime: the heads of the reference's own YoloV3 (models/detector/yolov3.py:71-91, darknet19
    features, random init, eval mode) on torch.rand images, seed 0.  Every head value is ~1e-5 .. 1e-9, so every score
    is 0.5 to within a few ulps (exactly 0.5 on the coarse scales): the ordering is decided by ties -> box index."""
    for name in ("torchinfo", "timm", "pytorch_lightning"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    if not hasattr(sys.modules["torchinfo"], "summary"):
        sys.modules["torchinfo"].summary = lambda *a, **k: None
    from utils.module_select import get_model
    from models.detector.yolov3 import YoloV3
    torch.manual_seed(0)
    backbone = get_model("darknet19")(pretrained="", features_only=True, out_indices=[3, 4, 5])
    model = YoloV3(backbone, 20, 9).eval()
    x = torch.rand(2, 3, 224, 224)
    with torch.no_grad():
        heads = [h.contiguous() for h in model(x)]
    nc, conf, S = 20, 0.25, 224
    rows = torch.cat([ref_decode_v3(h, nc, synth.YOLOV3_ANCHORS[3 * i: 3 * i + 3], S) for i, h in enumerate(heads)], dim=1)
    rows_np = rows.numpy()
    orows = O.decode_multi_scale([h.numpy() for h in heads], nc, synth.YOLOV3_ANCHORS, S, 3)
    rel = np.abs(orows[..., :5] - rows_np[..., :5]) / np.maximum(np.abs(rows_np[..., :5]), 1e-30)
    check("model_tie:decode", rel.max() <= 1e-5, "max rel %.2e, scores == 0.5: %.1f%%, distinct scores %d" % (
        rel.max(), 100 * float((rows_np[..., 4] == 0.5).mean()), len(np.unique(rows_np[..., 4]))))
    keeps, counts = [], []
    ok = True
    for b in range(2):
        k, _ = ref_nms_indices(rows[b], conf)
        om, um = cpu_modes(int((rows_np[b, :, 4] > conf).sum()))
        ok_, _ = O.nms_v3_indices(rows_np[b], conf, 0.45, offset_mode=om, union_mode=um)
        ref = k if om == O.OFFSET_TRICK else canon(k, rows_np[b])
        ok &= np.array_equal(ok_, ref)
        keeps.append(ref); counts.append(len(ref))
    check("model_tie:nms", ok, "kept %s of %d" % (counts, rows_np.shape[1]))
    keep_arr = np.full((2, max(counts)), -1, np.int32)
    for b, k in enumerate(keeps):
        keep_arr[b, :len(k)] = k
    np.savez_compressed(os.path.join(HERE, "model_tie.npz"), heads0=heads[0].numpy(), heads1=heads[1].numpy(),
                        heads2=heads[2].numpy(), rows=rows_np, keep=keep_arr, keep_count=np.array(counts, np.int32),
                        nc=nc, conf=np.float32(conf), input_size=S)


if __name__ == "__main__":
    O.build(force=True)
    if len(sys.argv) > 1:                      # regenerate only the named cases, e.g. `make_golden.py writers`
        for nm in sys.argv[1:]:
            globals()["case_" + nm]()
        print("FAILED: %s" % FAIL if FAIL else "all golden cases agree with the oracle")
        sys.exit(1 if FAIL else 0)
    case_decode_edges()
    case_adversarial()
    # C1: YOLOv2 416 VOC batch 8 dense (trick path, n<=1000 per image)
    case_heads("c1_yolov2_dense", "v2", 20, [13], 5, 8, 416, 0.25, 0.0, g=1)
    # C2 shape, 2 images, dense (vanilla path on CPU) and sparse (trick path)
    case_heads("c2_yolov3_dense_b2", "v3", 20, [52, 26, 13], 3, 2, 416, 0.25, 0.0, g=2, store_rows_stride=16)
    case_heads("c2_yolov3_sparse_b4", "v3", 20, [52, 26, 13], 3, 4, 416, 0.25, -5.0, g=3, store_rows_stride=16)
    # C3 shape (single class), 1 image dense
    case_heads("c3_person_dense_b1", "v3", 1, [52, 26, 13], 3, 1, 416, 0.25, 0.0, g=4, store_rows_stride=16)
    # BASELINE config 3 as written: the canonical 2-scale YOLOv4-tiny head (26/13), single class, 2 images dense
    case_heads("c3_person2s_dense_b2", "v3", 1, [26, 13], 3, 2, 416, 0.25, 0.0, g=6, store_rows_stride=4)
    # C4 shape, 1 image, conf .001
    case_heads("c4_coco608_b1", "v3", 80, [76, 38, 19], 3, 1, 608, 0.001, 0.0, g=5, store_rows_stride=32)
    case_wrappers()
    case_map()
    case_iou()
    case_writers()
    case_ultralytics()
    case_yolov1()
    case_confusion()
    case_metrics_ap()
    case_objcls()
    case_model_tie()
    print("FAILED: %s" % FAIL if FAIL else "all golden cases agree with the oracle")
    sys.exit(1 if FAIL else 0)
