"""CPU: the oracle (oracle/ypp_oracle.c) against the reference's own outputs (tests/golden/).

These fixtures were produced by tests/golden/make_golden.py importing the UNMODIFIED reference;
this is what pins the oracle.  Tolerances: decoded coordinates/scores 1e-5 relative (fp32,
different libm), everything index-like bit-exact.
"""
import numpy as np
import pytest

import util


@pytest.mark.parametrize("name", util.HEAD_CASES)
def test_decode_and_nms_against_reference(oracle, name):
    g = util.load(name)
    heads = [h.numpy() for h in util.case_heads(g)]
    nc, A, S = int(g["nc"]), int(g["A"]), int(g["input_size"])
    anchors = util.case_anchors(g)
    if str(g["kind"]) == "v2":
        rows = oracle.decode_predictions_v2(heads[0], nc, anchors, S)
    else:
        rows = oracle.decode_multi_scale(heads, nc, anchors, S, A)
    stride = int(g["rows_stride"])
    ref = g["rows"]
    sub = rows[:, ::stride]
    assert sub.shape == ref.shape
    assert util.rel_err(sub[..., :5], ref[..., :5]).max() <= 1e-5          # tolerance: 1e-5 relative
    assert (sub[..., 5] != ref[..., 5]).sum() <= 2                          # argmax near-ties only
    conf = float(g["conf"])
    for b in range(int(g["batch"])):
        ncand = int(g["cand_count"][b])
        keep, onc = oracle.nms_v3_indices(rows[b], conf, 0.45, offset_mode=oracle.OFFSET_AUTO_CPU,
                                          union_mode=oracle.UNION_SEPARATE)
        k_ref = g["keep"][b, : int(g["keep_count"][b])]
        # end-to-end on the oracle's own decoded rows: identical unless a score sits within an ulp
        # of a decision; require exact equality here (holds for these seeds)
        assert onc == ncand
        assert np.array_equal(util.canon(keep, rows[b]), util.canon(k_ref, rows[b]))


def test_adversarial_nms(oracle):
    g = util.load("nms_adversarial")
    names = sorted({k.rsplit("__", 1)[0] for k in g.files})
    assert len(names) >= 15
    for n in names:
        rows, keep = g[n + "__rows"], g[n + "__keep"]
        conf, iou = g[n + "__conf_iou"]
        om, um = g[n + "__mode"]
        got, _ = oracle.nms_v3_indices(rows, conf, iou, offset_mode=int(om), union_mode=int(um))
        assert np.array_equal(got, keep), n
        # and through the auto heuristic the CPU build of torchvision uses
        got2, _ = oracle.nms_v3_indices(rows, conf, iou, offset_mode=oracle.OFFSET_AUTO_CPU, union_mode=int(um))
        assert np.array_equal(got2, keep), n


def test_trick_vs_vanilla_differ_on_cross_class(oracle):
    g = util.load("nms_adversarial")
    rows = g["trick_cross_class__rows"]
    t, _ = oracle.nms_v3_indices(rows, .25, .45, offset_mode=oracle.OFFSET_TRICK)
    v, _ = oracle.nms_v3_indices(rows, .25, .45, offset_mode=oracle.OFFSET_NONE)
    assert len(t) == 2 and len(v) == 3


def test_decode_edges(oracle):
    g = util.load("decode_edges")
    rows = oracle.decode_predictions_v3(g["head"], int(g["nc"]), g["anchors"].tolist(), int(g["input_size"]))
    ref = g["rows"]
    assert np.array_equal(np.isinf(rows), np.isinf(ref))
    assert util.rel_err(rows[..., :5], ref[..., :5]).max() <= 1e-5
    assert np.array_equal(rows[..., 5], ref[..., 5])


def test_wrappers(oracle):
    from ypp import synth
    g = util.load("wrappers")
    heads = synth.make_heads(1, 20, [52, 26, 13], g=5, mu_obj=-3.0)
    assert [util.crc(h.numpy()) for h in heads] == [int(x) for x in g["v3_head_crc"]]
    rows = oracle.decode_multi_scale([h.numpy() for h in heads], 20, synth.YOLOV3_ANCHORS, 416)[0]
    out = oracle.nms_v3(rows, 0.25)
    assert out.shape == g["v3_rows"].shape and np.allclose(out, g["v3_rows"], rtol=1e-5, atol=0)
    h2 = synth.make_heads(1, 20, [13], A=5, g=6, mu_obj=-1.0)
    out2 = oracle.nms_v3(oracle.decode_predictions_v2(h2[0].numpy(), 20, synth.YOLOV2_SCALED_ANCHORS, 416)[0], 0.25)
    assert out2.shape == g["v2_rows"].shape and np.allclose(out2, g["v2_rows"], rtol=1e-5, atol=0)


def test_map_against_reference(oracle):
    g = util.load("map_cases")
    annots = g["acc_annots"]
    gt = np.concatenate([oracle.get_target_rows(a, 416, 8 * i) for i, a in enumerate(annots)])
    assert np.array_equal(gt, g["acc_gt_rows"])
    m, tp, pc = oracle.mean_average_precision(g["acc_gt_rows"], g["acc_det_rows"], 20, 0.5, True)
    assert abs(float(m) - float(g["acc_map"])) <= 1e-6                     # tolerance: 1e-6 absolute
    assert np.allclose(pc, g["acc_per_class"], atol=1e-6)
    m, tp, pc = oracle.mean_average_precision(g["syn_gt"], g["syn_det"], int(g["syn_nc"]), 0.5, True)
    assert abs(float(m) - float(g["syn_map"])) <= 1e-6
    assert np.allclose(pc, g["syn_per_class"], atol=1e-6)
    assert np.array_equal(pc[:, 1:], g["syn_per_class"][:, 1:])            # TP/FP/FN counts exact
    assert np.array_equal(tp, g["syn_tp_oracle"])
    assert float(oracle.mean_average_precision(g["syn_gt"], np.zeros((0, 7), np.float32), int(g["syn_nc"]))) == 0.0


def test_iou_helpers(oracle):
    g = util.load("iou_cases")
    with np.errstate(invalid="ignore"):
        assert np.array_equal(oracle.box_iou(g["a"], g["b"]), g["box_iou"], equal_nan=True)   # bit-exact
    assert np.array_equal(oracle.bbox_iou_cxcywh(g["c1"], g["c2"]), g["bbox_iou"])


def test_effective_threshold(oracle):
    # .45 rounds DOWN in fp32, .3 rounds UP: only the latter changes under the CPU kernel's double compare
    f45, f3 = np.float32(0.45), np.float32(0.3)
    assert oracle.effective_thr(0.45, True) == float(f45) == oracle.effective_thr(0.45, False)
    assert oracle.effective_thr(0.3, False) == float(f3)
    assert oracle.effective_thr(0.3, True) == float(np.nextafter(f3, np.float32(-1)))


def test_detection_writers(oracle):
    """Pred-file lines / COCO bbox ints vs what the reference's writer loop bodies produced on torch tensors."""
    g = util.load("writers")
    names = [str(x) for x in g["names"]]
    assert np.array_equal(oracle.detection_ints(g["boxes"], oracle.WRITER_PRED_FILE), g["pred_ints"])
    assert oracle.pred_file_lines(g["boxes"], names) == [str(x) for x in g["lines"]]
    rows7 = np.concatenate([(g["img_of"] + 7).astype(np.float32)[:, None], g["boxes"]], 1)
    S = float(g["input_size"])
    scale = [[w / S, h / S] for w, h in g["sizes"].tolist()]
    assert np.array_equal(oracle.detection_ints(rows7, oracle.WRITER_COCO, scale, img_idx_base=7), g["coco_ints"])


# ---------------------------------------------------------------- alternate semantics
ULTRA_VARIANTS = [dict(), dict(multi_label=True), dict(agnostic=True), dict(classes=[1, 3]),
                  dict(multi_label=True, max_det=9), dict(conf_thres=0.6, iou_thres=0.3),
                  dict(multi_label=True, agnostic=True, classes=[0, 2, 4]), dict(max_det=5000),
                  dict(multi_label=True, max_det=5000, conf_thres=0.1)]


def _split(rows, counts):
    off = np.concatenate([[0], np.cumsum(counts)])
    return [rows[off[i]: off[i + 1]] for i in range(len(counts))]


def test_ultralytics_nms(oracle):
    """utils/general.py:640-731 as run by the reference (CPU), every option combination of the fixture."""
    g = util.load("ultralytics")
    for tag in ("near", "far"):
        for vi, kw in enumerate(ULTRA_VARIANTS):
            want = _split(g["%s_%d_rows" % (tag, vi)], g["%s_%d_counts" % (tag, vi)])
            got = oracle.non_max_suppression(g[tag + "_pred"], **kw)
            assert [len(x) for x in got] == [len(x) for x in want], (tag, kw)
            assert all(np.array_equal(a, b) for a, b in zip(got, want)), (tag, kw)
    got = oracle.non_max_suppression(g["nc1_pred"], multi_label=True)
    assert all(np.array_equal(a, b) for a, b in zip(got, _split(g["nc1_rows"], g["nc1_counts"])))


def test_yolov1_decode(oracle):
    g = util.load("yolov1")
    for tag, nc, nb in (("voc_b2", 20, 2), ("nc3_b3", 3, 3)):
        got = oracle.decode_predictions_v1(g[tag + "_in"], nc, nb, 448)
        want = g[tag + "_out"]
        assert np.array_equal(got[..., 5], want[..., 5])
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)      # sigmoid differs by an ulp between libms


def test_confusion_matrix(oracle):
    g = util.load("confusion")
    nc = int(g["nc"])
    m = np.zeros((nc + 1, nc + 1))
    do, lo = g["det_off"], g["lab_off"]
    for i in range(len(do) - 1):
        oracle.confusion_process_batch(m, g["det"][do[i]: do[i + 1]], g["lab"][lo[i]: lo[i + 1]], nc, float(g["conf"]), float(g["iou_thres"]))
    assert np.array_equal(m, g["matrix"])
