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


def test_metrics_ap_side_against_reference(oracle):
    """utils/metrics.py: ap_per_class (:21-83), compute_ap (:86-111), bbox_iou (:192-230) - the oracle's plain-loop
    restatements AND the product's host implementations (ypp.compat.metrics: float64 numpy by design)
    against the reference's stored outputs."""
    import importlib
    g = util.load("metrics_ap")
    M = importlib.import_module("ypp.compat.metrics")
    for tag in ("t1", "t10", "few"):
        args = [g["%s_%s" % (tag, k)] for k in ("tp", "conf", "pred_cls", "target_cls")]
        want = [g["%s_%s" % (tag, k)] for k in ("p", "r", "ap", "f1", "classes")]
        for impl in (oracle.ap_per_class, M.ap_per_class):
            got = impl(*args)
            for a, b in zip(got, want):
                assert np.array_equal(np.asarray(a), b), (tag, impl.__module__)
            assert got[4].dtype == np.int32
    for impl in (oracle.compute_ap, M.compute_ap):
        ap, mpre, mrec = impl(g["ca_recall"], g["ca_precision"])
        assert ap == float(g["ca_ap"]) and np.array_equal(mpre, g["ca_mpre"]) and np.array_equal(mrec, g["ca_mrec"])
    b1, b2 = g["bi_box1"], g["bi_box2"]
    for xywh in (True, False):
        A, B = b1.copy(), b2.copy()
        if not xywh:
            A[:, 2:] += A[:, :2]
            B[:, 2:] += B[:, :2]
        for mode, kw in (("iou", {}), ("giou", {"GIoU": True}), ("diou", {"DIoU": True}), ("ciou", {"CIoU": True})):
            for bc, a_in in (("nn", A), ("1n", A[:1])):
                want = g["bi_%s_%s_%s" % ("xywh" if xywh else "xyxy", mode, bc)]
                got = oracle.bbox_iou_metrics(a_in, B, xywh=xywh, **kw)
                if mode == "ciou":
                    assert np.allclose(got, want, rtol=0, atol=1e-6, equal_nan=True)      # atan differs by an ulp between libms
                else:
                    assert np.array_equal(got, want, equal_nan=True), (xywh, mode, bc)


def test_objectness_times_class_scoring(oracle):
    """YPP_SCORE_OBJ_TIMES_CLS semantics = the scoring lines of the vendored NMS (utils/general.py:645,669,679-680):
    oracle.best_class_scores and the torch-op restatement the GPU test uses, against what the reference function
    itself returned for the decoded fixture heads (iou_thres = 1: its output is the scored candidate list)."""
    import torch
    g, heads = util.objcls_heads()
    from ypp import synth
    pred = util.torch_prediction(heads, int(g["nc"]), synth.YOLOV3_ANCHORS, 416)
    assert int(g["n_collisions"]) >= 10
    for b in range(2):
        idx, sc, cl = oracle.best_class_scores(pred[b].numpy(), float(g["conf"]))
        ti, ts, tc = util.ultra_best_class(pred[b], float(g["conf"]))
        assert np.array_equal(ti.numpy(), idx) and np.array_equal(ts.numpy(), sc) and np.array_equal(tc.numpy(), cl)
        # against what the reference function returned when the fixture was made (its own decode fed it; here the
        # decode is restated, and a vectorised CPU sigmoid may differ by an ulp with the memory layout or the machine):
        # same candidates up to the threshold's edge, scores to 1e-6, classes equal outside the 208 crafted cells
        # (whose winner depends on the last bit by construction)
        assert len(set(idx.tolist()) ^ set(g["idx_%d" % b].tolist())) <= 2
        common, ia, ib = np.intersect1d(idx, g["idx_%d" % b], return_indices=True)
        assert util.rel_err(sc[ia], g["conf_%d" % b][ib]).max() <= 1e-6
        crafted = (common >= 8112) & (common < 8112 + 208) if b == 1 else np.zeros(len(common), bool)
        assert int(((cl[ia] != g["cls_%d" % b][ib]) & ~crafted).sum()) <= 2
        if b == 1:
            assert set(cl[ia][crafted].tolist()) <= {2, 7} and set(g["cls_1"][ib][crafted].tolist()) == {2, 7}


def test_random_init_model_heads_tie_regime(oracle):
    """Heads of the random-init eval-mode reference model: scores 0.5 +- a few ulps, order decided by ties."""
    g = util.load("model_tie")
    from ypp import synth
    heads = [g["heads%d" % i] for i in range(3)]
    rows = oracle.decode_multi_scale(heads, int(g["nc"]), synth.YOLOV3_ANCHORS, int(g["input_size"]), 3)
    assert util.rel_err(rows[..., :5], g["rows"][..., :5]).max() <= 1e-5
    assert (rows[..., 4] == 0.5).mean() > 0.1
    for b in range(2):
        keep, _ = oracle.nms_v3_indices(g["rows"][b], float(g["conf"]), 0.45, offset_mode=oracle.OFFSET_AUTO_CPU,
                                        union_mode=oracle.UNION_SEPARATE)
        assert np.array_equal(util.canon(keep, g["rows"][b]), util.canon(g["keep"][b, : int(g["keep_count"][b])], g["rows"][b]))
