"""GPU: utils/metrics.py's AP side and bbox_iou, the objectness-times-class score mode, and the heads of the
random-init eval-mode reference model (scores tied at 0.5)."""
import numpy as np
import pytest
import torch

import util

pytestmark = pytest.mark.gpu


def test_bbox_iou_metrics_flavour(cuda):
    """ypp.compat.metrics.bbox_iou (utils/metrics.py:192-230) against the reference's stored outputs: IoU, GIoU and
    DIoU bit-exact (pure IEEE add/sub/mul/div/min/max), CIoU within 1e-6 absolute (atan)."""
    from ypp.compat.metrics import bbox_iou
    g = util.load("metrics_ap")
    b1, b2 = g["bi_box1"], g["bi_box2"]
    for xywh in (True, False):
        A, B = b1.copy(), b2.copy()
        if not xywh:
            A[:, 2:] += A[:, :2]
            B[:, 2:] += B[:, :2]
        for mode, kw in (("iou", {}), ("giou", {"GIoU": True}), ("diou", {"DIoU": True}), ("ciou", {"CIoU": True})):
            for bc, a_in in (("nn", A), ("1n", A[:1])):
                want = g["bi_%s_%s_%s" % ("xywh" if xywh else "xyxy", mode, bc)]
                got = bbox_iou(torch.tensor(a_in).to(cuda), torch.tensor(B).to(cuda), xywh=xywh, **kw).cpu().numpy()
                assert got.shape == want.shape
                if mode == "ciou":
                    assert np.allclose(got, want, rtol=0, atol=1e-6, equal_nan=True)
                else:
                    assert np.array_equal(got, want, equal_nan=True), (xywh, mode, bc)
    assert bbox_iou(torch.zeros((0, 4), device=cuda), torch.zeros((0, 4), device=cuda)).shape == (0, 1)
    with pytest.raises(ValueError):
        bbox_iou(torch.zeros(3, 4), torch.zeros(3, 4))                       # no CPU path


def test_ap_per_class_fed_by_gpu_tp_flags(oracle, cuda):
    """ap_per_class_from_rows: TP flags from ypp_map_match (GPU) -> ap_per_class (host float64) equals the oracle's
    flags -> the oracle's ap_per_class; and device tensors go through ap_per_class unchanged."""
    from ypp.compat import metrics as M
    m = util.load("map_cases")
    gt, det, nc = m["syn_gt"], m["syn_det"], int(m["syn_nc"])
    got = M.ap_per_class_from_rows(torch.tensor(gt).to(cuda), torch.tensor(det).to(cuda), nc)
    _, otp, _ = oracle.mean_average_precision(gt, det, nc, 0.5, True)
    want = oracle.ap_per_class((otp == 1).astype(np.float64)[:, None], det[:, 5].astype(np.float64), det[:, 6], gt[:, 6])
    for a, b in zip(got, want):
        assert np.array_equal(np.asarray(a), np.asarray(b))
    assert got[2].shape == (len(np.unique(gt[:, 6])), 1) and float(got[2].mean()) > 0.05
    g = util.load("metrics_ap")
    args = [g["t10_%s" % k] for k in ("tp", "conf", "pred_cls", "target_cls")]
    dev_args = [torch.tensor(a).to(cuda) for a in args]
    for a, b in zip(M.ap_per_class(*dev_args), [g["t10_%s" % k] for k in ("p", "r", "ap", "f1", "classes")]):
        assert np.array_equal(a, b)


def test_obj_times_cls_bitwise_against_torch_ops_on_device(cuda):
    """score_mode="obj_times_cls": candidates, scores and classes of D1 equal, bit for bit, the scoring lines of the
    vendored NMS (utils/general.py:645,669,679-680; tests/util.ultra_best_class is pinned to the reference function by
    make_golden.py:case_objcls) evaluated with torch ops on the same device on the same decoded values - including
    the cells whose products collide after rounding, where the winner is NOT argmax(sigmoid(cls))."""
    from ypp import engine, synth
    g, heads = util.objcls_heads()
    nc, conf = int(g["nc"]), float(g["conf"])
    dh = [h.to(cuda) for h in heads]
    pred = util.torch_prediction(dh, nc, synth.YOLOV3_ANCHORS, 416)
    pp = engine.PostProcessor(nc, synth.YOLOV3_ANCHORS, 416, conf, iou_threshold=1.0, score_mode="obj_times_cls",
                              offset_mode="none")
    res = pp(dh, want_keep_idx=True, row_width=6)           # IoU > 1 never holds: every candidate is kept
    dense = pp.decode(dh)                                    # dense decode in the same mode: score and class columns
    torch.cuda.synchronize()
    rows = res.rows()
    off, collisions = 0, 0
    for b in range(2):
        ti, ts, tc = util.ultra_best_class(pred[b], conf)
        k = int(res.keep_count[b])
        assert int(res.cand_count[b]) == k == ti.numel()
        kidx = res.keep_idx[b, :k].long()
        order = torch.argsort(kidx)
        assert torch.equal(kidx[order], ti)
        r = rows[off: off + k][order]
        assert torch.equal(r[:, 4], ts), "scores differ bitwise in %d boxes" % int((r[:, 4] != ts).sum())
        assert torch.equal(r[:, 5].long(), tc), "classes differ in %d boxes" % int((r[:, 5].long() != tc).sum())
        assert torch.equal(r[:, :4], pred[b][ti][:, :4])
        assert torch.equal(dense[b][ti][:, 4], ts) and torch.equal(dense[b][ti][:, 5].long(), tc)
        collisions += int((pred[b][ti][:, 5:].argmax(1) != tc).sum())
        # rows come out in (score desc, box index asc) order
        s, i = rows[off: off + k][:, 4], kidx
        assert bool(((s[:-1] > s[1:]) | ((s[:-1] == s[1:]) & (i[:-1] < i[1:]))).all())
        off += k
    assert collisions >= 3, "the fixture's colliding products did not collide on this device"


def test_obj_times_cls_against_reference_fixture_and_oracle_nms(oracle, cuda):
    """The same mode end to end: (1) against the candidate list the reference function itself produced on the CPU
    (scores 1e-5 relative - another libm -, classes equal outside the crafted last-bit cells); (2) keep sets at
    iou .45 bit-exact against the oracle's per-class NMS fed OUR scores and classes."""
    from ypp import engine, synth, _lib
    g, heads = util.objcls_heads()
    nc, conf = int(g["nc"]), float(g["conf"])
    dh = [h.to(cuda) for h in heads]
    pp1 = engine.PostProcessor(nc, synth.YOLOV3_ANCHORS, 416, conf, iou_threshold=1.0, score_mode="obj_times_cls", offset_mode="none")
    res = pp1(dh, want_keep_idx=True, row_width=6)
    pp = engine.PostProcessor(nc, synth.YOLOV3_ANCHORS, 416, conf, score_mode="obj_times_cls", offset_mode="none", union_mode="separate")
    nms = pp(dh, want_keep_idx=True, row_width=6)
    dense = pp.decode(dh).cpu().numpy()
    torch.cuda.synchronize()
    rows = res.rows().cpu().numpy()
    off = 0
    for b in range(2):
        k = int(res.keep_count[b])
        kidx = res.keep_idx[b, :k].cpu().numpy().astype(np.int64)
        r = rows[off: off + k]
        off += k
        ref_idx, ref_sc, ref_cl = g["idx_%d" % b], g["conf_%d" % b], g["cls_%d" % b]
        assert len(set(kidx.tolist()) ^ set(ref_idx.tolist())) <= 2
        common, ia, ib = np.intersect1d(kidx, ref_idx, return_indices=True)
        assert util.rel_err(r[ia, 4], ref_sc[ib]).max() <= 1e-5                     # tolerance: 1e-5 relative
        crafted = (common >= 8112) & (common < 8112 + 208) if b == 1 else np.zeros(len(common), bool)
        assert int(((r[ia, 5].astype(np.int64) != ref_cl[ib]) & ~crafted).sum()) <= 2
        # (2) the oracle's NMS on our dense rows (score = obj x cls, class = first maximum of the products)
        keep, ncand = oracle.nms_v3_indices(dense[b], conf, 0.45, offset_mode=_lib.OFFSET_NONE, union_mode=_lib.UNION_SEPARATE)
        kk = int(nms.keep_count[b])
        assert ncand == int(nms.cand_count[b]) == k
        assert np.array_equal(nms.keep_idx[b, :kk].cpu().numpy(), util.canon(keep, dense[b]))


@pytest.mark.parametrize("offset_mode,union_mode", [("auto_cpu", "separate"), ("auto_cuda", "fma"), ("none", "separate")])
def test_random_init_model_heads(oracle, cuda, offset_mode, union_mode):
    """SURVEY 8(d) third regime: heads of the random-init eval-mode reference model (every value ~1e-5 .. 1e-9, every
    score 0.5 +- a few ulps, a fifth of them exactly 0.5): decode within 1e-5, keep sets bit-exact against the oracle
    fed our rows, and equal to what the unmodified reference kept."""
    from ypp import engine, synth, _lib
    g = util.load("model_tie")
    heads = [torch.tensor(g["heads%d" % i]).to(cuda) for i in range(3)]
    S, nc, conf = int(g["input_size"]), int(g["nc"]), float(g["conf"])
    pp = engine.PostProcessor(nc, synth.YOLOV3_ANCHORS, S, conf, offset_mode=offset_mode, union_mode=union_mode)
    rows = pp.decode(heads).cpu().numpy()
    assert util.rel_err(rows[..., :5], g["rows"][..., :5]).max() <= 1e-5             # tolerance: 1e-5 relative
    assert (rows[..., 4] == 0.5).mean() > 0.1
    res = pp(heads, want_keep_idx=True)
    torch.cuda.synchronize()
    for b in range(2):
        k = int(res.keep_count[b])
        keep, ncand = oracle.nms_v3_indices(rows[b], conf, 0.45, offset_mode=_lib.OFFSET_MODES[offset_mode],
                                            union_mode=_lib.UNION_MODES[union_mode])
        assert ncand == int(res.cand_count[b]) == rows.shape[1]
        assert np.array_equal(res.keep_idx[b, :k].cpu().numpy(), util.canon(keep, rows[b]))
        if offset_mode == "auto_cpu":
            # end to end against the unmodified reference (CPU torch + torchvision): scores that differ in the last
            # bit between the two libms reorder ties, so compare as sets with a small allowance
            ref = set(g["keep"][b, : int(g["keep_count"][b])].tolist())
            got = set(res.keep_idx[b, :k].cpu().numpy().tolist())
            assert len(ref ^ got) <= max(4, len(ref) // 50), len(ref ^ got)
