"""GPU: the CUDA path (through the C ABI) against the CPU oracle and the golden fixtures.

Bars: decoded coordinates / scores within 1e-5 relative (fp32; tolerance written below);
candidate sets, keep indices, counts and output rows bit-exact (stage-wise: the oracle NMS is
fed the rows OUR decode produced, so libm differences cannot leak into an index comparison).
"""
import numpy as np
import pytest
import torch

import util

pytestmark = pytest.mark.gpu

DECODE_RTOL = 1e-5


def _pp(g, dev, **kw):
    from ypp import engine
    kind = str(g["kind"])
    return engine.PostProcessor(int(g["nc"]), util.case_anchors(g), int(g["input_size"]),
                                conf_threshold=float(g["conf"]), kind=kind,
                                anchors_per_scale=int(g["A"]), **kw)


def _oracle_rows(oracle, g, heads):
    nc, A, S = int(g["nc"]), int(g["A"]), int(g["input_size"])
    hs = [h.numpy() for h in heads]
    if str(g["kind"]) == "v2":
        return oracle.decode_predictions_v2(hs[0], nc, util.case_anchors(g), S)
    return oracle.decode_multi_scale(hs, nc, util.case_anchors(g), S, A)


def _check_decode(ours, ref, heads_cls_logits=None):
    assert ours.shape == ref.shape
    assert util.rel_err(ours[..., :5], ref[..., :5]).max() <= DECODE_RTOL
    mism = int((ours[..., 5] != ref[..., 5]).sum())
    assert mism <= max(2, ours.shape[0] * ours.shape[1] // 100000), "class arg-max mismatches: %d" % mism


@pytest.mark.parametrize("name", util.HEAD_CASES)
def test_dense_decode_matches_oracle_and_reference(oracle, cuda, name):
    g = util.load(name)
    heads = util.case_heads(g)
    pp = _pp(g, cuda)
    ours = pp.decode([h.to(cuda) for h in heads]).cpu().numpy()
    _check_decode(ours, _oracle_rows(oracle, g, heads))
    stride = int(g["rows_stride"])
    _check_decode(ours[:, ::stride], g["rows"])           # rows the unmodified reference produced


MODES = [("auto_cpu", "separate"), ("auto_cuda", "fma"), ("none", "separate"), ("trick", "fma"), ("trick", "separate")]


@pytest.mark.parametrize("name", util.HEAD_CASES)
@pytest.mark.parametrize("offset_mode,union_mode", MODES)
def test_chain_keep_sets_bit_exact(oracle, cuda, name, offset_mode, union_mode):
    from ypp import _lib
    g = util.load(name)
    heads = util.case_heads(g)
    dev_heads = [h.to(cuda) for h in heads]
    pp = _pp(g, cuda, offset_mode=offset_mode, union_mode=union_mode)
    rows = pp.decode(dev_heads).cpu().numpy()              # OUR decoded rows feed the oracle NMS
    res = pp(dev_heads, img_idx_base=5, row_width=7, want_keep_idx=True)
    torch.cuda.synchronize()
    kc = res.keep_count.cpu().numpy()
    cc = res.cand_count.cpu().numpy()
    kidx = res.keep_idx.cpu().numpy()
    det = res.rows().cpu().numpy()
    assert det.shape[0] == kc.sum()
    conf, B = float(g["conf"]), int(g["batch"])
    off = 0
    for b in range(B):
        keep, ncand = oracle.nms_v3_indices(rows[b], conf, 0.45, offset_mode=_lib.OFFSET_MODES[offset_mode],
                                            union_mode=_lib.UNION_MODES[union_mode])
        keep = util.canon(keep, rows[b])
        assert cc[b] == ncand == int((rows[b, :, 4] > np.float32(conf)).sum())
        assert kc[b] == len(keep), (name, b, kc[b], len(keep))
        assert np.array_equal(kidx[b, : kc[b]], keep)                       # bit-exact indices, NMS order
        d = det[off: off + kc[b]]
        assert np.all(d[:, 0] == 5 + b)
        assert np.array_equal(d[:, 1:], rows[b][keep])                      # rows = decoded rows, bit for bit
        off += kc[b]


@pytest.mark.parametrize("name", util.HEAD_CASES)
def test_chain_against_reference_keep_sets(cuda, name):
    """End to end against what the unmodified reference kept (CPU torch + torchvision): equal up to
    boxes whose score/IoU sits within an ulp of a decision under a different libm."""
    g = util.load(name)
    heads = util.case_heads(g)
    pp = _pp(g, cuda, offset_mode="auto_cpu", union_mode="separate")
    res = pp([h.to(cuda) for h in heads], want_keep_idx=True)
    kc = res.keep_count.cpu().numpy()
    kidx = res.keep_idx.cpu().numpy()
    cc = res.cand_count.cpu().numpy()
    tot, diff = 0, 0
    for b in range(int(g["batch"])):
        ref = set(g["keep"][b, : int(g["keep_count"][b])].tolist())
        got = set(kidx[b, : kc[b]].tolist())
        diff += len(ref ^ got)
        tot += len(ref)
        assert abs(int(cc[b]) - int(g["cand_count"][b])) <= 1
    assert diff <= max(2, tot // 2000), "keep sets differ from the reference in %d of %d boxes" % (diff, tot)


def test_staged_entry_points_equal_fused_call(cuda):
    from ypp import engine, synth
    heads = [h.to(cuda) for h in synth.make_heads(4, 20, [52, 26, 13], g=9)]
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    a = pp(heads, want_keep_idx=True)
    b = pp.staged(heads, want_keep_idx=True)
    assert torch.equal(a.keep_count, b.keep_count) and torch.equal(a.cand_count, b.cand_count)
    assert torch.equal(a.rows(), b.rows())
    n = int(a.keep_count.max())
    assert torch.equal(a.keep_idx[:, :n] * (torch.arange(n, device=cuda)[None] < a.keep_count[:, None]),
                       b.keep_idx[:, :n] * (torch.arange(n, device=cuda)[None] < b.keep_count[:, None]))


def test_decode_edges_match_reference(cuda):
    from ypp import engine
    g = util.load("decode_edges")
    pp = engine.PostProcessor(int(g["nc"]), g["anchors"].tolist(), int(g["input_size"]), kind="v3",
                              anchors_per_scale=int(g["A"]))
    ours = pp.decode([torch.tensor(g["head"]).to(cuda)]).cpu().numpy()
    ref = g["rows"]
    assert np.array_equal(np.isinf(ours), np.isinf(ref))
    assert util.rel_err(ours[..., :5], ref[..., :5]).max() <= DECODE_RTOL
    assert np.array_equal(ours[..., 5], ref[..., 5])      # saturated / collided class logits -> lowest index


def test_all_scores_tied(oracle, cuda):
    """Objectness logit exactly 0 everywhere -> every score is 0.5: ordering is 100% tie-break."""
    from ypp import engine, synth
    heads = synth.make_heads(2, 20, [26, 13], g=11)
    for h in heads:
        h.view(2, 3, 25, h.shape[2], h.shape[3])[:, :, 4] = 0.0
    anchors = synth.YOLOV3_ANCHORS[-6:]
    for om, um in (("auto_cuda", "fma"), ("none", "separate")):
        pp = engine.PostProcessor(20, anchors, 416, 0.25, offset_mode=om, union_mode=um)
        dev = [h.to(cuda) for h in heads]
        rows = pp.decode(dev).cpu().numpy()
        assert np.all(rows[..., 4] == 0.5)
        res = pp(dev, want_keep_idx=True)
        kc, kidx = res.keep_count.cpu().numpy(), res.keep_idx.cpu().numpy()
        from ypp import _lib
        for b in range(2):
            keep, _ = oracle.nms_v3_indices(rows[b], 0.25, 0.45, offset_mode=_lib.OFFSET_MODES[om],
                                            union_mode=_lib.UNION_MODES[um])
            assert np.array_equal(kidx[b, : kc[b]], util.canon(keep, rows[b]))
            assert np.all(np.diff(kidx[b, : kc[b]]) > 0)          # ties resolved by lower box index


def test_scores_tied_in_crowds(oracle, cuda):
    """A handful of distinct scores shared by thousands of boxes: the bucket+rank sort hands over to the
    stable LSD passes; a few hundred distinct scores stay on the bucket path with fat buckets."""
    from ypp import engine, _lib
    rng = np.random.RandomState(3)
    for levels in (7, 400):
        n = 6000
        rows = np.concatenate([rng.uniform(0, 416, (n, 2)), rng.uniform(8, 90, (n, 2)),
                               0.3 + 0.6 * rng.randint(0, levels, (n, 1)) / levels, rng.randint(0, 20, (n, 1))], 1).astype(np.float32)
        pp = engine.PostProcessor(20, [[1, 1]], 1, 0.25, kind="v2", anchors_per_scale=1, offset_mode="none", union_mode="separate")
        res = pp.nms_rows(torch.tensor(rows).to(cuda), want_keep_idx=True)
        k = res.keep_idx[0, : int(res.keep_count[0])].cpu().numpy()
        ref, _ = oracle.nms_v3_indices(rows, .25, .45, offset_mode=_lib.OFFSET_MODES["none"], union_mode=_lib.UNION_MODES["separate"])
        assert np.array_equal(k, util.canon(ref, rows)), levels


def test_score_distributions_of_both_bucket_functions(oracle, cuda):
    """The sort's two bucket functions and the hand-over between them: scores spread evenly over their range
    (value-linear buckets), spread over many binades (the linear buckets overflow, score-word buckets take over),
    one huge outlier (a linear range that is all outlier), an infinite score (no finite range at all)."""
    from ypp import engine, _lib
    rng = np.random.RandomState(17)
    n = 9000
    base = np.concatenate([rng.uniform(0, 416, (n, 2)), rng.uniform(8, 60, (n, 2)), np.zeros((n, 1)),
                           rng.randint(0, 20, (n, 1))], 1).astype(np.float32)
    cases = {"even": rng.uniform(0.001, 1.0, n), "binades": np.exp(-rng.uniform(0, 12, n)),
             "outlier": rng.uniform(0.001, 1.0, n), "infinite": rng.uniform(0.001, 1.0, n)}
    cases["outlier"][123] = 1e30
    cases["infinite"][77] = np.inf
    for name, sc in cases.items():
        rows = base.copy()
        rows[:, 4] = sc.astype(np.float32)
        pp = engine.PostProcessor(20, [[1, 1]], 1, 0.0005, kind="v2", anchors_per_scale=1, offset_mode="none", union_mode="fma")
        res = pp.nms_rows(torch.tensor(rows).to(cuda), want_keep_idx=True)
        k = res.keep_idx[0, : int(res.keep_count[0])].cpu().numpy()
        ref, _ = oracle.nms_v3_indices(rows, .0005, .45, offset_mode=_lib.OFFSET_MODES["none"], union_mode=_lib.UNION_MODES["fma"])
        assert np.array_equal(k, util.canon(ref, rows)), name


def test_empty_and_single_image(cuda):
    from ypp import engine, synth
    heads = [h.to(cuda) for h in synth.make_heads(1, 20, [52, 26, 13], g=12, mu_obj=-60.0)]
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    res = pp(heads, row_width=6)
    assert int(res.det_total.item()) == 0 and res.rows().shape == (0, 6)
    assert int(res.cand_count.item()) == 0


def test_one_process_two_devices(cuda):
    """One process driving two GPUs: function attributes and occupancy are per device (launcher caches too)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from ypp import engine, synth
    heads = synth.make_heads(4, 20, [52, 26, 13], g=21)
    out = []
    for d in (1, 0, 1):
        dev = torch.device("cuda", d)
        pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
        res = pp([h.to(dev) for h in heads], want_keep_idx=True)
        torch.cuda.synchronize(dev)
        out.append((res.keep_count.cpu(), res.rows().cpu()))
    for kc, rows in out[1:]:
        assert torch.equal(kc, out[0][0]) and torch.equal(rows, out[0][1])
