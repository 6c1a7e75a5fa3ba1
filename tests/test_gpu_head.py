"""GPU: the head's 1x1 convolution fused with the decode (tcgen05 TF32 kernel).

Two statements, as DESIGN.md explains: (1) the head values the kernel decodes equal the exact
convolution within the TF32 product bound 2^-9 * sum|x||w| (what cuDNN computes for this layer under
torch's default allow_tf32 obeys the same bound); (2) everything after the accumulator is D1 itself:
the unfused chain fed the kernel's own head tensors returns bit-identical counts, indices and rows."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ANCH9 = [[10, 13], [16, 30], [33, 23], [30, 61], [62, 45], [59, 119], [116, 90], [156, 198], [373, 326]]


def _inputs(B, cins, grids, AC, seed, channels_last=False):
    g = torch.Generator("cpu").manual_seed(seed)
    feats, weights = [], []
    for cin, hw in zip(cins, grids):
        x = torch.randn(B, cin, hw, hw, generator=g).relu_()             # the conv input is a ReLU output
        w = torch.randn(AC, cin, 1, 1, generator=g) * (1.6 / cin ** 0.5)
        feats.append(x)
        weights.append(w)
    return feats, weights


def _run(cuda, oracle, nc, A, anchors, kind, input_size, conf, B, cins, grids, seed, channels_last=False):
    from ypp import engine
    AC = A * (5 + nc)
    feats, weights = _inputs(B, cins, grids, AC, seed)
    pp = engine.PostProcessor(nc, anchors, input_size, conf, kind=kind, anchors_per_scale=A)
    df = [f.to(cuda) for f in feats]
    if channels_last:
        df = [f.contiguous(memory_format=torch.channels_last) for f in df]
    dw = [w.to(cuda) for w in weights]
    res, heads = pp.from_features(df, dw, want_keep_idx=True, want_heads=True)
    torch.cuda.synchronize()
    for x, w, h in zip(feats, weights, heads):
        exact, bound = oracle.head_conv1x1(x.numpy(), w.numpy())
        err = np.abs(h.cpu().numpy().astype(np.float64) - exact)
        tol = 2.0 ** -9 * bound * 1.01 + 1e-6                            # TF32 products + fp32 accumulation
        assert (err <= tol).all(), float((err / np.maximum(tol, 1e-30)).max())
        assert err.max() > 0                                             # it is TF32, not an fp32 copy of the answer
    ref = pp([h.contiguous() for h in heads], want_keep_idx=True)
    torch.cuda.synchronize()
    assert torch.equal(res.cand_count, ref.cand_count)
    assert torch.equal(res.keep_count, ref.keep_count)
    assert int(res.det_total) == int(ref.det_total) and int(res.det_total) > 0
    n = int(res.det_total)
    assert torch.equal(res.det_rows[:n], ref.det_rows[:n])
    kc = res.keep_count.cpu().numpy()
    for b in range(B):
        assert torch.equal(res.keep_idx[b, : kc[b]], ref.keep_idx[b, : kc[b]])
    return res


def test_fused_head_yolov3_three_scales(oracle, cuda):
    _run(cuda, oracle, 20, 3, ANCH9, "v3", 416, 0.25, 3, [64, 96, 128], [52, 26, 13], seed=1)


def test_fused_head_yolov3_real_channel_counts_channels_last(oracle, cuda):
    _run(cuda, oracle, 20, 3, ANCH9, "v3", 416, 0.25, 2, [256, 512, 1024], [52, 26, 13], seed=2, channels_last=True)


def test_fused_head_coco80_wide_tile_and_odd_planes(oracle, cuda):
    _run(cuda, oracle, 80, 3, ANCH9, "v3", 608, 0.05, 1, [72, 40, 36], [76, 38, 19], seed=3)


def test_fused_head_yolov2_five_anchors(oracle, cuda):
    from ypp import synth
    _run(cuda, oracle, 20, 5, synth.YOLOV2_SCALED_ANCHORS, "v2", 416, 0.25, 8, [1024], [13], seed=4)


def test_fused_head_single_class(oracle, cuda):
    _run(cuda, oracle, 1, 3, ANCH9, "v3", 416, 0.25, 4, [128, 256, 512], [52, 26, 13], seed=5)


def test_fused_head_argument_errors(cuda):
    from ypp import engine, _lib
    pp = engine.PostProcessor(20, ANCH9, 416, 0.25)
    x = torch.randn(1, 30, 13, 13, device=cuda)                          # Cin not a multiple of 4
    w = torch.randn(75, 30, 1, 1, device=cuda)
    with pytest.raises(_lib.YppError):
        pp.from_features([x], [w])
    with pytest.raises(ValueError):
        pp.from_features([torch.randn(1, 32, 13, 13, device=cuda)], [torch.randn(70, 32, 1, 1, device=cuda)])


def test_compat_entry_points_for_features(oracle, cuda):
    """DecodeYoloV3.forward_features and MeanAveragePrecision.update_state_features equal forward / update_state
    fed the head tensors the fused kernel itself produced."""
    from ypp import engine
    from ypp.compat.yolov3_utils import DecodeYoloV3, MeanAveragePrecision
    from ypp import synth
    feats, weights = _inputs(1, [64, 96, 128], [52, 26, 13], 75, seed=9)
    df, dw = [f.to(cuda) for f in feats], [w.to(cuda) for w in weights]
    dec = DecodeYoloV3(20, ANCH9, 416, conf_threshold=0.25)
    pp = engine.PostProcessor(20, ANCH9, 416, 0.25)
    _, heads = pp.from_features(df, dw, want_heads=True)
    assert torch.equal(dec.forward_features(df, dw), dec(heads))
    feats, weights = _inputs(4, [64, 96, 128], [52, 26, 13], 75, seed=10)
    df, dw = [f.to(cuda) for f in feats], [w.to(cuda) for w in weights]
    _, heads = pp.from_features(df, dw, want_heads=True)
    annot = synth.make_annots(4, 20, g=2).to(cuda)
    m1, m2 = MeanAveragePrecision(20, ANCH9, 416, 0.25), MeanAveragePrecision(20, ANCH9, 416, 0.25)
    m1.update_state_features(annot, df, dw)
    m2.update_state(annot, heads)
    assert torch.equal(m1.all_pred_boxes_variable, m2.all_pred_boxes_variable)
    assert torch.equal(m1.all_true_boxes_variable, m2.all_true_boxes_variable)
    assert float(m1.result()) == float(m2.result())
