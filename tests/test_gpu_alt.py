"""GPU: the alternate semantics of SURVEY 8(f) #4 through the C ABI -- the vendored Ultralytics
non_max_suppression, ConfusionMatrix.process_batch and the YOLO v1 decode -- against the fixtures the
reference itself produced (tests/golden/make_golden.py) and against the oracle on fresh inputs."""
import numpy as np
import pytest
import torch

import util
from test_oracle_golden import ULTRA_VARIANTS, _split

pytestmark = pytest.mark.gpu


def _lists(res_list):
    return [r.cpu().numpy() for r in res_list]


@pytest.mark.parametrize("tag", ["near", "far"])
def test_ultralytics_matches_reference(cuda, tag):
    from ypp.compat.general import non_max_suppression
    g = util.load("ultralytics")
    pred = torch.from_numpy(g[tag + "_pred"]).to(cuda)
    for vi, kw in enumerate(ULTRA_VARIANTS):
        want = _split(g["%s_%d_rows" % (tag, vi)], g["%s_%d_counts" % (tag, vi)])
        got = _lists(non_max_suppression(pred, union_mode="separate", **kw))     # the fixture ran torchvision's CPU nms
        assert [len(x) for x in got] == [len(x) for x in want], (tag, kw)
        for a, b in zip(got, want):
            assert np.array_equal(a, b), (tag, kw)


def test_ultralytics_single_class_and_empty(cuda):
    from ypp.compat.general import non_max_suppression
    g = util.load("ultralytics")
    pred = torch.from_numpy(g["nc1_pred"]).to(cuda)
    got = _lists(non_max_suppression(pred, multi_label=True, union_mode="separate"))
    for a, b in zip(got, _split(g["nc1_rows"], g["nc1_counts"])):
        assert np.array_equal(a, b)
    none = non_max_suppression(pred, conf_thres=1.0)
    assert [tuple(x.shape) for x in none] == [(0, 6), (0, 6)]
    with pytest.raises(NotImplementedError):
        non_max_suppression(pred, labels=[torch.zeros((1, 5))])


@pytest.mark.parametrize("kw", [dict(), dict(multi_label=True, conf_thres=0.05), dict(agnostic=True),
                                dict(max_nms=700, max_det=0, multi_label=True, conf_thres=0.05)])
def test_ultralytics_against_oracle_larger(oracle, cuda, kw):
    """COCO-like shape (80 classes), keep indices bit for bit, default device semantics (fma union)."""
    from ypp import engine, _lib
    rng = np.random.RandomState(5)
    B, n, nc = 2, 3000, 80
    p = np.zeros((B, n, 5 + nc), np.float32)
    p[..., 0:2] = rng.uniform(0, 640, (B, n, 2))
    p[..., 2:4] = rng.uniform(8, 200, (B, n, 2))
    p[..., 4] = rng.uniform(0, 1, (B, n)) ** 2
    p[..., 5:] = rng.uniform(0, 1, (B, n, nc)) ** 6
    okw = dict(kw)
    okw.setdefault("max_det", 300)
    want, want_idx = oracle.non_max_suppression(p, union_mode=_lib.UNION_FMA, return_index=True, **okw)
    res = engine.nms_ultralytics(torch.from_numpy(p).to(cuda), want_keep_idx=True, **okw)
    kc = res.keep_count.cpu().numpy()
    assert kc.tolist() == [len(x) for x in want]
    rows = res.rows().cpu().numpy()
    assert np.array_equal(rows, np.concatenate(want, 0))
    for b in range(B):
        assert np.array_equal(res.keep_idx[b, : kc[b]].cpu().numpy(), want_idx[b])


def test_yolov1_decode_matches_reference(cuda):
    from ypp.compat.yolov1_utils import decode_predictions, DecodeYoloV1
    g = util.load("yolov1")
    for tag, nc, nb in (("voc_b2", 20, 2), ("nc3_b3", 3, 3)):
        x = torch.from_numpy(g[tag + "_in"]).to(cuda)
        got = decode_predictions(x, nc, nb, 448).cpu().numpy()
        want = g[tag + "_out"]
        assert got.shape == want.shape
        assert np.array_equal(got[..., 5], want[..., 5])
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)
    boxes = DecodeYoloV1(20, 2, 448, conf_threshold=0.5)(torch.from_numpy(g["voc_b2_in"][:1]).to(cuda))
    assert boxes.dim() == 2 and boxes.shape[1] == 6 and bool((boxes[:, 4] > 0.5).all())


def test_confusion_matrix_matches_reference(cuda):
    from ypp.compat.metrics import ConfusionMatrix
    g = util.load("confusion")
    nc = int(g["nc"])
    det, lab = torch.from_numpy(g["det"]).to(cuda), torch.from_numpy(g["lab"]).to(cuda)
    do, lo = g["det_off"], g["lab_off"]
    one = ConfusionMatrix(nc, conf=float(g["conf"]), iou_thres=float(g["iou_thres"]))
    for i in range(len(do) - 1):                       # the reference's calling pattern: image by image
        one.process_batch(det[do[i]: do[i + 1]], lab[lo[i]: lo[i + 1]])
    assert np.array_equal(one.matrix, g["matrix"])
    allc = ConfusionMatrix(nc, conf=float(g["conf"]), iou_thres=float(g["iou_thres"]))
    allc.process_images(det, torch.from_numpy(do), lab, torch.from_numpy(lo))       # the whole set in one launch
    assert np.array_equal(allc.matrix, g["matrix"])
