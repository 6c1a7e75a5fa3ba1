"""GPU: the sweep-and-prune path of the per-class NMS kernel (segments of 192..512 boxes) and every
way out of it - edge-list overflow, deep dependency chains, boxes that are not ordinary, one x1 for
all - against the oracle's plain greedy NMS, keep indices bit-exact."""
import numpy as np
import pytest
import torch

import util

pytestmark = pytest.mark.gpu


def _rows(cxcywh, score, cls):
    n = len(score)
    return np.concatenate([np.asarray(cxcywh, np.float64).reshape(n, 4), np.asarray(score, np.float64).reshape(n, 1),
                           np.asarray(cls, np.float64).reshape(n, 1)], 1).astype(np.float32)


def _cases():
    rng = np.random.RandomState(42)
    out = {}
    n = 450
    sc = rng.permutation(n) / (2.0 * n) + 0.3                       # distinct scores
    # few edges: small boxes
    out["small_boxes"] = _rows(np.c_[rng.uniform(0, 416, (n, 2)), rng.uniform(5, 30, (n, 2))], sc, np.zeros(n))
    # hundreds of edges, several rounds of the fix-point
    out["mixed_sizes"] = _rows(np.c_[rng.uniform(0, 416, (n, 2)), rng.uniform(8, 90, (n, 2))], sc, np.zeros(n))
    # clusters around a few objects: the edge list overflows
    ctr = rng.uniform(60, 350, (6, 2))[rng.randint(0, 6, n)]
    out["clusters"] = _rows(np.c_[ctr + rng.normal(0, 6, (n, 2)), 60 + rng.normal(0, 5, (n, 2))], sc, np.zeros(n))
    # identical boxes
    out["identical"] = _rows(np.tile([100., 100., 50., 50.], (n, 1)), sc, np.zeros(n))
    # a chain: every box overlaps only its predecessor (IoU 0.6), scores descending along it
    m = 300
    out["chain"] = _rows(np.c_[20 + 10.0 * np.arange(m), np.full(m, 50.), np.full(m, 40.), np.full(m, 40.)],
                         0.9 - 0.001 * np.arange(m), np.zeros(m))
    # same chain, scores ascending along it (the last box wins, alternating back)
    out["chain_rev"] = _rows(np.c_[20 + 10.0 * np.arange(m), np.full(m, 50.), np.full(m, 40.), np.full(m, 40.)],
                             0.5 + 0.001 * np.arange(m), np.zeros(m))
    # one x1 for every box (one bin), spread in y
    w = rng.uniform(10, 40, n)
    out["same_x1"] = _rows(np.c_[100 + w / 2, rng.uniform(0, 2000, n), w, rng.uniform(10, 40, n)], sc, np.zeros(n))
    # boxes that are not ordinary: inverted, zero-area, huge
    r = _rows(np.c_[rng.uniform(0, 416, (n, 2)), rng.uniform(8, 90, (n, 2))], sc, np.zeros(n))
    r[5, 2] = -30; r[17, 3] = -12; r[40, 2] = 0; r[41, 2:4] = 0; r[60, 2:4] = 1e20; r[61, 0] = 3e38
    out["odd_boxes"] = r
    # huge but ordinary coordinates: areas overflow, every pair goes to the exact division
    r = _rows(np.c_[rng.uniform(0, 416, (n, 2)), rng.uniform(8, 90, (n, 2))], sc, np.zeros(n))
    out["scaled_1e18"] = (r * np.array([1e18, 1e18, 1e18, 1e18, 1, 1], np.float32)).astype(np.float32)
    # three classes of 350: per-class segments, class offsets in trick mode
    n3 = 1050
    out["three_classes"] = _rows(np.c_[rng.uniform(0, 416, (n3, 2)), rng.uniform(8, 90, (n3, 2))],
                                 rng.permutation(n3) / (2.0 * n3) + 0.3, rng.randint(0, 3, n3))
    # ties in score and in x1
    r = _rows(np.c_[np.round(rng.uniform(0, 416, (n, 2)) / 8) * 8, np.round(rng.uniform(8, 60, (n, 2)) / 8) * 8],
              np.round(rng.uniform(.3, 1, n) * 16) / 16, np.zeros(n))
    out["quantised"] = r
    return out


@pytest.mark.parametrize("iou", [0.45, 0.1, 0.9])
@pytest.mark.parametrize("om,um", [("trick", "fma"), ("none", "separate"), ("none", "fma")])
def test_sweep_and_prune_path_is_bit_exact(oracle, cuda, om, um, iou):
    from ypp import engine
    for name, rows in _cases().items():
        nc = int(rows[:, 5].max()) + 1
        pp = engine.PostProcessor(nc, [[1, 1]], 1, 0.25, iou_threshold=iou, kind="v2", anchors_per_scale=1,
                                  offset_mode=om, union_mode=um)
        res = pp.nms_rows(torch.tensor(rows).to(cuda), want_keep_idx=True)
        k = res.keep_idx[0, : int(res.keep_count[0])].cpu().numpy()
        ref, _ = oracle.nms_v3_indices(rows, .25, iou, offset_mode=oracle.OFFSET_TRICK if om == "trick" else oracle.OFFSET_NONE,
                                       union_mode=oracle.UNION_FMA if um == "fma" else oracle.UNION_SEPARATE)
        assert np.array_equal(k, util.canon(ref, rows)), (name, om, um, iou, len(k), len(ref))


def test_sweep_and_prune_batch_of_random_segments(oracle, cuda):
    """64 images x 4 classes, segment lengths spread over 150..500, box sizes from sparse to crowded."""
    from ypp import engine
    rng = np.random.RandomState(7)
    B, N, nc = 64, 1600, 4
    rows = np.zeros((B, N, 6), np.float32)
    for b in range(B):
        hi = rng.choice([20, 40, 80, 160])
        rows[b] = _rows(np.c_[rng.uniform(0, 416, (N, 2)), rng.uniform(6, hi, (N, 2))],
                        rng.uniform(0.05, 1, N), rng.randint(0, nc, N))
    pp = engine.PostProcessor(nc, [[1, 1]], 1, 0.25, kind="v2", anchors_per_scale=1)
    res = pp.nms_rows(torch.tensor(rows).to(cuda), want_keep_idx=True)
    kc = res.keep_count.cpu().numpy()
    ki = res.keep_idx.cpu().numpy()
    for b in range(B):
        ref, _ = oracle.nms_v3_indices(rows[b], .25, .45, offset_mode=oracle.OFFSET_AUTO_CUDA, union_mode=oracle.UNION_FMA)
        assert np.array_equal(ki[b, : kc[b]], util.canon(ref, rows[b])), b


def test_sweep_and_prune_against_torchvision_on_the_same_device(cuda):
    """Not the oracle: torchvision.ops.batched_nms on CUDA (what nms_v3 calls, utils/yolo_utils.py:346) on the
    same hand-made inputs; defaults (auto_cuda + fma) are the arithmetic of that kernel."""
    torchvision = pytest.importorskip("torchvision")
    from ypp import engine
    for name, rows in _cases().items():
        nc = int(rows[:, 5].max()) + 1
        t = torch.tensor(rows).to(cuda)
        pp = engine.PostProcessor(nc, [[1, 1]], 1, 0.25, kind="v2", anchors_per_scale=1)
        res = pp.nms_rows(t, want_keep_idx=True)
        k = res.keep_idx[0, : int(res.keep_count[0])].long()
        cand = torch.where(t[:, 4] > 0.25)[0]
        c = t[cand]
        xyxy = torch.stack([c[:, 0] - c[:, 2] / 2, c[:, 1] - c[:, 3] / 2, c[:, 0] + c[:, 2] / 2, c[:, 1] + c[:, 3] / 2], 1)
        keep = cand[torchvision.ops.batched_nms(xyxy, c[:, 4], c[:, 5], 0.45)]
        assert torch.equal(k, keep), (name, len(k), len(keep))
