"""Images with at most 256 candidates are finished inside S1 (rank by counting, every pair through the exact test, greedy
order as a fix-point: organise.cu, organise_small); larger ones go through the sort / segments / N1 machinery.  Both
state the same thing - torchvision.ops.batched_nms as nms_v3 calls it (utils/yolo_utils.py:346) - so they must agree bit
for bit.  The library reads YPP_SMALL once, so the general path runs in a child process with YPP_SMALL=0 on the same
inputs and the parent (default: small path on) compares everything G emits.  The general path is the one the oracle
and golden tests pin at sizes above 256; here it is the checker for the small one, and a few cases go to the CPU oracle
directly as well."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COUNTS = [0, 1, 2, 31, 32, 33, 64, 100, 200, 255, 256, 257, 300]


def _rows(rng, n_pass, n_total, nc, kind):
    """[n_total, 6] decoded rows cx,cy,w,h,score,cls of which exactly n_pass have score > 0.25."""
    r = np.zeros((n_total, 6), dtype=np.float32)
    if kind == "clustered":                 # a dozen centres, heavy overlap
        ctr = rng.uniform(40, 380, size=(12, 2))
        pick = rng.integers(0, 12, size=n_total)
        r[:, 0:2] = ctr[pick] + rng.normal(0, 12, size=(n_total, 2))
        r[:, 2:4] = rng.uniform(20, 160, size=(n_total, 2))
    elif kind == "corner":                  # boxes hanging over the top-left corner: negative x1, y1 under the class offsets
        r[:, 0:2] = rng.uniform(-30, 60, size=(n_total, 2))
        r[:, 2:4] = rng.uniform(10, 120, size=(n_total, 2))
        far = rng.integers(0, n_total, size=max(n_total // 8, 1))
        r[far, 0:2] = rng.uniform(300, 420, size=(len(far), 2))
    elif kind == "ties":                    # few distinct boxes and scores
        base = rng.uniform(50, 350, size=(6, 4)).astype(np.float32)
        base[:, 2:4] = rng.uniform(30, 90, size=(6, 2))
        r[:, 0:4] = base[rng.integers(0, 6, size=n_total)]
    else:                                   # "odd": zero-area, inverted, non-finite boxes among ordinary ones
        r[:, 0:2] = rng.uniform(0, 416, size=(n_total, 2))
        r[:, 2:4] = rng.uniform(5, 200, size=(n_total, 2))
        k = max(n_total // 10, 1)
        r[rng.integers(0, n_total, size=k), 2] = 0.0
        r[rng.integers(0, n_total, size=k), 3] = -rng.uniform(1, 50, size=k)
        r[rng.integers(0, n_total, size=max(k // 2, 1)), 0] = np.nan
        r[rng.integers(0, n_total, size=max(k // 2, 1)), 1] = np.inf
    r[:, 5] = rng.integers(0, nc, size=n_total)
    score = rng.uniform(0.26, 0.99, size=n_total).astype(np.float32)
    if kind == "ties":
        score = rng.choice(np.array([0.3, 0.5, 0.5, 0.9], dtype=np.float32), size=n_total)
    lose = rng.permutation(n_total)[: n_total - n_pass]
    score[lose] = rng.uniform(0.0, 0.25, size=len(lose)).astype(np.float32)
    r[:, 4] = score
    return r


def _batches():
    rng = np.random.default_rng(20260511)
    out = []
    for kind in ("clustered", "corner", "ties", "odd"):
        for nc in (1, 20):
            n_total = 384
            out.append((kind, nc, np.stack([_rows(rng, c, n_total, nc, kind) for c in COUNTS])))
    return out


def _compute(path):
    """Everything G emits for every batch / mode, written to an .npz (runs in the parent and in the child)."""
    import torch
    sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
    from ypp import engine, synth
    dev = torch.device("cuda:0")
    res = {}
    for kind, nc, rows in _batches():
        t = torch.from_numpy(rows).to(dev)
        for om in ("trick", "none", "auto_cuda", "auto_cpu"):
            for um in ("fma", "separate"):
                for thr in (0.45, 0.0):
                    pp = engine.PostProcessor(nc, synth.YOLOV3_ANCHORS, 416, 0.25, iou_threshold=thr, offset_mode=om, union_mode=um)
                    r = pp.nms_rows(t, want_keep_idx=True)
                    torch.cuda.synchronize()
                    tag = "%s_%d_%s_%s_%g" % (kind, nc, om, um, thr)
                    kc = r.keep_count.cpu().numpy()
                    ki = r.keep_idx.cpu().numpy()
                    res[tag + "_cand"] = r.cand_count.cpu().numpy()
                    res[tag + "_kc"] = kc
                    res[tag + "_ki"] = np.concatenate([ki[b, : kc[b]] for b in range(len(kc))]) if kc.sum() else np.zeros(0, np.int32)
                    res[tag + "_rows"] = r.rows().cpu().numpy()
    # the Ultralytics entry: obj * cls scores, multi_label rows, agnostic, max_nms / max_det cuts
    rng = np.random.default_rng(7)
    pred = np.zeros((6, 300, 5 + 4), dtype=np.float32)
    pred[..., 0:2] = rng.uniform(20, 400, size=pred.shape[:2] + (2,))
    pred[..., 2:4] = rng.uniform(20, 150, size=pred.shape[:2] + (2,))
    pred[..., 4] = rng.uniform(0, 1, size=pred.shape[:2])
    pred[..., 5:] = rng.uniform(0, 1, size=pred.shape[:2] + (4,))
    for b, keep in enumerate((0, 3, 40, 120, 250, 300)):
        pred[b, keep:, 4] = 0.0
    tp = torch.from_numpy(pred).to(dev)
    for kw in (dict(), dict(agnostic=True), dict(multi_label=True), dict(max_nms=50), dict(max_det=7), dict(classes=[1, 3])):
        r = engine.nms_ultralytics(tp, conf_thres=0.25, iou_thres=0.45, want_keep_idx=True, **kw)
        torch.cuda.synchronize()
        tag = "ultra_" + "_".join("%s%s" % (k, v) for k, v in kw.items())
        kc = r.keep_count.cpu().numpy()
        ki = r.keep_idx.cpu().numpy()
        res[tag + "_kc"] = kc
        res[tag + "_ki"] = np.concatenate([ki[b, : kc[b]] for b in range(len(kc))]) if kc.sum() else np.zeros(0, np.int32)
        res[tag + "_rows"] = r.rows().cpu().numpy()
    np.savez(path, **res)


def test_small_path_equals_general_path(cuda):
    with tempfile.TemporaryDirectory() as d:
        mine, ref = os.path.join(d, "small.npz"), os.path.join(d, "general.npz")
        assert os.environ.get("YPP_SMALL") in (None, "256"), "run the suite with the default YPP_SMALL"
        _compute(mine)
        env = dict(os.environ)
        env["YPP_SMALL"] = "0"
        out = subprocess.run([sys.executable, os.path.abspath(__file__), ref], env=env, capture_output=True, text=True, timeout=900)
        assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
        a, b = np.load(mine), np.load(ref)
        assert sorted(a.files) == sorted(b.files) and len(a.files) > 300
        n_small = 0
        for k in a.files:
            assert a[k].shape == b[k].shape, k
            assert np.array_equal(a[k], b[k], equal_nan=True), k
            if k.endswith("_cand"):
                n_small += int((a[k] <= 256).sum())
        assert n_small > 1000                    # the cases really are on the small side of the threshold


def test_small_path_against_the_oracle(cuda):
    """A slice of the same inputs against the CPU restatement directly (indices in NMS order, bit-exact)."""
    import torch
    from oracle import oracle as O
    from ypp import engine, synth, _lib
    dev = torch.device("cuda:0")
    for kind, nc, rows in _batches():
        if kind == "odd":
            continue                            # non-finite boxes: covered by the equality test and the golden edge cases
        pp = engine.PostProcessor(nc, synth.YOLOV3_ANCHORS, 416, 0.25, offset_mode="auto_cuda", union_mode="fma")
        r = pp.nms_rows(torch.from_numpy(rows).to(dev), want_keep_idx=True)
        kc, ki = r.keep_count.cpu().numpy(), r.keep_idx.cpu().numpy()
        for b in range(rows.shape[0]):
            keep, _ = O.nms_v3_indices(rows[b], 0.25, 0.45, offset_mode=_lib.OFFSET_AUTO_CUDA, union_mode=_lib.UNION_FMA)
            k = np.asarray(keep, dtype=np.int64)
            k = k[np.lexsort((k, -rows[b][k, 4].astype(np.float64)))]
            assert kc[b] == len(k) and np.array_equal(ki[b, : kc[b]], k), (kind, nc, b)


_ADAPTIVE = r"""
import sys
sys.path[:0] = [%r, %r]
import numpy as np, torch
from oracle import oracle as O
from ypp import engine, synth, _lib
dev = torch.device("cuda:0")
rng = np.random.default_rng(5)
B, N, nc = 3, 4000, 20                       # N / nc = 200: a many-class configuration, the long-segment kernel comes last
rows = np.zeros((B, N, 6), dtype=np.float32)
rows[..., 0:2] = rng.uniform(30, 380, size=(B, N, 2))
rows[..., 2:4] = rng.uniform(20, 120, size=(B, N, 2))
rows[..., 4] = rng.uniform(0.0, 1.0, size=(B, N))
rows[..., 5] = rng.integers(0, nc, size=(B, N))
rows[:, :1800, 5] = 3                        # ~1350 candidates of one class per image: segments > 512
pp = engine.PostProcessor(nc, synth.YOLOV3_ANCHORS, 416, 0.25)
t = torch.from_numpy(rows).to(dev)
outs = []
for call in range(3):                        # fresh process: the first call launches the kernel narrow (if the hint is on)
    r = pp.nms_rows(t, want_keep_idx=True)
    torch.cuda.synchronize()
    outs.append((r.keep_count.cpu().numpy().copy(), r.keep_idx.cpu().numpy().copy()))
for kc, ki in outs[1:]:
    assert np.array_equal(kc, outs[0][0])
    for b in range(B):
        assert np.array_equal(ki[b, : kc[b]], outs[0][1][b, : kc[b]])
kc, ki = outs[0]
for b in range(B):
    keep, _ = O.nms_v3_indices(rows[b], 0.25, 0.45, offset_mode=_lib.OFFSET_AUTO_CUDA, union_mode=_lib.UNION_FMA)
    k = np.asarray(keep, dtype=np.int64)
    k = k[np.lexsort((k, -rows[b][k, 4].astype(np.float64)))]
    assert kc[b] == len(k) and np.array_equal(ki[b, : kc[b]], k), b
print("adaptive ok", int(kc.sum()))
"""


@pytest.mark.parametrize("hint", ["1", "0"])
def test_launch_width_hint_does_not_change_results(cuda, hint):
    """ypp_adaptive_launch (include/ypp.h): until a device has seen a long segment the long-segment kernel of a
    many-class configuration is launched 8 CTAs wide.  Fresh process, long segments in the very first batch: the narrow
    launch, the full-width launches after it and the library without the hint all give the oracle's keep sets."""
    env = dict(os.environ)
    env["YPP_ADAPTIVE"] = hint
    out = subprocess.run([sys.executable, "-c", _ADAPTIVE % (ROOT, os.path.join(ROOT, "pytorch-object-detection_b200"))],
                         env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert "adaptive ok" in out.stdout


if __name__ == "__main__":
    _compute(sys.argv[1])
