"""GPU: the BASELINE.json configurations at their FULL sizes, through size-independent properties.

The oracle finishes a full-size image in tens of milliseconds but not a full batch of hundreds, so
at full size the chain is checked by properties that hold for any input: counts add up, output
order is (score desc, box index asc), kept boxes are candidates, NMS is idempotent on its own
output, a batch equals its images processed one by one, and a sample of images is compared
bit for bit with the oracle and with torchvision on the same device.
"""
import numpy as np
import pytest
import torch

import util

pytestmark = pytest.mark.gpu

FULL = ["yolov2_416_voc_b8", "yolov3_416_voc_b64", "yolov4tiny_416_person_b256", "yolov4tiny2s_416_person_b256",
        "yolov3_608_coco_b32"]


def _sample(B, n=8):
    """n images spread over the batch (first and last included)."""
    return sorted({int(round(i * (B - 1) / max(n - 1, 1))) for i in range(min(n, B))})


def _setup(name, dev, regime="dense", **kw):
    from ypp import engine, synth
    w = synth.WORKLOADS[name]
    heads = [h.to(dev) for h in synth.make_heads(w["batch"], w["nc"], w["grids"], A=w["A"], g=77,
                                                 mu_obj=0.0 if regime == "dense" else -5.0)]
    pp = engine.PostProcessor(w["nc"], synth.workload_anchors(w), w["input_size"], w["conf"], kind=w["kind"],
                              anchors_per_scale=w["A"], **kw)
    return w, heads, pp


@pytest.mark.parametrize("name", FULL)
@pytest.mark.parametrize("regime", ["dense", "sparse"])
def test_full_size_properties(oracle, cuda, name, regime):
    from ypp import _lib
    w, heads, pp = _setup(name, cuda, regime, offset_mode="none", union_mode="separate")
    B = w["batch"]
    dec = pp.decode(heads)                                   # [B, N, 6]
    N = dec.shape[1]
    res = pp(heads, want_keep_idx=True, row_width=7)
    torch.cuda.synchronize()
    kc, cc = res.keep_count.long(), res.cand_count.long()
    # counts
    assert int(res.det_total.item()) == int(kc.sum().item())
    assert torch.equal(cc, (dec[..., 4] > w["conf"]).sum(1))
    assert bool((kc <= cc).all())
    rows = res.rows()
    off = torch.cumsum(kc, 0) - kc
    img_col = torch.repeat_interleave(torch.arange(B, device=cuda), kc)
    assert torch.equal(rows[:, 0].long(), img_col)
    # every output row is the decoded row of its keep index, bit for bit
    pos = torch.arange(rows.shape[0], device=cuda) - off[img_col]
    kidx = res.keep_idx.long()[img_col, pos]
    assert torch.equal(rows[:, 1:], dec[img_col, kidx])
    assert bool((dec[img_col, kidx, 4] > w["conf"]).all())
    # order inside an image: score descending, ties by increasing box index; indices unique
    same = img_col[1:] == img_col[:-1]
    s0, s1 = rows[:-1, 5], rows[1:, 5]
    assert bool(((s0 > s1) | ((s0 == s1) & (kidx[:-1] < kidx[1:])) | ~same).all())
    flat = img_col * N + kidx
    assert flat.unique().numel() == flat.numel()
    # idempotence: NMS of the kept rows keeps all of them, in the same order (vanilla per-class mode)
    kmax = int(kc.max().item())
    if kmax > 0:
        padded = torch.zeros((B, kmax, 6), device=cuda)
        padded[img_col, pos] = rows[:, 1:]
        again = pp.nms_rows(padded, want_keep_idx=True)
        assert torch.equal(again.keep_count.long(), kc)
        assert torch.equal(again.rows(), rows[:, 1:])
    # a batch equals its images processed one by one (first, middle, last)
    for b in sorted({0, B // 2, B - 1}):
        one = pp([h[b:b + 1].contiguous() for h in heads], want_keep_idx=True, row_width=7, img_idx_base=b)
        assert int(one.keep_count.item()) == int(kc[b].item())
        assert torch.equal(one.keep_idx[0, : int(kc[b])], res.keep_idx[b, : int(kc[b])])
        assert torch.equal(one.rows(), rows[off[b]: off[b] + kc[b]])
    # eight images spread over the batch against the oracle fed OUR decoded rows (bit-exact indices)
    for b in _sample(B, 8):
        r = dec[b].cpu().numpy()
        keep, ncand = oracle.nms_v3_indices(r, w["conf"], 0.45, offset_mode=_lib.OFFSET_MODES["none"],
                                            union_mode=_lib.UNION_MODES["separate"])
        assert ncand == int(cc[b])
        assert np.array_equal(res.keep_idx[b, : int(kc[b])].cpu().numpy(), util.canon(keep, r)), (name, regime, b)


@pytest.mark.parametrize("name", FULL)
def test_full_size_default_modes_match_torchvision_on_device(cuda, name):
    """Default modes (auto_cuda, fma) = what the reference computes on a CUDA device: compare whole
    images with torchvision.ops.batched_nms run on the same GPU on our decoded rows."""
    tv = pytest.importorskip("torchvision")
    w, heads, pp = _setup(name, cuda, "dense")
    dec = pp.decode(heads)
    res = pp(heads, want_keep_idx=True)
    kc = res.keep_count.long()
    B = w["batch"]
    for b in _sample(B, 12):
        r = dec[b]
        idx = torch.where(r[:, 4] > w["conf"])[0]
        c = r[idx]
        xyxy = torch.stack([c[:, 0] - c[:, 2] / 2, c[:, 1] - c[:, 3] / 2, c[:, 0] + c[:, 2] / 2, c[:, 1] + c[:, 3] / 2], 1)
        keep = tv.ops.batched_nms(xyxy, c[:, 4], c[:, 5], 0.45)
        want = idx[keep]
        # the vanilla path's final score sort is not stable: canonicalise to (score desc, index asc)
        order = torch.tensor(np.lexsort((want.cpu().numpy(), -r[want, 4].cpu().numpy())), device=cuda)
        assert torch.equal(res.keep_idx[b, : int(kc[b])].long(), want[order]), (name, b)


@pytest.mark.parametrize("n", [25000, 25001, 60000])
def test_large_single_image_and_the_cuda_heuristic_flip(oracle, cuda, n):
    """torchvision's CUDA heuristic takes the coordinate trick up to 25 000 candidates (4 n <= 100 000) and the
    per-class loop above; 60 000 boxes no longer fit the sort's shared-memory key buffers.  Every box passes the
    threshold, 80 classes, one cross-class pair that only the trick suppresses."""
    from ypp import engine, _lib
    rng = np.random.RandomState(n)
    rows = np.concatenate([rng.uniform(0, 1200, (n, 2)), rng.uniform(8, 90, (n, 2)),
                           rng.permutation(n).reshape(n, 1) / (2.0 * n) + 0.3, rng.randint(0, 80, (n, 1))], 1).astype(np.float32)
    rows[0] = [1290, 1290, 20, 20, .99, 0]          # max coordinate 1300
    rows[1] = [-9, -9, 20, 20, .98, 1]              # class 1 at negative coordinates: lands on it under the trick
    pp = engine.PostProcessor(80, [[1, 1]], 1, 0.25, kind="v2", anchors_per_scale=1)      # auto_cuda + fma
    res = pp.nms_rows(torch.tensor(rows).to(cuda), want_keep_idx=True)
    k = res.keep_idx[0, : int(res.keep_count[0])].cpu().numpy()
    ref, ncand = oracle.nms_v3_indices(rows, .25, .45, offset_mode=_lib.OFFSET_MODES["auto_cuda"], union_mode=_lib.UNION_MODES["fma"])
    assert ncand == n == int(res.cand_count[0])
    assert np.array_equal(k, util.canon(ref, rows))
    assert (1 in set(k.tolist())) == (n > 25000)    # suppressed by the trick only
