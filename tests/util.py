"""Helpers shared by the test modules: golden fixtures, regenerated inputs, canonical orders."""
import os
import zlib

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
HEAD_CASES = ["c1_yolov2_dense", "c2_yolov3_dense_b2", "c2_yolov3_sparse_b4", "c3_person_dense_b1", "c3_person2s_dense_b2",
              "c4_coco608_b1"]


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)


def case_heads(g):
    """Regenerate the heads of a golden head case from its seed and verify the crc."""
    from ypp import synth
    heads = synth.make_heads(int(g["batch"]), int(g["nc"]), [int(x) for x in g["grids"]], A=int(g["A"]),
                             g=int(g["g"]), mu_obj=float(g["mu_obj"]))
    got = [crc(h.numpy()) for h in heads]
    assert got == [int(x) for x in g["head_crc"]], "regenerated heads differ from the fixture (torch version?)"
    return heads


def case_anchors(g):
    from ypp import synth
    if str(g["kind"]) == "v2":
        return synth.YOLOV2_SCALED_ANCHORS
    n = int(g["A"]) * len(g["grids"])
    return synth.YOLOV3_ANCHORS[:n] if n == 9 else synth.YOLOV3_ANCHORS[-n:]


def canon(keep, rows):
    """(score desc, index asc)."""
    k = np.asarray(keep, dtype=np.int64)
    order = np.lexsort((k, -rows[k, 4].astype(np.float64)))
    return k[order].astype(np.int32)


def rel_err(a, b):
    with np.errstate(invalid="ignore", divide="ignore"):
        fin = np.isfinite(b)
        return np.where(fin, np.abs(a - b) / np.maximum(np.abs(b), 1e-30), 0.0)


def ultra_best_class(pred, conf_thres):
    """The scoring lines of the vendored Ultralytics NMS (utils/general.py:645,669,679-680) in torch ops, for ONE image
    [n, 5+nc] on any device: obj > conf -> conf_j = cls_j * obj -> (conf, j) = max over j (first maximum) -> conf > conf.
    Returns (box numbers, conf, j).  Pinned to the reference function by tests/golden/make_golden.py:case_objcls."""
    import torch
    sel = torch.nonzero(pred[:, 4] > conf_thres)[:, 0]
    x = pred[sel].clone()
    x[:, 5:] *= x[:, 4:5]
    conf, j = x[:, 5:].max(1)
    m = conf > conf_thres
    return sel[m], conf[m], j[m]


def torch_prediction(heads, nc, anchors, input_size, A=3):
    """decode_predictions (utils/yolov3_utils.py:27-59) restated in plain torch ops on whatever device the heads are
    on, with the class PROBABILITIES kept: [B, n, 5+nc] rows cx,cy,w,h,sigmoid(obj),sigmoid(cls_0).. in the reference's
    box order (scale, anchor, y, x) - what the vendored non_max_suppression expects as `prediction`."""
    import torch
    out = []
    for i, h in enumerate(heads):
        B, _, H, W = h.shape
        p = h.view(B, A, 5 + nc, H, W).permute(0, 1, 3, 4, 2)
        stride = input_size / H
        an = torch.tensor([[a / stride, b / stride] for a, b in anchors[A * i: A * i + A]], device=h.device)
        gx = torch.arange(W, device=h.device).view(1, 1, 1, W).float().expand(B, A, H, W)
        gy = torch.arange(H, device=h.device).view(1, 1, H, 1).float().expand(B, A, H, W)
        xy = torch.stack([torch.sigmoid(p[..., 0]) + gx, torch.sigmoid(p[..., 1]) + gy], -1) * stride
        wh = torch.exp(p[..., 2:4]) * an.view(1, A, 1, 1, 2) * stride
        out.append(torch.cat([xy, wh, torch.sigmoid(p[..., 4:5]), torch.sigmoid(p[..., 5:])], -1).reshape(B, -1, 5 + nc))
    return torch.cat(out, 1)


def objcls_heads():
    """The heads of tests/golden/objcls.npz (seeded + the edits of make_golden.py:case_objcls), crc-checked."""
    import torch
    from ypp import synth
    g = load("objcls")
    nc = int(g["nc"])
    heads = synth.make_heads(2, nc, [52, 26, 13], g=int(g["g"]), mu_obj=float(g["mu_obj"]))
    heads[0][0, 5:25, 0, 0] = 20.0
    heads[0][0, 4, 0, 0] = 2.0
    gen = torch.Generator("cpu").manual_seed(71)
    v = torch.rand(8, 26, generator=gen) * 3 - 1
    delta = 2.0 ** -23 * (1 + torch.exp(v)) * (0.3 + 0.9 * torch.rand(8, 26, generator=gen))
    heads[1][1, 5:25, :8, :] = -3.0
    heads[1][1, 5 + 7, :8, :] = v
    heads[1][1, 5 + 2, :8, :] = v - delta
    heads[1][1, 4, :8, :] = torch.rand(8, 26, generator=gen) * 2
    assert [crc(h.numpy()) for h in heads] == [int(x) for x in g["head_crc"]], "regenerated objcls heads differ from the fixture"
    return g, heads


def ap_from_flags_numpy(triples, gt_per_class, nc, c_lo=0, c_hi=None, seg_count=None):
    """What ypp_map_ap_from_flags computes, in numpy with one fp32 rounding per operation (utils/yolo_utils.py:432-452
    on detections that already carry their TP flag): triples [n,3] or [n_seg, stride, 3] (score, cls, tp), valid rows
    per segment in seg_count; stable order by score descending inside a class -> per_class [nc,4]."""
    f = np.float32
    t = np.asarray(triples, dtype=np.float32)
    if t.ndim == 2:
        t = t[None]
    segs = [t[s, : (int(seg_count[s]) if seg_count is not None else t.shape[1])] for s in range(t.shape[0])]
    rows = np.concatenate(segs, 0) if segs else np.zeros((0, 3), np.float32)
    c_hi = nc if c_hi is None else c_hi
    out = np.zeros((nc, 4), np.float32)
    for c in range(c_lo, c_hi):
        ng = int(gt_per_class[c])
        if ng == 0:
            continue
        r = rows[rows[:, 1] == c]
        r = r[np.argsort(-r[:, 0].astype(np.float64), kind="stable")]
        tp = (r[:, 2] != 0).astype(np.int64)
        tpc = np.cumsum(tp)
        fpc = np.cumsum(1 - tp)
        prec = (tpc.astype(f) / ((tpc.astype(f) + fpc.astype(f)).astype(f) + f(1e-6)).astype(f)).astype(f)
        env = np.maximum.accumulate(prec[::-1])[::-1] if len(prec) else prec
        rec_den = f(np.float64(ng) + 1e-6)
        ap = f(0)
        for k, p in enumerate(np.nonzero(tp)[0], start=1):
            ap = f(ap + f(f(f(k) / rec_den - f(k - 1) / rec_den) * env[p]))
        n_tp = int(tp.sum())
        out[c] = [ap, n_tp, len(r) - n_tp, ng - n_tp]
    return out
