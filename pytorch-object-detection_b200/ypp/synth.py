"""Synthetic head tensors / ground truth of the benchmark shapes.

Deterministic on any machine: everything is drawn from a seeded CPU generator and
then moved to the device by the caller.  No datasets, no checkpoints.
"""
import torch

# anchors of configs/yolov3_voc.yaml:9-19 (pixels; the COCO yolov3 anchors)
YOLOV3_ANCHORS = [[10, 13], [16, 30], [33, 23], [30, 61], [62, 45], [59, 119],
                  [116, 90], [156, 198], [373, 326]]
# configs/yolov2_voc.yaml:8 (grid units)
YOLOV2_SCALED_ANCHORS = [[1.3221, 1.73145], [3.19275, 4.00944], [5.05587, 8.09892],
                         [9.47112, 4.84053], [11.2364, 10.0071]]

# name -> (kind, input_size, num_classes, grid sizes, anchors/scale, batch, conf)
WORKLOADS = {
    "yolov2_416_voc_b8":        dict(kind="v2", input_size=416, nc=20, grids=[13], A=5, batch=8, conf=0.25),
    "yolov3_416_voc_b64":       dict(kind="v3", input_size=416, nc=20, grids=[52, 26, 13], A=3, batch=64, conf=0.25),
    "yolov4tiny_416_person_b256": dict(kind="v3", input_size=416, nc=1, grids=[52, 26, 13], A=3, batch=256, conf=0.25),
    "yolov4tiny2s_416_person_b256": dict(kind="v3", input_size=416, nc=1, grids=[26, 13], A=3, batch=256, conf=0.25),
    "yolov3_608_coco_b32":      dict(kind="v3", input_size=608, nc=80, grids=[76, 38, 19], A=3, batch=32, conf=0.001),
}


def workload_anchors(w):
    if w["kind"] == "v2":
        return YOLOV2_SCALED_ANCHORS
    n = w["A"] * len(w["grids"])
    # the reference indexes anchors[3*idx:3*idx+3] in head order (utils/yolov3_utils.py:77-78)
    return YOLOV3_ANCHORS[:n] if n == 9 else YOLOV3_ANCHORS[-n:]


def make_heads(batch, nc, grids, A=3, g=0, mu_obj=0.0, wh_scale=0.5, obj_scale=1.5):
    """Heads of one global batch g: list of [B, A*(5+nc), H, H] fp32 CPU tensors.

    dense regime mu_obj=0 (about 77% of boxes pass conf .25), sparse mu_obj=-5
    (about 0.5%).  Seeds: 20260000 + 16*g + scale index.
    """
    heads = []
    for s, H in enumerate(grids):
        gen = torch.Generator("cpu").manual_seed(20260000 + 16 * g + s)
        t = torch.randn(batch, A, 5 + nc, H, H, generator=gen)
        t[:, :, 2:4] *= wh_scale
        t[:, :, 4] = t[:, :, 4] * obj_scale + mu_obj
        heads.append(t.reshape(batch, A * (5 + nc), H, H).contiguous())
    return heads


def make_annots(batch, nc, g=0, max_g=8):
    """Padded GT [B, max_g, 5] = cx,cy,w,h,cid normalised, -1 padded like collate_fn
    (dataset/detection/datasets.py:108-123).  Image i of batch g has 1 + (i mod 8) boxes."""
    gen = torch.Generator("cpu").manual_seed(777 + g)
    out = torch.full((batch, max_g, 5), -1.0)
    for i in range(batch):
        n = 1 + ((g * batch + i) % max_g)
        out[i, :n, 0:2] = torch.rand(n, 2, generator=gen) * 0.8 + 0.1
        out[i, :n, 2:4] = torch.rand(n, 2, generator=gen) * 0.45 + 0.05
        out[i, :n, 4] = torch.randint(0, nc, (n,), generator=gen).float()
    return out


def head_bytes_per_image(nc, grids, A=3):
    return 4 * sum(A * (5 + nc) * H * H for H in grids)


def boxes_per_image(grids, A=3):
    return sum(A * H * H for H in grids)


def plant_annots(heads, annots, nc, anchors, input_size, A=3, g=0, p_detect=0.85, p_dup=0.5, p_wrong=0.1,
                 jitter=0.06, obj_logit=2.5):
    """Write detections DERIVED FROM the ground truth into (sparse-regime) heads, in place: for every valid row of
    annots [B,T,5] (cx,cy,w,h normalised, class) one cell of a random scale / anchor gets the logits that decode to
    that box (jittered), a confident objectness and its class (a wrong one with probability p_wrong); with
    probability p_dup a second anchor of the same cell gets a weaker copy (a duplicate: false positive after the
    first claims the ground truth).  The evaluation of such heads has an mAP of order 0.5 instead of the ~1e-7 of
    random boxes against random ground truth, so a matcher that returns nothing cannot pass for a correct one."""
    import math
    rng = torch.Generator("cpu").manual_seed(9000 + g)
    B, T, _ = annots.shape
    u = torch.rand((B, T, 8), generator=rng)
    n = torch.randn((B, T, 8), generator=rng)
    for b in range(B):
        for t in range(T):
            row = annots[b, t]
            if float(row.sum()) <= 0 or u[b, t, 0] > p_detect:
                continue
            si = int(u[b, t, 1] * len(heads)) % len(heads)
            H = heads[si].shape[2]
            stride = input_size / H
            a0 = int(u[b, t, 2] * A) % A
            cls = int(row[4]) if u[b, t, 3] > p_wrong else int(u[b, t, 4] * nc) % nc
            cx, cy = float(row[0]) * input_size / stride, float(row[1]) * input_size / stride
            gx, gy = min(int(cx), H - 1), min(int(cy), H - 1)
            for k, a in enumerate((a0, (a0 + 1) % A)):
                if k == 1 and u[b, t, 5] > p_dup:
                    break
                aw, ah = anchors[A * si + a]
                fx = min(max(cx - gx + jitter * float(n[b, t, 2 * k]), 0.03), 0.97)
                fy = min(max(cy - gy + jitter * float(n[b, t, 2 * k + 1]), 0.03), 0.97)
                w = float(row[2]) * input_size * math.exp(jitter * float(n[b, t, 4 + k]))
                h = float(row[3]) * input_size * math.exp(jitter * float(n[b, t, 6 + k]))
                base = a * (5 + nc)
                hd = heads[si]
                hd[b, base + 0, gy, gx] = math.log(fx / (1 - fx))
                hd[b, base + 1, gy, gx] = math.log(fy / (1 - fy))
                hd[b, base + 2, gy, gx] = math.log(max(w, 1e-3) / aw)
                hd[b, base + 3, gy, gx] = math.log(max(h, 1e-3) / ah)
                hd[b, base + 4, gy, gx] = obj_logit - 1.2 * k + 0.8 * float(u[b, t, 6 + k])
                hd[b, base + 5: base + 5 + nc, gy, gx] = -3.0
                hd[b, base + 5 + cls, gy, gx] = 4.0
    return heads
