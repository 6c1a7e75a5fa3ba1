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
