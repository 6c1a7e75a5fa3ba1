"""GPU: the drop-in API (ypp.compat.*) against the golden fixtures and the oracle; library
cross-checks against torch / torchvision running on the same device."""
import numpy as np
import pytest
import torch

import util

pytestmark = pytest.mark.gpu


def _adv_cases():
    g = util.load("nms_adversarial")
    return g, sorted({k.rsplit("__", 1)[0] for k in g.files})


def test_nms_v3_adversarial_cases_match_reference(cuda):
    """Hand-made cases whose expected output came from the unmodified reference (CPU torchvision):
    ties, IoU exactly at the threshold, zero-area / inverted boxes, empty candidate set,
    cross-class overlap under the coordinate trick, the 1000/1001 heuristic flip."""
    from ypp.compat.yolo_utils import nms_v3
    g, names = _adv_cases()
    for n in names:
        rows = torch.tensor(g[n + "__rows"]).to(cuda)
        conf, iou = (float(x) for x in g[n + "__conf_iou"])
        keep = g[n + "__keep"]
        out = nms_v3(rows, conf, iou, num_classes=20, offset_mode="auto_cpu", union_mode="separate")
        assert out.shape == (len(keep), 6), n
        assert np.array_equal(out.cpu().numpy(), g[n + "__rows"][keep]), n      # rows of the input, bit-exact


def test_nms_v3_empty_returns_empty(cuda):
    from ypp.compat.yolo_utils import nms_v3
    out = nms_v3(torch.tensor([[5, 5, 4, 4, .2, 0.]], device=cuda))
    assert out.shape == (0, 6)
    out = nms_v3(torch.zeros((0, 6), device=cuda))
    assert out.shape == (0, 6)


def test_cross_class_trick_edge_is_reproduced(oracle, cuda):
    """Class-1 box at negative coordinates lands on the class-0 band under the trick: suppressed in
    trick mode, kept in vanilla mode (SURVEY s.8(c))."""
    from ypp import engine
    g, _ = _adv_cases()
    for name in ("trick_cross_class", "trick_cross_class2"):
        rows = torch.tensor(g[name + "__rows"]).to(cuda)
        for om, want in (("trick", oracle.OFFSET_TRICK), ("none", oracle.OFFSET_NONE)):
            pp = engine.PostProcessor(20, [[1, 1]], 1, 0.25, kind="v2", anchors_per_scale=1, offset_mode=om,
                                      union_mode="separate")
            res = pp.nms_rows(rows, want_keep_idx=True)
            k = res.keep_idx[0, : int(res.keep_count[0])].cpu().numpy()
            ref, _ = oracle.nms_v3_indices(g[name + "__rows"], .25, .45, offset_mode=want)
            assert np.array_equal(k, util.canon(ref, g[name + "__rows"])), (name, om)


def test_big_single_class_segment_and_whole_image_fallback(oracle, cuda):
    """> 2048 boxes of one class exercises the persistent big-segment kernel; trick mode with a
    forced cross-class edge exercises the class-agnostic whole-image segment at that size."""
    from ypp import engine
    rng = np.random.RandomState(5)
    n = 5000
    rows = np.concatenate([rng.uniform(0, 416, (n, 2)), rng.uniform(8, 120, (n, 2)),
                           rng.uniform(.26, 1, (n, 1)), np.zeros((n, 1))], 1).astype(np.float32)
    rows[n // 2:, 5] = 1
    rows[0] = [1000, 1000, 20, 20, .99, 0]        # near max_coord
    rows[n - 1] = [-9, -9, 20, 20, .98, 1]         # shifted onto it under the trick
    t = torch.tensor(rows).to(cuda)
    for om, um in (("trick", "fma"), ("none", "separate"), ("trick", "separate")):
        pp = engine.PostProcessor(2, [[1, 1]], 1, 0.25, kind="v2", anchors_per_scale=1, offset_mode=om, union_mode=um)
        res = pp.nms_rows(t, want_keep_idx=True)
        k = res.keep_idx[0, : int(res.keep_count[0])].cpu().numpy()
        ref, _ = oracle.nms_v3_indices(rows, .25, .45, offset_mode=oracle.OFFSET_TRICK if om == "trick" else oracle.OFFSET_NONE,
                                       union_mode=oracle.UNION_FMA if um == "fma" else oracle.UNION_SEPARATE)
        assert np.array_equal(k, util.canon(ref, rows)), (om, um)
        if om == "trick":
            assert (n - 1) not in set(k.tolist())
        else:
            assert (n - 1) in set(k.tolist())


def test_coupled_classes_share_a_segment(oracle, cuda):
    """Trick mode with cross-class suppression: only the classes an edge couples are merged.  Many boxes
    near max_coord overflow one detector list, which takes the every-box sweep; a dense synthetic image
    known to hold such an edge (workload seed 64, image 44) goes through the whole chain."""
    from ypp import engine, synth, _lib
    rng = np.random.RandomState(11)
    n, nc = 3000, 6
    rows = np.concatenate([rng.uniform(0, 416, (n, 2)), rng.uniform(8, 120, (n, 2)),
                           rng.uniform(.26, 1, (n, 1)), rng.randint(0, nc, (n, 1))], 1).astype(np.float32)
    rows[:200, 0:2] = rng.uniform(990, 1000, (200, 2))      # a crowd near max_coord, all classes
    rows[:200, 2:4] = rng.uniform(8, 20, (200, 2))
    rows[:200, 4] = rng.uniform(.26, .9, 200)
    rows[0] = [1000, 1000, 20, 20, .99, 1]
    rows[n - 1] = [-9, -9, 20, 20, .98, 2]                  # class 2 shifted onto class 1's band
    rows[n - 2] = [-8, -10, 20, 20, .97, 5]                 # class 5 onto class 4's band
    rows[1] = [1001, 999, 20, 20, .995, 4]
    t = torch.tensor(rows).to(cuda)
    for um in ("fma", "separate"):
        pp = engine.PostProcessor(nc, [[1, 1]], 1, 0.25, kind="v2", anchors_per_scale=1, offset_mode="trick", union_mode=um)
        res = pp.nms_rows(t, want_keep_idx=True)
        k = res.keep_idx[0, : int(res.keep_count[0])].cpu().numpy()
        ref, _ = oracle.nms_v3_indices(rows, .25, .45, offset_mode=oracle.OFFSET_TRICK,
                                       union_mode=oracle.UNION_FMA if um == "fma" else oracle.UNION_SEPARATE)
        assert np.array_equal(k, util.canon(ref, rows)), um
        assert (n - 1) not in set(k.tolist()) and (n - 2) not in set(k.tolist())
    w = synth.WORKLOADS["yolov3_416_voc_b64"]
    heads = [h[44:46].contiguous().to(cuda) for h in synth.make_heads(w["batch"], w["nc"], w["grids"], A=w["A"], g=64)]
    pp = engine.PostProcessor(w["nc"], synth.workload_anchors(w), w["input_size"], w["conf"], kind=w["kind"],
                              anchors_per_scale=w["A"], offset_mode="auto_cuda", union_mode="fma")
    dec = pp.decode(heads).cpu().numpy()
    res = pp(heads, want_keep_idx=True)
    for b in range(2):
        k = res.keep_idx[b, : int(res.keep_count[b])].cpu().numpy()
        ref, _ = oracle.nms_v3_indices(dec[b], w["conf"], .45, offset_mode=_lib.OFFSET_MODES["auto_cuda"],
                                       union_mode=_lib.UNION_MODES["fma"])
        assert np.array_equal(k, util.canon(ref, dec[b])), b


def test_wrappers_match_reference(cuda):
    from ypp import synth
    from ypp.compat.yolov3_utils import DecodeYoloV3, decode_predictions as dec3
    from ypp.compat.yolov2_utils import DecodeYoloV2, decode_predictions as dec2
    g = util.load("wrappers")
    heads = [h.to(cuda) for h in synth.make_heads(1, 20, [52, 26, 13], g=5, mu_obj=-3.0)]
    out = DecodeYoloV3(20, synth.YOLOV3_ANCHORS, 416, 0.25, offset_mode="auto_cpu", union_mode="separate")(heads)
    assert out.shape == g["v3_rows"].shape
    assert np.allclose(out.cpu().numpy(), g["v3_rows"], rtol=1e-5, atol=0)          # tolerance 1e-5 relative
    h2 = synth.make_heads(1, 20, [13], A=5, g=6, mu_obj=-1.0)[0].to(cuda)
    out2 = DecodeYoloV2(20, synth.YOLOV2_SCALED_ANCHORS, 416, 0.25, offset_mode="auto_cpu", union_mode="separate")(h2)
    assert out2.shape == g["v2_rows"].shape
    assert np.allclose(out2.cpu().numpy(), g["v2_rows"], rtol=1e-5, atol=0)
    with pytest.raises(AssertionError):                   # batch != 1, utils/yolov3_utils.py:74
        DecodeYoloV3(20, synth.YOLOV3_ANCHORS, 416)([torch.cat([h, h]) for h in heads])
    d = dec3(heads[2], 20, synth.YOLOV3_ANCHORS[6:9], 416)
    assert d.shape == (1, 507, 6)
    assert dec2(h2, 20, synth.YOLOV2_SCALED_ANCHORS, 416).shape == (1, 845, 6)


def test_map_accumulator_matches_reference(cuda):
    from ypp import synth
    from ypp.compat.yolov3_utils import MeanAveragePrecision
    from ypp.compat.yolo_utils import metrics_per_class
    g = util.load("map_cases")
    m = MeanAveragePrecision(20, synth.YOLOV3_ANCHORS, 416, 0.25, offset_mode="auto_cpu", union_mode="separate")
    assert m.all_pred_boxes_variable == 0
    for i in range(2):
        heads = [h.to(cuda) for h in synth.make_heads(8, 20, [52, 26, 13], g=40 + i, mu_obj=-4.0)]
        m.update_state(torch.tensor(g["acc_annots"][i]), heads)
    assert m.img_idx == 16
    assert np.array_equal(m.all_true_boxes_variable.cpu().numpy(), g["acc_gt_rows"])
    det = m.all_pred_boxes_variable.cpu().numpy()
    assert det.shape == g["acc_det_rows"].shape
    assert np.allclose(det, g["acc_det_rows"], rtol=1e-5, atol=0)
    assert abs(float(m.result()) - float(g["acc_map"])) <= 1e-6          # tolerance 1e-6 absolute
    pc = metrics_per_class(m.all_true_boxes_variable, m.all_pred_boxes_variable, 20).cpu().numpy()
    assert np.allclose(pc, g["acc_per_class"], atol=1e-6)
    m.reset_states()
    assert m.img_idx == 0 and m.all_true_boxes_variable == 0


def test_map_matching_against_reference_and_oracle(oracle, cuda):
    from ypp import engine
    g = util.load("map_cases")
    gt, det, nc = g["syn_gt"], g["syn_det"], int(g["syn_nc"])
    m, pc, tp = engine.map_match(torch.tensor(gt).to(cuda), torch.tensor(det).to(cuda), nc)
    assert abs(float(m) - float(g["syn_map"])) <= 1e-6
    assert np.array_equal(pc.cpu().numpy()[:, 1:], g["syn_per_class"][:, 1:])      # TP/FP/FN exact
    assert np.allclose(pc.cpu().numpy()[:, 0], g["syn_per_class"][:, 0], atol=1e-6)
    assert np.array_equal(tp.cpu().numpy(), g["syn_tp_oracle"])                    # per-detection flags
    # no detections / no ground truth
    z = torch.zeros((0, 7), device=cuda)
    assert float(engine.map_match(torch.tensor(gt).to(cuda), z, nc)[0]) == 0.0
    m0, pc0, tp0 = engine.map_match(z, torch.tensor(det).to(cuda), nc)
    assert float(m0) == 0.0 and bool((tp0 == 255).all())


def test_map_matching_large_random(oracle, cuda):
    """Many images, duplicates and ties in the matching: TP flags bit-exact versus the oracle."""
    from ypp import engine
    rng = np.random.RandomState(21)
    n_img, nc = 700, 20
    gts, dets = [], []
    for i in range(n_img):
        ng = rng.randint(0, 7)
        gt = np.concatenate([np.full((ng, 1), i), rng.uniform(50, 360, (ng, 2)), rng.uniform(20, 150, (ng, 2)),
                             np.ones((ng, 1)), rng.randint(0, nc, (ng, 1))], 1)
        gts.append(gt)
        for _ in range(rng.randint(0, 25)):
            if ng and rng.rand() < .75:
                s = gt[rng.randint(ng)]
                box = s[1:5] * rng.uniform(.85, 1.2, 4)
                c = s[6] if rng.rand() < .9 else rng.randint(nc)
            else:
                box = np.concatenate([rng.uniform(0, 416, 2), rng.uniform(10, 150, 2)])
                c = rng.randint(nc)
            dets.append(np.concatenate([[i], box, [rng.uniform(.25, 1)], [c]]))
    gt = np.concatenate(gts).astype(np.float32)
    det = np.array(dets, dtype=np.float32)
    det = det[rng.permutation(len(det))]                       # arbitrary row order is allowed
    om, otp, opc = oracle.mean_average_precision(gt, det, nc, 0.5, True)
    m, pc, tp = engine.map_match(torch.tensor(gt).to(cuda), torch.tensor(det).to(cuda), nc)
    assert np.array_equal(tp.cpu().numpy(), otp)
    assert np.array_equal(pc.cpu().numpy()[:, 1:], opc[:, 1:])
    assert np.allclose(pc.cpu().numpy()[:, 0], opc[:, 0], atol=1e-6)
    assert abs(float(m) - float(om)) <= 1e-6
    assert opc[:, 1].sum() > 500


def test_box_iou_and_target_rows(oracle, cuda):
    from ypp import engine
    from ypp.compat.metrics import box_iou
    from ypp.compat.yolo_utils import get_target_boxes_for_map
    g = util.load("iou_cases")
    out = box_iou(torch.tensor(g["a"]).to(cuda), torch.tensor(g["b"]).to(cuda)).cpu().numpy()
    assert np.array_equal(out, g["box_iou"], equal_nan=True)                       # bit-exact
    m = util.load("map_cases")
    an = torch.tensor(m["acc_annots"][0]).to(cuda)
    rows, cnt = engine.target_rows(an, 416, 0)
    ref = oracle.get_target_rows(m["acc_annots"][0], 416, 0)
    assert int(cnt) == len(ref) and np.array_equal(rows[: int(cnt)].cpu().numpy(), ref)
    d = get_target_boxes_for_map(an, 416)
    assert sorted(d.keys()) == list(range(8)) and sum(v.shape[0] for v in d.values()) == len(ref)


def test_library_cross_checks_torch_and_torchvision(cuda):
    """Not the oracle -- the same-device library ops: decode against a plain torch fp32 restatement
    (expected bit-identical: same expf / IEEE division), keep sets against
    torchvision.ops.batched_nms on CUDA (coordinate trick + fma-contracted union)."""
    torchvision = pytest.importorskip("torchvision")
    from ypp import engine, synth
    heads = [h.to(cuda) for h in synth.make_heads(3, 20, [52, 26, 13], g=13)]
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)            # defaults: auto_cuda + fma
    ours = pp.decode(heads)
    ref = []
    for i, h in enumerate(heads):
        B, _, H, W = h.shape
        p = h.view(B, 3, 25, H, W).permute(0, 1, 3, 4, 2)
        stride = 416 / H
        an = torch.tensor([[a / stride, b / stride] for a, b in synth.YOLOV3_ANCHORS[3 * i: 3 * i + 3]], device=cuda)
        gx = torch.arange(W, device=cuda).view(1, 1, 1, W).float().expand(B, 3, H, W)
        gy = torch.arange(H, device=cuda).view(1, 1, H, 1).float().expand(B, 3, H, W)
        xy = torch.stack([torch.sigmoid(p[..., 0]) + gx, torch.sigmoid(p[..., 1]) + gy], -1) * stride
        wh = torch.exp(p[..., 2:4]) * an.view(1, 3, 1, 1, 2) * stride
        sc = torch.sigmoid(p[..., 4:5])
        cl = torch.argmax(torch.sigmoid(p[..., 5:]), -1, keepdim=True).float()
        ref.append(torch.cat([xy, wh, sc, cl], -1).reshape(B, -1, 6))
    ref = torch.cat(ref, 1)
    assert torch.equal(ours, ref), "decode differs bitwise from torch CUDA ops in %d values" % int((ours != ref).sum())
    res = pp(heads, want_keep_idx=True)
    for b in range(3):
        r = ref[b]
        cand = torch.where(r[:, 4] > 0.25)[0]
        c = r[cand]
        xyxy = torch.stack([c[:, 0] - c[:, 2] / 2, c[:, 1] - c[:, 3] / 2, c[:, 0] + c[:, 2] / 2, c[:, 1] + c[:, 3] / 2], 1)
        keep = cand[torchvision.ops.batched_nms(xyxy, c[:, 4], c[:, 5], 0.45)]
        k = res.keep_idx[b, : int(res.keep_count[b])].long()
        assert torch.equal(k, keep), "image %d: %d vs %d kept" % (b, len(k), len(keep))


def test_detection_writers_match_reference(oracle, cuda, tmp_path):
    """ypp_detection_ints + the host formatting vs the stored outputs of the reference's writer loops
    (make_pred_file_yolov3.py:72-86, yolo2coco_pred_file_yolov3.py:83-100) and vs the oracle on a big random set."""
    import json
    from ypp import engine
    from ypp.compat import writers
    g = util.load("writers")
    names = [str(x) for x in g["names"]]
    boxes = torch.tensor(g["boxes"]).to(cuda)
    assert np.array_equal(engine.detection_ints(boxes, "pred_file").cpu().numpy(), g["pred_ints"])
    assert writers.pred_file_lines(boxes, names) == [str(x) for x in g["lines"]]
    S = int(g["input_size"])
    sizes = g["sizes"].tolist()
    rows7 = torch.cat([torch.tensor((g["img_of"] + 7).astype(np.float32))[:, None].to(cuda), boxes], 1)
    cats = {nm: int(c) for nm, c in zip(names, g["cat_ids"])}
    res = writers.coco_results(rows7, [11, 12, 13], [s[0] for s in sizes], [s[1] for s in sizes], S, cats, names, img_idx_base=7)
    assert np.array_equal(np.array([r["bbox"] for r in res]), g["coco_ints"])
    assert [r["image_id"] for r in res] == [11 + int(i) for i in g["img_of"]]
    assert [r["category_id"] for r in res] == [cats[names[int(c)]] for c in g["boxes"][:, 5]]
    assert [r["score"] for r in res] == [float(x) for x in g["boxes"][:, 4]]
    path = writers.write_coco_results(res, str(tmp_path / "results.json"))
    assert json.load(open(path)) == res
    # per-image pred files of a packed batch
    order = np.argsort(g["img_of"], kind="stable")
    counts = np.bincount(g["img_of"], minlength=3)
    paths = writers.write_pred_files(rows7[torch.tensor(order).to(cuda)], counts, names, str(tmp_path / "pred"))
    want = [str(x) for x in g["lines"][order]]
    assert "".join(open(p).read() for p in paths) == "".join(want)
    assert [p[-9:] for p in paths] == ["00001.txt", "00002.txt", "00003.txt"]
    # large random set against the oracle (both modes), 6- and 7-column rows
    rng = np.random.RandomState(8)
    n = 200000
    big = np.concatenate([rng.randint(0, 5, (n, 1)), rng.uniform(-50, 700, (n, 2)), rng.uniform(0, 500, (n, 2)),
                          rng.uniform(0, 1, (n, 1)), rng.randint(0, 80, (n, 1))], 1).astype(np.float32)
    sc = [[rng.uniform(0.3, 4), rng.uniform(0.3, 4)] for _ in range(5)]
    tb = torch.tensor(big).to(cuda)
    assert np.array_equal(engine.detection_ints(tb, "pred_file").cpu().numpy(), oracle.detection_ints(big, oracle.WRITER_PRED_FILE))
    assert np.array_equal(engine.detection_ints(tb, "coco", scale_xy=torch.tensor(sc, dtype=torch.float64)).cpu().numpy(),
                          oracle.detection_ints(big, oracle.WRITER_COCO, sc))
    assert np.array_equal(engine.detection_ints(tb[:, 1:].contiguous(), "coco", scale_xy=torch.tensor(sc[:1], dtype=torch.float64)).cpu().numpy(),
                          oracle.detection_ints(big[:, 1:], oracle.WRITER_COCO, sc[:1]))
    assert engine.detection_ints(tb[:0], "pred_file").shape == (0, 4)
