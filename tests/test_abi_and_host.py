"""CPU: the C-ABI library loads and exports every symbol include/ypp.h declares; host-side
logic (parameter blocks, synthetic workloads, error behaviour) -- no kernel is launched."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "ypp.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ypp_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from ypp import _lib
    assert os.path.exists(_lib.LIB_PATH), "run __graft_entry__.build() first"
    L = ctypes.CDLL(_lib.LIB_PATH)
    syms = _declared_symbols()
    assert len(syms) >= 16
    for s in syms:
        assert hasattr(L, s), "libypp.so does not export %s" % s
    assert sorted(_lib.EXPORTS) == syms


def test_host_only_entry_points(oracle):
    from ypp import _lib
    L = _lib.lib()
    assert L.ypp_version() >= 100
    assert L.ypp_workspace_bytes(64, 10647, 20) > 64 * 10647 * (8 + 16 + 16 + 4 + 1)
    assert L.ypp_workspace_bytes(0, 10647, 20) == 0
    assert L.ypp_map_workspace_bytes(100, 1000, 20) > 0
    for thr in (0.45, 0.3, 0.5, 0.7, 0.1):
        for dbl in (0, 1):
            assert float(L.ypp_effective_iou_thr(thr, dbl)) == oracle.effective_thr(thr, dbl)


def test_struct_layout_matches_header():
    from ypp import _lib
    # ypp_scale: pointer + 3 ints + 16 floats + 2 floats, 8-byte aligned; ypp_cfg: 8 x 4 bytes
    assert ctypes.sizeof(_lib.YppScale) == 96
    assert ctypes.sizeof(_lib.YppCfg) == 32


def test_errors_without_gpu():
    from ypp import engine
    pp = engine.PostProcessor(20, [[10, 13], [16, 30], [33, 23]], 416)
    with pytest.raises(ValueError):
        pp([torch.zeros(1, 75, 13, 13)])                 # CPU tensor: there is no CPU path
    with pytest.raises(ValueError):
        pp.decode([torch.zeros(75, 13, 13)])
    with pytest.raises(ValueError):
        engine.box_iou(torch.zeros(2, 4), torch.zeros(3, 4))


def test_missing_extension_fails_loudly(monkeypatch):
    from ypp import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libypp.so")
    with pytest.raises(ImportError):
        _lib.lib()


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "pytorch-object-detection_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(d, f)).read()
                assert "oracle" not in text.lower(), os.path.join(d, f)


def test_synth_is_deterministic_and_shaped():
    from ypp import synth
    w = synth.WORKLOADS["yolov3_416_voc_b64"]
    h1 = synth.make_heads(2, w["nc"], w["grids"], g=3)
    h2 = synth.make_heads(2, w["nc"], w["grids"], g=3)
    assert [tuple(h.shape) for h in h1] == [(2, 75, 52, 52), (2, 75, 26, 26), (2, 75, 13, 13)]
    assert all(torch.equal(a, b) for a, b in zip(h1, h2))
    assert synth.boxes_per_image(w["grids"]) == 10647
    assert synth.head_bytes_per_image(w["nc"], w["grids"]) == 1064700
    assert synth.head_bytes_per_image(80, [76, 38, 19]) == 7732620
    a = synth.make_annots(8, 20, g=1)
    assert a.shape == (8, 8, 5) and (a[0, 2:] == -1).all()
    dense = torch.sigmoid(h1[0].view(2, 3, 25, 52, 52)[:, :, 4]) > .25
    assert 0.7 < dense.float().mean() < 0.85


def test_scale_parameters_follow_the_reference_rounding(oracle):
    from ypp import engine, synth
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416)
    # same python-float division + single fp32 rounding as utils/yolov3_utils.py:29-31,40
    for i, H in enumerate([52, 26, 13]):
        aw, ah, sw, sh = oracle.scale_params_v3(synth.YOLOV3_ANCHORS[3 * i: 3 * i + 3], 416, H, H)
        for a, (w_, h_) in enumerate(synth.YOLOV3_ANCHORS[3 * i: 3 * i + 3]):
            assert ctypes.c_float(w_ / (416 / H)).value == float(aw[a])
            assert ctypes.c_float(h_ / (416 / H)).value == float(ah[a])
        assert float(sw) == 416 / H


def test_head_entry_points_check_their_arguments_before_any_launch():
    """ypp_head_decode_compact / ypp_head_postprocess (the fused head): struct layout and the
    argument errors that are raised before CUDA is touched."""
    from ypp import _lib
    L = _lib.lib()
    assert ctypes.sizeof(_lib.YppHeadSrc) == 32            # two pointers, int (+ padding), pointer
    cfg = _lib.YppCfg(20, 2, 1, 0.25, 0.45, 0, 3, 1)
    sc = (_lib.YppScale * 1)()
    sc[0].A, sc[0].H, sc[0].W = 3, 13, 13
    assert L.ypp_head_decode_compact(ctypes.byref(cfg), sc, None, None, 0, None) == -1        # YPP_E_BADARG: no sources
    assert b"null" in L.ypp_last_error()
    src = (_lib.YppHeadSrc * 1)()
    assert L.ypp_head_decode_compact(ctypes.byref(cfg), sc, src, None, 0, None) == -1         # feature / weight missing
    src[0].feature, src[0].weight, src[0].cin = 256, 512, 64
    assert L.ypp_head_decode_compact(ctypes.byref(cfg), sc, src, None, 0, None) == -3         # YPP_E_WORKSPACE
    assert L.ypp_head_postprocess(ctypes.byref(cfg), sc, src, None, 0, None, None, None, None, 7, 0, None, 0.0, None) == -3


def test_append_entry_points_check_their_arguments_before_any_launch():
    """ypp_nms_append needs the accumulator and two DIFFERENT cursor words; ypp_target_rows_append its cursor."""
    from ypp import _lib
    L = _lib.lib()
    cfg = _lib.YppCfg(20, 2, 1, 0.25, 0.45, 0, 3, 1)
    c = ctypes.byref(cfg)
    assert L.ypp_nms_append(c, 507, None, 0, None, None, None, 7, 0, 64, 72, 0.0, None, None) == -1       # no det_rows
    assert L.ypp_nms_append(c, 507, None, 0, None, None, 1024, 7, 0, 64, 64, 0.0, None, None) == -1       # one word twice
    assert b"cursor" in L.ypp_last_error()
    assert L.ypp_nms_append(c, 507, None, 0, None, None, 1024, 7, 0, 64, 72, 0.0, None, None) == -3       # YPP_E_WORKSPACE
    sc = (_lib.YppScale * 1)()
    sc[0].head, sc[0].A, sc[0].H, sc[0].W = 4096, 3, 13, 13
    sc[0].stride_w = sc[0].stride_h = 32.0
    assert L.ypp_postprocess_append(c, sc, None, 0, None, None, 1024, 7, 0, 64, 64, 0.0, None, None) == -1   # one word twice
    assert L.ypp_postprocess_append(c, sc, None, 0, None, None, 1024, 7, 0, 64, 72, 0.0, None, None) == -3   # YPP_E_WORKSPACE
    assert L.ypp_target_rows_append(1024, 2, 8, 416.0, 0.0, 2048, 16, None, None, None) == -1             # no cursor
    assert L.ypp_target_rows_append(1024, 0, 8, 416.0, 0.0, 2048, 16, 64, None, None) == -1               # B = 0


def test_oracle_head_conv_is_the_reference_layer(oracle):
    """oracle.head_conv1x1 against the layer itself, nn.Conv2d(Cin, A*(5+nc), 1, 1, bias=False)
    (models/detector/yolov3.py:56) run by torch on the CPU in fp32: equal within fp32 accumulation error."""
    g = torch.Generator().manual_seed(0)
    conv = torch.nn.Conv2d(48, 75, 1, 1, bias=False)
    x = torch.randn(2, 48, 13, 13, generator=g).relu_()
    with torch.no_grad():
        y = conv(x).numpy()
    exact, bound = oracle.head_conv1x1(x.numpy(), conv.weight.detach().numpy())
    assert exact.shape == y.shape
    assert (np.abs(y - exact) <= 2.0 ** -20 * bound + 1e-7).all()


def test_probe_word_arithmetic_survives_the_24_bit_wrap():
    """RowAccumulator reads (count << 24 | append number mod 2^24) from pinned memory: the append number is
    reconstructed relative to the number of appends issued, across the wrap."""
    from ypp.engine import RowAccumulator
    acc = RowAccumulator.__new__(RowAccumulator)     # the arithmetic only: no device, no pinned memory
    acc.bound = 10 ** 12
    for issued, reported in ((5, 3), (2 ** 24 + 7, 2 ** 24 + 2), (2 ** 24 + 3, 2 ** 24 - 4), (3 * 2 ** 24 + 1, 3 * 2 ** 24 + 1)):
        acc._seq = issued
        acc._host_np = [(12345 << 24) | (reported & 0xFFFFFF)]
        assert RowAccumulator._arrived(acc) == (reported, 12345)
        acc._booked = 1000
        acc._booked_at = [0] * RowAccumulator.RING
        acc._booked_at[reported % RowAccumulator.RING] = 400
        acc.bound = 10 ** 12
        RowAccumulator._refresh(acc)
        assert acc.bound == 12345 + 600              # rows after that append + the worst case of the appends since
    acc._seq = 200
    acc._host_np = [(1 << 24) | 100]                 # 100 appends behind: older than the ring remembers -> ignored
    acc.bound = 77
    RowAccumulator._refresh(acc)
    assert acc.bound == 77
    acc._host_np = [0]
    assert RowAccumulator._arrived(acc)[0] in (0, 200 - (200 & 0xFFFFFF))
