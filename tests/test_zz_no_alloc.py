"""GPU: ypp_postprocess_static - the chain with its scratch in a device array built into the library, for callers
that cannot allocate (no counterpart in the reference).  Called through
the C ABI (ctypes) and compared with ypp_postprocess on a caller workspace: counts, keep indices and rows bit for bit
(the tolerance of every index / row comparison in this suite), on two calls in a row with different inputs, and once
more with the first input (the array keeps the previous call's contents: nothing in it has to start at zero).
Collected last on purpose."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _call(L, fn_static, cfg, arr, B, N, ws, dev):
    out = dict(cand=torch.full((B,), -1, dtype=torch.int32, device=dev), keep=torch.full((B,), -1, dtype=torch.int32, device=dev),
               idx=torch.full((B, N), -1, dtype=torch.int32, device=dev), rows=torch.full((B * N, 7), -1.0, device=dev),
               total=torch.full((1,), -1, dtype=torch.int32, device=dev))
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    if fn_static:
        rc = L.ypp_postprocess_static(ctypes.byref(cfg), arr, p(out["cand"]), p(out["keep"]), p(out["idx"]), p(out["rows"]), 7,
                                      B * N, p(out["total"]), 3.0, st)
    else:
        rc = L.ypp_postprocess(ctypes.byref(cfg), arr, p(ws), ws.numel(), p(out["cand"]), p(out["keep"]), p(out["idx"]),
                               p(out["rows"]), 7, B * N, p(out["total"]), 3.0, st)
    assert rc == 0, L.ypp_last_error().decode()
    torch.cuda.synchronize()
    return out


def _same(a, b):
    assert torch.equal(a["cand"], b["cand"]) and torch.equal(a["keep"], b["keep"]) and torch.equal(a["total"], b["total"])
    n = int(a["total"])
    assert n > 0
    assert torch.equal(a["rows"][:n], b["rows"][:n])
    for i in range(a["keep"].numel()):
        k = int(a["keep"][i])
        assert torch.equal(a["idx"][i, :k], b["idx"][i, :k])


def test_postprocess_static_equals_postprocess(cuda):
    from ypp import _lib, engine, synth
    L = _lib.lib()
    B, nc = 2, 20
    pp = engine.PostProcessor(nc, synth.YOLOV3_ANCHORS, 416, 0.25)
    inputs = [[h.to(cuda) for h in synth.make_heads(B, nc, [52, 26, 13], g=950 + i, mu_obj=mu)] for i, mu in enumerate((0.0, -3.0))]
    N = 10647
    nbytes = L.ypp_workspace_bytes(B, N, nc)
    assert nbytes <= _lib.YPP_STATIC_WS_BYTES
    ws = torch.empty(nbytes, dtype=torch.uint8, device=cuda)
    want, got = [], []
    for heads in inputs + inputs[:1]:                      # dense, sparse, dense again: two calls in a row and a return
        arr, _, b_, n_ = pp._scales(heads)
        assert (b_, n_) == (B, N)
        cfg = pp._cfg(B, len(heads))
        got.append(_call(L, True, cfg, arr, B, N, None, cuda))
        want.append(_call(L, False, cfg, arr, B, N, ws, cuda))
    for a, b in zip(got, want):
        _same(a, b)
    assert int(got[0]["total"]) != int(got[1]["total"])    # the two inputs really differ
    _same(got[0], got[2])


def test_postprocess_static_rejects_sizes_beyond_the_array(cuda):
    from ypp import _lib, engine, synth
    L = _lib.lib()
    B, nc = 8, 20
    assert L.ypp_workspace_bytes(B, 10647, nc) > _lib.YPP_STATIC_WS_BYTES
    pp = engine.PostProcessor(nc, synth.YOLOV3_ANCHORS, 416, 0.25)
    heads = [torch.zeros((B, 75, g, g), device=cuda) for g in (52, 26, 13)]
    arr, _, _, N = pp._scales(heads)
    cfg = pp._cfg(B, 3)
    keep = torch.zeros(B, dtype=torch.int32, device=cuda)
    rc = L.ypp_postprocess_static(ctypes.byref(cfg), arr, None, ctypes.c_void_p(keep.data_ptr()), None, None, 7, 0, None, 0.0,
                                  ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -3                                         # YPP_E_WORKSPACE
    msg = L.ypp_last_error().decode()
    assert "postprocess_static" in msg and str(_lib.YPP_STATIC_WS_BYTES) in msg
