"""GPU: the checks compute-sanitizer would have made, done by hand (the tool is closed on this pool: it leaves GPUs
needing a reset).  (1) memcheck's job: the workspace and every output of the chain sit between 64 KiB guard zones
filled with a pattern; after the chain has run on every BASELINE shape the guards are untouched and the results
equal the ones computed in ordinary buffers - the workspace being EXACTLY ypp_workspace_bytes long, a kernel that
indexes past its array lands in a guard.  (2) racecheck's job: a shared-memory or cursor race shows as a result that
changes from run to run; the dense chain (every segment through the sweep-and-prune path, S1's bucket sort, D1's
mbarrier pipeline) and the append path are repeated and compared bit for bit, also with the PDL mask that launches
every kernel of the chain early."""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GUARD = 65536
SHAPES = [("v2", 20, [13], 5, 8, 0.25, 0.0), ("v3", 20, [52, 26, 13], 3, 4, 0.25, 0.0), ("v3", 20, [52, 26, 13], 3, 4, 0.25, -5.0),
          ("v3", 1, [52, 26, 13], 3, 2, 0.25, 0.0), ("v3", 1, [26, 13], 3, 3, 0.25, 0.0), ("v3", 80, [76, 38, 19], 3, 1, 0.001, 0.0),
          ("v3", 7, [11, 7], 3, 3, 0.25, 0.0)]


class Guarded:
    """`nbytes` usable bytes between two guard zones; view(dtype, shape) gives the usable part."""

    def __init__(self, nbytes, dev):
        self.n = int(nbytes)
        pad = (-self.n) % 256
        self.raw = torch.full((GUARD + self.n + pad + GUARD,), 0xA5, dtype=torch.uint8, device=dev)
        self.pad = pad

    def ptr(self):
        return ctypes.c_void_p(self.raw.data_ptr() + GUARD)

    def view(self, dtype, shape):
        return self.raw[GUARD: GUARD + self.n].view(dtype).view(shape)

    def intact(self):
        return bool((self.raw[:GUARD] == 0xA5).all()) and bool((self.raw[GUARD + self.n:] == 0xA5).all())


@pytest.mark.parametrize("kind,nc,grids,A,B,conf,mu", SHAPES)
def test_chain_stays_inside_its_buffers(cuda, kind, nc, grids, A, B, conf, mu):
    from ypp import _lib, engine, synth
    L = _lib.lib()
    anchors = synth.YOLOV2_SCALED_ANCHORS if kind == "v2" else (synth.YOLOV3_ANCHORS[:A * len(grids)] if len(grids) == 3 else synth.YOLOV3_ANCHORS[-A * len(grids):])
    size = 608 if grids[0] == 76 else 416
    pp = engine.PostProcessor(nc, anchors, size, conf, kind=kind, anchors_per_scale=A)
    heads = [h.to(cuda) for h in synth.make_heads(B, nc, grids, A=A, g=31, mu_obj=mu)]
    want = pp(heads, want_keep_idx=True)
    arr, _, _, N = pp._scales(heads)
    cfg = pp._cfg(B, len(heads))
    ws = Guarded(L.ypp_workspace_bytes(B, N, nc), cuda)
    assert ws.raw.data_ptr() % 256 == 0
    cand, keep, total = Guarded(4 * B, cuda), Guarded(4 * B, cuda), Guarded(4, cuda)
    idx, rows = Guarded(4 * B * N, cuda), Guarded(28 * B * N, cuda)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for rep in range(2):
        rc = L.ypp_postprocess(ctypes.byref(cfg), arr, ws.ptr(), ws.n, cand.ptr(), keep.ptr(), idx.ptr(), rows.ptr(), 7, B * N,
                               total.ptr(), 0.0, st)
        assert rc == 0, L.ypp_last_error().decode()
    torch.cuda.synchronize()
    for name, g in (("workspace", ws), ("cand_count", cand), ("keep_count", keep), ("det_total", total), ("keep_idx", idx), ("det_rows", rows)):
        assert g.intact(), "%s: a guard zone was written" % name
    n = int(want.det_total)
    assert int(total.view(torch.int32, (1,))) == n
    assert torch.equal(keep.view(torch.int32, (B,)), want.keep_count) and torch.equal(cand.view(torch.int32, (B,)), want.cand_count)
    assert torch.equal(rows.view(torch.float32, (B * N, 7))[:n], want.det_rows[:n])
    # rows past det_total are never written
    assert bool((rows.raw[GUARD + 28 * n: GUARD + rows.n] == 0xA5).all())
    # the append entry with a capacity smaller than the result: nothing past the capacity is touched
    cap = max(n // 2, 1)
    small = Guarded(28 * cap, cuda)
    cur = torch.zeros(2, dtype=torch.int64, device=cuda)
    rc = L.ypp_postprocess_append(ctypes.byref(cfg), arr, ws.ptr(), ws.n, cand.ptr(), keep.ptr(), small.ptr(), 7, cap,
                                  ctypes.c_void_p(cur.data_ptr()), ctypes.c_void_p(cur.data_ptr() + 8), 0.0, st, None)
    assert rc == 0, L.ypp_last_error().decode()
    torch.cuda.synchronize()
    assert small.intact() and ws.intact() and cur.tolist() == [0, n]
    assert torch.equal(small.view(torch.float32, (cap, 7))[: min(cap, n)], want.det_rows[: min(cap, n)])


def test_map_kernels_stay_inside_their_workspace(cuda):
    from ypp import _lib
    import util
    L = _lib.lib()
    m = util.load("map_cases")
    gt, det, nc = torch.tensor(m["syn_gt"]).to(cuda), torch.tensor(m["syn_det"]).to(cuda), int(m["syn_nc"])
    G, K = gt.shape[0], det.shape[0]
    ws = Guarded(L.ypp_map_workspace_bytes(G, K, nc), cuda)
    tp, pc, mp = Guarded(K, cuda), Guarded(16 * nc, cuda), Guarded(4, cuda)
    tr, gpc = Guarded(12 * K, cuda), Guarded(8 * nc, cuda)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    assert L.ypp_map_match(p(gt), G, p(det), K, nc, 0.5, ws.ptr(), ws.n, tp.ptr(), pc.ptr(), mp.ptr(), st) == 0
    torch.cuda.synchronize()
    assert ws.intact() and tp.intact() and pc.intact() and mp.intact()
    assert abs(float(mp.view(torch.float32, (1,))) - float(m["syn_map"])) <= 1e-6
    assert L.ypp_map_flags(p(gt), G, p(det), K, nc, 0.5, ws.ptr(), ws.n, tp.ptr(), tr.ptr(), gpc.ptr(), st) == 0
    pc2 = Guarded(16 * nc, cuda)
    pc2.view(torch.float32, (nc, 4)).zero_()
    assert L.ypp_map_ap_from_flags(tr.ptr(), 1, K, None, gpc.ptr(), nc, 0, nc, ws.ptr(), ws.n, pc2.ptr(), st) == 0
    torch.cuda.synchronize()
    assert ws.intact() and tp.intact() and tr.intact() and gpc.intact() and pc2.intact()
    assert torch.equal(pc2.view(torch.float32, (nc, 4)), pc.view(torch.float32, (nc, 4)))


_REPEAT = r"""
import sys, torch
sys.path[:0] = [%r, %r]
from ypp import engine, synth
dev = torch.device("cuda:0")
pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
heads = [h.to(dev) for h in synth.make_heads(16, 20, [52, 26, 13], g=41)]
one = [h.to(dev) for h in synth.make_heads(8, 1, [52, 26, 13], g=42)]
pp1 = engine.PostProcessor(1, synth.YOLOV3_ANCHORS, 416, 0.25)
first = first1 = None
for rep in range(40):
    r = pp(heads, want_keep_idx=True)
    r1 = pp1(one, want_keep_idx=True)
    valid = torch.arange(r.keep_idx.shape[1], device=dev)[None] < r.keep_count[:, None]        # the tail of keep_idx is not written
    cur = (r.keep_count.clone(), torch.where(valid, r.keep_idx, -1), r.rows().clone())
    cur1 = (r1.keep_count.clone(), r1.rows().clone())
    if first is None:
        first, first1 = cur, cur1
    assert all(torch.equal(a, b) for a, b in zip(cur, first)), "run %%d differs" %% rep
    assert all(torch.equal(a, b) for a, b in zip(cur1, first1)), "single-class run %%d differs" %% rep
acc_rows = None
for rep in range(6):
    p3 = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25, pipeline_depth=3)
    acc = engine.RowAccumulator(dev, 2)
    for i in range(7):
        p3.accumulate(acc, heads, img_idx_base=16 * i)
    v = acc.view().clone()
    if acc_rows is None:
        acc_rows = v
    assert torch.equal(v, acc_rows), "append run %%d differs" %% rep
print("repeat ok", int(first[0].sum()), int(first1[0].sum()), acc_rows.shape[0])
"""


@pytest.mark.parametrize("pdl", ["default", "0", "31"])
def test_results_do_not_change_from_run_to_run(cuda, pdl):
    """40 repetitions of a dense 16-image batch (320 segments on the sweep-and-prune path), of a single-class batch (the
    long-segment kernel) and 6 of a 7-batch pipelined append, each compared bit for bit with the first; in a fresh
    process because the PDL mask is read once per process."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ)
    env.pop("YPP_PDL", None)
    if pdl != "default":
        env["YPP_PDL"] = pdl
    out = subprocess.run([sys.executable, "-c", _REPEAT % (root, os.path.join(root, "pytorch-object-detection_b200"))],
                         env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert "repeat ok" in out.stdout
