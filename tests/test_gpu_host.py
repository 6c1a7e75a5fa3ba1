"""GPU: the host layer's ownership rules - scratch memory per (processor, stream), cached parameter blocks that
hold no tensors, several batches in flight in the accumulate path, thresholds that stay live attributes."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _heads(cuda, B, g, mu=0.0, grids=(52, 26, 13)):
    from ypp import synth
    return [h.to(cuda) for h in synth.make_heads(B, 20, list(grids), g=g, mu_obj=mu)]


def test_one_processor_on_two_streams_does_not_share_scratch(cuda):
    """Two dense batches through ONE processor from two streams at the same time: each stream has its own
    workspace, so both results equal the ones computed one after the other (a shared buffer - what a
    (device, tag) cache gave - mixes the two images' candidates)."""
    from ypp import engine, synth
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    a, b = _heads(cuda, 16, 900), _heads(cuda, 16, 901)
    want = [pp(h, want_keep_idx=True) for h in (a, b)]
    want = [(r.rows().clone(), r.keep_count.clone()) for r in want]
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    for rep in range(8):
        outs = []
        for st, h in ((s1, a), (s2, b)):
            with torch.cuda.stream(st):
                outs.append(pp(h, want_keep_idx=True))
        torch.cuda.synchronize()
        for (rows, kc), r in zip(want, outs):
            assert torch.equal(r.keep_count, kc)
            assert torch.equal(r.det_rows[: rows.shape[0]], rows)
    with torch.cuda.stream(s1):
        w1 = pp.workspace(1024, cuda)
    with torch.cuda.stream(s2):
        w2 = pp.workspace(1024, cuda)
    assert w1.data_ptr() != w2.data_ptr()
    other = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    with torch.cuda.stream(s1):
        assert other.workspace(1024, cuda).data_ptr() != w1.data_ptr()


def test_fresh_head_tensors_are_not_retained(cuda):
    """A validation loop hands over fresh head tensors every batch: the cached parameter blocks must not keep them
    alive (80 batches of 3 heads stayed resident before)."""
    from ypp import engine, synth
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    acc = engine.RowAccumulator(cuda)
    base = _heads(cuda, 8, 910, mu=-5.0)
    pp.accumulate(acc, [h.clone() for h in base])
    pp(base)
    torch.cuda.synchronize()
    m0 = torch.cuda.memory_allocated()
    per_batch = sum(h.numel() * 4 for h in base)
    for i in range(80):
        fresh = [h.clone() + 0.0 for h in base]
        pp.accumulate(acc, fresh, img_idx_base=8 * i)
        pp(fresh)
        del fresh
    acc.view()
    torch.cuda.synchronize()
    grown = torch.cuda.memory_allocated() - m0
    assert grown < 4 * per_batch, "memory grew by %.1f MB over 80 fresh batches" % (grown / 1e6)
    for pl in pp._plans.values():
        for name in pl.__slots__:
            assert not isinstance(getattr(pl, name), (torch.Tensor, list, tuple))


@pytest.mark.parametrize("depth", [1, 2, 3, 5])
def test_batches_in_flight_append_in_batch_order(cuda, depth):
    """accumulate with `depth` batches in flight on side streams: rows, order and counts equal the batches
    processed one by one; growth of the accumulator in the middle included (sync_every=1 forces it)."""
    from ypp import engine, synth
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25, pipeline_depth=depth)
    seq = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25, pipeline_depth=1)
    batches = [_heads(cuda, 4, 920 + i, mu=(0.0 if i % 3 == 0 else -3.0)) for i in range(9)]
    want = torch.cat([seq(h, img_idx_base=4 * i).rows().clone() for i, h in enumerate(batches)])
    for sync_every in (1, 16):
        acc = engine.RowAccumulator(cuda, sync_every)
        counts = []
        for i, h in enumerate(batches):
            sc = pp.accumulate(acc, [x.clone() for x in h], img_idx_base=4 * i)     # fresh tensors, dropped at once
            counts.append(sc)
        got = acc.view()
        assert got.shape == want.shape
        assert torch.equal(got, want)
        assert int(counts[-1].keep_count.sum()) == int(seq(batches[-1]).det_total)


def test_thresholds_are_live_attributes(cuda):
    """The reference reads self.conf_threshold on every call (utils/yolov3_utils.py:88,125); assigning it after
    construction must change the next call here too."""
    from ypp import engine, synth
    from ypp.compat.yolov3_utils import DecodeYoloV3, MeanAveragePrecision
    h = _heads(cuda, 1, 930, mu=-2.0)
    dec = DecodeYoloV3(20, synth.YOLOV3_ANCHORS, 416, 0.5)
    n_hi = dec(h).shape[0]
    dec.conf_threshold = 0.25
    got = dec(h)
    want = DecodeYoloV3(20, synth.YOLOV3_ANCHORS, 416, 0.25)(h)
    assert got.shape[0] > n_hi and torch.equal(got, want)
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.5)
    a = int(pp(h).det_total)
    pp.conf_threshold = 0.25
    assert int(pp(h).det_total) == want.shape[0] and a == n_hi
    pp.iou_threshold = 0.9
    assert int(pp(h).det_total) >= want.shape[0]
    m = MeanAveragePrecision(20, synth.YOLOV3_ANCHORS, 416, 0.5)
    annot = synth.make_annots(1, 20, g=0).to(cuda)
    m.update_state(annot, h)
    k0 = m.all_pred_boxes_variable.shape[0]
    m.reset_states()
    m.conf_threshold = 0.25
    m.update_state(annot, h)
    assert m.all_pred_boxes_variable.shape[0] == want.shape[0] and k0 == n_hi


def test_graph_keeps_its_own_workspace(cuda):
    """A captured chain bakes in the address of its workspace: a later, larger request through the same processor
    must not free it under the graph."""
    from ypp import engine, synth
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    small = _heads(cuda, 2, 940)
    g, res = pp.graphed(small, want_keep_idx=True)
    want = pp(small, want_keep_idx=True)
    want_rows, want_kc = want.rows().clone(), want.keep_count.clone()
    big = _heads(cuda, 32, 941)
    pp(big)                                         # grows the processor's own scratch
    junk = [torch.full((int(res.ws.numel()) // 4,), float("nan"), device=cuda) for _ in range(4)]   # would land on a freed buffer
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(res.keep_count, want_kc)
    assert torch.equal(res.det_rows[: want_rows.shape[0]], want_rows)
    del junk


def test_pipeline_c_abi_directly(cuda):
    """ypp_pipeline_* through ctypes, as a binding in another host language would use it: five batches submitted to a
    3-deep pipeline (own workspace per slot, two alternating cursor words, a probe word in pinned memory), joined on the
    caller's stream; rows, counts, cursor, tickets and the probe word against ypp_postprocess_append batch by batch."""
    import ctypes
    from ypp import _lib, engine, synth
    L = _lib.lib()
    B, nc, N = 4, 20, 10647
    pp = engine.PostProcessor(nc, synth.YOLOV3_ANCHORS, 416, 0.25)
    batches = [_heads(cuda, B, 960 + i, mu=(0.0 if i == 2 else -3.0)) for i in range(5)]
    want = torch.cat([pp(h, img_idx_base=B * i).rows().clone() for i, h in enumerate(batches)])
    h = ctypes.c_void_p()
    assert L.ypp_pipeline_create(0, ctypes.byref(h)) == -1 and L.ypp_pipeline_create(9, ctypes.byref(h)) == -1
    assert L.ypp_pipeline_create(3, ctypes.byref(h)) == 0
    nbytes = L.ypp_workspace_bytes(B, N, nc)
    ws = [torch.empty(nbytes, dtype=torch.uint8, device=cuda) for _ in range(3)]
    keep = [torch.zeros(B, dtype=torch.int32, device=cuda) for _ in range(3)]
    rows = torch.full((want.shape[0] + 64, 7), -7.0, device=cuda)
    cur = torch.zeros(2, dtype=torch.int64, device=cuda)
    word = torch.zeros(1, dtype=torch.int64).pin_memory()
    probe = _lib.YppProbe(word.data_ptr(), 0)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ticket = ctypes.c_int64(-1)
    keepalive = []
    for i, heads in enumerate(batches):
        slot = L.ypp_pipeline_next_slot(h)
        assert slot == i % 3 and L.ypp_pipeline_stream(h, slot)
        arr, hc, _, n_boxes = pp._scales(heads)
        keepalive.append((arr, hc))
        cfg = pp._cfg(B, 3)
        probe.seq = i + 1
        rc = L.ypp_pipeline_submit(h, ctypes.byref(cfg), arr, ws[slot].data_ptr(), nbytes, None, keep[slot].data_ptr(),
                                   rows.data_ptr(), 7, rows.shape[0], cur.data_ptr() + 8 * (i & 1), cur.data_ptr() + 8 * ((i + 1) & 1),
                                   ctypes.byref(probe), float(B * i), st, ctypes.byref(ticket))
        assert rc == 0, L.ypp_last_error().decode()
        assert ticket.value == i
    assert L.ypp_pipeline_query(h, 99) == 0
    assert L.ypp_pipeline_wait_ticket(h, 2) == 0 and L.ypp_pipeline_query(h, 2) == 1
    assert L.ypp_pipeline_join(h, st) == 0
    torch.cuda.synchronize()
    n = want.shape[0]
    assert cur[len(batches) & 1].item() == n
    assert torch.equal(rows[:n], want) and bool((rows[n:] == -7).all())
    assert int(word[0]) == (n << 24) | len(batches)
    assert int(keep[(len(batches) - 1) % 3].sum()) == int(pp(batches[-1]).det_total)
    assert all(L.ypp_pipeline_query(h, t) == 1 for t in range(5))
    # argument errors: one cursor word twice, missing rows
    arr, hc, _, _ = pp._scales(batches[0])
    cfg = pp._cfg(B, 3)
    assert L.ypp_pipeline_submit(h, ctypes.byref(cfg), arr, ws[0].data_ptr(), nbytes, None, None, rows.data_ptr(), 7, 8, cur.data_ptr(),
                                 cur.data_ptr(), None, 0.0, st, None) == -1
    assert L.ypp_pipeline_submit(h, ctypes.byref(cfg), arr, ws[0].data_ptr(), nbytes, None, None, None, 7, 8, cur.data_ptr(),
                                 cur.data_ptr() + 8, None, 0.0, st, None) == -1
    assert L.ypp_pipeline_destroy(h) == 0


def test_batches_in_flight_with_changing_batch_sizes(cuda):
    """The last batch of an epoch is smaller, and a loader may mix shapes: per-slot count buffers and workspaces are
    re-made while earlier batches are still in flight; rows and order still equal the one-by-one result."""
    from ypp import engine, synth
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25, pipeline_depth=3)
    seq = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25, pipeline_depth=1)
    shapes = [(4, (52, 26, 13)), (4, (52, 26, 13)), (2, (52, 26, 13)), (6, (26, 13)), (1, (52, 26, 13)), (4, (52, 26, 13)), (3, (13,))]
    batches, base = [], 0
    for i, (B, grids) in enumerate(shapes):
        heads = [h.to(cuda) for h in synth.make_heads(B, 20, list(grids), g=980 + i, mu_obj=-2.5)]
        anchors = synth.YOLOV3_ANCHORS[:3 * len(grids)] if len(grids) == 3 else synth.YOLOV3_ANCHORS[-3 * len(grids):]
        batches.append((heads, anchors, base))
        base += B
    want = []
    for heads, anchors, b0 in batches:
        seq.anchors = [list(map(float, a)) for a in anchors]
        seq._plans.clear()
        want.append(seq(heads, img_idx_base=b0).rows().clone())
    want = torch.cat(want)
    for rep in range(2):
        acc = engine.RowAccumulator(cuda, 2)
        for heads, anchors, b0 in batches:
            pp.anchors = [list(map(float, a)) for a in anchors]
            pp._plans.clear()
            pp.accumulate(acc, [h.clone() for h in heads], img_idx_base=b0)
        got = acc.view()
        assert got.shape == want.shape and want.shape[0] > 50
        assert torch.equal(got, want)


def test_default_depth_settles_on_two_for_sparse_output(cuda):
    """pipeline_depth not given: three batches in flight at first; once eight appends have reported fewer than 64 rows
    per image through the probe word the processor goes down to two (the host is the limit there).  The switch happens
    in the middle of the accumulation (old ring joined, new ring behind it): rows and order must not notice."""
    from ypp import engine, synth
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    seq = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25, pipeline_depth=1)
    batches = [_heads(cuda, 4, 1100 + i, mu=-5.0) for i in range(20)]           # 4 x 10647 boxes: above the pipelining limit
    want = torch.cat([seq(h, img_idx_base=4 * i).rows().clone() for i, h in enumerate(batches)])
    assert 0 < want.shape[0] < 64 * 4 * len(batches)
    acc = engine.RowAccumulator(cuda, 4)
    for i, h in enumerate(batches):
        if i == 12:
            torch.cuda.synchronize()                                            # every probe so far has arrived
        if i <= 8 or i > 12:                                                    # in between it depends on how far the GPU is
            assert pp._depth(4, 10647, None) == (3 if i <= 8 else 2)
        pp.accumulate(acc, [x.clone() for x in h], img_idx_base=4 * i)
    assert pp._auto_fixed and pp._auto_depth == 2
    got = acc.view()
    assert got.shape == want.shape and torch.equal(got, want)
    # dense output keeps three
    pd = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    accd = engine.RowAccumulator(cuda, 4)
    dense = [_heads(cuda, 4, 1200 + i, mu=0.0) for i in range(3)]
    for i in range(12):
        if i == 10:
            torch.cuda.synchronize()
        pd.accumulate(accd, dense[i % 3], img_idx_base=4 * i)
    assert pd._auto_fixed and pd._auto_depth == 3
    accd.view()
