"""GPU: append mode of the row gather and of the ground-truth ingest (ypp_nms_append / ypp_target_rows_append,
the torch.cat accumulation of utils/yolov3_utils.py:139-146 done behind a device cursor) against the
per-batch path: same rows, same order, bit for bit, through buffer growth and cursor reads."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _batches(cuda, n, B=4, mu=-3.0, grids=(52, 26, 13)):
    from ypp import synth
    return [[h.to(cuda) for h in synth.make_heads(B, 20, list(grids), g=300 + i, mu_obj=mu)] for i in range(n)]


@pytest.mark.parametrize("sync_every", [1, 2, 16])
def test_accumulated_rows_equal_concatenated_batches(cuda, sync_every):
    from ypp import engine, synth
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    batches = _batches(cuda, 7)
    want = torch.cat([pp(h, img_idx_base=4 * i).rows().clone() for i, h in enumerate(batches)])
    acc = engine.RowAccumulator(cuda, sync_every)
    for i, h in enumerate(batches):
        sc = pp.accumulate(acc, h, img_idx_base=4 * i)
    got = acc.view()
    assert got.shape == want.shape and want.shape[0] > 100
    assert torch.equal(got, want)
    # the counts of the last batch still come back
    assert int(sc.keep_count.sum()) == int(pp(batches[-1]).det_total)
    if sync_every == 16:
        assert acc.syncs == 1                       # only the view() above read the cursor


def test_growth_keeps_rows_when_batches_are_dense(cuda):
    """Dense heads (most boxes pass): every reserve() finds the bound at the capacity, reads the cursor, and
    grows when the rows really do not fit; earlier rows survive each move."""
    from ypp import engine, synth
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    batches = _batches(cuda, 5, B=2, mu=0.0, grids=(26, 13))
    want = torch.cat([pp(h, img_idx_base=2 * i).rows().clone() for i, h in enumerate(batches)])
    acc = engine.RowAccumulator(cuda, 1)
    caps = set()
    for i, h in enumerate(batches):
        pp.accumulate(acc, h, img_idx_base=2 * i)
        caps.add(acc.buf.shape[0])
    assert len(caps) > 1                            # it did grow
    assert torch.equal(acc.view(), want)


def test_overflow_drops_rows_but_counts_them(cuda):
    """The C entry point with a buffer that is too small: rows past det_capacity are not written, the cursor
    still counts them, nothing outside the buffer is touched."""
    from ypp import engine, synth, _lib
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    heads = _batches(cuda, 1)[0]
    full = pp(heads).rows().clone()
    n = full.shape[0]
    cap = n // 2
    sc = pp.staged(heads, stages=("decode", "sort"))
    guard = torch.full((cap + 64, 7), -7.0, device=cuda)
    cur = torch.tensor([3, 0], dtype=torch.int64, device=cuda)         # 3 rows "already there"
    L = _lib.lib()
    word = torch.zeros(1, dtype=torch.int64).pin_memory()
    probe = _lib.YppProbe(word.data_ptr(), 77)
    cfg = pp._cfg(4, 3)
    ws = pp.workspace(L.ypp_workspace_bytes(4, sc.n_boxes, 20), cuda)
    _lib.check(L.ypp_nms_append(ctypes.byref(cfg), sc.n_boxes, ws.data_ptr(), ws.numel(), None, sc.keep_count.data_ptr(),
                                guard.data_ptr(), 7, cap, cur.data_ptr(), cur.data_ptr() + 8, 0.0,
                                torch.cuda.current_stream().cuda_stream, ctypes.byref(probe)))
    assert cur.tolist() == [3, 3 + n]
    assert int(word[0]) == ((3 + n) << 24) | 77        # the kernel reported the cursor and the append number to pinned memory
    assert torch.equal(guard[3:cap], full[: cap - 3])
    assert bool((guard[:3] == -7).all()) and bool((guard[cap:] == -7).all())


def test_ground_truth_rows_append(cuda):
    from ypp import engine, synth
    acc = engine.RowAccumulator(cuda, 2)
    want = []
    for i in range(6):
        an = synth.make_annots(8, 20, g=i).to(cuda)
        rows, cnt = engine.target_rows(an, 416, 8 * i)
        want.append(rows[: int(cnt)].clone())
        engine.target_rows_append(acc, an, 416, 8 * i)
    assert torch.equal(acc.view(), torch.cat(want))
    # all-padding batch: nothing appended
    before = acc.view().shape[0]
    engine.target_rows_append(acc, torch.full((3, 5, 5), -1.0, device=cuda), 416, 48)
    assert acc.view().shape[0] == before


def test_accumulator_from_features_and_result(cuda):
    """update_state_features over several batches == rows of the per-batch fused-head path; result() is the mAP
    of exactly those rows."""
    from ypp import engine, synth
    from ypp.compat.yolov3_utils import MeanAveragePrecision
    gen = torch.Generator("cpu").manual_seed(5)
    cins, grids, B = (64, 96, 128), (52, 26, 13), 2
    ws = [(torch.randn(75, c, 1, 1, generator=gen) * (1.2 / c ** 0.5)).to(cuda) for c in cins]
    m = MeanAveragePrecision(20, synth.YOLOV3_ANCHORS, 416, 0.25, sync_every=2)
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    dets, gts = [], []
    for i in range(5):
        feats = [torch.randn(B, c, g, g, generator=gen).to(cuda).contiguous(memory_format=torch.channels_last)
                 for c, g in zip(cins, grids)]
        an = synth.make_annots(B, 20, g=i)
        m.update_state_features(an, feats, ws)
        dets.append(pp.from_features(feats, ws, img_idx_base=B * i).rows().clone())
        r, c = engine.target_rows(an.to(cuda), 416, B * i)
        gts.append(r[: int(c)].clone())
    assert torch.equal(m.all_pred_boxes_variable, torch.cat(dets))
    assert torch.equal(m.all_true_boxes_variable, torch.cat(gts))
    want = engine.map_match(torch.cat(gts), torch.cat(dets), 20)[0]
    assert float(m.result()) == float(want)


def test_append_on_a_device_that_is_not_current(cuda):
    """One process, two GPUs: the accumulators live on cuda:1 while cuda:0 is current."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from ypp import engine, synth
    d1 = torch.device("cuda", 1)
    pp = engine.PostProcessor(20, synth.YOLOV3_ANCHORS, 416, 0.25)
    batches = [[h.to(d1) for h in hs] for hs in _batches(cuda, 3)]
    want = torch.cat([pp(h, img_idx_base=4 * i).rows().clone() for i, h in enumerate(batches)])
    acc, gt = engine.RowAccumulator(d1, 2), engine.RowAccumulator(d1)
    for i, h in enumerate(batches):
        pp.accumulate(acc, h, img_idx_base=4 * i)
        engine.target_rows_append(gt, synth.make_annots(4, 20, g=i).to(d1), 416, 4 * i)
    assert torch.cuda.current_device() == 0
    assert torch.equal(acc.view(), want) and want.device == d1
    assert gt.view().shape[0] == sum(int((synth.make_annots(4, 20, g=i)[..., 0] >= 0).sum()) for i in range(3))
