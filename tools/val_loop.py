"""The validation loop's accumulator (MeanAveragePrecision.update_state, utils/yolov3_utils.py:108-146) two ways:
  sync    per batch: chain into a per-call buffer, read the two counts on the host, copy the rows behind the
          accumulated ones (what ypp.compat did before append mode);
  append  the gather and the ground-truth ingest write behind a device cursor, nothing is read back.
Wall clock over `--batches` batches with a synchronize at both ends, heads resident in HBM (ring of 6).
usage: python tools/val_loop.py [workload] [dense|sparse] [--batch B] [--batches K]"""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
import torch
from ypp import engine, synth
from ypp.compat._accum import _GrowBuf, MeanAveragePrecisionBase

ap = argparse.ArgumentParser()
ap.add_argument("workload", nargs="?", default="yolov3_416_voc_b64")
ap.add_argument("regime", nargs="?", default="sparse")
ap.add_argument("--batch", type=int, default=0)
ap.add_argument("--batches", type=int, default=200)
a = ap.parse_args()
w = synth.WORKLOADS[a.workload]
B = a.batch or w["batch"]
dev = torch.device("cuda:0")
mu = 0.0 if a.regime == "dense" else -5.0
ring = [[h.to(dev) for h in synth.make_heads(B, w["nc"], w["grids"], A=w["A"], g=i, mu_obj=mu)] for i in range(6)]
annots = [synth.make_annots(B, w["nc"], g=i).to(dev) for i in range(6)]
anchors = synth.workload_anchors(w)


def run_sync(K):
    pp = engine.PostProcessor(w["nc"], anchors, w["input_size"], w["conf"], kind=w["kind"], anchors_per_scale=w["A"])
    pred, true, img = _GrowBuf(), _GrowBuf(), 0
    for i in range(K):
        res = pp(ring[i % 6], img_idx_base=img, row_width=7)
        g, c = engine.target_rows(annots[i % 6], w["input_size"], img)
        n_det, n_gt = torch.cat([res.det_total, c]).tolist()
        pred.append(res.det_rows[:n_det]); true.append(g[:n_gt])
        img += B
    return pred.n, true.n


def run_append(K):
    m = MeanAveragePrecisionBase(w["nc"], anchors, w["input_size"], w["conf"], w["kind"], anchors_per_scale=w["A"])
    for i in range(K):
        m._update(annots[i % 6], ring[i % 6])
    n = (m._pred.view().shape[0], m._true.view().shape[0])
    return n + (m._pred.syncs,)


for name, fn in (("sync", run_sync), ("append", run_append)):
    fn(12)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = fn(a.batches)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("%-6s B=%d %s: %.3f ms/batch, %.0f images/s, rows %s" % (name, B, a.regime, 1e3 * dt / a.batches, B * a.batches / dt, out))
