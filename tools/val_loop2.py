"""MeanAveragePrecision.update_state loop, dense and sparse, per pipeline depth (wall clock, heads resident)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
import torch
from ypp import synth
from ypp.compat.yolov3_utils import MeanAveragePrecision
w = synth.WORKLOADS["yolov3_416_voc_b64"]
dev = torch.device("cuda:0")
B, nc = w["batch"], w["nc"]
n_b = int(sys.argv[1]) if len(sys.argv) > 1 else 96
for regime, mu in (("dense", 0.0), ("sparse", -5.0)):
    ring = [[h.to(dev) for h in synth.make_heads(B, nc, w["grids"], g=i, mu_obj=mu)] for i in range(6)]
    ann = [synth.make_annots(B, nc, g=i).to(dev) for i in range(6)]
    for depth in (1, 2, 3, 4):
        best, syncs = None, None
        m = MeanAveragePrecision(nc, synth.YOLOV3_ANCHORS, 416, 0.25, pipeline_depth=depth)
        for rep in range(4):
            m.reset_states()                  # epochs of ONE metric object, as module/yolov3_detector.py:43 does
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for i in range(n_b):
                m.update_state(ann[i % 6], ring[i % 6])
            t1 = time.perf_counter()
            n = m.all_pred_boxes_variable.shape[0]
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) * 1e3 / n_b
            best = dt if best is None else min(best, dt)
            syncs = m._pred.syncs
            host = (t1 - t0) * 1e3 / n_b
        print("%s depth %d: %.4f ms/batch (%.0f images/s), host loop %.4f ms/batch, rows %d, cursor reads %d, buffer %d rows"
              % (regime, depth, best, B / best * 1e3, host, n, syncs, m._pred.buf.shape[0]))
    del ring
