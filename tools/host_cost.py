"""Host cost of one MeanAveragePrecision.update_state call: the loop on tiny batches (the GPU is never the limit),
wall clock per call without any synchronisation, and a cProfile of it.
usage: python tools/host_cost.py [depth]"""
import cProfile, os, pstats, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
import torch
from ypp import synth
from ypp.compat.yolov3_utils import MeanAveragePrecision

depth = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dev = torch.device("cuda:0")
B = 1
ring = [[h.to(dev) for h in synth.make_heads(B, 20, [52, 26, 13], g=i, mu_obj=-5.0)] for i in range(6)]
ann = [synth.make_annots(B, 20, g=i).to(dev) for i in range(6)]
for d in (1, depth):
    m = MeanAveragePrecision(20, synth.YOLOV3_ANCHORS, 416, 0.25, pipeline_depth=d)
    for i in range(200):
        m.update_state(ann[i % 6], ring[i % 6])
    torch.cuda.synchronize()
    m.reset_states()
    n = 3000
    t0 = time.perf_counter()
    for i in range(n):
        m.update_state(ann[i % 6], ring[i % 6])
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print("depth %d: host %.1f us per update_state (queue drained %.1f us later per call)" % (d, 1e6 * (t1 - t0) / n, 1e6 * (t2 - t1) / n))
    if d == depth:
        m.reset_states()
        pr = cProfile.Profile()
        pr.enable()
        for i in range(n):
            m.update_state(ann[i % 6], ring[i % 6])
        pr.disable()
        torch.cuda.synchronize()
        pstats.Stats(pr).sort_stats("tottime").print_stats(18)
