"""Time D1 (decode + threshold + compaction) alone: CUDA events over K launches, ring of inputs > L2."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
import torch  # noqa: E402
from ypp import engine, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="yolov3_416_voc_b64")
ap.add_argument("--regime", default="dense")
ap.add_argument("--steps", type=int, default=50)
ap.add_argument("--ring", type=int, default=4)
ap.add_argument("--stages", default="decode")
ap.add_argument("--grids", default="", help="override the grid sizes, e.g. 52,26,12 (tuning experiments)")
a = ap.parse_args()
w = dict(synth.WORKLOADS[a.workload])
if a.grids:
    w["grids"] = [int(x) for x in a.grids.split(",")]
dev = torch.device("cuda:0")
mu = 0.0 if a.regime == "dense" else -5.0
ring = [[h.to(dev) for h in synth.make_heads(w["batch"], w["nc"], w["grids"], A=w["A"], g=i, mu_obj=mu)] for i in range(a.ring)]
pp = engine.PostProcessor(w["nc"], synth.workload_anchors(w), w["input_size"], w["conf"], kind=w["kind"], anchors_per_scale=w["A"])
res = pp(ring[0])
cand = res.cand_count.float().mean().item()
stages = tuple(a.stages.split(","))
for hs in ring:
    pp.staged(hs, stages=stages)
torch.cuda.synchronize()
REPS = max(1, a.steps // a.ring)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for _ in range(REPS):
        for hs in ring:
            pp.staged(hs, stages=stages)
for i in range(3):
    g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best = 1e9
for rep in range(5):
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / (REPS * a.ring))
hb = synth.head_bytes_per_image(w["nc"], w["grids"], w["A"])
alg = w["batch"] * (hb + 28 * cand)
print("%s %s lib=%s stages=%s: %.2f us/launch, %.0f GB/s algorithmic (%.1f%% of 6456), cand/img %.0f" % (
    a.workload, a.regime, os.environ.get("YPP_LIB", "default"), a.stages, best * 1e3, alg / best / 1e6,
    100 * alg / best / 1e6 / 6456.2, cand))
