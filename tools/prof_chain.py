"""Short driver for ncu: a few steps of the fused chain on one workload (no timing claims)."""
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
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--ring", type=int, default=2)
ap.add_argument("--events", action="store_true", help="print per-step CUDA-event times")
a = ap.parse_args()
w = synth.WORKLOADS[a.workload]
dev = torch.device("cuda:0")
mu = 0.0 if a.regime == "dense" else -5.0
ring = [[h.to(dev) for h in synth.make_heads(w["batch"], w["nc"], w["grids"], A=w["A"], g=i, mu_obj=mu)] for i in range(a.ring)]
pp = engine.PostProcessor(w["nc"], synth.workload_anchors(w), w["input_size"], w["conf"], kind=w["kind"], anchors_per_scale=w["A"])
res = pp(ring[0])
torch.cuda.synchronize()
evs = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps + 1)]
evs[0].record()
for i in range(a.steps):
    pp(ring[i % a.ring], out=res)
    evs[i + 1].record()
torch.cuda.synchronize()
if a.events:
    print("step ms:", ["%.3f" % evs[i].elapsed_time(evs[i + 1]) for i in range(a.steps)])
print("cand/img %.0f kept/img %.0f" % (res.cand_count.float().mean().item(), res.keep_count.float().mean().item()))
