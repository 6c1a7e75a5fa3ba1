"""Tuning build (-DYPP_SP_STATS): how the sweep-and-prune path of N1 fares on a workload (attempts, ways out,
pair tests, edges, rounds).  The phase clocks include the counters' own global atomics: read them as shares at most."""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
out = os.path.join(ROOT, "build_variants", "libypp_spstats.so")
if "YPP_LIB" not in os.environ:
    os.makedirs(os.path.dirname(out), exist_ok=True)
    from ypp import _lib
    _lib.build(defines=["YPP_SP_STATS"], out=out)
    os.environ["YPP_LIB"] = out
    sys.exit(subprocess.call([sys.executable] + sys.argv, env=os.environ))
import torch  # noqa: E402
from ypp import engine, synth, _lib  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "yolov3_416_voc_b64"
w = synth.WORKLOADS[wl]
dev = torch.device("cuda:0")
heads = [h.to(dev) for h in synth.make_heads(w["batch"], w["nc"], w["grids"], A=w["A"], g=0)]
pp = engine.PostProcessor(w["nc"], synth.workload_anchors(w), w["input_size"], w["conf"], kind=w["kind"], anchors_per_scale=w["A"])
L = _lib.lib()
buf = (ctypes.c_ulonglong * 16)()
pp(heads)
L.ypp_debug_sp_stats(buf, 1)
pp(heads)
L.ypp_debug_sp_stats(buf, 1)
names = ["attempts", "done", "not_ordinary", "edge_overflow", "too_deep", "pair_tests", "edges", "rounds",
         "clk_load", "clk_sort", "clk_sweep", "clk_resolve"]
v = list(buf)
print({n: v[i] for i, n in enumerate(names)})
if v[1]:
    print("per finished segment: tests %.0f edges %.1f rounds %.2f | cycles load %.0f sort %.0f sweep %.0f resolve %.0f"
          % (v[5] / v[0], v[6] / v[1], v[7] / v[1], v[8] / v[1], v[9] / v[1], v[10] / v[1], v[11] / v[1]))
