"""Phase timing inside S1 (CTA 0): build a -DYPP_PROFILE_S1 library, run the chain, read clock64 marks."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
os.environ["YPP_LIB"] = os.path.join(ROOT, "build_variants", "libypp_prof.so")
import torch
from ypp import engine, synth, _lib
wname = sys.argv[1] if len(sys.argv) > 1 else "yolov3_416_voc_b64"
regime = sys.argv[2] if len(sys.argv) > 2 else "dense"
w = synth.WORKLOADS[wname]
dev = torch.device("cuda:0")
heads = [h.to(dev) for h in synth.make_heads(w["batch"], w["nc"], w["grids"], A=w["A"], g=0, mu_obj=0.0 if regime == "dense" else -5.0)]
pp = engine.PostProcessor(w["nc"], synth.workload_anchors(w), w["input_size"], w["conf"], kind=w["kind"], anchors_per_scale=w["A"])
buf = (ctypes.c_longlong * 16)()
L = _lib.lib()
L.ypp_debug_s1_clocks.argtypes = [ctypes.c_void_p]
for _ in range(3):
    pp(heads)
torch.cuda.synchronize()
L.ypp_debug_s1_clocks(buf)          # reads and resets the marks
pp(heads)
torch.cuda.synchronize()
rc = L.ypp_debug_s1_clocks(buf)
v = list(buf)
names = ["tile scan", "tile gather", "detector", "radix sort", "skeys copy", "class count+seg", "rank", "flag check+out", "-", "-"]
print("rc", rc, "total cycles", v[8] - v[0])
for i in range(8):
    print("%-12s %8d cycles  %6.2f us @1.965GHz" % (names[i], v[i + 1] - v[i], (v[i + 1] - v[i]) / 1965.0))
print("longest CTA %.2f us; first CTA start -> last CTA end %.2f us" % (v[9] / 1965.0, (v[11] - v[10]) / 1e3))
if (v[7] - v[3]) < 20000 and v[12] == 0:
    print('(small-image path: the rows radix sort .. rank above are its rank / pairs / fix-point / outputs phases; marks cost ~1 us each at this scale)')
elif v[12]:
    print("detector split: pass 1 %.2f, bin scan %.2f, pass 2 %.2f, pair tests %.2f us" % (
        (v[12] - v[2]) / 1965.0, (v[13] - v[12]) / 1965.0, (v[14] - v[13]) / 1965.0, (v[3] - v[14]) / 1965.0))
