"""Short driver for ncu: a few fused-head (F1) chains at the C2 shape (no timing claims)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
import torch  # noqa: E402
from ypp import engine, synth  # noqa: E402

dev = torch.device("cuda:0")
B, nc, A = 64, 20, 3
cins, grids = [256, 512, 1024], [52, 26, 13]
AC = A * (5 + nc)
g = torch.Generator("cpu").manual_seed(4242)
feats = [torch.randn(B, c, h, h, generator=g).to(dev).contiguous(memory_format=torch.channels_last) for c, h in zip(cins, grids)]
row_std = torch.tensor(([1.0, 1.0, 0.5, 0.5, 1.5] + [1.0] * nc) * A).reshape(AC, 1, 1, 1)
weights = [(torch.randn(AC, c, 1, 1, generator=g) * row_std / c ** 0.5).to(dev) for c in cins]
pp = engine.PostProcessor(nc, synth.YOLOV3_ANCHORS, 416, 0.25)
res = pp.from_features(feats, weights)
for _ in range(3):
    pp.from_features(feats, weights, out=res)
torch.cuda.synchronize()
print("cand/img %.0f kept/img %.0f" % (res.cand_count.float().mean().item(), res.keep_count.float().mean().item()))
