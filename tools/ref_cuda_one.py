"""The reference's own path for ONE image on the GPU (eager decode_predictions op list per scale + nms_v3 ->
torchvision.ops.batched_nms), for an ncu launch list: which kernels the shipped sm_100 build launches and how long each
takes.  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python tools/ref_cuda_one.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
import torch
import torchvision as tv
sys.argv = [sys.argv[0]]
import bench
from ypp import synth
dev = torch.device("cuda:0")
w = synth.WORKLOADS["yolov3_416_voc_b64"]
anchors = synth.workload_anchors(w)
heads = [h[:1].contiguous().to(dev) for h in synth.make_heads(2, w["nc"], w["grids"], g=4000, mu_obj=0.0)]


def one():
    rows = torch.cat([bench._ref_decode_eager(torch, h, w["nc"], anchors[3 * i: 3 * i + 3], 416) for i, h in enumerate(heads)], dim=1)
    return bench._ref_nms_v3_eager(torch, tv, rows[0], 0.25, 0.45)


for _ in range(3):
    k = one()
torch.cuda.synchronize()
torch.cuda.nvtx.range_push("measured")
k = one()
torch.cuda.synchronize()
torch.cuda.nvtx.range_pop()
print("kept", k.shape[0])
