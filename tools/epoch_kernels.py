"""The mAP kernels of one epoch end on ONE GPU, sized like a 2-rank run (for an ncu launch list):
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/map_launches.csv python tools/epoch_kernels.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
import torch
sys.argv = [sys.argv[0]]
import bench
from ypp import engine, synth
from ypp.compat.yolov3_utils import MeanAveragePrecision
dev = torch.device("cuda:0")
w = synth.WORKLOADS["yolov3_416_voc_b64"]
nc, B = w["nc"], w["batch"]
pairs = [([h.to(dev) for h in hs], an.to(dev)) for hs, an in bench.epoch_data(synth, w, 0, 6)]
m = MeanAveragePrecision(nc, synth.workload_anchors(w), 416, 0.25)
for i in range(32):
    m.update_state(pairs[i % 6][1], pairs[i % 6][0])
gt, det = m.all_true_boxes_variable, m.all_pred_boxes_variable
torch.cuda.synchronize()
for rep in range(2):
    tr, gpc, _ = engine.map_flags(gt, det, nc)
    recv = torch.stack([tr, tr])                                  # what two ranks would have gathered
    cnt = torch.tensor([tr.shape[0]] * 2, dtype=torch.int64, device=dev)
    pc = engine.map_ap_from_flags(recv, gpc * 2, nc, 0, nc // 2, seg_count=cnt)
    engine.map_match(torch.cat([gt, gt]), torch.cat([det, det]), nc, want_tp=False)
    torch.cuda.synchronize()
print("done", det.shape, gt.shape)
