"""Throughput of the chain with 1, 2 or 3 steps in flight (one stream + workspace each)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
import torch  # noqa: E402
from ypp import engine, synth  # noqa: E402

w = synth.WORKLOADS["yolov3_416_voc_b64"]
dev = torch.device("cuda:0")
ring = [[h.to(dev) for h in synth.make_heads(w["batch"], w["nc"], w["grids"], A=w["A"], g=i)] for i in range(12)]
for ns in (1, 2, 3, 4, 6):
    pps = [engine.PostProcessor(w["nc"], synth.workload_anchors(w), w["input_size"], w["conf"], kind=w["kind"],
                                anchors_per_scale=w["A"]) for k in range(ns)]
    streams = [torch.cuda.Stream() for _ in range(ns)]
    graphs = [pps[i % ns].graphed(ring[i]) for i in range(12)]
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = 120
    main = torch.cuda.current_stream()
    for rep in range(2):
        e0.record(main)
        for s in streams:
            s.wait_stream(main)
        for i in range(K):
            with torch.cuda.stream(streams[i % ns]):
                graphs[i % 12][0].replay()
        for s in streams:
            main.wait_stream(s)
        e1.record(main)
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    print("streams %d: %.4f ms/step  %.0f images/s" % (ns, ms, w["batch"] / ms * 1e3))
