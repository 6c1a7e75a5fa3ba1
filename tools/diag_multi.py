"""Per-rank diagnostics for the N>1 bench: host cores, chain time after short/long warm-up, H2D rate."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from ypp import engine, synth  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
print(rank, "cpus", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)), "OMP", os.environ.get("OMP_NUM_THREADS"), flush=True)
w = synth.WORKLOADS["yolov3_416_voc_b64"]
pp = engine.PostProcessor(w["nc"], synth.workload_anchors(w), w["input_size"], w["conf"], kind=w["kind"], anchors_per_scale=w["A"])
host = [synth.make_heads(w["batch"], w["nc"], w["grids"], A=w["A"], g=rank * 64 + i, mu_obj=0.0) for i in range(4)]
ring = [[h.to(dev) for h in hs] for hs in host]
graphs = [pp.graphed(hs) for hs in ring]
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def timed(tag, K=20):
    e0.record()
    for i in range(K):
        graphs[i % 4][0].replay()
    e1.record()
    torch.cuda.synchronize()
    print(rank, tag, "%.3f ms/step" % (e0.elapsed_time(e1) / K), flush=True)


for i in range(3):
    graphs[i % 4][0].replay()
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
timed("after 3 warm-up steps")
timed("again")
t = time.time()
while time.time() - t < 0.5:
    graphs[0][0].replay()
torch.cuda.synchronize()
timed("after 0.5 s warm-up")
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
timed("after barrier")
time.sleep(0.2)
timed("after 0.2 s idle")
pinned = [h.pin_memory() for h in host[0]]
dst = [torch.empty_like(h, device=dev) for h in host[0]]
nbytes = sum(h.numel() * 4 for h in host[0])
for tag in ("h2d", "h2d (all ranks together)"):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for i in range(20):
        for d, h in zip(dst, pinned):
            d.copy_(h, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    print(rank, tag, "%.1f GB/s" % (20 * nbytes / e0.elapsed_time(e1) / 1e6), flush=True)
if world > 1:
    dist.destroy_process_group()
