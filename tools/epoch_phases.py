"""Where the epoch end goes (flags mode), phase by phase with a synchronize + barrier between phases (so the sum is
larger than the real epoch end, but the split is readable).  Run under torchrun:
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/epoch_phases.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
import torch, torch.distributed as dist
sys.argv = [sys.argv[0]]
import bench
from ypp import engine, synth
from ypp.dist import ShardedMeanAveragePrecision, gather_padded, class_range, gather_rows

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    import datetime
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
w = synth.WORKLOADS["yolov3_416_voc_b64"]
anchors = synth.workload_anchors(w)
nc, B = w["nc"], w["batch"]
pairs = [([h.to(dev) for h in hs], an.to(dev)) for hs, an in bench.epoch_data(synth, w, rank, 6)]
m = ShardedMeanAveragePrecision(nc, anchors, 416, 0.25, world * 32 * B)
m.warm()


def sync():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def phase(name, fn, acc):
    sync()
    t0 = time.perf_counter()
    r = fn()
    torch.cuda.synchronize()
    acc.setdefault(name, []).append(1e3 * (time.perf_counter() - t0))
    return r


acc = {}
for rep in range(4):
    m.reset_states()
    for i in range(32):
        m.update_state(pairs[i % 6][1], pairs[i % 6][0])
    gt, det = phase("1 local_rows (2 cursor reads)", m.local_rows, acc)
    tr, gpc, _ = phase("2 map_flags (M1)", lambda: engine.map_flags(gt, det, nc), acc)
    recv, counts_t, counts, extras = phase("3 gather_padded triples+gt counts", lambda: gather_padded(tr, None, extra=gpc), acc)
    g = phase("4 extras.sum", lambda: extras.sum(0), acc)
    lo, hi = class_range(nc, rank, world)
    pc = phase("5 map_ap_from_flags (M2, my classes)", lambda: engine.map_ap_from_flags(recv, g, nc, lo, hi, seg_count=counts_t), acc)

    def gather_pc():
        if world == 1:
            return pc
        cmax = -(-nc // world)
        send = torch.zeros((cmax, 4), dtype=torch.float32, device=dev)
        send[: hi - lo] = pc[lo:hi]
        got = torch.empty((world, cmax, 4), dtype=torch.float32, device=dev)
        dist.all_gather_into_tensor(got.view(-1), send.view(-1))
        return torch.cat([got[r, : class_range(nc, r, world)[1] - class_range(nc, r, world)[0]] for r in range(world)], dim=0)
    pcs = phase("6 all-gather per-class + cat", gather_pc, acc)
    phase("7 map_mean + item", lambda: float(engine.map_mean(pcs)[0].item()), acc)
    phase("R rows mode: 2 gather_rows", lambda: (gather_rows(gt), gather_rows(det)), acc)
    ga, da = gather_rows(gt), gather_rows(det)
    phase("R rows mode: map_match", lambda: engine.map_match(ga, da, nc, want_tp=False), acc)
    sync()
    t0 = time.perf_counter()
    v = m.per_class("flags")[0].item()
    acc.setdefault("whole epoch end, flags (no syncs between)", []).append(1e3 * (time.perf_counter() - t0))
    sync()
    t0 = time.perf_counter()
    v = m.per_class("rows")[0].item()
    acc.setdefault("whole epoch end, rows (no syncs between)", []).append(1e3 * (time.perf_counter() - t0))
engine.launch_count(True)
m.per_class("flags")                       # collectives: every rank
n_flags = engine.launch_count(True)
m.per_class("rows")
n_rows = engine.launch_count(True)
if rank == 0:
    print("world %d, %d detections / %d ground truths per rank" % (world, det.shape[0], gt.shape[0]))
    for k, v in acc.items():
        print("%-48s %.3f ms (min of %d)" % (k, min(v), len(v)))
    print("kernel launches of ours: flags mode %d, rows mode %d" % (n_flags, n_rows))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
