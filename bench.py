#!/usr/bin/env python
"""bench.py -- post-processed images/sec (decode + NMS) of the YOLO post-processing hot path.

    python bench.py --gpus N --steps K --warmup W                 (ours; N>1 under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W  (the reference algorithm on the host cores)

A step = one pass of the hot path over one batch of synthetic heads (default workload:
YOLOv3-416 VOC, 3 scales, batch 64, dense regime: ~77% of the 10,647 boxes/image pass conf .25).
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "post-processed images/sec (decode+NMS)"
UNIT = "images/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="yolov3_416_voc_b64")
    ap.add_argument("--regime", default="dense", choices=["dense", "sparse"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the sparse-regime and mAP epoch extras")
    ap.add_argument("--no-ref-cuda", action="store_true", help="skip the reference-path-on-this-GPU comparison")
    ap.add_argument("--ring", type=int, default=6, help="distinct input batches cycled through (> L2 in total)")
    ap.add_argument("--in-flight", type=int, default=3,
                    help="steps in flight: each on its own stream with its own workspace and outputs (must divide --ring)")
    return ap.parse_args()


def workload_config(args, w, n_gpus):
    from ypp import synth
    return {
        "workload": args.workload, "regime": args.regime,
        "heads": ["[%d,%d,%d,%d]" % (w["batch"], w["A"] * (5 + w["nc"]), g, g) for g in w["grids"]],
        "batch_per_gpu": w["batch"], "global_batch": w["batch"] * n_gpus,
        "boxes_per_image": synth.boxes_per_image(w["grids"], w["A"]), "num_classes": w["nc"],
        "conf_threshold": w["conf"], "iou_threshold": 0.45,
        "parallelism": "image-sharded x%d, no data-path collective" % n_gpus,
        "steps_in_flight": max(1, args.in_flight),
        "l2_policy": "inputs larger than L2: ring of %d distinct batches (%.0f MB) cycled" % (
            args.ring, args.ring * w["batch"] * synth.head_bytes_per_image(w["nc"], w["grids"], w["A"]) / 1e6),
    }


# ------------------------------------------------------------------------------------ clocks

class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.idx = gpu_index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm_sorted = sorted(sm)
        return {"sm_mhz": sm_sorted[len(sm_sorted) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw)}


# ------------------------------------------------------------------------------------ reference arm

def oracle_step(O, heads_np, w, anchors):
    """decode_predictions (+cat) and per-image nms_v3 with torchvision-CPU semantics, all host threads."""
    if w["kind"] == "v2":
        rows = O.decode_predictions_v2(heads_np[0], w["nc"], anchors, w["input_size"])
    else:
        rows = O.decode_multi_scale(heads_np, w["nc"], anchors, w["input_size"], w["A"])
    keep, kc, cc = O.nms_v3_batch(rows, w["conf"], 0.45, offset_mode=O.OFFSET_AUTO_CPU, union_mode=O.UNION_SEPARATE)
    return int(kc.sum()), int(cc.sum())


def cpu_baseline(args, w, anchors, sample_images, budget_s=12.0, max_reps=3):
    """The oracle port of the reference algorithm timed on this box's host cores, bounded sample."""
    from oracle import oracle as O
    from ypp import synth
    O.build()
    O.set_num_threads(len(os.sched_getaffinity(0)))
    mu = 0.0 if args.regime == "dense" else -5.0
    heads = [h.numpy() for h in synth.make_heads(sample_images, w["nc"], w["grids"], A=w["A"], g=1000, mu_obj=mu)]
    oracle_step(O, [h[:2] for h in heads], w, anchors)        # warm the thread pool
    best, reps, t_all = None, 0, time.perf_counter()
    while reps < max_reps and (time.perf_counter() - t_all) < budget_s:
        t0 = time.perf_counter()
        oracle_step(O, heads, w, anchors)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
        reps += 1
    return {"value": sample_images / best, "unit": UNIT, "cores": O.num_threads(), "kind": "port",
            "sample": "%d images of %s (%s regime), oracle/ypp_oracle.c decode + per-image nms_v3 "
                      "(torchvision-CPU semantics), OpenMP over images, best of %d" % (
                          sample_images, args.workload, args.regime, reps)}


def run_reference(args):
    from oracle import oracle as O
    from ypp import synth
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    O.build()
    # every host thread this process may run on (torchrun exports OMP_NUM_THREADS=1 to its workers)
    cores = len(os.sched_getaffinity(0))
    O.set_num_threads(cores)
    w = synth.WORKLOADS[args.workload]
    anchors = synth.workload_anchors(w)
    sample = w["batch"]                         # one step = one batch of the workload, as on the GPU arm
    mu = 0.0 if args.regime == "dense" else -5.0
    ring = [[h.numpy() for h in synth.make_heads(sample, w["nc"], w["grids"], A=w["A"], g=2000 + i, mu_obj=mu)]
            for i in range(2)]
    for i in range(args.warmup):
        oracle_step(O, ring[i % 2], w, anchors)
    t0 = time.perf_counter()
    for i in range(args.steps):
        oracle_step(O, ring[i % 2], w, anchors)
    dt = time.perf_counter() - t0
    val = sample * args.steps / dt
    desc = ("%d images of %s (%s regime) per step; oracle/ypp_oracle.c (C restatement of decode_predictions + "
            "nms_v3/torchvision-CPU; the Python reference cannot travel to this box), OpenMP over images" % (
                sample, args.workload, args.regime))
    cfg = workload_config(args, w, args.gpus)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg, "gpu_launches": 0,
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": O.num_threads(), "kind": "port", "sample": desc},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------ our arm

def _bind_to_gpu_numa_node(gpu_index):
    """Run this rank (and first-touch its pinned host buffers) on the CPUs NVML reports as local to its GPU,
    so that with several ranks per box the H2D copies of the e2e path do not cross sockets.  Returns what
    happened, for the JSON line (affinity is an optimisation, but whether it took place explains the e2e curve)."""
    info = {"bound": False, "cpus_before": len(os.sched_getaffinity(0))}
    try:
        import pynvml
    except Exception as e:                          # nvidia-ml-py not importable
        info["why"] = "pynvml import failed: %s" % type(e).__name__
        return info
    try:
        pynvml.nvmlInit()
        try:                                        # NVML ignores CUDA_VISIBLE_DEVICES: go through the PCI address
            import torch
            pr = torch.cuda.get_device_properties(gpu_index)
            h = pynvml.nvmlDeviceGetHandleByPciBusId("%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id))
            info["handle"] = "pci"
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            info["handle"] = "index"
        n = (os.cpu_count() + 63) // 64
        words = pynvml.nvmlDeviceGetCpuAffinity(h, n)
        cpus = {64 * i + b for i, wd in enumerate(words) for b in range(64) if (wd >> b) & 1}
        info["gpu_local_cpus"] = len(cpus)
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            info["bound"] = True
            info["cpus_after"] = len(cpus)
        else:
            info["why"] = "no GPU-local CPU inside this process's affinity mask"
    except Exception as e:
        info["why"] = "%s: %s" % (type(e).__name__, str(e)[:80])
    return info



def run_ours(args):
    import torch
    import torch.distributed as dist
    from ypp import engine, synth
    from ypp import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = _bind_to_gpu_numa_node(local)
    if world > 1:
        import datetime
        # a rank that dies or disagrees must not cost ten minutes of eight GPUs: collectives give up after two
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
    if not os.path.exists(_lib.LIB_PATH):       # a snapshot taken before __graft_entry__.build(): compile, never fall back
        if local == 0:
            _lib.build()
    if world > 1:
        dist.barrier()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    w = synth.WORKLOADS[args.workload]
    anchors = synth.workload_anchors(w)
    B, nc, grids, A = w["batch"], w["nc"], w["grids"], w["A"]
    mu = 0.0 if args.regime == "dense" else -5.0
    pp = engine.PostProcessor(nc, anchors, w["input_size"], w["conf"], kind=w["kind"], anchors_per_scale=A)
    host_ring = [synth.make_heads(B, nc, grids, A=A, g=rank * 64 + i, mu_obj=mu) for i in range(args.ring)]
    ring = [[h.to(dev) for h in hs] for hs in host_ring]
    n_boxes = synth.boxes_per_image(grids, A)
    head_bytes = synth.head_bytes_per_image(nc, grids, A)
    K, W = args.steps, max(args.warmup, 3)

    # ---- value: K steps, inputs resident in HBM, CUDA events on the launching stream.
    # Each step replays the CUDA graph of the chain captured on that ring entry (PostProcessor.graphed):
    # the chain is 6 short launches, so eager launches would time the host, not the GPU.
    # NS steps are in flight at a time, each on its own stream with its own workspace and output buffers: S1 is one
    # CTA per image (64 of the 148 SMs at batch 64), so the neighbouring step's kernels fill the rest of the machine.
    NS = max(1, args.in_flight)
    if args.ring % NS != 0:
        raise SystemExit("bench.py: --in-flight must divide --ring")
    engine.launch_count(True)
    graphs = [pp.graphed(hs) for hs in ring]              # every captured chain owns its workspace and outputs
    launches_per_step = engine.launch_count(True) // (2 * len(ring))      # warm-up call + captured call
    res = graphs[0][1]
    streams = [torch.cuda.Stream(device=dev) for _ in range(NS)]
    main = torch.cuda.current_stream()

    def run_steps(n):
        """n steps, step i on stream i % NS (ring entry i % ring was captured with that stream's workspace)."""
        for st in streams:
            st.wait_stream(main)
        for i in range(n):
            with torch.cuda.stream(streams[i % NS]):
                graphs[i % args.ring][0].replay()
        for st in streams:
            main.wait_stream(st)

    run_steps(W)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed_windows(fn, est_ms, min_total_ms=120.0, lo=5, hi=400):
        """fn() = one window of exactly K steps.  Repeated until >= min_total_ms of device time are covered (a single
        window of 40 steps is 3 ms: its run-to-run noise is as large as the effects being reported); every window is
        bracketed by the barrier + synchronize and timed with CUDA events, max over ranks.  Returns the list of ms."""
        est_ms = max_over_ranks(est_ms)             # every rank must run the SAME number of windows (each holds a barrier)
        reps = int(min(hi, max(lo, -(-min_total_ms // max(est_ms, 1e-3)))))
        out = []
        for _ in range(reps):
            barrier()
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            out.append(e0.elapsed_time(e1))
        if world > 1:
            t = torch.tensor(out, dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            out = t.tolist()
        return out

    def spread(xs):
        xs = sorted(xs)
        return {"median": xs[len(xs) // 2], "min": xs[0], "max": xs[-1], "p10": xs[len(xs) // 10], "p90": xs[(9 * len(xs)) // 10],
                "windows": len(xs)}

    # one step at a time on one stream (the latency of a step), before the clocks are sampled
    def one_at_a_time():
        for i in range(K):
            graphs[i % args.ring][0].replay()

    one_at_a_time()
    torch.cuda.synchronize()
    e0.record(); one_at_a_time(); e1.record(); torch.cuda.synchronize()
    single_windows = timed_windows(one_at_a_time, e0.elapsed_time(e1))
    ms_single = spread(single_windows)["median"] / K
    sampler = ClockSampler(local) if rank == 0 else None
    e0.record(); run_steps(K); e1.record(); torch.cuda.synchronize()
    est = e0.elapsed_time(e1)
    if sampler:
        sampler.start()
    windows = timed_windows(lambda: run_steps(K), est)
    launches = launches_per_step * K
    win = spread(windows)
    ms_total = win["median"]
    # keep the GPU busy a little longer so that the 200 ms sampler sees clocks under load
    t_end = time.time() + 1.0
    while time.time() < t_end:
        graphs[0][0].replay()
    torch.cuda.synchronize()
    clocks = sampler.stop() if sampler else None
    value = world * B * K / (ms_total * 1e-3)
    cand_per_img = float(res.cand_count.float().mean().item())
    kept_per_img = float(res.keep_count.float().mean().item())

    # ---- roofline of the dominant HBM kernel: D1 (decode + threshold + compaction) alone:
    # one graph holding D1 on every ring entry, REPS times over; replayed and timed with CUDA events
    REPS = 8
    for hs in ring:
        pp.staged(hs, stages=("decode",))
    torch.cuda.synchronize()
    g_d1 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g_d1):
        for _ in range(REPS):
            for hs in ring:
                pp.staged(hs, stages=("decode",))
    for i in range(3):
        g_d1.replay()
    torch.cuda.synchronize()
    d1_list = []
    for rep in range(120):                          # 120 x 48 launches x ~19 us = ~110 ms of device time
        e0.record()
        g_d1.replay()
        e1.record()
        torch.cuda.synchronize()
        d1_list.append(e0.elapsed_time(e1) / (REPS * len(ring)))
    d1_sp = spread(d1_list)
    d1_ms = d1_sp["median"]
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json, burst copy)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    alg_bytes = B * (head_bytes + 28.0 * cand_per_img)
    achieved = alg_bytes / (d1_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "d1_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(args.workload + ":" + args.regime)
        except Exception:
            traffic = None
    dram_bytes = None
    if isinstance(traffic, dict):
        dram_bytes = traffic.get("dram_bytes_per_launch")
    elif isinstance(traffic, (int, float)):
        dram_bytes = float(traffic)
    roofline = {"kernel": "ypp_decode_compact_kernel (D1)", "bound": "hbm", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": dram_bytes, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "ms_per_launch": d1_ms, "ms_per_launch_spread": d1_sp,
                "frac_on_dram_bytes": (dram_bytes / (d1_ms * 1e-3) / 1e9 / peak) if dram_bytes else None,
                "traffic_source": "profiles/d1_traffic.json (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of one launch)"
                                  if dram_bytes else None,
                "bytes_per_image": "4*sum_s A(5+nc)HW + 28*n_cand = %d + 28*%.0f" % (head_bytes, cand_per_img),
                "how": "CUDA events around a graph of %d back-to-back D1 launches over the %d-batch ring (inputs > L2), median of 120 replays; "
                       "frac = algorithmic bytes / time / peak (SURVEY 8(d)); frac_on_dram_bytes = the bytes ncu saw reach HBM / the same time / peak "
                       "(most of the 28-byte record writes are still in L2 when the kernel ends)" % (REPS * len(ring), len(ring))}

    # ---- stage split: each stage captured 8 times into a CUDA graph on the same (decoded, sorted) state and replayed,
    # so the figures are device time without host gaps (repeating S1 or N1+G on the same state costs the same work)
    stage_ms = {}
    for name, stages in (("sort", ("sort",)), ("nms_gather", ("nms",))):
        g_st, _ = pp.graphed(ring[0], stages=stages, repeat=8)
        pp.staged(ring[0], stages=("decode", "sort"))
        g_st.replay()
        torch.cuda.synchronize()
        best = 1e9
        for rep in range(3):
            pp.staged(ring[0], stages=("decode", "sort"))
            torch.cuda.synchronize()
            e0.record()
            g_st.replay()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 8)
        stage_ms[name] = best
    stage_ms["decode_compact"] = d1_ms

    # ---- e2e: host (pinned) heads -> device, chain, counts + rows back to the host, every step
    pinned = [[h.pin_memory() for h in hs] for hs in host_ring[:2]]
    dbuf = [[torch.empty_like(h, device=dev) for h in hs] for hs in host_ring[:2]]
    outs = [pp(ring[0]), pp(ring[0])]
    host_cnt = torch.empty(B + 1, dtype=torch.int32).pin_memory()
    host_rows = torch.empty((B * n_boxes, 7), dtype=torch.float32).pin_memory()
    copy_stream = torch.cuda.Stream(device=dev)
    main = torch.cuda.current_stream()
    copied = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    d2h_bytes = 0

    def e2e_loop(n):
        nonlocal d2h_bytes
        d2h_bytes = 0
        for s in (0,):
            with torch.cuda.stream(copy_stream):
                for hd, hp in zip(dbuf[0], pinned[0]):
                    hd.copy_(hp, non_blocking=True)
                copied[0].record(copy_stream)
        for i in range(n):
            cur, nxt = i & 1, (i + 1) & 1
            if i + 1 < n:                           # prefetch the next batch while this one computes
                with torch.cuda.stream(copy_stream):
                    if i >= 1:
                        copy_stream.wait_event(consumed[nxt])
                    for hd, hp in zip(dbuf[nxt], pinned[nxt]):
                        hd.copy_(hp, non_blocking=True)
                    copied[nxt].record(copy_stream)
            main.wait_event(copied[cur])
            r = pp(dbuf[cur], out=outs[cur])
            consumed[cur].record(main)
            host_cnt[:B].copy_(r.keep_count, non_blocking=True)
            host_cnt[B:].copy_(r.det_total, non_blocking=True)
            main.synchronize()                      # the user needs the count to size the row read
            tot = int(host_cnt[B])
            host_rows[:tot].copy_(r.det_rows[:tot], non_blocking=True)
            main.synchronize()
            d2h_bytes += 4 * (B + 1) + 28 * tot

    e2e_loop(W)
    e2e_est = None
    e0.record(); e2e_loop(K); e1.record(); torch.cuda.synchronize()
    e2e_est = e0.elapsed_time(e1)
    e2e_windows = timed_windows(lambda: e2e_loop(K), e2e_est, min_total_ms=150.0, lo=3, hi=20)
    e2e_sp = spread(e2e_windows)
    e2e_ms = e2e_sp["median"]
    # the ceiling of that number: this rank's H2D copies alone (same pinned buffers, same copy stream), every rank at once
    def h2d_only(n):
        with torch.cuda.stream(copy_stream):
            for i in range(n):
                for hd, hp in zip(dbuf[i & 1], pinned[i & 1]):
                    hd.copy_(hp, non_blocking=True)
        main.wait_stream(copy_stream)
    h2d_only(2)
    torch.cuda.synchronize()
    barrier()
    e0.record(); h2d_only(8); e1.record(); torch.cuda.synchronize()
    h2d_ms = e0.elapsed_time(e1) / 8
    h2d_gbs = B * head_bytes / (h2d_ms * 1e-3) / 1e9
    per_rank = [{"rank": rank, "h2d_gb_per_s": h2d_gbs, "numa": numa}]
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, per_rank[0])
        per_rank = gathered
    e2e = {"value": world * B * K / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": B * head_bytes,
           "d2h_bytes_per_step": d2h_bytes // K, "ms_per_step": e2e_ms / K, "window_ms": e2e_sp,
           "h2d_only_ms_per_step": h2d_ms, "h2d_gb_per_s_this_rank": h2d_gbs,
           "h2d_bound_images_per_s": world * B / (max(r["h2d_gb_per_s"] and B * head_bytes / (r["h2d_gb_per_s"] * 1e9) for r in per_rank)),
           "per_rank": per_rank,
           "how": "pinned host heads -> H2D on a copy stream (double buffered) -> ypp_postprocess -> "
                  "D2H of keep counts + kept rows [K,7]; all inside the timed region; h2d_only = the same copies with nothing "
                  "else running, all ranks at once: the ceiling the host links put on e2e at this N"}

    extras = {}
    if not args.no_extras:
        extras = run_extras(args, torch, dist, engine, synth, dev, rank, world, w, anchors, pp, barrier, max_over_ranks,
                            ring_dense=ring[0], ring_all=ring)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(args, w, anchors, sample_images=min(B, 64))

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": workload_config(args, w, world),
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
                "cpu_baseline": cpu, "stage_ms": stage_ms,
                "timing": {"what": "the K-step window repeated; value / ms_per_step are the MEDIAN window (max over ranks per window)",
                           "window_ms": win, "total_timed_ms": sum(windows),
                           "value_spread": {"best": world * B * K / (win["min"] * 1e-3), "worst": world * B * K / (win["max"] * 1e-3)}},
                "single_stream": {"ms_per_step": ms_single, "value": world * B / (ms_single * 1e-3), "unit": UNIT,
                                  "window_ms": spread(single_windows),
                                  "what": "the same K steps one at a time on one stream: the latency of a step"},
                "per_image": {"candidates": cand_per_img, "kept": kept_per_img}}
        line.update(extras)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_extras(args, torch, dist, engine, synth, dev, rank, world, w, anchors, pp, barrier, max_over_ranks, ring_dense=None, ring_all=None):
    """Reported beside the headline: the sparse (deployment-like) regime, and the mAP epoch end
    (all-gather of detections over NCCL + GPU TP/FP matching) on sparse-regime detections."""
    B, nc, grids, A = w["batch"], w["nc"], w["grids"], w["A"]
    out = {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_ring = args.ring
    if rank == 0 and not args.no_ref_cuda:
        cases = [(args.workload, "dense"), (args.workload, "sparse")]
        if args.workload != "yolov4tiny_416_person_b256" and world == 1:
            cases.append(("yolov4tiny_416_person_b256", "dense"))          # one class: the O(n^2) stress of the mask kernel
        out["ref_cuda"] = [ref_cuda_compare(torch, engine, synth, dev, nm, rg) for nm, rg in cases]
    if world > 1:
        dist.barrier()
    sparse = [[h.to(dev) for h in synth.make_heads(B, nc, grids, A=A, g=500 + rank * 64 + i, mu_obj=-5.0)]
              for i in range(n_ring)]
    # same protocol as the headline: NS steps in flight, each stream with its own processor (workspace) and outputs
    NS = max(1, args.in_flight)
    sg = [pp.graphed(hs) for hs in sparse]
    res = sg[0][1]
    streams = [torch.cuda.Stream(device=dev) for _ in range(NS)]
    main = torch.cuda.current_stream()

    def run_sparse(n):
        for st in streams:
            st.wait_stream(main)
        for i in range(n):
            with torch.cuda.stream(streams[i % NS]):
                sg[i % n_ring][0].replay()
        for st in streams:
            main.wait_stream(st)

    run_sparse(3)
    torch.cuda.synchronize()
    e0.record()
    for i in range(args.steps):
        sg[i % n_ring][0].replay()
    e1.record()
    torch.cuda.synchronize()
    ms_one = e0.elapsed_time(e1) / args.steps
    barrier()
    e0.record()
    run_sparse(args.steps)
    e1.record()
    torch.cuda.synchronize()
    ms = max_over_ranks(e0.elapsed_time(e1))
    out["sparse_regime"] = {"value": world * B * args.steps / (ms * 1e-3), "unit": UNIT,
                            "candidates_per_image": float(res.cand_count.float().mean().item()),
                            "ms_per_step": ms / args.steps, "steps_in_flight": NS, "single_stream_ms_per_step": ms_one}
    out["map_epoch"] = map_epoch_extra(args, torch, dist, engine, synth, dev, rank, world, w, anchors, barrier, max_over_ranks)
    # the drop-in accumulator on DENSE batches (MeanAveragePrecision.update_state, heads resident): what a caller of
    # ypp.compat gets per batch, one batch at a time vs several in flight inside the product (PostProcessor.accumulate)
    if w["kind"] == "v3" and ring_all is not None:
        from ypp.compat.yolov3_utils import MeanAveragePrecision as MAPv3
        ann = [synth.make_annots(B, nc, g=900 + i).to(dev) for i in range(len(ring_all))]
        vl = {}
        for depth in (1, 2, 3, 4, None):
            mm = MAPv3(nc, anchors, w["input_size"], w["conf"], anchors_per_scale=A, pipeline_depth=depth)
            best = None
            for rep in range(4):                     # epochs of ONE metric object (module/yolov3_detector.py:43): from the
                mm.reset_states()                    # second on the accumulators are allocated at their final size at once
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                for i in range(96):
                    mm.update_state(ann[i % len(ann)], ring_all[i % len(ring_all)])
                n_rows = mm.all_pred_boxes_variable.shape[0]
                torch.cuda.synchronize()
                dt = (time.perf_counter() - t0) * 1e3 / 96
                best = dt if best is None else min(best, dt)
            vl["depth_%d_ms_per_batch" % depth if depth else "auto_ms_per_batch"] = best
            if depth is None:
                vl["auto_depth_chosen"] = int(mm._pp._auto_depth)
            vl["rows"] = int(n_rows)
            del mm
        vl["default_depth"] = "auto: 3; 2 once eight batches have reported fewer than 64 rows per image"
        vl["images_per_s_default"] = B / (vl["auto_ms_per_batch"] * 1e-3)
        vl["what"] = "96 dense batches per epoch through ypp.compat.yolov3_utils.MeanAveragePrecision.update_state, wall clock incl. the final row count read; best of 4 epochs"
        out["val_loop_dense"] = vl
    # detection writers: integer box columns of one dense step's kept rows
    rows = pp(ring_dense).rows()
    scale = torch.tensor([[500 / w["input_size"], 375 / w["input_size"]]] * B, dtype=torch.float64).float().to(dev)
    engine.detection_ints(rows, "coco", scale_xy=scale)
    torch.cuda.synchronize()
    e0.record()
    for i in range(10):
        engine.detection_ints(rows, "coco", scale_xy=scale)
    e1.record()
    torch.cuda.synchronize()
    ms_w = e0.elapsed_time(e1) / 10
    out["writers"] = {"rows": int(rows.shape[0]), "ms": ms_w, "rows_per_s": rows.shape[0] / (ms_w * 1e-3),
                      "gb_per_s": rows.shape[0] * (4 * rows.shape[1] + 16) / (ms_w * 1e-3) / 1e9,
                      "what": "ypp_detection_ints, COCO mode: 28 B row read + 16 B int4 write per detection"}
    # alternate semantics: the vendored Ultralytics NMS on a YOLOv5-640-shaped batch
    gen = torch.Generator(device="cpu").manual_seed(1234)
    ub, un, unc = 32, 25200, 80
    pred = torch.rand((ub, un, 5 + unc), generator=gen)
    pred[..., 0:2] *= 640
    pred[..., 2:4] = pred[..., 2:4] * 150 + 8
    pred[..., 4] = pred[..., 4] ** 8                     # ~16 % of the boxes pass obj > 0.25
    pred = pred.to(dev)
    r = engine.nms_ultralytics(pred)
    torch.cuda.synchronize()
    e0.record()
    for i in range(10):
        r = engine.nms_ultralytics(pred)
    e1.record()
    torch.cuda.synchronize()
    ms_u = e0.elapsed_time(e1) / 10
    out["ultralytics_nms"] = {"images_per_s": ub / (ms_u * 1e-3), "ms": ms_u, "batch": ub, "boxes_per_image": un, "classes": unc,
                              "candidates_per_image": float(r.cand_count.float().mean().item()),
                              "kept_per_image": float(r.keep_count.float().mean().item()),
                              "gb_per_s": pred.numel() * 4 / (ms_u * 1e-3) / 1e9,
                              "what": "ypp_nms_ultralytics, conf .25 / iou .45 / max_det 300, best class, input resident in HBM"}
    del pred
    if len(grids) == 3 and w["kind"] == "v3":
        # the reference's own calling shape: DecodeYoloV3.forward on ONE image (utils/yolov3_utils.py:73-90), the sized
        # [K,6] tensor returned to the caller (so each call ends with a host read of K); wall clock, heads resident
        from ypp.compat.yolov3_utils import DecodeYoloV3
        dec = DecodeYoloV3(nc, anchors, w["input_size"], w["conf"])
        one = [[h[:1].contiguous() for h in hs] for hs in sparse]
        for i in range(20):
            k = dec(one[i % n_ring]).shape[0]
        torch.cuda.synchronize()
        n_calls = 300
        t0 = time.perf_counter()
        for i in range(n_calls):
            k = dec(one[i % n_ring]).shape[0]
        torch.cuda.synchronize()
        ms_d = (time.perf_counter() - t0) * 1e3 / n_calls
        out["decode_module_b1"] = {"ms_per_call": ms_d, "images_per_s": 1e3 / ms_d, "kept_last": int(k),
                                   "what": "ypp.compat.yolov3_utils.DecodeYoloV3.forward, batch 1, sparse regime, result sized on the host"}
        out["fused_head"] = fused_head_extra(torch, engine, dev, w, anchors)
    return out


def fused_head_extra(torch, engine, dev, w, anchors, cins=(256, 512, 1024)):
    """the heads' 1x1 convolutions fused with decode+compaction (tcgen05 TF32 kernel F1) on the
    reference's channel counts (models/detector/yolov3.py:56,62,68), against cuDNN's convolution + the unfused chain.
    Features channels-last and resident, 3 distinct batches cycled (> L2), weights scaled to the workload's head statistics."""
    B, nc, grids, A = w["batch"], w["nc"], w["grids"], w["A"]
    AC = A * (5 + nc)
    g = torch.Generator("cpu").manual_seed(4242)
    ring = [[torch.randn(B, c, h, h, generator=g).to(dev).contiguous(memory_format=torch.channels_last)
             for c, h in zip(cins, grids)] for _ in range(3)]
    row_std = torch.tensor(([1.0, 1.0, 0.5, 0.5, 1.5] + [1.0] * nc) * A).reshape(AC, 1, 1, 1)
    weights = [(torch.randn(AC, c, 1, 1, generator=g) * row_std / c ** 0.5).to(dev) for c in cins]
    wcl = [x.contiguous(memory_format=torch.channels_last) for x in weights]
    pp = engine.PostProcessor(nc, anchors, w["input_size"], w["conf"], kind=w["kind"], anchors_per_scale=A)
    res = pp.from_features(ring[0], weights)
    hcl = [torch.empty((B, AC, h, h), device=dev).contiguous(memory_format=torch.channels_last) for h in grids]
    hnchw = [torch.empty((B, AC, h, h), device=dev) for h in grids]
    state = {"i": 0}

    def fused(stages):
        def f():
            state["i"] += 1
            pp.from_features(ring[state["i"] % 3], weights, stages=stages, out=res)
        return f

    def cudnn():
        state["i"] += 1
        for x, ww, o, oc in zip(ring[state["i"] % 3], wcl, hcl, hnchw):
            torch.ops.aten.convolution.out(x, ww, None, [1, 1], [0, 0], [1, 1], False, [0, 0], 1, out=o)
            oc.copy_(o)                      # the contiguous NCHW head the reference's decode (and D1) reads

    def unfused():
        cudnn()
        pp(hnchw, out=res)

    def timed(fn, per_graph=6, n=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for _ in range(per_graph):
                fn()
        gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            gr.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / (n * per_graph)

    t_f1, t_chain = timed(fused(("decode",))), timed(fused(("decode", "sort", "nms")))
    t_conv, t_unf = timed(cudnn), timed(unfused)
    feat_bytes = sum(B * c * h * h * 4 for c, h in zip(cins, grids))
    cand = float(res.cand_count.float().mean().item())
    alg = feat_bytes + sum(AC * c * 4 for c in cins) + 28.0 * cand * B
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak = float(json.load(open(peaks_path))["hbm_gbs"]) if os.path.exists(peaks_path) else 6650.0
    return {"kernel": "ypp_head_decode_kernel (F1: TMA + tcgen05 kind::tf32 + TMEM epilogue)", "cin": list(cins),
            "f1_ms": t_f1, "f1_gb_per_s": alg / (t_f1 * 1e-3) / 1e9, "f1_frac_of_hbm_peak": alg / (t_f1 * 1e-3) / 1e9 / peak,
            "algorithmic_bytes": alg, "fused_chain_ms": t_chain, "fused_chain_images_per_s": B / (t_chain * 1e-3),
            "cudnn_conv_plus_nchw_copy_ms": t_conv, "cudnn_plus_unfused_chain_ms": t_unf,
            "cudnn_plus_unfused_images_per_s": B / (t_unf * 1e-3), "candidates_per_image": cand,
            "allow_tf32": bool(torch.backends.cudnn.allow_tf32),
            "what": "input of the heads' 1x1 convs [B,H,W,Cin] fp32 resident in HBM -> kept rows; CUDA-graph timed"}


# ------------------------------------------------------------------------------------ C5: the evaluation epoch

def epoch_data(synth, w, rank, n_distinct):
    """n_distinct (heads, annots) pairs of rank `rank`: sparse-regime heads with detections derived from that batch's
    ground truth planted in (synth.plant_annots), so the epoch's mAP is of order 0.7 and says something."""
    B, nc, grids, A = w["batch"], w["nc"], w["grids"], w["A"]
    anchors = synth.workload_anchors(w)
    pairs = []
    for i in range(n_distinct):
        g = 7000 + rank * 64 + i
        heads = synth.make_heads(B, nc, grids, A=A, g=g, mu_obj=-5.0)
        an = synth.make_annots(B, nc, g=g)
        if w["kind"] == "v3":
            synth.plant_annots(heads, an, nc, anchors, w["input_size"], A=A, g=g)
        pairs.append((heads, an))
    return pairs


def map_epoch_extra(args, torch, dist, engine, synth, dev, rank, world, w, anchors, barrier, max_over_ranks):
    """BASELINE config 5: every rank accumulates n_batches x B images through the drop-in accumulator
    (ShardedMeanAveragePrecision.update_state = MeanAveragePrecision.update_state with global image ids; batches in
    flight inside the product), then the epoch end in both exchange modes:
      rows   all-gather of [K,7] detections + [G,7] ground truths (NCCL), whole matching on every rank (north star);
      flags  local TP flags, all-gather of (score, cls, tp) + per-class ground-truth counts, class-sharded AP.
    Wall clock between barriers (the host drives the loop), max over ranks; second epoch reported (the first warms
    NCCL and the allocator).  Rank 0 then recomputes everything in ONE process over all ranks' images and compares."""
    from ypp.dist import ShardedMeanAveragePrecision
    from ypp.compat.yolov3_utils import MeanAveragePrecision as MAPv3
    from ypp.compat.yolov2_utils import MeanAveragePrecision as MAPv2
    B, nc = w["batch"], w["nc"]
    n_batches, n_distinct = 32, min(args.ring, 6)
    n_total = world * n_batches * B
    pairs = [([h.to(dev) for h in hs], an.to(dev)) for hs, an in epoch_data(synth, w, rank, n_distinct)]
    m = ShardedMeanAveragePrecision(nc, anchors, w["input_size"], w["conf"], n_total, kind=w["kind"],
                                    anchors_per_scale=w["A"])
    m.warm()

    def feed(metric):
        for i in range(n_batches):
            hs, an = pairs[i % n_distinct]
            metric.update_state(an, hs if w["kind"] == "v3" else hs[0])

    def epoch(mode):
        """-> (ms accumulate, ms epoch end, mAP, per_class) with the accumulation queued behind nothing."""
        m.reset_states()
        barrier()
        t0 = time.perf_counter()
        feed(m)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        mp, pc = m.per_class(mode)
        val = float(mp[0].item())                 # the read of the result ends the epoch
        t2 = time.perf_counter()
        return 1e3 * (t1 - t0), 1e3 * (t2 - t1), val, pc

    res = {}
    for mode in ("flags", "rows"):
        epoch(mode)
        acc_ms, end_ms, val, pc = epoch(mode)
        # the epoch is ONE interval per rank (barrier -> result read), max over ranks; the two parts are maxed separately
        # for information only (a rank that finishes accumulating early waits inside the first collective: its "end" is
        # longer by the skew, so max(accumulate) + max(end) counts the skew twice)
        tot_ms = max_over_ranks(acc_ms + end_ms)
        acc_ms, end_ms = max_over_ranks(acc_ms), max_over_ranks(end_ms)
        res[mode] = {"accumulate_ms": acc_ms, "epoch_end_ms": end_ms, "epoch_ms": tot_ms, "mAP": val,
                     "images_per_s": n_total / (tot_ms * 1e-3), "per_class": pc}
    gt_l, det_l = m.local_rows()
    # the same epoch on one GPU, in this run: this rank's own images, no exchange
    Single = MAPv3 if w["kind"] == "v3" else MAPv2
    one = Single(nc, anchors, w["input_size"], w["conf"], anchors_per_scale=w["A"]) if w["kind"] == "v3" else \
        Single(nc, anchors, w["input_size"], w["conf"])
    one_ms = None
    for _ in range(2):
        one.reset_states()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        feed(one)
        float(one.result().item())
        one_ms = 1e3 * (time.perf_counter() - t0)
    one_ms = max_over_ranks(one_ms)
    out = {"images": n_total, "images_per_rank": n_batches * B, "detections_per_rank": int(det_l.shape[0]),
           "ground_truths_per_rank": int(gt_l.shape[0]),
           "flags": {k: v for k, v in res["flags"].items() if k != "per_class"},
           "rows": {k: v for k, v in res["rows"].items() if k != "per_class"},
           "one_gpu_epoch_ms": one_ms, "one_gpu_images_per_s": n_batches * B / (one_ms * 1e-3),
           "epoch_efficiency_flags": one_ms / res["flags"]["epoch_ms"], "epoch_efficiency_rows": one_ms / res["rows"]["epoch_ms"],
           "mAP": res["flags"]["mAP"],
           "what": "n_batches=%d x B=%d images per rank; detections derived from the ground truth; epoch = update_state loop + result(), "
                   "wall clock between barriers, max over ranks; efficiency = the same epoch on one GPU (this run) / the N-GPU epoch" % (n_batches, B)}
    # ---- hardware parity: ONE process over every rank's images (rank 0 regenerates the other ranks' data)
    same_modes = bool(torch.equal(res["flags"]["per_class"], res["rows"]["per_class"])) and res["flags"]["mAP"] == res["rows"]["mAP"]
    gt_all, det_all = m.gathered()
    if rank == 0:
        ref = Single(nc, anchors, w["input_size"], w["conf"], anchors_per_scale=w["A"]) if w["kind"] == "v3" else \
            Single(nc, anchors, w["input_size"], w["conf"])
        ref.reset_states()
        for r in range(world):
            pr = pairs if r == 0 else [([h.to(dev) for h in hs], an.to(dev)) for hs, an in epoch_data(synth, w, r, n_distinct)]
            for i in range(n_batches):
                hs, an = pr[i % n_distinct]
                ref.update_state(an, hs if w["kind"] == "v3" else hs[0])
            del pr
        r_gt, r_det = ref.all_true_boxes_variable, ref.all_pred_boxes_variable
        r_map, r_pc, r_tp = engine.map_match(r_gt, r_det, nc, want_tp=True)
        _, _, g_tp = engine.map_match(gt_all, det_all, nc, want_tp=True)
        checks = {"gathered_detections_equal": bool(det_all.shape == r_det.shape and torch.equal(det_all, r_det)),
                  "gathered_ground_truths_equal": bool(gt_all.shape == r_gt.shape and torch.equal(gt_all, r_gt)),
                  "tp_flags_equal": bool(g_tp.shape == r_tp.shape and torch.equal(g_tp, r_tp)),
                  "per_class_rows_mode_equal": bool(torch.equal(res["rows"]["per_class"], r_pc)),
                  "per_class_flags_mode_equal": bool(torch.equal(res["flags"]["per_class"], r_pc)),
                  "mAP_abs_diff": abs(res["flags"]["mAP"] - float(r_map[0].item())), "modes_agree": same_modes}
        ok = all(v for k, v in checks.items() if k != "mAP_abs_diff") and checks["mAP_abs_diff"] <= 1e-6
        out["parity"] = "ok" if ok else "FAILED"
        out["parity_checks"] = checks
        out["single_process_mAP"] = float(r_map[0].item())
        out["true_positives"] = int((r_tp == 1).sum().item())
    barrier()
    return out


# ------------------------------------------------------------------------------------ the shipped sm_100 kernels to beat

def _ref_decode_eager(torch, inp, num_classes, anchors, input_size):
    """decode_predictions of the reference (utils/yolov3_utils.py:27-59) as the same sequence of eager torch ops on
    the heads' own device: view -> permute -> contiguous copy, materialised anchor / grid tensors, sigmoid / exp,
    *= stride, argmax of the class sigmoids, cat.  (The reference itself cannot travel to this box; this is its op
    list, so the launches and the HBM traffic are the reference's.)"""
    B, _, H, W = inp.shape
    A = len(anchors)
    sh, sw = input_size / H, input_size / W
    dev = inp.device
    pred = inp.view(B, A, -1, H, W).permute(0, 1, 3, 4, 2).contiguous()
    stride = torch.tensor([sw, sh] * 2, dtype=torch.float32, device=dev)
    sa = torch.tensor([[aw / sw, ah / sh] for aw, ah in anchors], dtype=torch.float32, device=dev).unsqueeze(0)
    sa = sa.repeat(B, 1, 1).repeat(1, 1, H * W).view(B, A, H, W, 2)
    gx = torch.arange(0, W, device=dev).repeat(H, 1).repeat(B * A, 1, 1).view(B, A, H, W, 1).float()
    gy = torch.arange(0, W, device=dev).repeat(H, 1).t().repeat(B * A, 1, 1).view(B, A, H, W, 1).float()
    pxy = torch.sigmoid(pred[..., 0:2]) + torch.cat([gx, gy], dim=-1)
    pwh = torch.exp(pred[..., 2:4]) * sa
    pbox = torch.cat([pxy, pwh], dim=-1).view(B, -1, 4)
    pbox *= stride
    pconf = torch.sigmoid(pred[..., 4:5]).view(B, -1, 1)
    pcls = torch.argmax(torch.sigmoid(pred[..., 5:]).view(B, -1, num_classes), dim=-1, keepdim=True)
    return torch.cat((pbox, pconf, pcls), -1)


def _ref_nms_v3_eager(torch, tv, boxes, conf, iou):
    """nms_v3 (utils/yolo_utils.py:334-348) on one image's rows, on their device."""
    boxes = boxes[torch.where(boxes[..., 4] > conf)[0]]
    if boxes.shape[0] == 0:
        return boxes
    x1 = (boxes[..., 0] - (boxes[..., 2] / 2)).unsqueeze(dim=-1)
    y1 = (boxes[..., 1] - (boxes[..., 3] / 2)).unsqueeze(dim=-1)
    x2 = (boxes[..., 0] + (boxes[..., 2] / 2)).unsqueeze(dim=-1)
    y2 = (boxes[..., 1] + (boxes[..., 3] / 2)).unsqueeze(dim=-1)
    keep = tv.ops.batched_nms(torch.cat([x1, y1, x2, y2], dim=-1), boxes[..., 4], boxes[..., 5], iou)
    return boxes[keep]


def ref_cuda_compare(torch, engine, synth, dev, name, regime, reps=7):
    """SURVEY 2a K6/K7, BASELINE.md 4: the reference's own path on THIS GPU - eager decode + per-image nms_v3 ->
    torchvision.ops.batched_nms (its shipped sm_100 nms_kernel_impl<float> + gather_keep_from_mask) - against our chain
    on the same resident heads; and batched_nms ALONE on our compacted candidates (candidate selection and the xyxy
    conversion done beforehand, outside the timed region) against our S1 + N1 + N1b + G on the same candidates.
    CUDA events, 3 warm-ups, median of `reps`; the per-image Python loop and its host syncs are part of the reference."""
    try:
        import torchvision as tv
    except Exception as e:
        return {"unavailable": "torchvision: %s" % type(e).__name__}
    w = synth.WORKLOADS[name]
    anchors = synth.workload_anchors(w)
    B, nc, grids, A = w["batch"], w["nc"], w["grids"], w["A"]
    heads = [h.to(dev) for h in synth.make_heads(B, nc, grids, A=A, g=4000, mu_obj=0.0 if regime == "dense" else -5.0)]
    pp = engine.PostProcessor(nc, anchors, w["input_size"], w["conf"], kind=w["kind"], anchors_per_scale=A)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def med(fn, n=reps, warm=3):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(n):
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        return ts[len(ts) // 2]

    kept_ref = [0]

    def ref_chain():
        if w["kind"] == "v2":
            rows = _ref_decode_eager(torch, heads[0], nc, anchors, w["input_size"])
        else:
            rows = torch.cat([_ref_decode_eager(torch, h, nc, anchors[A * i: A * i + A], w["input_size"]) for i, h in enumerate(heads)], dim=1)
        k = 0
        for b in range(B):
            k += _ref_nms_v3_eager(torch, tv, rows[b], w["conf"], 0.45).shape[0]
        kept_ref[0] = k
        return rows

    def ref_decode_only():
        if w["kind"] == "v2":
            return _ref_decode_eager(torch, heads[0], nc, anchors, w["input_size"])
        return torch.cat([_ref_decode_eager(torch, h, nc, anchors[A * i: A * i + A], w["input_size"]) for i, h in enumerate(heads)], dim=1)

    g_all, res = pp.graphed(heads)
    ours_ms = med(g_all.replay, n=25)
    ours_eager_ms = med(lambda: pp(heads, out=res), n=25)
    ref_ms = med(ref_chain)
    ref_dec_ms = med(ref_decode_only)
    kept_ours = int(res.keep_count.sum().item())
    # batched_nms alone on OUR candidates
    rows = pp.decode(heads)
    cands = []
    for b in range(B):
        c = rows[b][rows[b][:, 4] > w["conf"]]
        xyxy = torch.stack([c[:, 0] - c[:, 2] / 2, c[:, 1] - c[:, 3] / 2, c[:, 0] + c[:, 2] / 2, c[:, 1] + c[:, 3] / 2], 1).contiguous()
        cands.append((xyxy, c[:, 4].contiguous(), c[:, 5].contiguous()))

    def tv_nms_only():
        for xyxy, sc, cl in cands:
            if xyxy.shape[0]:
                tv.ops.batched_nms(xyxy, sc, cl, 0.45)

    tv_ms = med(tv_nms_only)
    pp.staged(heads, stages=("decode",), out=res, ws=res.ws)
    g_sn = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(g_sn):
        pp.staged(heads, stages=("sort", "nms"), out=res, ws=res.ws)
    ours_sn_ms = med(g_sn.replay, n=25)
    n_cand = float(res.cand_count.float().mean().item())
    return {"workload": name, "regime": regime, "batch": B, "candidates_per_image": n_cand,
            "kept_total_ours": kept_ours, "kept_total_reference_path": kept_ref[0],
            "ours_chain_ms": ours_ms, "ours_chain_eager_launch_ms": ours_eager_ms,
            "ref_cuda_chain_ms": ref_ms, "ref_cuda_decode_ms": ref_dec_ms,
            "ours_images_per_s": B / (ours_ms * 1e-3), "ref_cuda_images_per_s": B / (ref_ms * 1e-3),
            "chain_speedup": ref_ms / ours_ms,
            "tv_batched_nms_only_ms": tv_ms, "ours_sort_nms_rows_ms": ours_sn_ms, "nms_speedup": tv_ms / ours_sn_ms,
            "what": "ref_cuda = eager decode_predictions + per-image nms_v3 -> torchvision.ops.batched_nms on this GPU (heads resident); "
                    "tv_batched_nms_only = %d batched_nms calls on our compacted candidates; ours = CUDA-graph replay of the same work" % B}


if __name__ == "__main__":
    a = parse_args()
    sys.exit(run_reference(a) if a.impl == "reference" else run_ours(a))
