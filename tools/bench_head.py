"""F1 timing: fused 1x1 conv + decode + compaction (+ sort/NMS) against cuDNN conv2d + the unfused chain."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402
from ypp import engine, synth  # noqa: E402


def timed(fn, n=20, warm=5, per_graph=6):
    """fn() is captured per_graph times into one CUDA graph (no host gaps), replayed n times."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(per_graph):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (n * per_graph)


def main():
    dev = torch.device("cuda:0")
    B, nc, A = 64, 20, 3
    cins, grids = [256, 512, 1024], [52, 26, 13]
    AC = A * (5 + nc)
    g = torch.Generator("cpu").manual_seed(0)
    ring = []
    for r in range(3):                                              # 3 x 310 MB of features: larger than L2
        feats = [torch.randn(B, c, h, h, generator=g).to(dev).contiguous(memory_format=torch.channels_last)
                 for c, h in zip(cins, grids)]
        ring.append(feats)
    # head statistics of the synthetic workload (ypp/synth.py): xy 1.0, wh 0.5, objectness 1.5, classes 1.0
    row_std = torch.tensor(([1.0, 1.0, 0.5, 0.5, 1.5] + [1.0] * nc) * A).reshape(AC, 1, 1, 1)
    weights = [(torch.randn(AC, c, 1, 1, generator=g) * row_std / c ** 0.5).to(dev) for c in cins]
    wcl = [w.contiguous(memory_format=torch.channels_last) for w in weights]
    pp = engine.PostProcessor(nc, synth.YOLOV3_ANCHORS, 416, 0.25)
    res = pp.from_features(ring[0], weights)
    torch.cuda.synchronize()
    state = {"i": 0}

    def fused(stages):
        def f():
            state["i"] += 1
            pp.from_features(ring[state["i"] % 3], weights, stages=stages, out=res)
        return f

    hbuf = [[torch.empty((B, AC, h, h), device=dev).contiguous(memory_format=torch.channels_last) for h in grids],
            [torch.empty((B, AC, h, h), device=dev) for h in grids]]

    def cudnn_only():
        state["i"] += 1
        for x, w, o, oc in zip(ring[state["i"] % 3], wcl, hbuf[0], hbuf[1]):
            torch.ops.aten.convolution.out(x, w, None, [1, 1], [0, 0], [1, 1], False, [0, 0], 1, out=o)

    def cudnn_nchw():
        cudnn_only()
        for o, oc in zip(hbuf[0], hbuf[1]):
            oc.copy_(o)                                               # the NCHW head tensor D1 reads

    def unfused():
        cudnn_nchw()
        pp(hbuf[1], out=res)

    feat_bytes = sum(B * c * h * h * 4 for c, h in zip(cins, grids))
    t_f1 = timed(fused(("decode",)))
    t_all = timed(fused(("decode", "sort", "nms")))
    t_conv = timed(cudnn_only)
    t_conv2 = timed(cudnn_nchw)
    t_unf = timed(unfused)
    print("cand/img %.0f kept/img %.0f" % (res.cand_count.float().mean().item(), res.keep_count.float().mean().item()))
    print("F1 (3 launches: conv+decode+compact)  %.3f ms  = %.0f GB/s of feature bytes (%.0f MB)" % (t_f1, feat_bytes / t_f1 / 1e6, feat_bytes / 1e6))
    print("fused chain F1+S1+N1+G                %.3f ms  = %.0f images/s" % (t_all, B / t_all * 1e3))
    print("cuDNN conv2d x3 (TF32=%s, channels_last) %.3f ms" % (torch.backends.cudnn.allow_tf32, t_conv))
    print("  + channels_last -> NCHW head copy      %.3f ms" % t_conv2)
    print("cuDNN conv2d x3 + unfused chain       %.3f ms  = %.0f images/s" % (t_unf, B / t_unf * 1e3))


if __name__ == "__main__":
    main()
