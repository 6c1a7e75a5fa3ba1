"""Chain and stage times for a given seed group / device (single process)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-object-detection_b200")]
import torch  # noqa: E402
from ypp import engine, synth  # noqa: E402

g0, devi = int(sys.argv[1]), int(sys.argv[2])
torch.cuda.set_device(devi)
dev = torch.device("cuda", devi)
w = synth.WORKLOADS["yolov3_416_voc_b64"]
pp = engine.PostProcessor(w["nc"], synth.workload_anchors(w), w["input_size"], w["conf"], kind=w["kind"], anchors_per_scale=w["A"])
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for g in range(g0, g0 + 4):
    hs = [h.to(dev) for h in synth.make_heads(w["batch"], w["nc"], w["grids"], A=w["A"], g=g, mu_obj=0.0)]
    r = pp(hs)
    torch.cuda.synchronize()
    out = []
    for stages in (("decode",), ("sort",), ("nms",)):
        pp.staged(hs, stages=("decode", "sort"))
        torch.cuda.synchronize()
        e0.record()
        for i in range(5):
            pp.staged(hs, stages=stages)
        e1.record()
        torch.cuda.synchronize()
        out.append("%s %.3f" % (stages[0], e0.elapsed_time(e1) / 5))
    print("dev", devi, "g", g, "cand", int(r.cand_count.sum()), "kept", int(r.keep_count.sum()), "maxcand", int(r.cand_count.max()), " ".join(out), flush=True)
    # peek into the workspace (layout of ypp_make_ws, csrc/ypp_common.cuh)
    import numpy as np
    from ypp._lib import lib
    B, N, nc = w["batch"], synth.boxes_per_image(w["grids"], w["A"]), w["nc"]
    ws = pp.workspace(lib().ypp_workspace_bytes(B, N, nc), dev).cpu().numpy()
    off = [0]
    def take(nbytes):
        o = off[0]
        off[0] += (nbytes + 255) // 256 * 256
        return o
    o_cand, o_keep, o_flag, o_step, o_q, o_nt = take(4 * B), take(4 * B), take(4 * B), take(4 * B), take(4 * B), take(4)
    T = (N // 256 + 4 * 8 + 8) * 8
    take(4 * T); take(4 * B * T); take(16 * B * T)
    per = min(N // 512 + 1, nc)
    o_bigq = take(4 * B * per)
    o_seg = take(4 * B * (nc + 1))
    flag = ws[o_flag:o_flag + 4 * B].view(np.int32)
    q = ws[o_q:o_q + 4 * B].view(np.int32).sum()
    seg = ws[o_seg:o_seg + 4 * B * (nc + 1)].view(np.int32).reshape(B, nc + 1)
    sizes = np.diff(seg, axis=1)
    print("   flags", np.unique(flag & 3, return_counts=True), "n_low/n_high/edges of flagged", [(f >> 8 & 255, f >> 16 & 255, f >> 24 & 255) for f in flag if f & 1], "long segments", q, "max seg", sizes.max(), "segs>512", (sizes > 512).sum(),
          "flagged imgs", np.nonzero(flag & 1)[0], "their seg sizes", sizes[(flag & 1) != 0])
