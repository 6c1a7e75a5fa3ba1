"""CPU, world_size 2, gloo: the N>1 host logic (image sharding + variable-length all-gather)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import util


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path[:0] = [root, os.path.join(root, "pytorch-object-detection_b200"), os.path.join(root, "tests")]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from ypp.dist import gather_rows, shard_range
    from oracle import oracle as O
    g = util.load("map_cases")
    gt, det, nc = g["syn_gt"], g["syn_det"], int(g["syn_nc"])
    n_img = 48
    lo, hi = shard_range(n_img, rank, world)
    my_det = torch.tensor(det[(det[:, 0] >= lo) & (det[:, 0] < hi)])
    my_gt = torch.tensor(gt[(gt[:, 0] >= lo) & (gt[:, 0] < hi)])
    counts = []
    all_det = gather_rows(my_det, counts_out=counts)
    all_gt = gather_rows(my_gt)
    empty = gather_rows(torch.zeros((0, 7)))
    uneven = gather_rows(torch.full((rank * 3, 7), float(rank)))
    m = float(O.mean_average_precision(all_gt.numpy(), all_det.numpy(), nc))
    # "flags" mode: flag locally (the oracle stands in for ypp_map_flags), exchange (score, cls, tp) + per-class ground-truth
    # counts, AP of this rank's block of classes (numpy stands in for ypp_map_ap_from_flags), gather the per-class rows
    from ypp.dist import map_from_flags, class_range
    _, ltp, _ = O.mean_average_precision(my_gt.numpy(), my_det.numpy(), nc, 0.5, True)
    triples = torch.tensor(np.stack([my_det[:, 5].numpy(), my_det[:, 6].numpy(), (ltp == 1).astype(np.float32)], 1))
    gpc = torch.tensor(np.bincount(my_gt[:, 6].numpy().astype(np.int64), minlength=nc)[:nc])

    def ap_fn(recv, gpc_all, nc_, lo_, hi_, seg_count=None):
        return torch.tensor(util.ap_from_flags_numpy(recv.numpy(), gpc_all.numpy(), nc_, lo_, hi_, seg_count.numpy()))

    def mean_fn(pc):
        s_ = np.float32(0)
        for v in pc[:, 0].numpy():
            s_ = np.float32(s_ + v)
        return torch.tensor([np.float32(s_ / np.float32(pc.shape[0]))])

    fm, fpc = map_from_flags(triples, gpc, nc, ap_fn=ap_fn, mean_fn=mean_fn)
    q.put((rank, counts, all_det.numpy(), all_gt.numpy(), empty.shape[0], uneven.numpy(), m, float(fm[0]), fpc.numpy(),
           class_range(nc, rank, world)))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_range_partitions():
    from ypp.dist import shard_range
    for n, w in ((16384, 8), (10, 3), (2, 4), (0, 2)):
        r = [shard_range(n, k, w) for k in range(w)]
        assert r[0][0] == 0 and r[-1][1] == n
        assert all(r[k][1] == r[k + 1][0] for k in range(w - 1))
        sizes = [b - a for a, b in r]
        assert max(sizes) - min(sizes) <= 1
    assert shard_range(16384, 3, 8) == (6144, 8192)


def test_gather_rows_and_sharded_map_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=180) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    g = util.load("map_cases")
    det, gt = g["syn_det"], g["syn_gt"]
    for rank, counts, all_det, all_gt, n_empty, uneven, m, fm, fpc, crange in out:
        # flags mode: every rank ends with the single-process per-class table and mAP
        assert abs(fm - float(g["syn_map"])) <= 1e-6
        assert np.allclose(fpc[:, 0], g["syn_per_class"][:, 0], atol=1e-6) and np.array_equal(fpc[:, 1:], g["syn_per_class"][:, 1:])
        assert crange == ((0, 3), (3, 6))[rank]
        assert sum(counts) == len(det)
        # image-major fixture: rank-after-rank concatenation restores the original order
        assert np.array_equal(all_det, det) and np.array_equal(all_gt, gt)
        assert n_empty == 0
        assert uneven.shape == (3, 7) and np.all(uneven == 1.0)
        assert abs(m - float(g["syn_map"])) <= 1e-6          # gathered mAP == single-process mAP
