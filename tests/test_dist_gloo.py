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
    q.put((rank, counts, all_det.numpy(), all_gt.numpy(), empty.shape[0], uneven.numpy(), m))
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
    for rank, counts, all_det, all_gt, n_empty, uneven, m in out:
        assert sum(counts) == len(det)
        # image-major fixture: rank-after-rank concatenation restores the original order
        assert np.array_equal(all_det, det) and np.array_equal(all_gt, gt)
        assert n_empty == 0
        assert uneven.shape == (3, 7) and np.all(uneven == 1.0)
        assert abs(m - float(g["syn_map"])) <= 1e-6          # gathered mAP == single-process mAP
