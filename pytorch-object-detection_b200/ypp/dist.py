"""Multi-GPU plumbing: one process per GPU, images sharded in contiguous blocks, and the single
exchange step of the path at the end of an evaluation epoch.  decode -> sort -> NMS needs no
communication.  Two ways through the epoch end, same result bit for bit:

  "rows"   the north star's plain form: variable-length all-gather of the [K_r,7] detection and [G_r,7] ground-truth
           rows, then the whole matching (ypp_map_match) on every rank over the gathered set;
  "flags"  the matching of utils/yolo_utils.py:404-430 never looks across images, and an image lives on one rank:
           every rank flags ITS detections against ITS ground truths (ypp_map_flags), the ranks all-gather only
           (score, cls, tp) triples - 12 bytes per detection, no ground-truth rows, just per-class counts riding on the
           size exchange -, rank r computes the AP of its own contiguous block of classes (ypp_map_ap_from_flags) and
           a last all-gather collects the nc per-class results.  Work per rank shrinks with the number of ranks
           instead of growing with the gathered set.

The reference never gathers (its metric state is rank-local, module/yolov3_detector.py:45-47); parity for this
module is judged against the single-process result over all images.  Works with backend "nccl" (CUDA tensors,
NVLink/NVSwitch) and "gloo" (CPU tensors: the tests run the plumbing there with CPU stand-ins for the
kernels).
"""
from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of rank `rank`; the first n_items % world ranks get one extra."""
    base, extra = divmod(int(n_items), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def class_range(num_classes: int, rank: int, world: int) -> Tuple[int, int]:
    """The classes whose AP rank `rank` computes in "flags" mode: the same contiguous split as the images."""
    return shard_range(num_classes, rank, world)


def _world(group):
    if not (dist.is_available() and dist.is_initialized()):
        return 1, 0
    return dist.get_world_size(group), dist.get_rank(group)


def gather_padded(rows: torch.Tensor, group: Optional[dist.ProcessGroup] = None, extra: Optional[torch.Tensor] = None):
    """All-gather rows [k_r, C] of every rank WITHOUT repacking: returns (recv [world, kmax, C] of which the first
    counts[r] rows of segment r are valid, counts int64 [world] on the rows' device, the same counts as a list, and
    extras [world, E]: `extra` (int64 [E], e.g. per-class ground-truth counts) of every rank - it rides on the size
    exchange, so the whole thing is two collectives and one host sync (the sizes NCCL needs)."""
    world, _ = _world(group)
    rows = rows.contiguous()
    dev, C = rows.device, rows.shape[1]
    E = 0 if extra is None else int(extra.numel())
    meta = torch.empty(1 + E, dtype=torch.int64, device=dev)
    meta[0] = rows.shape[0]
    if E:
        meta[1:] = extra.to(device=dev, dtype=torch.int64).reshape(-1)
    if world == 1:
        return rows.unsqueeze(0), meta[:1], [rows.shape[0]], meta[1:].unsqueeze(0)
    all_meta = torch.empty((world, 1 + E), dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(all_meta.view(-1), meta, group=group)
    counts_t = all_meta[:, 0].contiguous()
    counts = counts_t.tolist()                      # the one host sync
    kmax = max(counts)
    recv = torch.empty((world, kmax, C), dtype=rows.dtype, device=dev)
    if kmax:
        send = rows
        if rows.shape[0] < kmax:
            send = torch.empty((kmax, C), dtype=rows.dtype, device=dev)
            send[: rows.shape[0]] = rows
        dist.all_gather_into_tensor(recv.view(world * kmax, C), send, group=group)
    return recv, counts_t, counts, all_meta[:, 1:]


def gather_rows(rows: torch.Tensor, group: Optional[dist.ProcessGroup] = None,
                counts_out: Optional[list] = None) -> torch.Tensor:
    """All-gather rows [k_r, C] of every rank into [sum k_r, C], rank after rank.

    Two collectives: the int64 counts, then the payload padded to the largest count (NCCL
    all-gather needs equal sizes; NVSwitch makes the padding cheap).  One host sync (the counts).
    """
    world, _ = _world(group)
    if world == 1:
        if counts_out is not None:
            counts_out[:] = [rows.shape[0]]
        return rows
    recv, _, counts, _ = gather_padded(rows, group)
    if counts_out is not None:
        counts_out[:] = counts
    kmax = recv.shape[1]
    if kmax == 0:
        return rows[:0]
    if all(c == kmax for c in counts):
        return recv.view(world * kmax, rows.shape[1])
    return torch.cat([recv[r, : counts[r]] for r in range(world)], dim=0)


def map_from_flags(triples: torch.Tensor, gt_per_class: torch.Tensor, num_classes: int,
                   group: Optional[dist.ProcessGroup] = None, ap_fn: Optional[Callable] = None,
                   mean_fn: Optional[Callable] = None):
    """The exchange + M2 of "flags" mode.  triples [K_r,3] (score, cls, tp) of THIS rank's detections in image order,
    gt_per_class int64 [nc] of THIS rank's ground truths.  Returns (mAP [1], per_class [nc,4]) of the whole set, the
    same on every rank.  ap_fn / mean_fn default to the CUDA kernels (the gloo tests pass CPU stand-ins)."""
    if ap_fn is None:
        from . import engine
        ap_fn, mean_fn = engine.map_ap_from_flags, engine.map_mean
    world, rank = _world(group)
    nc = int(num_classes)
    recv, counts_t, _, extras = gather_padded(triples, group, extra=gt_per_class)
    gpc = extras.sum(0)
    lo, hi = class_range(nc, rank, world)
    pc = ap_fn(recv, gpc, nc, lo, hi, seg_count=counts_t)                    # [nc,4]: rows lo..hi-1 filled, the rest zero
    if world > 1:
        # every row has exactly one owner and is zero everywhere else: the sum IS the gather (x + 0 + ... + 0 == x in
        # fp32, every value here is >= 0), in one collective and without repacking
        dist.all_reduce(pc, op=dist.ReduceOp.SUM, group=group)
    return mean_fn(pc), pc


class ShardedMeanAveragePrecision:
    """MeanAveragePrecision over an image-sharded evaluation set.

    Each rank feeds its own images to update_state (global image ids = shard offset + local
    index, stored in the fp32 img column like the reference, exact below 2**24); result() returns the mAP of the
    WHOLE set on every rank.  mode "flags" (default) or "rows": see the module text; both give the same bits.
    """

    def __init__(self, num_classes, anchors, input_size, conf_threshold, n_images_total, kind="v3",
                 group=None, mode="flags", **modes):
        from .compat._accum import MeanAveragePrecisionBase
        if mode not in ("flags", "rows"):
            raise ValueError("mode must be 'flags' or 'rows'")
        self.group = group
        self.mode = mode
        world, rank = _world(group)
        self.lo, self.hi = shard_range(n_images_total, rank, world)
        self._acc = MeanAveragePrecisionBase(num_classes, anchors, input_size, conf_threshold, kind, **modes)
        self.num_classes = num_classes
        self.reset_states()

    def warm(self):
        """NCCL sets a collective up on its first call (milliseconds) and the caching allocator has to find the
        receive buffers once: run the epoch end's three all-gathers on a few bytes now, outside any timed epoch."""
        world, _ = _world(self.group)
        if world > 1 and torch.cuda.is_available():
            dev = torch.device("cuda", torch.cuda.current_device())
            gather_padded(torch.zeros((4, 3), device=dev), self.group, extra=torch.zeros(self.num_classes, dtype=torch.int64, device=dev))
            gather_padded(torch.zeros((4, 7), device=dev), self.group)
            dist.all_reduce(torch.zeros((self.num_classes, 4), device=dev), group=self.group)

    def reset_states(self):
        self._acc.reset_states()
        self._acc.img_idx = self.lo

    def update_state(self, y_true, y_preds):
        self._acc._update(y_true, list(y_preds) if isinstance(y_preds, (list, tuple)) else [y_preds])

    def local_rows(self):
        from . import engine
        gt, det = engine.views(self._acc._true, self._acc._pred)           # one host read for both counts
        return gt, det

    def gathered(self):
        gt, det = self.local_rows()
        return gather_rows(gt, self.group), gather_rows(det, self.group)

    def _empty(self):
        return torch.zeros((0, 7), dtype=torch.float32, device=torch.device("cuda", torch.cuda.current_device()))

    def per_class(self, mode=None):
        """(mAP [1], per_class [nc,4] = AP, TP, FP, FN) of the whole set."""
        from . import engine
        mode = mode or self.mode
        if mode == "rows":
            gt, det = self.gathered()
            m, pc, _ = engine.map_match(gt, det, self.num_classes, want_tp=False)
            return m, pc
        gt, det = self.local_rows()
        triples, gpc, _ = engine.map_flags(gt, det, self.num_classes)
        return map_from_flags(triples, gpc, self.num_classes, self.group)

    def result(self, mode=None):
        return self.per_class(mode)[0][0]
