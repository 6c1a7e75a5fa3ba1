"""Multi-GPU plumbing: one process per GPU, images sharded in contiguous blocks, and the single
exchange step of the path -- a variable-length all-gather of detection / ground-truth rows for
the mAP@.5 matching.  decode -> sort -> NMS needs no communication.

The reference never gathers (its metric state is rank-local, module/yolov3_detector.py:45-47);
parity for this module is judged against the single-process result over all images.
Works with backend "nccl" (CUDA tensors, NVLink/NVSwitch) and "gloo" (CPU tensors, tests).
"""
from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of rank `rank`; the first n_items % world ranks get one extra."""
    base, extra = divmod(int(n_items), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_rows(rows: torch.Tensor, group: Optional[dist.ProcessGroup] = None,
                counts_out: Optional[list] = None) -> torch.Tensor:
    """All-gather rows [k_r, C] of every rank into [sum k_r, C], rank after rank.

    Two collectives: the int64 counts, then the payload padded to the largest count (NCCL
    all-gather needs equal sizes; NVSwitch makes the padding cheap).  One host sync (the counts).
    """
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        if counts_out is not None:
            counts_out[:] = [rows.shape[0]]
        return rows
    world = dist.get_world_size(group)
    rows = rows.contiguous()
    cnt = torch.tensor([rows.shape[0]], dtype=torch.int64, device=rows.device)
    all_cnt = torch.empty(world, dtype=torch.int64, device=rows.device)
    dist.all_gather_into_tensor(all_cnt, cnt, group=group)
    counts = all_cnt.tolist()
    if counts_out is not None:
        counts_out[:] = counts
    kmax = max(counts)
    C = rows.shape[1]
    if kmax == 0:
        return rows[:0]
    send = rows
    if rows.shape[0] < kmax:
        send = torch.empty((kmax, C), dtype=rows.dtype, device=rows.device)
        send[: rows.shape[0]] = rows
    recv = torch.empty((world, kmax, C), dtype=rows.dtype, device=rows.device)
    dist.all_gather_into_tensor(recv.view(world * kmax, C), send, group=group)
    if all(c == kmax for c in counts):
        return recv.view(world * kmax, C)
    return torch.cat([recv[r, : counts[r]] for r in range(world)], dim=0)


class ShardedMeanAveragePrecision:
    """MeanAveragePrecision over an image-sharded evaluation set.

    Each rank feeds its own images to update_state (global image ids = shard offset + local
    index, stored in the fp32 img column like the reference, exact below 2**24); result()
    all-gathers [K_r,7] detections and [G_r,7] ground truths and runs the GPU matching on the
    gathered set, so every rank returns the mAP of the WHOLE set.
    """

    def __init__(self, num_classes, anchors, input_size, conf_threshold, n_images_total, kind="v3",
                 group=None, **modes):
        from .compat._accum import MeanAveragePrecisionBase
        self.group = group
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.lo, self.hi = shard_range(n_images_total, rank, world)
        self._acc = MeanAveragePrecisionBase(num_classes, anchors, input_size, conf_threshold, kind, **modes)
        self.num_classes = num_classes
        self.reset_states()

    def reset_states(self):
        self._acc.reset_states()
        self._acc.img_idx = self.lo

    def update_state(self, y_true, y_preds):
        self._acc._update(y_true, list(y_preds) if isinstance(y_preds, (list, tuple)) else [y_preds])

    def gathered(self):
        det = gather_rows(self._acc._pred.view() if self._acc._pred is not None else self._empty(), self.group)
        gt = gather_rows(self._acc._true.view() if self._acc._true is not None else self._empty(), self.group)
        return gt, det

    def _empty(self):
        return torch.zeros((0, 7), dtype=torch.float32, device=torch.device("cuda", torch.cuda.current_device()))

    def result(self):
        from .compat.yolo_utils import mean_average_precision
        gt, det = self.gathered()
        return mean_average_precision(gt, det, self.num_classes)
