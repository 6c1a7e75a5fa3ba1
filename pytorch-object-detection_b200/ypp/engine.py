"""Host side of the hot path: torch tensors in, stream-ordered C-ABI calls, torch tensors out.

PyTorch is used for device memory (caching allocator), the current CUDA stream and, in
ypp.dist, torch.distributed.  All arithmetic happens in libypp.so (csrc/*.cu).
"""
import contextlib
import ctypes
import os
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch

from . import _lib
from ._lib import YppCfg, YppScale, check
from ._lib import lib as _load_lib

_hinted = False


def lib():
    """The loaded library; the first call on a machine with a GPU also turns on ypp_adaptive_launch (include/ypp.h: the
    long-segment kernel is launched narrow until the device has seen a long segment; a hint, results never depend on
    it; YPP_ADAPTIVE=0 leaves it off)."""
    global _hinted
    L = _load_lib()
    if not _hinted:
        _hinted = True
        if torch.cuda.is_available() and os.environ.get("YPP_ADAPTIVE", "1") != "0":
            L.ypp_adaptive_launch(1)            # a failure only means full-width launches, as before
    return L


# the per-call host cost is what bounds small batches: the current device and stream come from the two C getters
# torch.cuda's Python helpers wrap (an order of magnitude cheaper); fall back to the public API if they move
_cur_dev = getattr(torch._C, "_cuda_getDevice", None) or torch.cuda.current_device
_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_NULL_CTX = contextlib.nullcontext()


def _stream():
    if _raw_stream is not None:
        return ctypes.c_void_p(_raw_stream(_cur_dev()))
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _on(device):
    """Context that makes `device` current; nothing to do (and nothing to pay) when it already is."""
    idx = device.index
    if idx is None or idx == _cur_dev():
        return _NULL_CTX
    return torch.cuda.device(device)


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream_key(device):
    """(device ordinal, raw handle of the stream current on it)."""
    idx = device.index if device.index is not None else _cur_dev()
    if _raw_stream is not None:
        return idx, _raw_stream(idx)
    return idx, torch.cuda.current_stream(idx).cuda_stream


class _Scratch:
    """Scratch buffers of ONE owner (a PostProcessor, or the module-level helpers below), one per
    (device, stream, tag), grown on demand.  A buffer is only ever handed to work queued on the stream it is
    keyed on, so everything that touches it is stream-ordered; two streams (or two owners) never share one.
    A buffer that is replaced goes back to the caching allocator, which is stream-ordered on that same stream.
    Streams come and go (CUDA graph captures, per-call side streams): beyond `max_live` entries the oldest go."""

    def __init__(self, max_live=16):
        self._bufs = {}
        self.max_live = max_live

    def get(self, nbytes: int, device, tag: str = "post") -> torch.Tensor:
        key = _stream_key(device) + (tag,)
        buf = self._bufs.get(key)
        if buf is None or buf.numel() < nbytes:
            if buf is None and len(self._bufs) >= self.max_live:
                self._bufs.pop(next(iter(self._bufs)))
            buf = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
            self._bufs[key] = buf
        return buf


_module_scratch = _Scratch()


def _workspace(nbytes: int, device, tag: str) -> torch.Tensor:
    """Scratch of the module-level entry points (map_match, nms_ultralytics, confusion_update)."""
    return _module_scratch.get(nbytes, device, tag)


def _check_head(t: torch.Tensor, name="head") -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or t.dim() != 4:
        raise ValueError("%s must be a 4-d tensor [B, A*(5+nc), H, W]" % name)
    if not t.is_cuda:
        raise ValueError("%s must live on a CUDA device (ypp has no CPU path)" % name)
    if t.dtype != torch.float32:
        raise ValueError("%s must be float32, got %s" % (name, t.dtype))
    return t if t.is_contiguous() else t.contiguous()


@dataclass
class PostResult:
    """Device-side result of one batch; nothing here has synchronised with the host."""
    det_rows: torch.Tensor            # [capacity, row_width], first det_total rows valid
    det_total: torch.Tensor           # int32 [1]
    keep_count: torch.Tensor          # int32 [B]
    cand_count: torch.Tensor          # int32 [B]
    keep_idx: Optional[torch.Tensor]  # int32 [B, N] (first keep_count[b] entries valid) or None
    n_boxes: int

    def rows(self) -> torch.Tensor:
        """Valid rows (one host sync for the count)."""
        return self.det_rows[: int(self.det_total.item())]

    def per_image(self) -> List[torch.Tensor]:
        counts = self.keep_count.tolist()
        out, off = [], 0
        for c in counts:
            out.append(self.det_rows[off: off + c])
            off += c
        return out


class RowAccumulator:
    """[n, 7] rows growing on the device with the row count kept ON the device: what
    `self.all_pred_boxes_variable = torch.cat([...])` (utils/yolov3_utils.py:139-146) builds, without the host ever
    waiting for the device inside the loop.  The kernels append behind a device cursor (ypp_nms_append /
    ypp_target_rows_append).  The host needs the count only to size the buffer, and an upper bound does for that:
      - the kernel that advances the cursor also stores (count << 24 | append number) into ONE pinned host word
        (ypp_probe): no copy, no event, nothing between two chains on the stream;
      - bound = the newest count that has arrived + the worst case of the appends issued after that one;
      - when the bound does not fit, the buffer grows by copying `bound` rows on the device (rows past the real count
        are garbage nobody reads) - a wait on the GPU for the appends in flight, none on the host.  Only if the host
        has run more than STALE appends ahead of the device does it first wait for the append STALE back (newer
        batches stay queued, nothing drains): a bound that stale would grow the buffer for rows that never come.
    `sync_every` sizes the growth steps (worst cases booked per step).  Only count() / view() read the cursor itself.

    Appends may be queued on side streams (PostProcessor.accumulate with several batches in flight): everything
    that reads or moves the rows on the current stream (count, growth, view) first waits, on the device, for them."""

    STALE = 6
    RING = 64
    _SEQ = 0xFFFFFF

    def __init__(self, device, sync_every=16, capacity_hint=0):
        self.device = torch.device(device)
        self.sync_every = max(1, int(sync_every))
        self.capacity_hint = int(capacity_hint)     # rows the previous epoch ended with: the first allocation covers them
        self.cur = torch.zeros(2, dtype=torch.int64, device=self.device)   # two words: see ypp_nms_append
        self.k = 0                                                         # appends so far: the live word is k & 1
        self.buf = None
        self.bound = 0                                                     # rows <= bound, known without a sync
        self.syncs = 0                                                     # host reads of the cursor (count / view)
        self._ring = None                                                  # the pipeline the newest append went through
        self._streams = set()                                              # side streams that were told about buf / cur
        self._host = torch.zeros(1, dtype=torch.int64).pin_memory()        # the probe word
        self._host_np = self._host.numpy()
        self._probe = _lib.YppProbe(self._host.data_ptr(), 0)
        self._probe_ref = ctypes.byref(self._probe)
        self._seq = 0                                                      # appends issued (probe sequence numbers start at 1)
        self._booked = 0                                                   # sum of the worst cases of every append so far
        self._booked_at = [0] * self.RING                                  # ... as it was when append seq was issued

    def join(self):
        """The current stream waits (on the device) for every append queued on side streams so far."""
        if self._ring is not None:
            self._ring.join(_stream())
            self._ring = None

    def used_by(self, stream):
        """A side stream is about to append: the caching allocator must not hand the buffers to anyone else
        while work queued on that stream still touches them."""
        key = (stream.cuda_stream, 0 if self.buf is None else self.buf.data_ptr())
        if key not in self._streams:
            self._streams.add(key)
            self.cur.record_stream(stream)
            if self.buf is not None:
                self.buf.record_stream(stream)

    def next_probe(self, ring=None):
        """The ypp_probe* for the append about to be issued (after reserve()); ring: the pipeline it goes through."""
        self._seq += 1
        self._booked_at[self._seq % self.RING] = self._booked
        self._probe.seq = self._seq & self._SEQ
        if ring is not None:
            self._ring = ring
        return self._probe_ref

    def _arrived(self):
        """(append number, rows after it) of the newest append whose probe word has arrived; (0, 0) before the first."""
        v = int(self._host_np[0])
        low = v & self._SEQ
        seq = self._seq - ((self._seq - low) & self._SEQ)      # undo the 24-bit wrap
        return seq, v >> 24

    def _refresh(self):
        seq, n = self._arrived()
        if seq > 0 and self._seq - seq < self.RING:
            b = n + (self._booked - self._booked_at[seq % self.RING])
            if b < self.bound:
                self.bound = b

    def _wait_for(self, seq, timeout_s=30.0):
        """The host waits until append `seq` has reported (the device is busy with it or with earlier work)."""
        import time
        t0 = time.perf_counter()
        spins = 0
        while self._arrived()[0] < seq:
            spins += 1
            if spins & 1023 == 0 and time.perf_counter() - t0 > timeout_s:
                raise _lib.YppError("RowAccumulator: append %d did not report within %.0f s (device error?)" % (seq, timeout_s))

    def count(self) -> int:
        self.join()
        self.syncs += 1
        n = int(self.cur[self.k & 1].item())
        self.bound = n
        return n

    def reserve(self, worst: int):
        """Room for `worst` more rows.  Never drains the device."""
        cap = 0 if self.buf is None else self.buf.shape[0]
        if self.bound + worst > cap:
            self._refresh()
            if self.bound + worst > cap and self._seq > self.STALE:
                self._wait_for(self._seq - self.STALE)
                self._refresh()
            if self.bound + worst > cap:
                want = max(self.bound + worst * self.sync_every, 2 * cap, 1024)
                if self.buf is None and self.capacity_hint:
                    want = max(want, self.capacity_hint + self.capacity_hint // 8 + worst * self.STALE)
                nb = torch.empty((want, 7), dtype=torch.float32, device=self.device)
                if self.buf is not None and self.bound:
                    self.join()                                    # the appends in flight wrote (and sized) the old buffer
                    n = min(self.bound, cap)
                    nb[:n] = self.buf[:n]
                self.buf = nb
        self.bound += worst
        self._booked += worst

    def cursors(self, in_place=False):
        """(pointer to the live word, pointer to the word the next append writes)."""
        base = self.cur.data_ptr()
        if in_place:
            return base + 8 * (self.k & 1), base + 8 * (self.k & 1)
        a, b = base + 8 * (self.k & 1), base + 8 * ((self.k + 1) & 1)
        self.k += 1
        return a, b

    def view(self) -> torch.Tensor:
        if self.buf is None:
            return torch.empty((0, 7), dtype=torch.float32, device=self.device)
        n = self.count()
        # reserve() books the worst case of every append, so the kernels never had to drop a row
        assert n <= self.buf.shape[0], "RowAccumulator: %d rows were appended to a buffer of %d" % (n, self.buf.shape[0])
        return self.buf[:n]


def views(*accs):
    """view() of several accumulators with ONE host read of their cursors (each .item() is a device synchronisation)."""
    live = [a for a in accs if a is not None and a.buf is not None]
    for a in live:
        a.join()
    counts = torch.stack([a.cur[a.k & 1] for a in live]).tolist() if live else []
    out, it = [], iter(counts)
    for a in accs:
        if a is None or a.buf is None:
            dev = a.device if a is not None else torch.device("cuda", _cur_dev())
            out.append(torch.empty((0, 7), dtype=torch.float32, device=dev))
            continue
        n = int(next(it))
        a.syncs += 1
        a.bound = n
        assert n <= a.buf.shape[0], "RowAccumulator: %d rows were appended to a buffer of %d" % (n, a.buf.shape[0])
        out.append(a.buf[:n])
    return out


class _Plan:
    """Parameter blocks of one set of head buffers.  Holds NO tensor: the key (pointer, shape, dtype, device of
    every head) guarantees the pointers inside `arr` are those of whatever tensors are passed with that key."""
    __slots__ = ("arr", "B", "n_boxes", "cfg", "nbytes", "dev")

    def __init__(self, arr, B, n_boxes, cfg, nbytes, dev):
        self.arr, self.B, self.n_boxes, self.cfg, self.nbytes, self.dev = arr, B, n_boxes, cfg, nbytes, dev


_pipeline_pool = {}      # (device index, depth) -> [(handle, streams)] not in use


class _Ring:
    """The batches in flight of one PostProcessor on one device: a ypp_pipeline (side streams + events, owned by the
    library) and, per slot, the workspace and count outputs this side owns.  The caching allocator remembers every
    stream a tensor was record_stream()ed on and touches it when the tensor is freed, so a pipeline's streams must
    outlive every tensor they have seen: pipelines are never destroyed, a ring that goes away hands its pipeline back
    to a per-(device, depth) pool and the next ring takes it over (their batches simply queue behind)."""

    def __init__(self, dev, depth):
        self.dev, self.depth = dev, depth
        self.key = (dev.index if dev.index is not None else _cur_dev(), depth)
        free = _pipeline_pool.setdefault(self.key, [])
        if free:
            self.handle, self.streams = free.pop()
        else:
            h = ctypes.c_void_p()
            with _on(dev):
                check(lib().ypp_pipeline_create(depth, ctypes.byref(h)))
            L = lib()
            self.handle = h
            self.streams = [torch.cuda.ExternalStream(L.ypp_pipeline_stream(h, i), device=dev) for i in range(depth)]
        self.ws = [None] * depth
        self.sc = [None] * depth

    def __del__(self):
        try:
            _pipeline_pool[self.key].append((self.handle, self.streams))
        except Exception:
            pass

    def slot(self, B, nbytes):
        i = lib().ypp_pipeline_next_slot(self.handle)
        st = self.streams[i]
        if self.ws[i] is None or self.ws[i].numel() < nbytes:
            self.ws[i] = torch.empty(int(nbytes), dtype=torch.uint8, device=self.dev)
            self.ws[i].record_stream(st)
        sc = self.sc[i]
        if sc is None or sc.keep_count.numel() != B:
            z = torch.empty
            sc = PostResult(z((0, 7), device=self.dev), z(1, dtype=torch.int32, device=self.dev), z(B, dtype=torch.int32, device=self.dev),
                            z(B, dtype=torch.int32, device=self.dev), None, 0)
            sc.keep_count.record_stream(st)
            sc.cand_count.record_stream(st)
            self.sc[i] = sc
        return i, st, self.ws[i], sc

    def join(self, raw_stream):
        check(lib().ypp_pipeline_join(self.handle, raw_stream))


class PostProcessor:
    """decode -> threshold/compact -> sort -> per-class NMS -> rows for a batch of heads.

    anchors: pixels for kind="v3" (divided by the stride in float64, rounded to fp32 once, as
    utils/yolov3_utils.py:31,40 does) or grid units for kind="v2" (utils/yolov2_utils.py:39).
    Heads are consumed in list order with anchors[A*i : A*i+A] (utils/yolov3_utils.py:76-78).
    Defaults reproduce what the reference computes when it runs on a CUDA device:
    torchvision's CUDA heuristic for the coordinate trick and its fma-contracted union.

    Scratch memory belongs to the processor and is keyed on the stream it is used from, so two processors, or
    one processor called from two streams, never share a buffer.  pipeline_depth: batches `accumulate` keeps
    in flight (each on its own side stream and workspace; only the row gathers, which append behind one
    device cursor, are chained in batch order).  None = 3 for batches of at least 32 768 boxes, else 1;
    the environment variable YPP_PIPELINE overrides the default.
    """

    def __init__(self, num_classes, anchors, input_size, conf_threshold=0.25, iou_threshold=0.45,
                 kind="v3", anchors_per_scale=None, offset_mode="auto_cuda", union_mode="fma",
                 score_mode="obj", compare_in_double=None, pipeline_depth=None):
        self.num_classes = int(num_classes)
        self.anchors = [list(map(float, a)) for a in anchors]
        self.input_size = input_size
        self._plans = {}
        self._scratch = {}
        self._ws = _Scratch()
        self._rings = {}
        self._batches = 0
        if pipeline_depth is None and os.environ.get("YPP_PIPELINE"):
            pipeline_depth = int(os.environ["YPP_PIPELINE"])
        self.pipeline_depth = pipeline_depth
        self._auto_depth, self._auto_fixed = 3, False
        self.kind = kind
        if anchors_per_scale is None:
            anchors_per_scale = len(self.anchors) if kind == "v2" else 3
        self.A = int(anchors_per_scale)
        self.offset_mode = _lib.OFFSET_MODES[offset_mode] if isinstance(offset_mode, str) else int(offset_mode)
        self.union_mode = _lib.UNION_MODES[union_mode] if isinstance(union_mode, str) else int(union_mode)
        self.score_mode = _lib.SCORE_MODES[score_mode] if isinstance(score_mode, str) else int(score_mode)
        if compare_in_double is None:
            compare_in_double = self.union_mode == _lib.UNION_SEPARATE   # CPU kernel semantics
        self._compare_in_double = bool(compare_in_double)
        self.conf_threshold = conf_threshold
        self.iou_threshold = iou_threshold

    # the thresholds are plain attributes in the reference (read on every call, utils/yolov3_utils.py:88,125):
    # assigning one takes effect on the next call here too (the cached parameter blocks are dropped)
    @property
    def conf_threshold(self):
        return self._conf

    @conf_threshold.setter
    def conf_threshold(self, v):
        self._conf = float(v)
        self._plans.clear()

    @property
    def iou_threshold(self):
        return self._iou

    @iou_threshold.setter
    def iou_threshold(self, v):
        self._iou = float(v)
        self.iou_thr_eff = float(lib().ypp_effective_iou_thr(self._iou, int(self._compare_in_double)))
        self._plans.clear()

    def workspace(self, nbytes, device) -> torch.Tensor:
        """This processor's scratch buffer for the stream current on `device`."""
        return self._ws.get(nbytes, device)

    # -- parameter blocks ------------------------------------------------------------------
    def _cfg(self, B, n_scales):
        return YppCfg(self.num_classes, int(B), int(n_scales), self._conf, self.iou_thr_eff,
                      self.score_mode, self.offset_mode, self.union_mode)

    def _scales(self, heads: Sequence[torch.Tensor]):
        n = len(heads)
        if n < 1 or n > _lib.YPP_MAX_SCALES:
            raise ValueError("1..%d heads supported, got %d" % (_lib.YPP_MAX_SCALES, n))
        arr = (YppScale * n)()
        heads = [_check_head(h, "head %d" % i) for i, h in enumerate(heads)]
        B = heads[0].shape[0]
        n_boxes = 0
        for i, h in enumerate(heads):
            if h.shape[0] != B or h.device != heads[0].device:
                raise ValueError("all heads must share batch size and device")
            _, C, H, W = h.shape
            an = self.anchors[self.A * i: self.A * i + self.A]
            if len(an) != self.A:
                raise ValueError("not enough anchors for head %d" % i)
            if C != self.A * (5 + self.num_classes):
                raise ValueError("head %d has %d channels, expected %d*(5+%d)" % (i, C, self.A, self.num_classes))
            stride_h = self.input_size / H
            stride_w = self.input_size / W
            s = arr[i]
            s.head = h.data_ptr()
            s.A, s.H, s.W = self.A, H, W
            for a, (w_, h_) in enumerate(an):
                if self.kind == "v2":
                    s.anchor_w[a], s.anchor_h[a] = w_, h_
                else:
                    s.anchor_w[a], s.anchor_h[a] = w_ / stride_w, h_ / stride_h
            s.stride_w, s.stride_h = stride_w, stride_h
            n_boxes += self.A * H * W
        return arr, heads, B, n_boxes

    # -- entry points ------------------------------------------------------------------------
    def decode(self, heads: Sequence[torch.Tensor]) -> torch.Tensor:
        """= decode_predictions per scale + torch.cat(dim=1): [B, n_boxes, 6]."""
        arr, heads, B, n_boxes = self._scales(heads)
        out = torch.empty((B, n_boxes, 6), dtype=torch.float32, device=heads[0].device)
        cfg = self._cfg(B, len(heads))
        with _on(heads[0].device):
            check(lib().ypp_decode_dense(ctypes.byref(cfg), arr, _ptr(out), _stream()))
        return out

    def _outputs(self, B, n_boxes, device, row_width, want_keep_idx, out):
        if out is not None:
            return out
        cap = B * n_boxes
        return PostResult(
            det_rows=torch.empty((cap, row_width), dtype=torch.float32, device=device),
            det_total=torch.empty(1, dtype=torch.int32, device=device),
            keep_count=torch.empty(B, dtype=torch.int32, device=device),
            cand_count=torch.empty(B, dtype=torch.int32, device=device),
            keep_idx=torch.empty((B, n_boxes), dtype=torch.int32, device=device) if want_keep_idx else None,
            n_boxes=n_boxes)

    def _plan(self, heads):
        """Parameter blocks for these exact head buffers, cached (the per-call host cost is what bounds the sparse
        regime): key = (pointer, shape, ...) of every head.  Returns (plan, heads actually read): the second is
        only there to keep contiguous copies of non-contiguous heads alive until the launches are queued."""
        key = tuple((h.data_ptr(), tuple(h.shape), h.is_contiguous(), h.dtype, h.device.index) for h in heads)
        plan = self._plans.get(key)
        if plan is not None:
            return plan, heads
        arr, heads_c, B, n_boxes = self._scales(heads)
        plan = _Plan(arr, B, n_boxes, self._cfg(B, len(heads)), lib().ypp_workspace_bytes(B, n_boxes, self.num_classes),
                     heads_c[0].device)
        if all(hc is h for hc, h in zip(heads_c, heads)):        # temporaries are not worth remembering
            if len(self._plans) > 64:
                self._plans.clear()
            self._plans[key] = plan
        return plan, heads_c

    def __call__(self, heads: Sequence[torch.Tensor], img_idx_base=0, row_width=7, want_keep_idx=False,
                 out: Optional[PostResult] = None, ws: Optional[torch.Tensor] = None) -> PostResult:
        plan, heads = self._plan(heads)
        dev = plan.dev
        if dev.index is not None and dev.index != _cur_dev():
            with _on(dev):
                return self.__call__(heads, img_idx_base, row_width, want_keep_idx, out, ws)
        res = self._outputs(plan.B, plan.n_boxes, dev, row_width, want_keep_idx, out)
        if ws is None:
            ws = self._ws.get(plan.nbytes, dev)
        check(lib().ypp_postprocess(ctypes.byref(plan.cfg), plan.arr, _ptr(ws), ws.numel(), _ptr(res.cand_count),
                                    _ptr(res.keep_count), _ptr(res.keep_idx), _ptr(res.det_rows),
                                    res.det_rows.shape[1], res.det_rows.shape[0], _ptr(res.det_total),
                                    float(img_idx_base), _stream()))
        return res

    def from_features(self, features: Sequence[torch.Tensor], weights: Sequence[torch.Tensor], img_idx_base=0,
                      row_width=7, want_keep_idx=False, want_heads=False, stages=("decode", "sort", "nms"),
                      out: Optional[PostResult] = None):
        """The chain fed by the INPUTS of the heads' 1x1 convolutions: features[i] is the
        [B, Cin, H, W] tensor the reference passes to nn.Conv2d(Cin, A*(5+nc), 1, bias=False)
        (models/detector/yolov3.py:56,62,68), weights[i] that layer's weight [A*(5+nc), Cin, 1, 1].  The
        head tensors are never written: one tcgen05 TF32 kernel per scale decodes the accumulator tiles.
        Features are read channels-last; a torch.channels_last tensor is used in place, anything else is
        converted first (one extra pass).  want_heads=True also returns the head tensors the kernel saw."""
        n = len(features)
        if n != len(weights) or n < 1 or n > _lib.YPP_MAX_SCALES:
            raise ValueError("features and weights must be equally long lists of 1..%d tensors" % _lib.YPP_MAX_SCALES)
        dev = features[0].device
        B = features[0].shape[0]
        arr = (YppScale * n)()
        src = (_lib.YppHeadSrc * n)()
        keep_alive, heads_out, n_boxes = [], [], 0
        AC = self.A * (5 + self.num_classes)
        for i, (x, w) in enumerate(zip(features, weights)):
            if x.dim() != 4 or not x.is_cuda or x.dtype != torch.float32 or x.shape[0] != B or x.device != dev:
                raise ValueError("feature %d must be a float32 CUDA tensor [B, Cin, H, W] on one device" % i)
            _, Cin, H, W = x.shape
            w2 = w.reshape(w.shape[0], -1)
            if w2.shape != (AC, Cin) or w.dtype != torch.float32 or w.device != dev:
                raise ValueError("weight %d must be float32 [%d, %d(,1,1)] on the features' device" % (i, AC, Cin))
            xl = x.permute(0, 2, 3, 1)
            if not xl.is_contiguous():
                xl = xl.contiguous()                       # NCHW input: one transposing pass
            w2 = w2.contiguous()
            keep_alive += [xl, w2]
            an = self.anchors[self.A * i: self.A * i + self.A]
            if len(an) != self.A:
                raise ValueError("not enough anchors for scale %d" % i)
            stride_h, stride_w = self.input_size / H, self.input_size / W
            s = arr[i]
            s.head = None
            s.A, s.H, s.W = self.A, H, W
            for a, (w_, h_) in enumerate(an):
                if self.kind == "v2":
                    s.anchor_w[a], s.anchor_h[a] = w_, h_
                else:
                    s.anchor_w[a], s.anchor_h[a] = w_ / stride_w, h_ / stride_h
            s.stride_w, s.stride_h = stride_w, stride_h
            src[i].feature, src[i].weight, src[i].cin = xl.data_ptr(), w2.data_ptr(), Cin
            if want_heads:
                heads_out.append(torch.empty((B, AC, H, W), dtype=torch.float32, device=dev))
                src[i].head_out = heads_out[-1].data_ptr()
            n_boxes += self.A * H * W
        res = self._outputs(B, n_boxes, dev, row_width, want_keep_idx, out)
        cfg = self._cfg(B, n)
        L = lib()
        with _on(dev):
            ws = self._ws.get(L.ypp_workspace_bytes(B, n_boxes, self.num_classes), dev)
            if tuple(stages) == ("decode", "sort", "nms"):
                check(L.ypp_head_postprocess(ctypes.byref(cfg), arr, src, _ptr(ws), ws.numel(), _ptr(res.cand_count),
                                             _ptr(res.keep_count), _ptr(res.keep_idx), _ptr(res.det_rows),
                                             res.det_rows.shape[1], res.det_rows.shape[0], _ptr(res.det_total),
                                             float(img_idx_base), _stream()))
            else:
                if "decode" in stages:
                    check(L.ypp_head_decode_compact(ctypes.byref(cfg), arr, src, _ptr(ws), ws.numel(), _stream()))
                if "sort" in stages:
                    check(L.ypp_sort_segments(ctypes.byref(cfg), n_boxes, _ptr(ws), ws.numel(), _stream()))
                if "nms" in stages:
                    check(L.ypp_nms(ctypes.byref(cfg), n_boxes, _ptr(ws), ws.numel(), _ptr(res.cand_count),
                                    _ptr(res.keep_count), _ptr(res.keep_idx), _ptr(res.det_rows),
                                    res.det_rows.shape[1], res.det_rows.shape[0], _ptr(res.det_total),
                                    float(img_idx_base), _stream()))
        return (res, heads_out) if want_heads else res

    def _depth(self, B, n_boxes, acc=None):
        """Batches in flight for this call.  Not given by the caller: 1 for small batches (nothing to overlap); else 3
        until eight appends of `acc` have reported their row count to the host (the pinned probe word, no sync), then
        - once, for the life of this processor - 2 if they held only a few rows per image (< 64: the host is the limit
        there and every side stream costs it an event pair: 0.0333 against 0.0355 ms per sparse batch), 3 otherwise
        (dense output: 3 and 4 measure the same, 0.077-0.079 ms per batch)."""
        d = self.pipeline_depth
        if d is None:
            if B * n_boxes < 32768:
                return 1
            d = self._auto_depth
            if not self._auto_fixed and acc is not None:
                seq, n = acc._arrived()
                if seq >= 8:
                    per_image = n / float(seq * B)
                    d = self._auto_depth = 2 if per_image < 64 else 3
                    self._auto_fixed = True
        return max(1, int(d))

    def _ring(self, dev, depth):
        ring = self._rings.get(dev.index)
        if ring is None or ring.depth != depth:
            if ring is not None:
                with _on(dev):
                    ring.join(_stream())
            ring = _Ring(dev, depth)
            self._rings[dev.index] = ring
        return ring

    def join(self):
        """The current stream of every device this processor has batches in flight on waits for them (on the device)."""
        for idx, ring in self._rings.items():
            with _on(ring.dev):
                ring.join(_stream())

    def accumulate(self, acc: RowAccumulator, heads: Sequence[torch.Tensor], weights=None, img_idx_base=0):
        """The chain with the gather appending to `acc` (7-column rows) instead of filling a per-call buffer:
        nothing is read back, so a validation loop queues batch after batch.  weights given: `heads` are the
        INPUTS of the heads' 1x1 convolutions (from_features).

        With a pipeline depth above 1 (see the class) batch i runs on side stream i mod depth: its decode, sort
        and suppression overlap the neighbouring batches' (the per-image sort alone leaves most SMs idle), its
        row gather waits for the previous batch's (the cursor chain) - so rows still land in batch order - and
        `acc` remembers the event behind it; acc.view() / count() / a later growth wait for that event.  The
        counts returned belong to this batch and are valid for the current stream after acc.join()."""
        L = lib()
        if weights is not None:                       # fused-head entry: current stream, one batch at a time
            B = heads[0].shape[0]
            n_boxes = sum(self.A * f.shape[2] * f.shape[3] for f in heads)
            dev = heads[0].device
            if acc.device != dev:
                raise ValueError("accumulator and heads are on different devices")
            sc = self._scratch.get((B, dev))
            if sc is None:
                z = torch.empty
                sc = PostResult(z((0, 7), device=dev), z(1, dtype=torch.int32, device=dev), z(B, dtype=torch.int32, device=dev),
                                z(B, dtype=torch.int32, device=dev), None, n_boxes)
                self._scratch[(B, dev)] = sc
            acc.join()
            self.from_features(heads, weights, stages=("decode", "sort"), out=sc)
            cfg = self._cfg(B, len(heads))
            with _on(dev):
                ws = self._ws.get(L.ypp_workspace_bytes(B, n_boxes, self.num_classes), dev)
                acc.reserve(B * n_boxes)
                cin, cout = acc.cursors()
                check(L.ypp_nms_append(ctypes.byref(cfg), n_boxes, _ptr(ws), ws.numel(), _ptr(sc.cand_count), _ptr(sc.keep_count),
                                       _ptr(acc.buf), 7, acc.buf.shape[0], cin, cout, float(img_idx_base), _stream(), acc.next_probe()))
            return sc
        plan, heads = self._plan(heads)
        dev, B, n_boxes = plan.dev, plan.B, plan.n_boxes
        if acc.device != dev:
            raise ValueError("accumulator and heads are on different devices")
        depth = self._depth(B, n_boxes, acc)
        with _on(dev):
            if depth == 1:
                acc.join()
                sc = self._scratch.get((B, dev))
                if sc is None:
                    z = torch.empty
                    sc = PostResult(z((0, 7), device=dev), z(1, dtype=torch.int32, device=dev), z(B, dtype=torch.int32, device=dev),
                                    z(B, dtype=torch.int32, device=dev), None, n_boxes)
                    self._scratch[(B, dev)] = sc
                ws = self._ws.get(plan.nbytes, dev)
                acc.reserve(B * n_boxes)
                cin, cout = acc.cursors()
                check(L.ypp_postprocess_append(ctypes.byref(plan.cfg), plan.arr, _ptr(ws), ws.numel(), _ptr(sc.cand_count),
                                               _ptr(sc.keep_count), _ptr(acc.buf), 7, acc.buf.shape[0], cin, cout,
                                               float(img_idx_base), _stream(), acc.next_probe()))
                return sc
            ring = self._ring(dev, depth)
            i, st, ws, sc = ring.slot(B, plan.nbytes)
            acc.reserve(B * n_boxes)                 # may move the rows: on the current stream, behind the batches in flight
            cin, cout = acc.cursors()
            acc.used_by(st)
            for h in heads:                          # the allocator must not reuse a head the side stream still reads
                h.record_stream(st)
            check(L.ypp_pipeline_submit(ring.handle, ctypes.byref(plan.cfg), plan.arr, _ptr(ws), ws.numel(), _ptr(sc.cand_count),
                                        _ptr(sc.keep_count), _ptr(acc.buf), 7, acc.buf.shape[0], cin, cout, acc.next_probe(ring),
                                        float(img_idx_base), _stream(), None))
        sc.n_boxes = n_boxes
        return sc

    def graphed(self, heads: Sequence[torch.Tensor], img_idx_base=0, row_width=7, want_keep_idx=False,
                repeat=1, stages=None, ws: Optional[torch.Tensor] = None):
        """Capture the chain on these (static) head buffers into a CUDA graph.  Returns
        (graph, PostResult); graph.replay() re-runs it with whatever the buffers then hold.
        `repeat` > 1 captures that many back-to-back invocations (timing loops); `stages`
        restricts the capture to some of ("decode", "sort", "nms").  The graph bakes in the address of its
        workspace: a private one is allocated here (or `ws` is used: graphs that are only ever replayed on one
        stream may share one) and lives as long as the result (`res.ws`)."""
        heads = [_check_head(h) for h in heads]
        arr, _, B, n_boxes = self._scales(heads)
        nbytes = lib().ypp_workspace_bytes(B, n_boxes, self.num_classes)
        if ws is None:
            ws = torch.empty(int(nbytes), dtype=torch.uint8, device=heads[0].device)
        elif ws.numel() < nbytes:
            raise ValueError("graphed: workspace of %d bytes, %d needed" % (ws.numel(), nbytes))
        res = self.staged(heads, img_idx_base, row_width, want_keep_idx, stages=("decode", "sort", "nms"), ws=ws)   # warm-up, allocations
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(repeat):
                if stages is None:
                    self(heads, img_idx_base, row_width, want_keep_idx, out=res, ws=ws)
                else:
                    self.staged(heads, img_idx_base, row_width, want_keep_idx, stages=stages, out=res, ws=ws)
        res.ws = ws
        return g, res

    def staged(self, heads: Sequence[torch.Tensor], img_idx_base=0, row_width=7, want_keep_idx=False,
               stages=("decode", "sort", "nms"), out: Optional[PostResult] = None,
               ws: Optional[torch.Tensor] = None) -> PostResult:
        """Same chain through the separate C-ABI stages (used to time / profile D1 alone); stages from
        ("decode", "sort", "nms" = "suppress" + "rows")."""
        arr, heads, B, n_boxes = self._scales(heads)
        dev = heads[0].device
        res = self._outputs(B, n_boxes, dev, row_width, want_keep_idx, out)
        cfg = self._cfg(B, len(heads))
        L = lib()
        mask = 0
        for st in stages:
            mask |= _lib.STAGES[st]
        with _on(dev):
            if ws is None:
                ws = self._ws.get(L.ypp_workspace_bytes(B, n_boxes, self.num_classes), dev)
            check(L.ypp_run_stages(ctypes.byref(cfg), arr, n_boxes, _ptr(ws), ws.numel(), mask, _ptr(res.cand_count),
                                   _ptr(res.keep_count), _ptr(res.keep_idx), _ptr(res.det_rows), res.det_rows.shape[1],
                                   res.det_rows.shape[0], _ptr(res.det_total), None, None, float(img_idx_base), _stream(), None))
        return res

    def nms_rows(self, rows: torch.Tensor, img_idx_base=0, row_width=6, want_keep_idx=False) -> PostResult:
        """nms_v3 on decoded rows [B, N, 6] (or [N, 6]); class ids must be integers in [0, nc)."""
        if rows.dim() == 2:
            rows = rows.unsqueeze(0)
        if rows.dim() != 3 or rows.shape[-1] != 6:
            raise ValueError("rows must be [B, N, 6] or [N, 6]")
        if not rows.is_cuda or rows.dtype != torch.float32:
            raise ValueError("rows must be a float32 CUDA tensor")
        rows = rows.contiguous()
        B, N, _ = rows.shape
        dev = rows.device
        if N == 0:
            z = torch.zeros
            return PostResult(z((0, row_width), device=dev), z(1, dtype=torch.int32, device=dev),
                              z(B, dtype=torch.int32, device=dev), z(B, dtype=torch.int32, device=dev),
                              z((B, 0), dtype=torch.int32, device=dev) if want_keep_idx else None, 0)
        res = self._outputs(B, N, dev, row_width, want_keep_idx, None)
        cfg = self._cfg(B, 1)
        L = lib()
        with _on(dev):
            ws = self._ws.get(L.ypp_workspace_bytes(B, N, self.num_classes), dev)
            check(L.ypp_nms_rows(ctypes.byref(cfg), _ptr(rows), N, _ptr(ws), ws.numel(), _ptr(res.cand_count),
                                 _ptr(res.keep_count), _ptr(res.keep_idx), _ptr(res.det_rows),
                                 res.det_rows.shape[1], res.det_rows.shape[0], _ptr(res.det_total),
                                 float(img_idx_base), _stream()))
        return res


# ------------------------------------------------------------------------------------ mAP side

def target_rows(annot: torch.Tensor, input_size, img_idx_base=0):
    """[B,T,5] padded annots -> ([B*T,7] buffer, int32 count[1]); get_target_boxes_for_map +
    the img_idx column (utils/yolo_utils.py:682-710, utils/yolov3_utils.py:134-137)."""
    if annot.dim() != 3 or annot.shape[-1] != 5:
        raise ValueError("annot must be [B, max_num_annots, 5]")
    if not annot.is_cuda:
        raise ValueError("annot must be a CUDA tensor")
    annot = annot.to(torch.float32).contiguous()
    B, T, _ = annot.shape
    out = torch.empty((B * T, 7), dtype=torch.float32, device=annot.device)
    cnt = torch.empty(1, dtype=torch.int32, device=annot.device)
    with _on(annot.device):
        check(lib().ypp_target_rows(_ptr(annot), B, T, float(input_size), float(img_idx_base), _ptr(out),
                                    _ptr(cnt), _stream()))
    return out, cnt


def target_rows_append(acc: RowAccumulator, annot: torch.Tensor, input_size, img_idx_base=0):
    """target_rows appending to `acc` behind its device cursor (no count comes back)."""
    if annot.dim() != 3 or annot.shape[-1] != 5:
        raise ValueError("annot must be [B, max_num_annots, 5]")
    if not annot.is_cuda or annot.device != acc.device:
        raise ValueError("annot must be a CUDA tensor on the accumulator's device")
    if annot.dtype != torch.float32 or not annot.is_contiguous():
        annot = annot.to(torch.float32).contiguous()
    B, T, _ = annot.shape
    with _on(annot.device):
        acc.join()
        acc.reserve(B * T)
        cin, _ = acc.cursors(in_place=True)
        check(lib().ypp_target_rows_append(_ptr(annot), B, T, float(input_size), float(img_idx_base), _ptr(acc.buf),
                                           acc.buf.shape[0], cin, _stream(), acc.next_probe()))


def box_iou(box1: torch.Tensor, box2: torch.Tensor) -> torch.Tensor:
    """utils/metrics.py:233-255: xyxy [N,4] x [M,4] -> [N,M]."""
    if not (box1.is_cuda and box2.is_cuda):
        raise ValueError("box_iou needs CUDA tensors")
    a = box1.to(torch.float32).contiguous()
    b = box2.to(torch.float32).contiguous()
    out = torch.empty((a.shape[0], b.shape[0]), dtype=torch.float32, device=a.device)
    with _on(a.device):
        check(lib().ypp_box_iou(_ptr(a), a.shape[0], _ptr(b), b.shape[0], _ptr(out), _stream()))
    return out


def bbox_iou(box1: torch.Tensor, box2: torch.Tensor, xywh=True, GIoU=False, DIoU=False, CIoU=False, eps=1e-7) -> torch.Tensor:
    """utils/metrics.py:192-230: box1 [1|n,4] against box2 [n,4] -> [n,1]."""
    if not (box1.is_cuda and box2.is_cuda):
        raise ValueError("bbox_iou needs CUDA tensors (ypp has no CPU path)")
    a = box1.to(torch.float32).reshape(-1, 4).contiguous()
    b = box2.to(torch.float32).reshape(-1, 4).contiguous()
    n = max(a.shape[0], b.shape[0]) if a.shape[0] and b.shape[0] else 0
    out = torch.empty((n, 1), dtype=torch.float32, device=a.device)
    mode = 3 if CIoU else 2 if DIoU else 1 if GIoU else 0
    with _on(a.device):
        check(lib().ypp_bbox_iou(_ptr(a), a.shape[0], _ptr(b), b.shape[0], int(bool(xywh)), mode, float(eps), _ptr(out), _stream()))
    return out


def map_match(true_boxes: torch.Tensor, pred_boxes: torch.Tensor, num_classes, iou_threshold=0.5,
              want_tp=True):
    """mean_average_precision / metrics_per_class on the GPU.
    Returns (map float32[1], per_class float32[nc,4] = AP,TP,FP,FN, tp uint8[K] or None)."""
    gt = true_boxes.reshape(-1, 7).to(torch.float32).contiguous()
    det = pred_boxes.reshape(-1, 7).to(torch.float32).contiguous()
    if not (gt.is_cuda and det.is_cuda):
        raise ValueError("map_match needs CUDA tensors")
    dev = det.device
    G, K = gt.shape[0], det.shape[0]
    L = lib()
    ws = _workspace(L.ypp_map_workspace_bytes(G, K, int(num_classes)), dev, "map")
    m = torch.empty(1, dtype=torch.float32, device=dev)
    pc = torch.empty((int(num_classes), 4), dtype=torch.float32, device=dev)
    tp = torch.empty(max(K, 1), dtype=torch.uint8, device=dev) if want_tp else None
    with _on(dev):
        check(L.ypp_map_match(_ptr(gt), G, _ptr(det), K, int(num_classes), float(iou_threshold), _ptr(ws),
                              ws.numel(), _ptr(tp), _ptr(pc), _ptr(m), _stream()))
    return m, pc, (tp[:K] if tp is not None else None)


def map_flags(true_boxes: torch.Tensor, pred_boxes: torch.Tensor, num_classes, iou_threshold=0.5, want_tp=False):
    """M1 alone (ypp_map_flags): the TP flag of every detection against the ground truths given, no AP.
    Returns (triples float32 [K,3] = score, cls, tp; gt_per_class int64 [nc]; tp uint8 [K] or None)."""
    gt = true_boxes.reshape(-1, 7).to(torch.float32).contiguous()
    det = pred_boxes.reshape(-1, 7).to(torch.float32).contiguous()
    if not (gt.is_cuda and det.is_cuda):
        raise ValueError("map_flags needs CUDA tensors")
    dev = det.device
    G, K, nc = gt.shape[0], det.shape[0], int(num_classes)
    L = lib()
    triples = torch.empty((K, 3), dtype=torch.float32, device=dev)
    gpc = torch.empty(nc, dtype=torch.int64, device=dev)
    tp = torch.empty(max(K, 1), dtype=torch.uint8, device=dev) if want_tp else None
    with _on(dev):
        ws = _workspace(L.ypp_map_workspace_bytes(G, K, nc), dev, "map")
        check(L.ypp_map_flags(_ptr(gt), G, _ptr(det), K, nc, float(iou_threshold), _ptr(ws), ws.numel(), _ptr(tp),
                              _ptr(triples), _ptr(gpc), _stream()))
    return triples, gpc, (tp[:K] if tp is not None else None)


def map_ap_from_flags(triples: torch.Tensor, gt_per_class: torch.Tensor, num_classes, c_lo=0, c_hi=None,
                      seg_count: Optional[torch.Tensor] = None, per_class: Optional[torch.Tensor] = None) -> torch.Tensor:
    """M2 (ypp_map_ap_from_flags) for the classes [c_lo, c_hi): triples [n,3] or the padded all-gather buffer
    [n_seg, seg_stride, 3] with seg_count int64 [n_seg] valid rows each; gt_per_class int64 [nc] (global counts).
    Returns per_class float32 [nc,4] with the rows of the owned classes filled in (AP, TP, FP, FN)."""
    nc = int(num_classes)
    c_hi = nc if c_hi is None else int(c_hi)
    if not triples.is_cuda or triples.dtype != torch.float32 or triples.shape[-1] != 3:
        raise ValueError("triples must be a float32 CUDA tensor [..., 3]")
    t = triples.contiguous()
    if t.dim() == 2:
        n_seg, stride = 1, t.shape[0]
    else:
        n_seg, stride = t.shape[0], t.shape[1]
    dev = t.device
    gpc = gt_per_class.to(device=dev, dtype=torch.int64).contiguous()
    sc = None if seg_count is None else seg_count.to(device=dev, dtype=torch.int64).contiguous()
    if per_class is None:
        per_class = torch.zeros((nc, 4), dtype=torch.float32, device=dev)
    L = lib()
    with _on(dev):
        ws = _workspace(L.ypp_map_workspace_bytes(0, n_seg * stride, nc), dev, "map")
        check(L.ypp_map_ap_from_flags(_ptr(t), n_seg, stride, _ptr(sc), _ptr(gpc), nc, int(c_lo), c_hi, _ptr(ws), ws.numel(),
                                      _ptr(per_class), _stream()))
    return per_class


def map_mean(per_class: torch.Tensor) -> torch.Tensor:
    """mean AP over all classes of per_class [nc,4], fp32 in class order -> float32 [1]."""
    pc = per_class.to(torch.float32).contiguous()
    m = torch.empty(1, dtype=torch.float32, device=pc.device)
    with _on(pc.device):
        check(lib().ypp_map_mean(_ptr(pc), pc.shape[0], _ptr(m), _stream()))
    return m


def detection_ints(rows: torch.Tensor, mode="pred_file", scale_xy: Optional[torch.Tensor] = None, img_idx_base=0) -> torch.Tensor:
    """Integer box columns of the reference's detection writers for rows [K,6|7] -> int32 [K,4].
    pred_file: xmin,ymin,xmax,ymax (make_pred_file_yolov3.py:73-86); coco: xmin,ymin,w,h after scaling by
    scale_xy[img] = fp32(width / input_size), fp32(height / input_size) (yolo2coco_pred_file_yolov3.py:83-100)."""
    from ._lib import WRITER_MODES
    if rows.dim() != 2 or rows.shape[1] not in (6, 7):
        raise ValueError("rows must be [K,6] or [K,7]")
    if not rows.is_cuda or rows.dtype != torch.float32:
        raise ValueError("rows must be a float32 CUDA tensor")
    rows = rows.contiguous()
    K = rows.shape[0]
    out = torch.empty((K, 4), dtype=torch.int32, device=rows.device)
    n_img = 0
    if scale_xy is not None:
        scale_xy = scale_xy.to(device=rows.device, dtype=torch.float32).reshape(-1, 2).contiguous()
        n_img = scale_xy.shape[0]
    with _on(rows.device):
        check(lib().ypp_detection_ints(_ptr(rows), K, rows.shape[1], WRITER_MODES[mode], _ptr(scale_xy), n_img,
                                       float(img_idx_base), _ptr(out), _stream()))
    return out


# ------------------------------------------------------------------------------------ alternate semantics

def nms_ultralytics(prediction: torch.Tensor, conf_thres=0.25, iou_thres=0.45, classes=None, agnostic=False,
                    multi_label=False, max_det=300, max_nms=30000, max_wh=4096, union_mode="fma",
                    compare_in_double=None, row_width=6, img_idx_base=0, want_keep_idx=False) -> PostResult:
    """The vendored Ultralytics non_max_suppression (utils/general.py:640-731) for a whole batch:
    prediction [B, n, 5+nc] -> packed rows x1,y1,x2,y2,conf,cls.  union_mode "fma" = what the reference
    computes on a CUDA device (torchvision's CUDA nms), "separate" = on the CPU."""
    from ._lib import YppUltraCfg
    if prediction.dim() != 3 or prediction.shape[2] < 6:
        raise ValueError("prediction must be [B, n, 5+nc]")
    if not prediction.is_cuda or prediction.dtype != torch.float32:
        raise ValueError("prediction must be a float32 CUDA tensor (ypp has no CPU path)")
    if not (0 <= conf_thres <= 1) or not (0 <= iou_thres <= 1):         # the reference's asserts (:648-649)
        raise AssertionError("conf_thres and iou_thres must be in [0, 1]")
    pred = prediction.contiguous()
    B, n, C = pred.shape
    nc = C - 5
    dev = pred.device
    um = _lib.UNION_MODES[union_mode] if isinstance(union_mode, str) else int(union_mode)
    if compare_in_double is None:
        compare_in_double = um == _lib.UNION_SEPARATE
    L = lib()
    rows = int(L.ypp_ultra_rows(n, nc, int(bool(multi_label)))) if n else 0
    z = torch.zeros
    if B == 0 or n == 0:
        return PostResult(z((0, row_width), device=dev), z(1, dtype=torch.int32, device=dev),
                          z(B, dtype=torch.int32, device=dev), z(B, dtype=torch.int32, device=dev),
                          z((B, 0), dtype=torch.int32, device=dev) if want_keep_idx else None, 0)
    allow = None
    if classes is not None:
        allow = torch.zeros(nc, dtype=torch.uint8)
        for c in classes:
            if float(c) == int(c) and 0 <= int(c) < nc:
                allow[int(c)] = 1
        allow = allow.to(dev)
    cfg = YppUltraCfg(nc, B, float(conf_thres), float(L.ypp_effective_iou_thr(float(iou_thres), int(bool(compare_in_double)))),
                      int(bool(multi_label)), int(bool(agnostic)), int(max_det), int(max_nms), float(max_wh),
                      None if allow is None else allow.data_ptr(), um)
    per_img = min(rows, int(max_det)) if max_det > 0 else rows
    res = PostResult(det_rows=torch.empty((B * per_img, row_width), dtype=torch.float32, device=dev),
                     det_total=torch.empty(1, dtype=torch.int32, device=dev),
                     keep_count=torch.empty(B, dtype=torch.int32, device=dev),
                     cand_count=torch.empty(B, dtype=torch.int32, device=dev),
                     keep_idx=torch.empty((B, rows), dtype=torch.int32, device=dev) if want_keep_idx else None,
                     n_boxes=rows)
    ws = _workspace(L.ypp_workspace_bytes(B, rows, nc), dev, "post")
    with _on(dev):
        check(L.ypp_nms_ultralytics(ctypes.byref(cfg), _ptr(pred), n, _ptr(ws), ws.numel(), _ptr(res.cand_count),
                                    _ptr(res.keep_count), _ptr(res.keep_idx), _ptr(res.det_rows), row_width,
                                    res.det_rows.shape[0], _ptr(res.det_total), float(img_idx_base), _stream()))
    res._keepalive = allow
    return res


def decode_yolov1(inp: torch.Tensor, num_classes, num_boxes, input_size, S=7) -> torch.Tensor:
    """utils/yolov1_utils.py:12-66: [B, S*S*(5*num_boxes+nc)] -> [B, S*S, 6] cx,cy,w,h,confidence,class."""
    if not inp.is_cuda or inp.dtype != torch.float32:
        raise ValueError("input must be a float32 CUDA tensor (ypp has no CPU path)")
    per = S * S * (5 * int(num_boxes) + int(num_classes))
    x = inp.contiguous().view(-1, per)
    B = x.shape[0]
    out = torch.empty((B, S * S, 6), dtype=torch.float32, device=x.device)
    if B:
        with _on(x.device):
            check(lib().ypp_decode_yolov1(_ptr(x), B, S, int(num_classes), int(num_boxes), float(input_size), _ptr(out), _stream()))
    return out


def confusion_update(matrix: torch.Tensor, detections: torch.Tensor, det_off: torch.Tensor, labels: torch.Tensor,
                     lab_off: torch.Tensor, nc, conf=0.25, iou_thres=0.45) -> torch.Tensor:
    """ConfusionMatrix.process_batch (utils/metrics.py:122-160) for len(det_off)-1 images in one launch.
    matrix int64 [(nc+1),(nc+1)] on the device, accumulated in place."""
    det = detections.to(torch.float32).contiguous()
    lab = labels.to(torch.float32).contiguous()
    dev = matrix.device
    if matrix.dtype != torch.int64 or not matrix.is_cuda or not matrix.is_contiguous() or matrix.numel() != (nc + 1) * (nc + 1):
        raise ValueError("matrix must be a contiguous int64 CUDA tensor [(nc+1),(nc+1)]")
    d_off = det_off.to(device=dev, dtype=torch.int32).contiguous()
    l_off = lab_off.to(device=dev, dtype=torch.int32).contiguous()
    n_img = d_off.numel() - 1
    K = det.shape[0]
    scratch = _workspace(max(K, 1) * 8, dev, "cm")
    with _on(dev):
        check(lib().ypp_confusion_update(_ptr(det), _ptr(d_off), _ptr(lab), _ptr(l_off), n_img, K, int(nc), float(conf),
                                         float(iou_thres), _ptr(scratch), scratch.numel(), _ptr(matrix), _stream()))
    return matrix


def launch_count(reset=False) -> int:
    return int(lib().ypp_launch_count(int(bool(reset))))
