"""Drop-in for utils/metrics.py.

    box_iou            utils/metrics.py:233-255    N x M kernel
    bbox_iou           utils/metrics.py:192-230    element-wise kernel (IoU / GIoU / DIoU / CIoU)
    ConfusionMatrix    utils/metrics.py:113-160    process_batch as a kernel
    compute_ap         utils/metrics.py:86-111     host, float64 (a curve of at most a few thousand points: SURVEY 8(a) X)
    ap_per_class       utils/metrics.py:21-83      host, float64; fed by the TP flags the GPU matching produces
    ap_per_class_from_rows                         the two together: ypp_map_match's flags -> ap_per_class
The plots (plot=True, save_dir, plot_pr_curve, ...) need matplotlib and a display and are not provided.
"""
import numpy as np
import torch

from .. import engine
from ..engine import box_iou, bbox_iou  # noqa: F401

_trapezoid = getattr(np, "trapezoid", None) or np.trapz


def _host(a):
    """numpy view of an array-like; CUDA tensors are copied to the host (one synchronising copy)."""
    if isinstance(a, torch.Tensor):
        return a.detach().cpu().numpy()
    return np.asarray(a)


def compute_ap(recall, precision):
    """Average precision of one recall / precision curve -> (ap, mpre, mrec): sentinel points (0, 1) and (1, 0),
    the precision replaced by its envelope (maximum of everything to the right), and the envelope integrated by
    the trapezoid rule on 101 equally spaced recall values (COCO's 101-point interpolation)."""
    mrec = np.concatenate(([0.0], _host(recall), [1.0]))
    mpre = np.concatenate(([1.0], _host(precision), [0.0]))
    mpre = np.maximum.accumulate(mpre[::-1])[::-1]
    grid = np.linspace(0, 1, 101)
    return _trapezoid(np.interp(grid, mrec, mpre), grid), mpre, mrec


def ap_per_class(tp, conf, pred_cls, target_cls, plot=False, save_dir=".", names=()):
    """tp [n, T] (T IoU thresholds; [n] is taken as T = 1), conf [n], pred_cls [n], target_cls [m] (numpy arrays or
    tensors on any device) -> (p, r, ap [nc, T], f1, classes int32) over the classes that occur in target_cls,
    p / r / f1 taken at the confidence that maximises the mean F1.  float64 on the host, as the reference."""
    if plot:
        raise NotImplementedError("ap_per_class(plot=True): the curve plots are not provided")
    tp, conf, pred_cls, target_cls = _host(tp), _host(conf), _host(pred_cls), _host(target_cls)
    if tp.ndim == 1:
        tp = tp[:, None]
    order = np.argsort(-conf)                         # numpy's default sort: ties land where the reference's land
    tp, conf, pred_cls = tp[order], conf[order], pred_cls[order]
    classes = np.unique(target_cls)
    n_cls, T = classes.shape[0], tp.shape[1]
    px = np.linspace(0, 1, 1000)
    ap, p, r = np.zeros((n_cls, T)), np.zeros((n_cls, 1000)), np.zeros((n_cls, 1000))
    labels_of = {c: int(np.count_nonzero(target_cls == c)) for c in classes}
    for ci, c in enumerate(classes):
        mine = pred_cls == c
        n_lab = labels_of[c]
        if n_lab == 0 or not mine.any():
            continue
        hits = tp[mine]
        tpc = hits.cumsum(0)
        fpc = (1 - hits).cumsum(0)
        recall = tpc / (n_lab + 1e-16)
        precision = tpc / (tpc + fpc)
        neg_conf = -conf[mine]                        # increasing, as np.interp needs
        r[ci] = np.interp(-px, neg_conf, recall[:, 0], left=0)
        p[ci] = np.interp(-px, neg_conf, precision[:, 0], left=1)
        for j in range(T):
            ap[ci, j] = compute_ap(recall[:, j], precision[:, j])[0]
    f1 = 2 * p * r / (p + r + 1e-16)
    best = f1.mean(0).argmax()
    return p[:, best], r[:, best], ap, f1[:, best], classes.astype("int32")


def ap_per_class_from_rows(true_boxes, pred_boxes, num_classes, iou_threshold=0.5):
    """ap_per_class fed by the GPU matching: true_boxes [G,7] / pred_boxes [K,7] rows img,cx,cy,w,h,score,cls (the
    accumulators of MeanAveragePrecision) -> TP flags from ypp_map_match (the reference's greedy rule,
    utils/yolo_utils.py:404-430, at one IoU threshold) -> (p, r, ap [nc,1], f1, classes).  Detections of classes
    without ground truth count as false positives, as they would coming out of a matcher."""
    _, _, tp = engine.map_match(true_boxes, pred_boxes, num_classes, iou_threshold, want_tp=True)
    det = pred_boxes.reshape(-1, 7)
    flags = (tp == 1).to(torch.float64)[:, None]
    return ap_per_class(flags, det[:, 5].to(torch.float64), det[:, 6], true_boxes.reshape(-1, 7)[:, 6])


class ConfusionMatrix:
    """Same constructor, process_batch(detections [N,6] xyxy,conf,cls; labels [M,5] cls,xyxy) and .matrix
    ((nc+1) x (nc+1), predicted x true, last index = background).  The counts live on the device and
    process_batch never synchronises; .matrix copies them out.  process_images() takes a whole batch."""

    def __init__(self, nc, conf=0.25, iou_thres=0.45, device=None):
        self.nc = nc
        self.conf = conf
        self.iou_thres = iou_thres
        self._dev = torch.device(device) if device is not None else None
        self._m = None

    def _matrix(self, dev):
        if self._m is None:
            self._dev = dev
            self._m = torch.zeros((self.nc + 1, self.nc + 1), dtype=torch.int64, device=dev)
        return self._m

    def process_batch(self, detections, labels):
        off_d = torch.tensor([0, detections.shape[0]], dtype=torch.int32)
        off_l = torch.tensor([0, labels.shape[0]], dtype=torch.int32)
        self.process_images(detections, off_d, labels, off_l)

    def process_images(self, detections, det_off, labels, lab_off):
        """detections [K,6] and labels [G,5] of several images, row offsets [n_img+1] each."""
        if not detections.is_cuda:
            raise ValueError("detections must live on a CUDA device (ypp has no CPU path)")
        m = self._matrix(detections.device)
        engine.confusion_update(m, detections, det_off, labels.to(detections.device), lab_off, self.nc, self.conf, self.iou_thres)

    @property
    def matrix(self):
        if self._m is None:
            return np.zeros((self.nc + 1, self.nc + 1))
        return self._m.cpu().numpy().astype(np.float64)
