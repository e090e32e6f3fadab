"""Box utilities and NMS with the reference's signatures (``utils/utils.py``).

``non_max_suppression`` keeps the reference's conventions — it rewrites
``prediction[..., :4]`` from xywh to xyxy *in place* (``utils/utils.py:244``) and
returns a python list with one ``[k,7]`` fp32 CPU tensor (or ``None``) per image —
but the work runs in the CUDA kernels of ``csrc/nms.cu`` and returns the same
bits as the reference's python loop.  ``non_max_suppression_device`` is the
allocation-free variant used by the fused pipeline.
"""
import ctypes

import torch

from .. import _lib


def to_cpu(tensor):
    """utils/utils.py:14-15."""
    return tensor.detach().cpu()


def load_classes(path):
    """utils/utils.py:18-24."""
    with open(path, "r") as fh:
        return fh.read().split("\n")[:-1]


def _device_of(t):
    if t.is_cuda:
        return t.device
    if not torch.cuda.is_available():
        raise _lib.EngineUnavailable("no CUDA device: the B200 kernels have no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def xywh2xyxy(x):
    """utils/utils.py:53-59 (pure tensor algebra, stays where ``x`` lives)."""
    y = x.new(x.shape)
    y[..., 0] = x[..., 0] - x[..., 2] / 2
    y[..., 1] = x[..., 1] - x[..., 3] / 2
    y[..., 2] = x[..., 0] + x[..., 2] / 2
    y[..., 3] = x[..., 1] + x[..., 3] / 2
    return y


def rescale_boxes(boxes, current_dim, original_shape):
    """utils/utils.py:36-50: in place on columns 0..3, returns its argument.
    Runs ``ay_rescale_merge`` with merging switched off."""
    dev = _device_of(boxes)
    k = boxes.shape[0]
    if k == 0:
        return boxes
    rows = torch.zeros((1, k, 7), dtype=torch.float32, device=dev)
    rows[0, :, :4] = boxes[:, :4].to(dev, torch.float32)
    counts = torch.tensor([k], dtype=torch.int32, device=dev)
    out = torch.empty_like(rows)
    out_counts = torch.empty_like(counts)
    lib, ctx = _lib.load(), _lib.context(dev.index)
    with torch.cuda.device(dev):
        _lib.check(lib.ay_rescale_merge(ctx, _ptr(rows), _ptr(counts), 1, k, int(current_dim),
                                        int(original_shape[0]), int(original_shape[1]), 0, _ptr(out),
                                        _ptr(out_counts), _stream(dev)))
        _lib.check(lib.ay_check(ctx, _stream(dev)))
    boxes[:, :4] = out[0, :, :4].to(boxes.device, boxes.dtype)
    return boxes


def bbox_iou(box1, box2, x1y1x2y2=True):
    """utils/utils.py:202-232 (``+1`` pixel convention), computed by ``ay_bbox_iou``."""
    if not x1y1x2y2:
        box1, box2 = xywh2xyxy(box1), xywh2xyxy(box2)
    dev = _device_of(box2)
    b1 = box1[:, :4].to(dev, torch.float32).contiguous()
    b2 = box2[:, :4].to(dev, torch.float32).contiguous()
    n1, n2 = b1.shape[0], b2.shape[0]
    if n2 == 1 and n1 > 1:
        b1, b2, n1, n2 = b2, b1, n2, n1            # IoU is symmetric
    out = torch.empty((n2,), dtype=torch.float32, device=dev)
    lib, ctx = _lib.load(), _lib.context(dev.index)
    with torch.cuda.device(dev):
        _lib.check(lib.ay_bbox_iou(ctx, _ptr(b1), n1, _ptr(b2), n2, _ptr(out), _stream(dev)))
    return out.to(box2.device)


def non_max_suppression_device(pred, conf_thres, nms_thres, cap=None):
    """pred: CUDA fp32 [B,N,5+C], rewritten to xyxy in place.
    -> (rows [B,cap,7], keep [B,cap] int32, counts [B] int32), all on the GPU."""
    if not (pred.is_cuda and pred.dtype == torch.float32 and pred.is_contiguous() and pred.dim() == 3):
        raise ValueError("expected a contiguous CUDA fp32 [B,N,5+C] tensor")
    B, N, cols = pred.shape
    cap = int(cap or N)
    dev = pred.device
    rows = torch.empty((B, cap, 7), dtype=torch.float32, device=dev)
    keep = torch.empty((B, cap), dtype=torch.int32, device=dev)
    counts = torch.empty((B,), dtype=torch.int32, device=dev)
    lib, ctx = _lib.load(), _lib.context(dev.index)
    with torch.cuda.device(dev):
        _lib.check(lib.ay_nms(ctx, _ptr(pred), B, N, cols - 5, float(conf_thres), float(nms_thres), _ptr(rows),
                              _ptr(keep), _ptr(counts), cap, _stream(dev)))
    return rows, keep, counts


def non_max_suppression(prediction, conf_thres=0.5, nms_thres=0.4, return_keep=False):
    """utils/utils.py:235-273.  Returns detections with shape
    (x1, y1, x2, y2, object_conf, class_score, class_pred) per image, ``None``
    where nothing passes ``conf_thres``; ``prediction[..., :4]`` becomes xyxy."""
    dev = _device_of(prediction)
    work = prediction.detach().to(dev, torch.float32).contiguous()
    if work.data_ptr() == prediction.data_ptr():
        work = work.clone()
    rows, keep, counts = non_max_suppression_device(work, conf_thres, nms_thres)
    lib, ctx = _lib.load(), _lib.context(dev.index)
    with torch.cuda.device(dev):
        _lib.check(lib.ay_check(ctx, _stream(dev)))
    prediction[..., :4] = work[..., :4].to(prediction.device, prediction.dtype)
    counts_h = counts.cpu().tolist()
    kmax = max(counts_h) if counts_h else 0
    rows_h = rows[:, :max(kmax, 1)].cpu()
    keep_h = keep[:, :max(kmax, 1)].cpu()
    output, kept = [], []
    for b, k in enumerate(counts_h):
        output.append(rows_h[b, :k].clone() if k > 0 else None)
        kept.append(keep_h[b, :k].to(torch.int64) if k > 0 else None)
    return (output, kept) if return_keep else output


# ----------------------------------------------------------------------------- evaluation statistics
def xywh2xyxy_np(x):
    """utils/utils.py:60-66."""
    import numpy as np
    y = np.zeros_like(x)
    half_w, half_h = x[..., 2] / 2, x[..., 3] / 2
    y[..., 0], y[..., 1] = x[..., 0] - half_w, x[..., 1] - half_h
    y[..., 2], y[..., 3] = x[..., 0] + half_w, x[..., 1] + half_h
    return y


def compute_ap(recall, precision):
    """Area under the precision envelope (utils/utils.py:123-149, py-faster-rcnn's AP): sentinels at both ends,
    running maximum from the right, sum of (delta recall) * precision where recall changes."""
    import numpy as np
    mrec = np.concatenate(([0.0], recall, [1.0]))
    mpre = np.concatenate(([0.0], precision, [0.0]))
    mpre = np.maximum.accumulate(mpre[::-1])[::-1]
    step = np.nonzero(mrec[1:] != mrec[:-1])[0]
    return np.sum((mrec[step + 1] - mrec[step]) * mpre[step + 1])


def ap_per_class(tp, conf, pred_cls, target_cls):
    """utils/utils.py:69-120: per class present in ``target_cls`` -> (precision, recall, AP, F1, classes) with the
    precision / recall of the LAST point of each curve, predictions ranked by descending ``conf``."""
    import numpy as np
    order = np.argsort(-conf)
    tp, conf, pred_cls = tp[order], conf[order], pred_cls[order]
    target_cls = np.asarray(target_cls)
    classes = np.unique(target_cls)
    ap, p, r = [], [], []
    for c in classes:
        mine = pred_cls == c
        n_gt, n_p = (target_cls == c).sum(), mine.sum()
        if n_p == 0 and n_gt == 0:
            continue
        if n_p == 0 or n_gt == 0:
            ap.append(0), r.append(0), p.append(0)
            continue
        tpc = tp[mine].cumsum()
        fpc = (1 - tp[mine]).cumsum()
        recall_curve = tpc / (n_gt + 1e-16)
        precision_curve = tpc / (tpc + fpc)
        r.append(recall_curve[-1])
        p.append(precision_curve[-1])
        ap.append(compute_ap(recall_curve, precision_curve))
    p, r, ap = np.array(p), np.array(r), np.array(ap)
    f1 = 2 * p * r / (p + r + 1e-16)
    return p, r, ap, f1, classes.astype("int32")


def get_batch_statistics(outputs, targets, iou_threshold):
    """utils/utils.py:152-190: per image with detections -> [true_positives, scores, labels].  A prediction (in NMS
    order) is a true positive when its label occurs among the image's targets and its best-IoU target (first index on
    ties, IoU >= iou_threshold) has not been claimed yet; the scan stops once every target is claimed.
    ``outputs``: list of [k,7] tensors / None in network-input pixels; ``targets``: [n,6] (sample, class, x1,y1,x2,y2).
    The IoUs of one image come from ONE ``ay_bbox_iou`` launch over all (prediction, target) pairs."""
    import numpy as np
    batch_metrics = []
    for sample_i, output in enumerate(outputs):
        if output is None:
            continue
        output = output.detach().cpu()
        pred_boxes, pred_scores, pred_labels = output[:, :4], output[:, 4], output[:, -1]
        true_positives = np.zeros(pred_boxes.shape[0])
        annotations = targets[targets[:, 0] == sample_i][:, 1:]
        if len(annotations):
            target_labels = annotations[:, 0]
            target_boxes = annotations[:, 1:].to(torch.float32)
            n_p, n_t = pred_boxes.shape[0], target_boxes.shape[0]
            iou = bbox_iou(pred_boxes.to(torch.float32).repeat_interleave(n_t, 0), target_boxes.repeat(n_p, 1)).view(n_p, n_t)
            best_iou, best_idx = iou.max(1)
            known = set(target_labels.tolist())
            claimed = set()
            for pred_i in range(n_p):
                if len(claimed) == n_t:
                    break
                if float(pred_labels[pred_i]) not in known:
                    continue
                j = int(best_idx[pred_i])
                if float(best_iou[pred_i]) >= iou_threshold and j not in claimed:
                    true_positives[pred_i] = 1
                    claimed.add(j)
        batch_metrics.append([true_positives, pred_scores, pred_labels])
    return batch_metrics
