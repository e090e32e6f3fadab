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
