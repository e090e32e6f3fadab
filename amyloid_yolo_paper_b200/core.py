"""``mergeDetections`` with the reference's signature (``core.py:366-423``).

Same-class (0 = CAA, 1 = Cored) overlapping boxes are replaced by their integer
hull until nothing changes; values and row order equal the reference's (python
``set`` order included, see ``csrc/merge.cu``).  The rest of the reference's
``core.py`` is paper analysis and out of scope.
"""
import ctypes

import torch

from . import _lib
from .utils.utils import _device_of, _ptr, _stream


def merge_detections_device(rows, counts, current_dim, original_shape, merge=True):
    """rows [B,cap,7] / counts [B] on the GPU (as produced by
    ``non_max_suppression_device``) -> rescaled (+merged) rows and counts."""
    B, cap, _ = rows.shape
    dev = rows.device
    out = torch.empty_like(rows)
    out_counts = torch.empty_like(counts)
    lib, ctx = _lib.load(), _lib.context(dev.index)
    with torch.cuda.device(dev):
        _lib.check(lib.ay_rescale_merge(ctx, _ptr(rows), _ptr(counts), B, cap, int(current_dim),
                                        int(original_shape[0]), int(original_shape[1]), 1 if merge else 0,
                                        _ptr(out), _ptr(out_counts), _stream(dev)))
    return out, out_counts


def mergeDetections(detections):
    """[k,7] tensor (x1,y1,x2,y2,conf,cls_conf,cls_pred) -> new [k',7] fp32 tensor."""
    dev = _device_of(detections)
    k = detections.shape[0]
    rows = detections.detach().to(dev, torch.float32).reshape(1, k, 7).contiguous()
    counts = torch.tensor([k], dtype=torch.int32, device=dev)
    # identity rescale: (x - 0) / 1 * 1 is exact in fp32
    out, out_counts = merge_detections_device(rows, counts, 1, (1, 1), merge=True)
    lib, ctx = _lib.load(), _lib.context(dev.index)
    with torch.cuda.device(dev):
        _lib.check(lib.ay_check(ctx, _stream(dev)))
    n = int(out_counts.item())
    return out[0, :n].to(detections.device).clone()
