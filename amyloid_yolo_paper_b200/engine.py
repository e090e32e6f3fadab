"""Host-side owner of one network instance inside the sm_100a engine.

Turns the cfg block list (``utils/parse_config.py``) and a state_dict in the
reference's naming (``models.py:31-44``) into the engine's layer plan
(``ay_net_build``), folds BatchNorm into the convolution weights
(``w' = w * gamma / sqrt(var + 1e-5)``, ``b' = beta - mean * gamma / sqrt(var + 1e-5)``,
the eval-mode algebra of ``models.py:40-45``) and uploads them, then runs
forwards on torch-allocated device buffers.  PyTorch is used for memory and
streams only; every kernel launched here lives in ``csrc/``.
"""
import ctypes

import numpy as np
import torch

from . import _lib


def _absolute(index, current):
    """Route/shortcut indices: negative = relative to the current layer
    (``layer_outputs[int(i)]`` at the time layer ``current`` runs, models.py:245-248)."""
    index = int(index)
    return current + index if index < 0 else index


def build_layer_descs(module_defs):
    descs = (_lib.LayerDesc * len(module_defs))()
    for i, d in enumerate(module_defs):
        L = descs[i]
        kind = d["type"]
        if kind == "convolutional":
            L.type = _lib.AY_CONV
            L.filters = int(d["filters"])
            L.size = int(d["size"])
            L.stride = int(d["stride"])
            L.leaky = 1 if d["activation"] == "leaky" else 0
            if d["activation"] not in ("leaky", "linear"):
                raise ValueError("layer %d: activation %r is not supported" % (i, d["activation"]))
        elif kind == "maxpool":
            L.type = _lib.AY_MAXPOOL
            L.size = int(d["size"])
            L.stride = int(d["stride"])
        elif kind == "upsample":
            L.type = _lib.AY_UPSAMPLE
            L.stride = int(d["stride"])
        elif kind == "route":
            L.type = _lib.AY_ROUTE
            srcs = [_absolute(v, i) for v in d["layers"].split(",")]
            L.n_src = len(srcs)
            for k, s in enumerate(srcs):
                L.src[k] = s
        elif kind == "shortcut":
            L.type = _lib.AY_SHORTCUT
            L.n_src = 1
            L.src[0] = _absolute(d["from"], i)
        elif kind == "yolo":
            L.type = _lib.AY_YOLO
            mask = [int(v) for v in d["mask"].split(",")]
            flat = [int(v) for v in d["anchors"].split(",")]
            pairs = [(flat[j], flat[j + 1]) for j in range(0, len(flat), 2)]
            L.n_anchors = len(mask)
            L.classes = int(d["classes"])
            for k, m in enumerate(mask):
                L.anchors[2 * k] = pairs[m][0]
                L.anchors[2 * k + 1] = pairs[m][1]
        else:
            raise ValueError("unsupported cfg block [%s] at layer %d" % (kind, i))
    return descs


def fold_batchnorm(module_defs, state, layer):
    """-> (weight OIHW fp32, bias fp32) of conv ``layer`` with BN folded in."""
    prefix = "module_list.%d." % layer
    w = state[prefix + "conv_%d.weight" % layer].detach().to("cpu", torch.float64)
    if int(module_defs[layer]["batch_normalize"]):
        bn = prefix + "batch_norm_%d." % layer
        gamma = state[bn + "weight"].detach().to("cpu", torch.float64)
        beta = state[bn + "bias"].detach().to("cpu", torch.float64)
        mean = state[bn + "running_mean"].detach().to("cpu", torch.float64)
        var = state[bn + "running_var"].detach().to("cpu", torch.float64)
        scale = gamma / torch.sqrt(var + 1e-5)
        w = w * scale.view(-1, 1, 1, 1)
        b = beta - mean * scale
    else:
        b = state[prefix + "conv_%d.bias" % layer].detach().to("cpu", torch.float64)
    return (np.ascontiguousarray(w.to(torch.float32).numpy()), np.ascontiguousarray(b.to(torch.float32).numpy()))


class Engine:
    def __init__(self, module_defs, hyperparams, state_dict, device, max_batch=8, precision="bf16", img_size=None):
        if not torch.cuda.is_available():
            raise _lib.EngineUnavailable("no CUDA device: the B200 engine has no CPU fallback")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.EngineUnavailable("the engine runs on CUDA devices only (got %s)" % device)
        self.index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", self.index)
        self.lib = _lib.load()
        self.module_defs = module_defs
        self.max_batch = int(max_batch)
        self.precision = precision
        # network input size: the cfg's [net] height unless the caller runs the network at another
        # multiple of 32 (the reference takes img_dim from the input tensor, models.py:238)
        self.img_size = int(img_size) if img_size is not None else int(hyperparams["height"])
        if self.img_size < 32 or self.img_size % 32 != 0:
            raise ValueError("network input size must be a positive multiple of 32 (got %d)" % self.img_size)
        self.in_channels = int(hyperparams["channels"])
        self.ctx = ctypes.c_void_p()
        code = self.lib.ay_create(ctypes.byref(self.ctx), self.index)
        if code != 0:
            raise _lib.EngineUnavailable(self.lib.ay_last_error().decode())
        descs = build_layer_descs(module_defs)
        dtype = {"bf16": _lib.AY_BF16, "fp16": _lib.AY_FP16, "fp32": _lib.AY_F32}[precision]
        with torch.cuda.device(self.index):
            _lib.check(self.lib.ay_net_build(self.ctx, descs, len(module_defs), self.in_channels, self.img_size,
                                             self.max_batch, dtype))
            for i, d in enumerate(module_defs):
                if d["type"] == "convolutional":
                    w, b = fold_batchnorm(module_defs, state_dict, i)
                    _lib.check(self.lib.ay_net_set_conv(self.ctx, i, w.ctypes.data_as(ctypes.c_void_p),
                                                        b.ctypes.data_as(ctypes.c_void_p)))
            _lib.check(self.lib.ay_net_finalize(self.ctx))
        rows, cols = ctypes.c_int(), ctypes.c_int()
        _lib.check(self.lib.ay_net_output_shape(self.ctx, ctypes.byref(rows), ctypes.byref(cols)))
        self.rows, self.cols = rows.value, cols.value

    # ---------------------------------------------------------------- running
    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _forward(self, inp, kind, src_h, src_w, out=None):
        batch = inp.shape[0]
        if out is None:
            out = torch.empty((batch, self.rows, self.cols), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.index):
            for lo in range(0, batch, self.max_batch):
                hi = min(batch, lo + self.max_batch)
                _lib.check(self.lib.ay_forward(self.ctx, ctypes.c_void_p(inp[lo:hi].data_ptr()), kind, hi - lo,
                                               src_h, src_w, ctypes.c_void_p(out[lo:hi].data_ptr()), self._stream()))
        return out

    def forward_nchw(self, x, out=None):
        """x: [B,3,S,S] float (any device) -> decoded [B,rows,cols] fp32 on the GPU."""
        if x.dim() != 4 or x.shape[1] != self.in_channels or x.shape[2] != self.img_size or x.shape[3] != self.img_size:
            raise ValueError("expected input [B,%d,%d,%d], got %s" % (self.in_channels, self.img_size, self.img_size,
                                                                     tuple(x.shape)))
        x = x.to(self.device, torch.float32, non_blocking=True).contiguous()
        return self._forward(x, _lib.AY_IN_F32_NCHW, self.img_size, self.img_size, out)

    def forward_u8(self, tiles, out=None):
        """tiles: [B,H,W,3] uint8 square tiles; ToTensor + nearest Resize are fused
        into the first convolution (utils/transforms.py:89-111)."""
        if tiles.dim() != 4 or tiles.shape[3] != 3 or tiles.dtype != torch.uint8 or tiles.shape[1] != tiles.shape[2]:
            raise ValueError("expected uint8 tiles [B,H,H,3], got %s %s" % (tuple(tiles.shape), tiles.dtype))
        tiles = tiles.to(self.device, non_blocking=True).contiguous()
        return self._forward(tiles, _lib.AY_IN_U8_HWC, tiles.shape[1], tiles.shape[2], out)

    def check(self):
        with torch.cuda.device(self.index):
            _lib.check(self.lib.ay_check(self.ctx, self._stream()))

    # ---------------------------------------------------------------- test hooks
    def layer_output(self, layer, batch):
        c, h, w = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        _lib.check(self.lib.ay_net_layer_shape(self.ctx, layer, ctypes.byref(c), ctypes.byref(h), ctypes.byref(w)))
        out = torch.empty((batch, c.value, h.value, w.value), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.index):
            _lib.check(self.lib.ay_net_layer_output(self.ctx, layer, batch, ctypes.c_void_p(out.data_ptr()),
                                                    self._stream()))
        return out

    def layer_impl(self, layer):
        return self.lib.ay_net_layer_impl(self.ctx, layer)

    def launch_count(self):
        return self.lib.ay_net_launch_count(self.ctx)

    def use_reference_conv(self, enable):
        _lib.check(self.lib.ay_net_use_reference_conv(self.ctx, 1 if enable else 0))

    def close(self):
        if getattr(self, "ctx", None) is not None and self.ctx.value:
            self.lib.ay_destroy(self.ctx)
            self.ctx = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
