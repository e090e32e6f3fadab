"""The consensus CAA filter of the reference on the sm_100a engine.

``filterDetectionsByCAAModel`` is the step right after ``mergeDetections`` in every production caller of the
reference (``validation.py:129,345``, ``prospective.py:87``): every distinct detection of a 1536^2 tile is cropped to
the 256^2 window centred on it (``get256Img``, ``core.py:109-159``), a small CNN (``Net``, ``core.py:161-205``: six
Conv3x3-BatchNorm-ReLU-MaxPool stages and one fully connected layer) classifies the crop, and a box labelled CAA
whose sigmoid CAA score is <= 0.5 is dropped (``core.py:425-452``).

Here the network runs inside the same engine as the detector: the convolutions are the tcgen05 implicit-GEMM kernel
with a ReLU epilogue (``leaky = 2``), the pooling is the engine's 2x2 max-pool, the classifier is the ``AY_LINEAR``
output layer, and all crops of a tile go through in one batch instead of one ``model(...)`` call per box.
The crops are cut and normalised by one gather kernel (``ay_crop_normalize``).  PyTorch only allocates the buffers;
there is no CPU or PyTorch fallback.
"""
import ctypes

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from .engine import Engine

CHANNELS = (3, 16, 32, 48, 64, 80, 96)
CROP = 256


class Net(nn.Module):
    """Drop-in for ``core.Net`` (``core.py:161-205``): same constructor, the same ``features`` / ``classifier``
    layout and parameter names, so the reference's pickled model's ``state_dict()`` loads unchanged.  The modules
    only hold the parameters; ``forward`` runs the engine and returns the logits ``[B, num_classes]``."""

    def __init__(self, fc_nodes=512, num_classes=3, dropout=0.5, precision="fp32"):
        super().__init__()
        self.drop = 0.2
        stages = []
        for cin, cout in zip(CHANNELS[:-1], CHANNELS[1:]):
            stages += [nn.Conv2d(cin, cout, 3, padding=1), nn.BatchNorm2d(cout), nn.ReLU(inplace=True),
                       nn.MaxPool2d(kernel_size=2, stride=2)]
        self.features = nn.Sequential(*stages)
        self.classifier = nn.Sequential(nn.Linear(96 * 4 * 4, num_classes))
        self.train_loss_curve, self.dev_loss_curve, self.train_auprc, self.dev_auprc = [], [], [], []
        self.num_classes = num_classes
        # "fp32" (default: the gate is a threshold on the logit; measured 1e-6 from the reference's logits), "fp16" (the fast
        # 16-bit tensor-core path at 1.3e-3), "bf16" (same speed, 1e-2)
        self.precision = precision
        self._engine = None
        self._engine_key = None
        self.eval()

    def _key(self, device, max_batch):
        return (str(device), int(max_batch), self.precision,
                sum(int(t._version) for t in self.state_dict().values()), id(self.features[0].weight))

    def engine(self, device=None, max_batch=64):
        """The engine instance holding this model's weights on ``device`` (rebuilt when parameters change)."""
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else "cuda"
        key = self._key(device, max_batch)
        if self._engine is None or self._engine_key != key:
            self._engine = CAAEngine(self.state_dict(), device, max_batch=max_batch, precision=self.precision,
                                     num_classes=self.num_classes)
            self._engine_key = key
        return self._engine

    def load_state_dict(self, *args, **kwargs):
        out = super().load_state_dict(*args, **kwargs)
        self._engine = None
        return out

    def forward(self, x):
        """x: [B,3,256,256] float, normalised as ``core.py:435`` does -> logits [B, num_classes] on x's device."""
        if self.training:
            raise NotImplementedError("training the CAA filter is out of scope; call .eval()")
        dev = x.device if x.is_cuda else None
        eng = self.engine(dev)
        out = eng.forward_nchw(x).reshape(x.shape[0], -1)
        eng.check()
        return out.to(x.device)


def fold_stage(state, k):
    """Conv2d (with bias) + eval-mode BatchNorm2d of stage ``k`` as one convolution, in float64:
    w' = w * g / sqrt(var + eps), b' = (b - mean) * g / sqrt(var + eps) + beta."""
    c, b = "features.%d." % (4 * k), "features.%d." % (4 * k + 1)
    f64 = lambda name: state[name].detach().to("cpu", torch.float64)          # noqa: E731
    scale = f64(b + "weight") / torch.sqrt(f64(b + "running_var") + 1e-5)
    w = f64(c + "weight") * scale.view(-1, 1, 1, 1)
    bias = (f64(c + "bias") - f64(b + "running_mean")) * scale + f64(b + "bias")
    return (np.ascontiguousarray(w.to(torch.float32).numpy()), np.ascontiguousarray(bias.to(torch.float32).numpy()))


class CAAEngine(Engine):
    """Layer plan of ``core.Net`` inside the engine: 6 x (AY_CONV 3x3 with ReLU, AY_MAXPOOL 2/2) + AY_LINEAR.
    ``forward_nchw`` returns the logits as ``[B, 1, num_classes]``."""

    def __init__(self, state_dict, device, max_batch=64, precision="fp32", num_classes=3):
        if not torch.cuda.is_available():
            raise _lib.EngineUnavailable("no CUDA device: the B200 engine has no CPU fallback")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.EngineUnavailable("the engine runs on CUDA devices only (got %s)" % device)
        self.index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", self.index)
        self.lib = _lib.load()
        self.module_defs = None
        self.max_batch = int(max_batch)
        self.precision = precision
        self.img_size = CROP
        self.in_channels = 3
        n = 13
        # fp32 mode: the tensor-core convolution (3 x tf32) needs channel counts that are multiples of 32; 16, 48 and 80
        # are padded with zero filters (ReLU(0 + 0) = 0 feeds zero weights of the next stage: the same values, and stages
        # 2-6 leave the CUDA-core kernel).  The 16-bit path pads its buffers to 32 channels by itself.
        chans = list(CHANNELS)
        if precision == "fp32":
            chans = [3] + [-(-c // 32) * 32 for c in CHANNELS[1:]]
        descs = (_lib.LayerDesc * n)()
        for k in range(6):
            conv, pool = descs[2 * k], descs[2 * k + 1]
            conv.type, conv.filters, conv.size, conv.stride, conv.leaky = _lib.AY_CONV, chans[k + 1], 3, 1, 2
            pool.type, pool.size, pool.stride = _lib.AY_MAXPOOL, 2, 2
        descs[12].type, descs[12].filters = _lib.AY_LINEAR, int(num_classes)
        self.ctx = ctypes.c_void_p()
        if self.lib.ay_create(ctypes.byref(self.ctx), self.index) != 0:
            raise _lib.EngineUnavailable(self.lib.ay_last_error().decode())
        dtype = {"bf16": _lib.AY_BF16, "fp16": _lib.AY_FP16, "fp32": _lib.AY_F32}[precision]
        P = lambda a: a.ctypes.data_as(ctypes.c_void_p)                        # noqa: E731
        with torch.cuda.device(self.index):
            _lib.check(self.lib.ay_net_build(self.ctx, descs, n, 3, CROP, self.max_batch, dtype))
            for k in range(6):
                w, b = fold_stage(state_dict, k)
                if w.shape[:2] != (chans[k + 1], chans[k]):
                    wp = np.zeros((chans[k + 1], chans[k], 3, 3), dtype=np.float32)
                    wp[:w.shape[0], :w.shape[1]] = w
                    bp = np.zeros((chans[k + 1],), dtype=np.float32)
                    bp[:b.shape[0]] = b
                    w, b = wp, bp
                _lib.check(self.lib.ay_net_set_conv(self.ctx, 2 * k, P(w), P(b)))
            w = np.ascontiguousarray(state_dict["classifier.0.weight"].detach().to("cpu", torch.float32).numpy())
            b = np.ascontiguousarray(state_dict["classifier.0.bias"].detach().to("cpu", torch.float32).numpy())
            if w.shape != (int(num_classes), 96 * 4 * 4):
                raise ValueError("classifier.0.weight has shape %s, expected (%d, 1536)" % (w.shape, num_classes))
            _lib.check(self.lib.ay_net_set_conv(self.ctx, 12, P(w), P(b)))
            _lib.check(self.lib.ay_net_finalize(self.ctx))
        rows, cols = ctypes.c_int(), ctypes.c_int()
        _lib.check(self.lib.ay_net_output_shape(self.ctx, ctypes.byref(rows), ctypes.byref(cols)))
        self.rows, self.cols = rows.value, cols.value


# ----------------------------------------------------------------------------- crops (core.py:109-159)
def _axis_window(centre, size):
    """One axis of ``get256Img``: 256 pixels centred on ``centre`` when they fit strictly inside (the reference's
    comparisons are ``centre - 128 > 0`` and ``centre + 128 < size``), else the window flush with the near edge."""
    if size < CROP:
        raise ValueError("a %d-pixel tile has no 256-pixel window" % size)
    low_ok, high_ok = centre - CROP // 2 > 0, centre + CROP // 2 < size
    if low_ok and high_ok:
        return centre - CROP // 2, centre + CROP // 2
    if high_ok:
        return 0, CROP
    if low_ok:
        return size - CROP, size
    raise ValueError("a %d-pixel tile has no 256-pixel window around %d" % (size, centre))


def crop_window(bbox_coord, size=1536):
    """(x_left, x_right, y_top, y_bottom) of the 256^2 crop for a box (top-left x, top-left y, width, height),
    identical to the slices of ``core.py:116-158`` for every tile wider than 256 pixels.  The reference's nine
    branches reduce to the same rule applied to each axis on its own."""
    cx = int(bbox_coord[0] + (bbox_coord[2] / 2))
    cy = int(bbox_coord[1] + (bbox_coord[3] / 2))
    xl, xr = _axis_window(cx, size)
    yt, yb = _axis_window(cy, size)
    return xl, xr, yt, yb


def get256Img(img, bbox_coord):
    """``core.py:109-159``: the 256 x 256 crop of a square tile ``img`` ([H,W,3] array or tensor) centred on the box."""
    xl, xr, yt, yb = crop_window(bbox_coord, img.shape[0])
    return img[yt:yb, xl:xr]


def normalised_crops(tile, windows, mean, std):
    """``core.py:214-217`` for a list of windows at once: uint8 RGB tile [H,W,3] on the GPU -> [n,3,256,256] fp32,
    ToTensor (``/255``) then Normalize (``- mean``, ``/ std``) with the same fp32 operation order as torchvision, so the
    network input carries the reference's bits.  One launch of ``ay_crop_normalize``; there is no CPU path."""
    if not (torch.is_tensor(tile) and tile.is_cuda and tile.dtype == torch.uint8 and tile.dim() == 3 and tile.shape[2] == 3
            and tile.stride(2) == 1 and tile.stride(1) == 3):
        raise ValueError("expected a CUDA uint8 tile [H,W,3] with packed pixels")
    dev = tile.device
    n = len(windows)
    for xl, xr, yt, yb in windows:
        if xr - xl != CROP or yb - yt != CROP:
            raise ValueError("window (%d, %d, %d, %d) is not %d x %d" % (xl, xr, yt, yb, CROP, CROP))
    win = torch.tensor(windows, dtype=torch.int32).reshape(n, 4).to(dev)
    out = torch.empty((n, 3, CROP, CROP), dtype=torch.float32, device=dev)
    m = (ctypes.c_float * 3)(*[float(v) for v in mean])
    s = (ctypes.c_float * 3)(*[float(v) for v in std])
    lib, ctx = _lib.load(), _lib.context(dev.index)
    with torch.cuda.device(dev):
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(lib.ay_crop_normalize(ctx, ctypes.c_void_p(tile.data_ptr()), tile.shape[0], tile.shape[1], tile.stride(0),
                                         ctypes.c_void_p(win.data_ptr()), n, CROP, m, s, ctypes.c_void_p(out.data_ptr()), st))
    return out


def class_preds(img_256, model, norm, device=None):
    """``getClassPreds`` (``core.py:210-221``) for one RGB 256^2 crop or a stack of them: sigmoid of the three logits,
    one (cored, diffuse, CAA) tuple per crop.  (The reference converts cv2's BGR to RGB first; crops passed here are
    RGB already.)"""
    eng = model.engine(device)
    crops = torch.as_tensor(np.asarray(img_256) if not torch.is_tensor(img_256) else img_256)
    single = crops.dim() == 3
    if single:
        crops = crops.unsqueeze(0)
    if crops.dtype != torch.uint8 or tuple(crops.shape[1:]) != (CROP, CROP, 3):
        raise ValueError("expected uint8 RGB crops [n,256,256,3], got %s %s" % (tuple(crops.shape), crops.dtype))
    crops = crops.to(eng.device).contiguous()
    tall = crops.reshape(-1, CROP, 3)                      # the stack as one tall tile: crop i is rows [256 i, 256 (i+1))
    x = normalised_crops(tall, [(0, CROP, CROP * i, CROP * (i + 1)) for i in range(len(crops))], norm["mean"], norm["std"])
    logits = eng.forward_nchw(x).reshape(len(crops), -1)
    eng.check()
    preds = [tuple(p) for p in torch.sigmoid(logits).cpu().tolist()]
    return preds[0] if single else preds


getClassPreds = class_preds


# ----------------------------------------------------------------------------- the filter (core.py:425-452)
MODEL_PICKLE = "pickles/model_all_fold_3_thresholding_2_l2.pkl"
NORM_FILE = "pickles/normalization.npy"


def load_reference_model(path=MODEL_PICKLE, precision="fp32"):
    """The reference pickles the whole ``core.Net`` module (``core.py:432``); its ``state_dict()`` is what this ``Net``
    needs.  Unpickling resolves ``core.Net`` in the caller's environment, exactly as ``torch.load`` does there."""
    obj = torch.load(path, map_location="cpu", weights_only=False)
    state = obj if isinstance(obj, dict) else obj.state_dict()
    net = Net(num_classes=int(state["classifier.0.weight"].shape[0]), precision=precision)
    net.load_state_dict(state)
    return net


def _load_tile(img):
    """Path -> RGB uint8 [H,W,3] (the reference reads BGR with cv2 and swaps to RGB in getClassPreds, core.py:214)."""
    if isinstance(img, str):
        from .utils.datasets import load_image
        return torch.from_numpy(np.ascontiguousarray(load_image(img)))
    return torch.as_tensor(img)


def filterDetectionsByCAAModel(img_name, detections, classes, model=None, norm=None, device=None, return_scores=False):
    """``core.py:425-452``.  ``img_name``: path of the 1536^2 tile, or the tile itself as an RGB uint8 [H,W,3] array /
    tensor (already on the GPU in the device-resident pipeline).  ``detections`` [n,7] as ``mergeDetections`` returns
    them.  Rows are deduplicated and ordered by Python's own ``set`` of row tuples, as in the reference; a row whose
    class is "CAA" and whose sigmoid CAA score is <= 0.5 is dropped.  ``model`` / ``norm`` default to the reference's
    pickles under ``pickles/`` (they are LFS objects the caller must provide).  Returns ``torch.as_tensor`` of the kept
    rows (an empty tensor when nothing is kept), plus every distinct row's (cored, diffuse, CAA) scores on request."""
    if model is None:
        model = load_reference_model()
    if norm is None:
        norm = np.load(NORM_FILE, allow_pickle=True).item()
    rows = list(set(tuple(entry) for entry in detections.tolist()))
    if not rows:
        empty = torch.as_tensor([])
        return (empty, []) if return_scores else empty
    eng = model.engine(device)
    tile = _load_tile(img_name)
    if tile.dim() != 3 or tile.shape[2] != 3 or tile.dtype != torch.uint8 or tile.shape[0] != tile.shape[1]:
        raise ValueError("expected a square RGB uint8 tile [H,H,3], got %s %s" % (tuple(tile.shape), tile.dtype))
    tile = tile.to(eng.device, non_blocking=True)
    size = tile.shape[0]
    windows = [crop_window((int(x1), int(y1), int(x2 - x1), int(y2 - y1)), size) for x1, y1, x2, y2, _, _, _ in rows]
    scores = []
    for lo in range(0, len(rows), eng.max_batch):               # one engine batch per <= max_batch distinct boxes
        x = normalised_crops(tile, windows[lo:lo + eng.max_batch], norm["mean"], norm["std"])
        logits = eng.forward_nchw(x).reshape(x.shape[0], -1)
        scores += torch.sigmoid(logits).cpu().tolist()
    eng.check()
    kept = [list(row) for row, s in zip(rows, scores) if not (classes[int(row[6])] == "CAA" and s[2] <= 0.5)]
    out = torch.as_tensor(kept)
    return (out, [tuple(s) for s in scores]) if return_scores else out


def filter_batch(tiles, detections_list, classes, model, norm, device=None):
    """``filterDetectionsByCAAModel`` for a batch of tiles in one pass: ``tiles`` uint8 ``[B,H,H,3]`` (host or device),
    ``detections_list`` the per-tile ``[k,7]`` tensors (or None) of ``TileDetector.detect``.  The distinct boxes of ALL
    tiles go through ``ay_crop_normalize`` and the network together (the batch is addressed as one tall tile), so the
    per-call cost of the single-tile form is paid once per batch.  Returns, per tile, what the single-tile call returns
    (same rows, same set order)."""
    eng = model.engine(device)
    tiles = torch.as_tensor(tiles)
    if tiles.dim() != 4 or tiles.shape[3] != 3 or tiles.dtype != torch.uint8 or tiles.shape[1] != tiles.shape[2]:
        raise ValueError("expected uint8 tiles [B,H,H,3], got %s %s" % (tuple(tiles.shape), tiles.dtype))
    if len(detections_list) != tiles.shape[0]:
        raise ValueError("%d detection lists for %d tiles" % (len(detections_list), tiles.shape[0]))
    size = tiles.shape[1]
    tall = tiles.to(eng.device, non_blocking=True).contiguous().reshape(-1, size, 3)
    rows_of, windows = [], []
    for b, det in enumerate(detections_list):
        rows = [] if det is None else list(set(tuple(entry) for entry in det.tolist()))
        rows_of.append(rows)
        for x1, y1, x2, y2, _, _, _ in rows:
            xl, xr, yt, yb = crop_window((int(x1), int(y1), int(x2 - x1), int(y2 - y1)), size)
            windows.append((xl, xr, yt + b * size, yb + b * size))
    scores = []
    for lo in range(0, len(windows), eng.max_batch):
        x = normalised_crops(tall, windows[lo:lo + eng.max_batch], norm["mean"], norm["std"])
        scores += torch.sigmoid(eng.forward_nchw(x).reshape(x.shape[0], -1)).cpu().tolist()
    eng.check()
    out, at = [], 0
    for rows in rows_of:
        s = scores[at:at + len(rows)]
        at += len(rows)
        out.append(torch.as_tensor([list(r) for r, sc in zip(rows, s) if not (classes[int(r[6])] == "CAA" and sc[2] <= 0.5)]))
    return out
