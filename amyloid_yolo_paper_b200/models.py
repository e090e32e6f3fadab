"""``Darknet`` — drop-in for the reference's ``models.py`` on the inference path.

Same constructor, attributes, parameter names and weight-file formats as the
reference (``models.py:225-336``), so ``load_state_dict(torch.load(ckpt))`` and
``load_darknet_weights(path)`` work on the reference's checkpoints unchanged.
The modules registered under ``module_list`` only *hold* the parameters: the
forward pass never calls them.  ``forward`` hands the batch to the sm_100a
engine behind the C-ABI library (``include/amyloid_yolo.h``), which runs the
whole network — tcgen05/TMEM implicit-GEMM convolutions with folded BatchNorm,
LeakyReLU, residual adds, concat and the YOLO decode fused in — and returns the
decoded ``[B, sum(A*G*G), 5+C]`` fp32 tensor on the CPU, like the reference
(``models.py:254``).  There is no CPU or PyTorch fallback: without the CUDA
library or a GPU, ``forward`` raises.

Training (``targets is not None``) is out of scope and raises.
"""
import numpy as np
import torch
import torch.nn as nn

from .utils.parse_config import parse_model_config


class Upsample(nn.Module):
    """Parameter-free marker for ``[upsample]`` (reference ``models.py:86-96``)."""

    def __init__(self, scale_factor, mode="nearest"):
        super().__init__()
        self.scale_factor = scale_factor
        self.mode = mode


class YOLOLayer(nn.Module):
    """Marker for ``[yolo]``: anchors / classes of one detection scale
    (reference ``models.py:98-113``); the decode itself (``models.py:137-169``)
    runs inside the head-convolution kernel's epilogue."""

    def __init__(self, anchors, num_classes, img_dim=416):
        super().__init__()
        self.anchors = anchors
        self.num_anchors = len(anchors)
        self.num_classes = num_classes
        self.img_dim = img_dim
        self.grid_size = 0
        self.metrics = {}


def create_modules(module_defs):
    """cfg blocks -> (hyperparams, ModuleList) with the reference's module and
    parameter names (``models.py:16-83``): ``conv_{i}``, ``batch_norm_{i}``,
    ``leaky_{i}``, ``maxpool_{i}``, ``upsample_{i}``, ``route_{i}``,
    ``shortcut_{i}``, ``yolo_{i}``.  Pops the ``[net]`` block off the list."""
    hyperparams = module_defs.pop(0)
    channels = [int(hyperparams["channels"])]
    module_list = nn.ModuleList()
    for i, d in enumerate(module_defs):
        seq = nn.Sequential()
        kind = d["type"]
        if kind == "convolutional":
            bn = int(d["batch_normalize"])
            filters = int(d["filters"])
            k = int(d["size"])
            seq.add_module("conv_%d" % i, nn.Conv2d(channels[-1], filters, k, int(d["stride"]),
                                                    (k - 1) // 2, bias=not bn))
            if bn:
                seq.add_module("batch_norm_%d" % i, nn.BatchNorm2d(filters, momentum=0.9, eps=1e-5))
            if d["activation"] == "leaky":
                seq.add_module("leaky_%d" % i, nn.LeakyReLU(0.1))
        elif kind == "maxpool":
            k, s = int(d["size"]), int(d["stride"])
            if k == 2 and s == 1:
                seq.add_module("_debug_padding_%d" % i, nn.ZeroPad2d((0, 1, 0, 1)))
            seq.add_module("maxpool_%d" % i, nn.MaxPool2d(k, s, (k - 1) // 2))
            filters = channels[-1]
        elif kind == "upsample":
            seq.add_module("upsample_%d" % i, Upsample(int(d["stride"]), "nearest"))
            filters = channels[-1]
        elif kind == "route":
            filters = sum(channels[1:][int(j)] for j in d["layers"].split(","))
            seq.add_module("route_%d" % i, nn.Sequential())
        elif kind == "shortcut":
            filters = channels[1:][int(d["from"])]
            seq.add_module("shortcut_%d" % i, nn.Sequential())
        elif kind == "yolo":
            mask = [int(v) for v in d["mask"].split(",")]
            flat = [int(v) for v in d["anchors"].split(",")]
            pairs = [(flat[j], flat[j + 1]) for j in range(0, len(flat), 2)]
            seq.add_module("yolo_%d" % i, YOLOLayer([pairs[j] for j in mask], int(d["classes"]),
                                                    int(hyperparams["height"])))
            filters = channels[-1]
        else:
            raise ValueError("unsupported cfg block [%s] at layer %d" % (kind, i))
        module_list.append(seq)
        channels.append(filters)
    return hyperparams, module_list


class Darknet(nn.Module):
    """YOLOv3 detector with the reference's API (``models.py:225-336``)."""

    def __init__(self, config_path, img_size=416):
        super().__init__()
        self.module_defs = parse_model_config(config_path)
        self.hyperparams, self.module_list = create_modules(self.module_defs)
        self.yolo_layers = [layer[0] for layer in self.module_list if isinstance(layer[0], YOLOLayer)]
        self.img_size = img_size
        self.seen = 0
        self.header_info = np.array([0, 0, 0, self.seen, 0], dtype=np.int32)
        self._engine = None           # built lazily on first forward (needs a GPU)
        self._engine_key = None
        self._weights_epoch = 0
        self._tensor_cache = None
        # "bf16" | "fp16": activation storage of the tensor-core path (2e-2 bar on the raw heads);
        # "fp32": the accuracy mode, fp32 storage + FFMA accumulation on the CUDA cores (1e-4 bar)
        self.precision = "bf16"

    # ------------------------------------------------------------------ engine
    def invalidate_engine(self):
        """Force the next forward to re-fold and re-upload the weights.  Needed only after writes the
        staleness key cannot see: ``param.data.copy_()/fill_()`` (``.data`` has its own version counter)."""
        self._weights_epoch += 1

    def _staleness_key(self):
        """Cheap per-call key: the tensors are listed once (no ``state_dict()`` walk per forward, which was
        first-order at batch 1) and only their identity and in-place version counters are read."""
        cache = self._tensor_cache
        if cache is None:
            cache = self._tensor_cache = tuple(self.parameters()) + tuple(self.buffers())
        v = 0
        for t in cache:
            v += t._version
        return (self._weights_epoch, v, cache[0].data_ptr() if cache else 0)

    def _apply(self, fn, *args, **kwargs):
        # .to() / .cuda() / .float(): parameters are re-created
        self._tensor_cache = None
        self._weights_epoch += 1
        return super()._apply(fn, *args, **kwargs)

    def apply(self, fn):
        # model.apply(weights_init_normal) writes through .data (utils/utils.py:27-33)
        out = super().apply(fn)
        self._tensor_cache = None
        self._weights_epoch += 1
        return out

    def load_state_dict(self, *args, **kwargs):
        out = super().load_state_dict(*args, **kwargs)
        self._tensor_cache = None
        self._weights_epoch += 1
        return out

    def engine(self, device=None, max_batch=None, img_size=None):
        """The sm_100a execution engine for this network (created on demand and re-created when the
        weights, device, precision, input size or batch capacity change).  A replaced engine is not
        destroyed here: it is released when its last user (e.g. a :class:`TileDetector`) lets go of it,
        so nobody is left holding a freed ``ay_ctx``."""
        from .engine import Engine
        if device is None:
            p = next(self.parameters())
            device = p.device if p.is_cuda else torch.device("cuda", torch.cuda.current_device()
                                                             if torch.cuda.is_available() else 0)
        device = torch.device(device)
        size = int(img_size) if img_size is not None else int(self.hyperparams["height"])
        key = (device.type, device.index, self.precision, size, self._staleness_key())
        eng = self._engine
        if eng is None or self._engine_key != key or (max_batch is not None and max_batch > eng.max_batch):
            cap = max(int(max_batch or 8), eng.max_batch if (eng is not None and self._engine_key == key) else 0)
            self._engine = None           # drop our reference first: frees the old buffers unless someone else holds it
            del eng
            self._engine = Engine(self.module_defs, self.hyperparams, self.state_dict(), device,
                                  max_batch=cap, precision=self.precision, img_size=size)
            self._engine_key = key
        return self._engine

    def forward(self, x, targets=None):
        """x: [B,3,S,S] float (S multiple of 32) -> CPU fp32 [B, sum(A*G*G), 5+C]
        (reference ``models.py:237-255``; like the reference the network runs at the input's own size,
        ``img_dim = x.shape[2]``, models.py:238)."""
        if targets is not None:
            raise NotImplementedError("training branch (models.py:173-222) is out of scope")
        if x.dim() != 4 or x.shape[2] != x.shape[3] or x.shape[2] % 32 != 0:
            raise ValueError("expected input [B,C,S,S] with S a multiple of 32, got %s" % (tuple(x.shape),))
        eng = self.engine(max_batch=x.shape[0], img_size=x.shape[2])
        out = eng.forward_nchw(x)
        eng.check()
        return out.cpu()

    # ------------------------------------------------------------ weight files
    def load_darknet_weights(self, weights_path):
        """Darknet ``.weights`` reader (reference ``models.py:257-308``): int32[5]
        header (``header[3]`` = images seen) then fp32 per conv:
        [bn.bias, bn.weight, bn.running_mean, bn.running_var] or [conv.bias],
        then conv.weight (OIHW).  ``darknet53.conv.74`` stops after layer 74."""
        with open(weights_path, "rb") as fh:
            header = np.fromfile(fh, dtype=np.int32, count=5)
            self.header_info = header
            self.seen = header[3]
            blob = np.fromfile(fh, dtype=np.float32)
        cutoff = 75 if "darknet53.conv.74" in weights_path else None
        pos = 0
        self._weights_epoch += 1

        def take(param):
            nonlocal pos
            n = param.numel()
            param.data.copy_(torch.from_numpy(blob[pos:pos + n]).view_as(param))
            pos += n

        for i, (d, module) in enumerate(zip(self.module_defs, self.module_list)):
            if i == cutoff:
                break
            if d["type"] != "convolutional":
                continue
            conv = module[0]
            if d["batch_normalize"]:
                bn = module[1]
                take(bn.bias)
                take(bn.weight)
                take(bn.running_mean)
                take(bn.running_var)
            else:
                take(conv.bias)
            take(conv.weight)

    def save_darknet_weights(self, path, cutoff=-1):
        """Inverse of :meth:`load_darknet_weights` (reference ``models.py:310-336``;
        like the reference, ``cutoff=-1`` slices off the last block, which is a
        ``[yolo]`` block and holds no weights)."""
        with open(path, "wb") as fh:
            self.header_info[3] = self.seen
            self.header_info.tofile(fh)
            for d, module in zip(self.module_defs[:cutoff], self.module_list[:cutoff]):
                if d["type"] != "convolutional":
                    continue
                conv = module[0]
                if d["batch_normalize"]:
                    bn = module[1]
                    for t in (bn.bias, bn.weight, bn.running_mean, bn.running_var):
                        t.data.cpu().numpy().tofile(fh)
                else:
                    conv.bias.data.cpu().numpy().tofile(fh)
                conv.weight.data.cpu().numpy().tofile(fh)
