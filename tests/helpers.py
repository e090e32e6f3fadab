"""Shared test helpers (CPU-only imports; GPU use is up to the caller)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MINI_CFG = """
[net]
batch=1
width=64
height=64
channels=3

[convolutional]
batch_normalize=1
filters=32
size=3
stride=1
pad=1
activation=leaky

[convolutional]
batch_normalize=1
filters=64
size=3
stride=2
pad=1
activation=leaky

[convolutional]
batch_normalize=1
filters=32
size=1
stride=1
pad=1
activation=leaky

[convolutional]
batch_normalize=1
filters=64
size=3
stride=1
pad=1
activation=leaky

[shortcut]
from=-3
activation=linear

[convolutional]
batch_normalize=1
filters=128
size=3
stride=2
pad=1
activation=leaky

[convolutional]
batch_normalize=1
filters=64
size=1
stride=1
pad=1
activation=leaky

[convolutional]
batch_normalize=1
filters=128
size=3
stride=1
pad=1
activation=leaky

[shortcut]
from=-3
activation=linear

[convolutional]
batch_normalize=1
filters=256
size=3
stride=2
pad=1
activation=leaky

[convolutional]
batch_normalize=1
filters=128
size=1
stride=1
pad=1
activation=leaky

[convolutional]
batch_normalize=1
filters=256
size=3
stride=1
pad=1
activation=leaky

[shortcut]
from=-3
activation=linear

[convolutional]
batch_normalize=1
filters=512
size=3
stride=2
pad=1
activation=leaky

[convolutional]
batch_normalize=1
filters=256
size=1
stride=1
pad=1
activation=leaky

[convolutional]
batch_normalize=1
filters=512
size=3
stride=1
pad=1
activation=leaky

[shortcut]
from=-3
activation=linear

[convolutional]
size=1
stride=1
pad=1
filters=21
activation=linear

[yolo]
mask = 6,7,8
anchors = 10,13,  16,30,  33,23,  30,61,  62,45,  59,119,  116,90,  156,198,  373,326
classes=2
num=9

[route]
layers = -3

[convolutional]
batch_normalize=1
filters=128
size=1
stride=1
pad=1
activation=leaky

[upsample]
stride=2

[route]
layers = -1, 12

[convolutional]
batch_normalize=1
filters=128
size=1
stride=1
pad=1
activation=leaky

[convolutional]
batch_normalize=1
filters=256
size=3
stride=1
pad=1
activation=leaky

[convolutional]
size=1
stride=1
pad=1
filters=21
activation=linear

[yolo]
mask = 3,4,5
anchors = 10,13,  16,30,  33,23,  30,61,  62,45,  59,119,  116,90,  156,198,  373,326
classes=2
num=9
"""


def write_mini_cfg(directory):
    path = os.path.join(str(directory), "mini.cfg")
    with open(path, "w") as fh:
        fh.write(MINI_CFG)
    return path


def rel_max(a, b):
    """max|a-b| / max|b| (the head-output parity metric of SURVEY.md §8d)."""
    denom = float(b.abs().max())
    return float((a - b).abs().max()) / (denom if denom > 0 else 1.0)


def decoded_to_raw(dec, blocks, img_dim=416):
    """Invert the YOLO decode (models.py:137-169) in float64: decoded [B, sum A*G*G, 5+C] -> list of raw head
    tensors [B, A, G, G, 5+C] in cfg order.  The north-star tolerance (2e-2 relative in bf16) is stated on the RAW
    head outputs; exp() in the decode amplifies a raw error of d into a relative box-size error of d, so the
    comparison is made here, before the exponential."""
    import torch
    from oracle import restate as R
    outs, off = [], 0
    d64 = dec.to(torch.float64)
    for b in blocks[1:]:
        if b["type"] != "yolo":
            continue
        anchors = R.yolo_anchors(b)
        A = len(anchors)
        # grid size from the rows left: grids come in cfg order and are recovered from the producing conv's size
        G = b["_grid"]
        n = A * G * G
        p = d64[:, off:off + n].view(dec.shape[0], A, G, G, -1).clone()
        off += n
        stride = img_dim / G
        gx = torch.arange(G, dtype=torch.float64).view(1, 1, 1, G)
        gy = torch.arange(G, dtype=torch.float64).view(1, 1, G, 1)

        def logit(v):
            v = v.clamp(1e-9, 1 - 1e-9)
            return torch.log(v) - torch.log1p(-v)

        raw = torch.empty_like(p)
        raw[..., 0] = logit(p[..., 0] / stride - gx)
        raw[..., 1] = logit(p[..., 1] / stride - gy)
        for a, (aw, ah) in enumerate(anchors):
            raw[:, a, :, :, 2] = torch.log(p[:, a, :, :, 2] / aw)
            raw[:, a, :, :, 3] = torch.log(p[:, a, :, :, 3] / ah)
        raw[..., 4:] = logit(p[..., 4:])
        outs.append(raw)
    return outs


def raw_head_error(got_dec, ref_raw_heads, blocks, img_dim=416):
    """max over heads of max|raw_got - raw_ref| / max|raw_ref|, raw_got recovered from the decoded output.
    ref_raw_heads: the oracle's raw head tensors [B, A*(5+C), G, G] (darknet_forward(return_raw_heads=True))."""
    import torch
    yolo = [b for b in blocks[1:] if b["type"] == "yolo"]
    for b, r in zip(yolo, ref_raw_heads):
        b["_grid"] = int(r.shape[-1])
    got = decoded_to_raw(got_dec, blocks, img_dim)
    worst = 0.0
    for g, r in zip(got, ref_raw_heads):
        B, _, G, _ = r.shape
        A = g.shape[1]
        want = r.to(torch.float64).view(B, A, -1, G, G).permute(0, 1, 3, 4, 2)
        # sigmoid saturates in fp32 beyond |t| ~ 15: the decoded value no longer determines t there
        ok = want.abs() < 12.0
        err = ((g - want).abs() * ok).max() / want.abs().max()
        worst = max(worst, float(err))
    return worst
