"""Seeded synthetic inputs for tests and benchmarks (no dataset, no network).

* ``synthetic_tiles``       — uint8 HWC "whole-slide" tiles (1536x1536x3 by default)
* ``synthetic_state_dict``  — a checkpoint in the reference's ``.pth`` key layout
* ``synthetic_predictions`` — decoded head output ``[B, N, 5+C]`` for NMS stress
                               
* ``synthetic_detections``  — post-NMS rows for the box-merging step

Only numpy/torch; used by the product-side bench and by tests alike.
"""
import numpy as np
import torch

GRIDS_416 = (13, 26, 52)
ANCHORS_V3 = {13: ((116, 90), (156, 198), (373, 326)),
              26: ((30, 61), (62, 45), (59, 119)),
              52: ((10, 13), (16, 30), (33, 23))}


def synthetic_tiles(batch, size=1536, seed=0xA11CE, first_tile=0):
    """uint8 [batch, size, size, 3]; tile t depends only on (seed, first_tile+t).
    Smooth low-frequency field plus blobs so BatchNorm statistics stay sane."""
    out = np.empty((batch, size, size, 3), dtype=np.uint8)
    yy, xx = np.meshgrid(np.arange(size, dtype=np.float32), np.arange(size, dtype=np.float32), indexing="ij")
    for b in range(batch):
        rng = np.random.default_rng([seed, first_tile + b])
        img = np.full((size, size, 3), 200.0, dtype=np.float32)
        img += 20.0 * np.sin(xx / rng.uniform(40, 200) + rng.uniform(0, 6))[..., None]
        img += 15.0 * np.cos(yy / rng.uniform(40, 200) + rng.uniform(0, 6))[..., None]
        for _ in range(int(rng.integers(4, 14))):
            cx, cy = rng.uniform(0, size, 2)
            r = rng.uniform(15, 120)
            blob = np.exp(-((xx - cx) ** 2 + (yy - cy) ** 2) / (2 * r * r))
            img -= blob[..., None] * rng.uniform(40, 140, 3).astype(np.float32)
        img += rng.normal(0, 6, img.shape).astype(np.float32)
        out[b] = np.clip(img, 0, 255).astype(np.uint8)
    return out


def synthetic_tiles_fast(batch, size=1536, seed=0xA11CE, first_tile=0):
    """Cheaper variant for large benchmark batches: uniform random bytes
    (FLOPs and bytes of the path do not depend on content)."""
    rng = np.random.default_rng([seed, first_tile, batch])
    return rng.integers(0, 256, size=(batch, size, size, 3), dtype=np.uint8)


def synthetic_state_dict(model, seed=0, conf_bias=None, wh_gain=None):
    """Fill ``model`` (a Darknet module) with the seeded synthetic checkpoint of
    SURVEY.md §8c and return its state_dict: conv N(0,0.02) scaled up so that
    activations neither die nor explode over 75 layers, BN gamma N(1,0.02),
    beta small, running stats randomised so BN folding is exercised, head
    objectness biases set from ``conf_bias`` so that a controllable fraction of
    rows clears the confidence threshold.  ``wh_gain`` (optional) shrinks the box-size outputs."""
    g = torch.Generator().manual_seed(seed)
    sd = model.state_dict()
    defs = model.module_defs
    for key, t in sd.items():
        leaf = key.rsplit(".", 1)[1]
        layer = int(key.split(".")[1])
        nxt = defs[layer + 1]["type"] if layer + 1 < len(defs) else ""
        if "conv_" in key and leaf == "weight":
            fan_in = t[0].numel()
            gain = 0.75 * 2.0 / 1.01                # slightly contracting for leaky(0.1) + random BN stats
            if nxt == "shortcut":
                gain *= 0.1                         # keep the residual stream from exploding
            if nxt == "yolo":
                gain = 1.5                          # raw head outputs of order 1
            t.copy_(torch.randn(t.shape, generator=g) * (gain / fan_in) ** 0.5)
        elif "conv_" in key and leaf == "bias":
            t.copy_(torch.randn(t.shape, generator=g) * 0.3)
        elif "batch_norm_" in key and leaf == "weight":
            t.copy_(1.0 + 0.02 * torch.randn(t.shape, generator=g))
        elif "batch_norm_" in key and leaf == "bias":
            t.copy_(0.05 * torch.randn(t.shape, generator=g))
        elif leaf == "running_mean":
            t.copy_(0.1 * torch.randn(t.shape, generator=g))
        elif leaf == "running_var":
            t.copy_(0.5 + torch.rand(t.shape, generator=g))
        elif leaf == "num_batches_tracked":
            t.fill_(105)
    if conf_bias is not None:
        head = 0                                     # heads in cfg order (grids 13, 26, 52 for yolov3-custom)
        for key, t in sd.items():
            if "conv_" in key and key.endswith(".bias"):
                bias = conf_bias[min(head, len(conf_bias) - 1)] if isinstance(conf_bias, (tuple, list)) else conf_bias
                head += 1
                per_anchor = t.numel() // 3
                for a in range(3):
                    t[a * per_anchor + 4] = bias
    if wh_gain is not None:
        # box-size channels (tw, th) of the heads scaled down: sizes stay within a factor ~2 of the anchors, like a
        # trained detector's, instead of exp(N(0,1.5)) * anchor (mergeDetections enumerates box pixels on the CPU path)
        for key, t in sd.items():
            layer = int(key.split(".")[1])
            if "conv_" in key and layer + 1 < len(defs) and defs[layer + 1]["type"] == "yolo":
                per_anchor = t.shape[0] // 3
                for a in range(3):
                    t[a * per_anchor + 2:a * per_anchor + 4] *= wh_gain
    model.load_state_dict(sd)
    return model.state_dict()


def synthetic_predictions(batch, num_classes=2, seed=5, img=416, grids=GRIDS_416,
                          clusters=(200, 400), anchors=None):
    """Decoded predictions laid out like the real head: for each grid G (coarse
    to fine) and anchor a, rows a*G*G + gy*G + gx; centres near the cell centre,
    sizes = anchor * exp(N(0,0.3)); a few hundred planted object clusters so
    that IoU>0.4 groups of 2..60 rows exist; conf ~ U(0.01,1), cls ~ U(0,1)^C;
    scores conf*max(cls) made distinct within a tile (reference tie order is
    unspecified)."""
    anchors = anchors or ANCHORS_V3
    rng = np.random.default_rng(seed)
    n_rows = sum(3 * g * g for g in grids)
    out = np.empty((batch, n_rows, 5 + num_classes), dtype=np.float32)
    for b in range(batch):
        rows = []
        for g in grids:
            stride = img / g
            gy, gx = np.meshgrid(np.arange(g), np.arange(g), indexing="ij")
            for a in range(3):
                aw, ah = anchors[g][a] if g in anchors else (stride * 2, stride * 2)
                cx = (gx.reshape(-1) + 0.5 + rng.uniform(-0.5, 0.5, g * g)) * stride
                cy = (gy.reshape(-1) + 0.5 + rng.uniform(-0.5, 0.5, g * g)) * stride
                w = aw * np.exp(rng.normal(0, 0.3, g * g))
                h = ah * np.exp(rng.normal(0, 0.3, g * g))
                rows.append(np.stack([cx, cy, w, h], 1))
        box = np.concatenate(rows, 0)
        n_clusters = int(rng.integers(clusters[0], clusters[1] + 1))
        for _ in range(n_clusters):
            size = int(rng.integers(2, 61))
            members = rng.choice(n_rows, size=size, replace=False)
            centre = rng.uniform(0, img, 2)
            wh = rng.uniform(12, 160, 2)
            box[members, 0:2] = centre + rng.normal(0, 0.04, (size, 2)) * wh
            box[members, 2:4] = wh * np.exp(rng.normal(0, 0.06, (size, 2)))
        conf = rng.uniform(0.01, 1.0, n_rows)
        cls = rng.uniform(0.0, 1.0, (n_rows, num_classes))
        p = np.concatenate([box, conf[:, None], cls], 1).astype(np.float32)
        for _ in range(64):
            score = p[:, 4] * p[:, 5:].max(1)
            _, first, counts = np.unique(score, return_index=True, return_counts=True)
            if (counts == 1).all():
                break
            dup = np.setdiff1d(np.arange(n_rows), first)
            p[dup, 4] = rng.uniform(0.01, 1.0, dup.size).astype(np.float32)
        out[b] = p
    return out


def synthetic_detections(count, seed=0, extent=1536.0, num_classes=2, overlap=0.5, wh=(20, 260), jitter=40):
    """Post-NMS rows [count,7] (x1,y1,x2,y2,conf,cls_conf,cls_pred) in original
    tile coordinates, with a tunable share of overlapping same-class pairs and
    some boxes hanging over the top/left edge (negative coordinates)."""
    rng = np.random.default_rng(seed)
    rows = np.zeros((count, 7), dtype=np.float32)
    for i in range(count):
        if i and rng.random() < overlap:
            j = int(rng.integers(0, i))
            cx = (rows[j, 0] + rows[j, 2]) / 2 + rng.normal(0, jitter)
            cy = (rows[j, 1] + rows[j, 3]) / 2 + rng.normal(0, jitter)
            cls = rows[j, 6] if rng.random() < 0.8 else float(rng.integers(0, num_classes))
        else:
            cx, cy = rng.uniform(-20, extent + 20, 2)
            cls = float(rng.integers(0, num_classes))
        w, h = rng.uniform(wh[0], wh[1], 2)
        rows[i] = (cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2,
                   rng.uniform(0.8, 1.0), rng.uniform(0.5, 1.0), cls)
    return rows


# ----------------------------------------------------------------------------- planted objects
# A seeded random checkpoint has a Gaussian objectness tail: every row that clears conf_thres sits within the 16-bit
# activations' noise of the threshold, so "final detections agree" cannot be tested on it in bf16.  A trained detector
# is bimodal instead (object / no object).  The checkpoint below gets that property by construction: three MARKER
# channels carry a sub-sampled copy of the input's R, G and B planes through the whole network unchanged (centre-tap
# unit weights, identity BatchNorm, zero contribution from the residual branches), and the objectness / class logits of
# head h (grid 13, 26, 52) read marker h only.  Setting one source pixel of a tile plants one confident detection at a
# chosen grid cell; every other pixel and every other channel stays random, so the box regressions of the planted rows
# still come through all 75 random convolutions.
PLANT_CONF_GAIN, PLANT_CONF_BIAS = 8.0, -2.5     # t_conf = 8 m - 2.5: m = 0 -> conf 0.076; m in [0.58, 1] -> 0.90 .. 0.996
PLANT_CLS_GAIN, PLANT_CLS_MID = 30.0, 0.794      # class 0 for m > 0.794 (pixel value > 202), class 1 below
PLANT_VALUES = {0: (215, 255), 1: (150, 190)}    # marker pixel values per class (distinct per object: distinct scores)


def planted_state_dict(model, seed=0, anchor=(0, 2, 2), xy_gain=0.3, wh_gain=0.05):
    """Seeded random checkpoint of ``model`` (yolov3-custom layout: no maxpool) with the marker channels described above.
    Head h fires on anchor ``anchor[h]`` only.  Returns the state_dict (also loaded into ``model``)."""
    sd = synthetic_state_dict(model, seed=seed, conf_bias=None, wh_gain=None)
    defs = model.module_defs
    marks = []                     # per layer: input-channel index of the three markers in that layer's OUTPUT (or None)
    cur = [0, 1, 2]                # the network input itself: R, G, B
    channels = [int(model.hyperparams["channels"])]
    head = 0
    for i, d in enumerate(defs):
        t = d["type"]
        pre = "module_list.%d." % i
        if t == "convolutional":
            w = sd[pre + "conv_%d.weight" % i]
            k = w.shape[2]
            is_head = i + 1 < len(defs) and defs[i + 1]["type"] == "yolo"
            if is_head:
                per = w.shape[0] // 3
                classes = per - 5
                if classes != 2:
                    raise ValueError("planted checkpoint is defined for 2 classes")
                for a in range(3):
                    base = a * per
                    w[base:base + 2] *= xy_gain
                    w[base + 2:base + 4] *= wh_gain
                    sd[pre + "conv_%d.bias" % i][base:base + 4] *= xy_gain
                    w[base + 4:base + per] = 0.0
                    b = sd[pre + "conv_%d.bias" % i]
                    if a == anchor[min(head, len(anchor) - 1)] and cur is not None:
                        m = cur[min(head, 2)]
                        w[base + 4, m, 0, 0] = PLANT_CONF_GAIN
                        b[base + 4] = PLANT_CONF_BIAS
                        w[base + 5, m, 0, 0] = PLANT_CLS_GAIN
                        b[base + 5] = -PLANT_CLS_GAIN * PLANT_CLS_MID
                        w[base + 6, m, 0, 0] = -PLANT_CLS_GAIN
                        b[base + 6] = PLANT_CLS_GAIN * PLANT_CLS_MID
                    else:
                        b[base + 4] = -8.0
                        b[base + 5:base + per] = 0.0
                head += 1
                nxt = None
            else:
                ends_branch = i + 1 < len(defs) and defs[i + 1]["type"] == "shortcut"
                for j in range(3):
                    w[j] = 0.0
                    if not ends_branch and cur is not None:
                        w[j, cur[j], k // 2, k // 2] = 1.0          # centre tap: out(y,x) = in(s*y, s*x)
                    bn = pre + "batch_norm_%d." % i
                    sd[bn + "weight"][j] = 1.0
                    sd[bn + "bias"][j] = 0.0
                    sd[bn + "running_mean"][j] = 0.0
                    sd[bn + "running_var"][j] = 1.0 - 1e-5          # gamma / sqrt(var + eps) = 1
                nxt = None if ends_branch else [0, 1, 2]
            channels.append(w.shape[0])
        elif t == "shortcut":
            nxt = marks[i + int(d["from"])]                           # branch contributes exactly 0 on the marker channels
            channels.append(channels[-1])
        elif t == "route":
            srcs = [int(v) for v in d["layers"].split(",")]
            srcs = [s if s >= 0 else i + s for s in srcs]
            off, nxt = 0, None
            for s in srcs:                                            # the last source that carries markers wins (the
                if marks[s] is not None:                              # non-upsampled one of [upsample, stage output])
                    nxt = [off + c for c in marks[s]]
                off += channels[s + 1]
            channels.append(off)
        elif t == "upsample":
            nxt = cur
            channels.append(channels[-1])
        elif t == "yolo":
            nxt = None
            channels.append(channels[-1])
        else:
            raise ValueError("planted checkpoint: unsupported block [%s]" % t)
        marks.append(nxt)
        cur = nxt
    model.load_state_dict(sd)
    return model.state_dict()


def planted_tiles(batch, size=1536, seed=7, img=416, objects=(10, 16), grids=(13, 26), pair_share=0.35):
    """uint8 [batch,size,size,3] random tiles with planted objects for :func:`planted_state_dict`.
    -> (tiles, truth) where truth[b] is a list of (grid, gy, gx, class, value).  Marker h (R, G, B for grids
    13, 26, 52) is sampled at network pixel (stride*gy, stride*gx), i.e. source pixel floor(p*size/img) under the
    nearest resize: those source pixels are zeroed everywhere and set to a distinct value per planted object.
    ``pair_share`` of the objects get a same-class neighbour in the adjacent cell (an IoU > 0.4 cluster for NMS,
    overlapping boxes for mergeDetections)."""
    rng = np.random.default_rng([seed, batch, size])
    tiles = rng.integers(0, 256, size=(batch, size, size, 3), dtype=np.uint8)
    scale = np.float32(size) / np.float32(img)
    src = np.minimum(np.floor(np.arange(img, dtype=np.float32) * scale).astype(np.int64), size - 1)
    all_grids = (13, 26, 52)
    truth = []
    for b in range(batch):
        for h, g in enumerate(all_grids):                 # background: marker planes read 0 at every sampled position
            pos = src[np.arange(g) * (img // g)]
            tiles[b][np.ix_(pos, pos, [h])] = 0
        used, objs = set(), []
        values = {c: list(rng.permutation(np.arange(lo, hi + 1))) for c, (lo, hi) in PLANT_VALUES.items()}
        n = int(rng.integers(objects[0], objects[1] + 1))
        while len(objs) < n:
            g = int(rng.choice(grids))
            gy, gx = (int(v) for v in rng.integers(1, g - 2, 2))
            cls = int(rng.integers(0, 2))
            cells = [(gy, gx)]
            if rng.random() < pair_share:
                cells.append((gy, gx + 1))
            if any((g, y + dy, x + dx) in used for y, x in cells for dy in (-1, 0, 1) for dx in (-1, 0, 1)):
                continue
            for y, x in cells:
                used.add((g, y, x))
                objs.append((g, y, x, cls, int(values[cls].pop())))
        for g, gy, gx, cls, val in objs:
            h = all_grids.index(g)
            stride = img // g
            tiles[b, src[stride * gy], src[stride * gx], h] = val
        truth.append(objs)
    return tiles, truth
