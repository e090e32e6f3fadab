"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's detection path.

This file is the *checker*: tests/, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` leg of ``bench.py`` may import it; nothing under
``amyloid_yolo_paper_b200/`` (the product) may.  Every function cites the
reference lines it restates (paths relative to the reference checkout).

Parity status: the reference pins almost nothing numerically on this path
(only three scalar IoU values, ``unit_test.py:137-150``), so the oracle is
pinned against *outputs of the reference itself run in the build container*
(``oracle/make_golden.py`` -> ``tests/golden/``) plus those IoU known answers.
See DESIGN.md "Oracle".

Two flavours are provided where the reference's cost structure matters:

* ``*_faithful``  — same algorithmic structure as the reference (python loop
  of tensor ops for NMS, pixel-set enumeration for box merging).  This is what
  the CPU baseline times.
* ``*_exact``     — closed-form / vectorised restatements that return the same
  bits (validated against the faithful ones and against the real reference)
  and finish quickly at stress sizes; they also return keep indices, which the
  reference does not expose.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

f32 = np.float32

# --------------------------------------------------------------------------
# cfg handling (models.py:16-83 create_modules, host side only)
# --------------------------------------------------------------------------


def parse_cfg(path):
    """utils/parse_config.py:3-21."""
    blocks = []
    for line in open(path).read().split("\n"):
        if not line or line.startswith("#"):
            continue
        line = line.strip()
        if line.startswith("["):
            blocks.append({"type": line[1:-1].rstrip()})
            if blocks[-1]["type"] == "convolutional":
                blocks[-1]["batch_normalize"] = 0
        else:
            k, v = line.split("=")
            blocks[-1][k.rstrip()] = v.strip()
    return blocks


def yolo_anchors(block):
    """models.py:68-72: anchors selected by mask."""
    idx = [int(v) for v in block["mask"].split(",")]
    flat = [int(v) for v in block["anchors"].split(",")]
    pairs = [(flat[i], flat[i + 1]) for i in range(0, len(flat), 2)]
    return [pairs[i] for i in idx]


# --------------------------------------------------------------------------
# forward + decode (models.py:237-255, 127-172)
# --------------------------------------------------------------------------


def decode_head(raw, anchors, num_classes, img_dim):
    """YOLOLayer inference branch, models.py:137-169 (fp32, torch CPU ops).

    raw: [B, A*(5+C), G, G] -> [B, A*G*G, 5+C], row = a*G*G + gy*G + gx."""
    B, _, G, _ = raw.shape
    A = len(anchors)
    p = raw.view(B, A, 5 + num_classes, G, G).permute(0, 1, 3, 4, 2).contiguous()
    stride = img_dim / G
    gx = torch.arange(G, dtype=torch.float32).repeat(G, 1).view(1, 1, G, G)
    gy = torch.arange(G, dtype=torch.float32).repeat(G, 1).t().view(1, 1, G, G)
    scaled = torch.tensor([(aw / stride, ah / stride) for aw, ah in anchors], dtype=torch.float32)
    aw = scaled[:, 0:1].view(1, A, 1, 1)
    ah = scaled[:, 1:2].view(1, A, 1, 1)
    boxes = torch.empty(B, A, G, G, 4, dtype=torch.float32)
    boxes[..., 0] = torch.sigmoid(p[..., 0]) + gx
    boxes[..., 1] = torch.sigmoid(p[..., 1]) + gy
    boxes[..., 2] = torch.exp(p[..., 2]) * aw
    boxes[..., 3] = torch.exp(p[..., 3]) * ah
    return torch.cat((boxes.view(B, -1, 4) * stride,
                      torch.sigmoid(p[..., 4]).view(B, -1, 1),
                      torch.sigmoid(p[..., 5:]).view(B, -1, num_classes)), -1)


def _conv_key(i, kind, leaf):
    return "module_list.%d.%s_%d.%s" % (i, kind, i, leaf)


def darknet_forward(blocks, state, x, emulate_bf16=False, return_layers=False, return_raw_heads=False):
    """Functional restatement of Darknet.forward (models.py:237-255) +
    create_modules (models.py:16-83) in eval mode.

    blocks: parse_cfg() output INCLUDING the [net] block; state: state_dict in
    the reference's key naming (models.py:31-44).  With ``emulate_bf16`` the
    arithmetic of the B200 pipeline is mimicked instead (BatchNorm folded into
    the weights in fp32, weights and every stored activation rounded to bf16,
    fp32 accumulation, fp32 head outputs) — used to derive tolerances and to
    localise GPU mismatches layer by layer."""
    hyper, defs = blocks[0], blocks[1:]
    img_dim = x.shape[2]
    outs, yolo, raws = [], [], []
    # a float64 input evaluates the whole network in float64 (tests use it to size the fp32 noise floor)
    ft = torch.float64 if (x.dtype == torch.float64 and not emulate_bf16) else torch.float32

    def q(t):
        return t.to(torch.bfloat16).to(torch.float32) if emulate_bf16 else t

    for i, d in enumerate(defs):
        t = d["type"]
        if t == "convolutional":
            k = int(d["size"])
            w = state[_conv_key(i, "conv", "weight")].to(ft)
            bn = int(d["batch_normalize"])
            if bn:
                g = state[_conv_key(i, "batch_norm", "weight")].to(ft)
                b = state[_conv_key(i, "batch_norm", "bias")].to(ft)
                mu = state[_conv_key(i, "batch_norm", "running_mean")].to(ft)
                var = state[_conv_key(i, "batch_norm", "running_var")].to(ft)
            else:
                cb = state[_conv_key(i, "conv", "bias")].to(ft)
            if emulate_bf16:
                if bn:
                    s = g / torch.sqrt(var + 1e-5)
                    w = w * s.view(-1, 1, 1, 1)
                    cb = b - mu * s
                is_head = defs[i + 1]["type"] == "yolo" if i + 1 < len(defs) else False
                # the engine's first layer (Cin=3) contracts pixel values 0..255 (exact in bf16 for uint8 sources)
                # with bf16(w / 255): ToTensor's scale is folded into the weights
                if i == 0 and w.shape[1] < 8:
                    x = q(x * 255.0)
                    wq = q((w.double() / 255.0).float())
                else:
                    wq = q(w)
                fused_add = (not is_head) and i + 1 < len(defs) and defs[i + 1]["type"] == "shortcut"
                x = F.conv2d(x, wq, None, int(d["stride"]), (k - 1) // 2) + cb.view(1, -1, 1, 1)
                if d["activation"] == "leaky":
                    x = F.leaky_relu(x, 0.1)
                if not is_head and not fused_add:
                    x = q(x)                    # fused shortcut: one rounding after the add
            else:
                x = F.conv2d(x, w, None if bn else cb, int(d["stride"]), (k - 1) // 2)
                if bn:
                    x = F.batch_norm(x, mu, var, g, b, False, 0.1, 1e-5)
                if d["activation"] == "leaky":
                    x = F.leaky_relu(x, 0.1)
        elif t == "maxpool":
            k, s = int(d["size"]), int(d["stride"])
            if k == 2 and s == 1:
                x = F.pad(x, (0, 1, 0, 1))
            x = F.max_pool2d(x, k, s, (k - 1) // 2)
        elif t == "upsample":
            x = F.interpolate(x, scale_factor=int(d["stride"]), mode="nearest")
        elif t == "route":
            x = torch.cat([outs[int(j)] for j in d["layers"].split(",")], 1)
        elif t == "shortcut":
            x = outs[-1] + outs[int(d["from"])]
            x = q(x)
        elif t == "yolo":
            raws.append(x)
            x = decode_head(x, yolo_anchors(d), int(d["classes"]), img_dim)
            yolo.append(x)
        outs.append(x)
    res = torch.cat(yolo, 1)
    if return_layers or return_raw_heads:
        return res, (outs if return_layers else raws)
    return res


def preprocess_u8(tile_hwc_u8, size):
    """utils/transforms.py:89-111 for a square RGB tile: ToTensor (/255, CHW)
    then F.interpolate(size, mode='nearest').  [H,W,3] uint8 -> [3,size,size]."""
    t = torch.from_numpy(np.ascontiguousarray(tile_hwc_u8)).permute(2, 0, 1).to(torch.float32).div(255)
    return F.interpolate(t.unsqueeze(0), size=size, mode="nearest").squeeze(0)


def nearest_src_index(dst_size, src_size):
    """Source index used by F.interpolate(mode='nearest') (ATen
    nearest_neighbor_compute_source_index: floor(dst * float(src/dst)))."""
    scale = f32(src_size) / f32(dst_size)
    return np.minimum(np.floor(np.arange(dst_size, dtype=f32) * scale).astype(np.int64), src_size - 1)


# --------------------------------------------------------------------------
# box utilities (utils/utils.py:36-59, 202-232)
# --------------------------------------------------------------------------


def xywh2xyxy(a):
    """utils/utils.py:53-59, numpy fp32: x1 = cx - w/2 ... (two roundings)."""
    a = np.asarray(a, dtype=f32)
    out = np.empty_like(a)
    half_w = a[..., 2] / f32(2)
    half_h = a[..., 3] / f32(2)
    out[..., 0] = a[..., 0] - half_w
    out[..., 1] = a[..., 1] - half_h
    out[..., 2] = a[..., 0] + half_w
    out[..., 3] = a[..., 1] + half_h
    return out


def bbox_iou_xyxy(head, boxes):
    """utils/utils.py:218-230 with x1y1x2y2=True, one fp32 rounding per op.
    head: [4], boxes: [n,4] -> [n] fp32."""
    head = np.asarray(head, dtype=f32)
    boxes = np.asarray(boxes, dtype=f32)
    one = f32(1)
    ix1 = np.maximum(head[0], boxes[:, 0])
    iy1 = np.maximum(head[1], boxes[:, 1])
    ix2 = np.minimum(head[2], boxes[:, 2])
    iy2 = np.minimum(head[3], boxes[:, 3])
    iw = np.maximum((ix2 - ix1) + one, f32(0))
    ih = np.maximum((iy2 - iy1) + one, f32(0))
    inter = iw * ih
    a1 = ((head[2] - head[0]) + one) * ((head[3] - head[1]) + one)
    a2 = ((boxes[:, 2] - boxes[:, 0]) + one) * ((boxes[:, 3] - boxes[:, 1]) + one)
    return (inter / (((a1 + a2) - inter) + f32(1e-16))).astype(f32)


def rescale_boxes(rows, current_dim, original_shape):
    """utils/utils.py:36-50 on a numpy [k,>=4] fp32 array (returns a copy).

    pad/unpad are python floats in the reference and meet the fp32 tensor as
    scalars, so they are rounded to fp32 at each op."""
    rows = np.array(rows, dtype=f32, copy=True)
    oh, ow = original_shape
    pad_x = max(oh - ow, 0) * (current_dim / max(original_shape))
    pad_y = max(ow - oh, 0) * (current_dim / max(original_shape))
    unpad_h = current_dim - pad_y
    unpad_w = current_dim - pad_x
    for col, pad, unpad, orig in ((0, pad_x, unpad_w, ow), (1, pad_y, unpad_h, oh),
                                  (2, pad_x, unpad_w, ow), (3, pad_y, unpad_h, oh)):
        rows[:, col] = ((rows[:, col] - f32(pad // 2)) / f32(unpad)) * f32(orig)
    return rows


# --------------------------------------------------------------------------
# ATen CPU float32 sum order (torch 2.11, aten/src/ATen/native/cpu/SumKernel.cpp)
# needed because the reference's NMS merges clusters with torch.sum
# (utils/utils.py:267).  V = 8 fp32 lanes is what the build container's ATen
# uses; tests/test_oracle_cpu.py re-checks it against torch.sum at start-up.
# --------------------------------------------------------------------------

ATEN_LANES = 8


def _multi_row_sum(rows):
    """rows: [size, n] fp32 -> [n]; 4-level cascade, one fp32 add per step."""
    size, n = rows.shape
    level_power = max(4, (max(size, 1) - 1).bit_length() // 4)
    step = 1 << level_power
    acc = np.zeros((4, n), dtype=f32)
    i = 0
    while i + step <= size:
        for _ in range(step):
            acc[0] = acc[0] + rows[i]
            i += 1
        for j in range(1, 4):
            acc[j] = acc[j] + acc[j - 1]
            acc[j - 1] = f32(0)
            if (i & ((step - 1) << (j * level_power))) != 0:
                break
    while i < size:
        acc[0] = acc[0] + rows[i]
        i += 1
    for j in range(1, 4):
        acc[0] = acc[0] + acc[j]
    return acc[0].copy()


def aten_sum_outer(mat):
    """torch.sum(mat, 0) for a contiguous [k, n<8] fp32 matrix."""
    return _multi_row_sum(np.ascontiguousarray(mat, dtype=f32))


def _row_sum(x):
    """x: [size, lanes] -> [lanes] (ATen row_sum: 4 interleaved partials)."""
    size, lanes = x.shape
    q = size // 4
    if q:
        ps = _multi_row_sum(x[:4 * q].reshape(q, 4 * lanes)).reshape(4, lanes)
    else:
        ps = np.zeros((4, lanes), dtype=f32)
    for i in range(4 * q, size):
        ps[0] = ps[0] + x[i]
    ps[0] = ps[0] + ps[1]
    ps[0] = ps[0] + ps[2]
    ps[0] = ps[0] + ps[3]
    return ps[0].copy()


def aten_sum_inner(vec, lanes=None):
    """torch.sum(vec) for k contiguous fp32 values."""
    lanes = lanes or ATEN_LANES
    x = np.ascontiguousarray(vec, dtype=f32).reshape(-1)
    k = x.shape[0]
    if k < lanes:
        return f32(_row_sum(x.reshape(k, 1))[0])
    nv = k // lanes
    vacc = _row_sum(x[:nv * lanes].reshape(nv, lanes))
    t = f32(0)
    for e in x[nv * lanes:]:
        t = f32(t + e)
    for l in range(lanes):
        t = f32(t + vacc[l])
    return t


# --------------------------------------------------------------------------
# non_max_suppression (utils/utils.py:235-273)
# --------------------------------------------------------------------------


def nms_faithful(prediction, conf_thres=0.5, nms_thres=0.4):
    """Same structure and tensor ops as utils/utils.py:235-273 (python while
    loop, one bbox_iou call per kept box, torch.sum cluster merge); mutates
    prediction[..., :4] in place like the reference.  This is the function the
    CPU baseline times."""
    xy = prediction[..., :4]
    conv = xy.new(xy.shape)
    conv[..., 0] = xy[..., 0] - xy[..., 2] / 2
    conv[..., 1] = xy[..., 1] - xy[..., 3] / 2
    conv[..., 2] = xy[..., 0] + xy[..., 2] / 2
    conv[..., 3] = xy[..., 1] + xy[..., 3] / 2
    prediction[..., :4] = conv
    result = [None] * len(prediction)
    for bi, pred in enumerate(prediction):
        pred = pred[pred[:, 4] >= conf_thres]
        if not pred.size(0):
            continue
        score = pred[:, 4] * pred[:, 5:].max(1)[0]
        pred = pred[(-score).argsort()]
        cconf, cidx = pred[:, 5:].max(1, keepdim=True)
        det = torch.cat((pred[:, :5], cconf.float(), cidx.float()), 1)
        kept = []
        while det.size(0):
            h = det[0, :4].unsqueeze(0)
            ix1 = torch.max(h[:, 0], det[:, 0]); iy1 = torch.max(h[:, 1], det[:, 1])
            ix2 = torch.min(h[:, 2], det[:, 2]); iy2 = torch.min(h[:, 3], det[:, 3])
            inter = torch.clamp(ix2 - ix1 + 1, min=0) * torch.clamp(iy2 - iy1 + 1, min=0)
            a1 = (h[:, 2] - h[:, 0] + 1) * (h[:, 3] - h[:, 1] + 1)
            a2 = (det[:, 2] - det[:, 0] + 1) * (det[:, 3] - det[:, 1] + 1)
            iou = inter / (a1 + a2 - inter + 1e-16)
            absorb = (iou > nms_thres) & (det[0, -1] == det[:, -1])
            w = det[absorb, 4:5]
            det[0, :4] = (w * det[absorb, :4]).sum(0) / w.sum()
            kept.append(det[0])
            det = det[~absorb]
        if kept:
            result[bi] = torch.stack(kept)
    return result


def nms_exact(image_pred, conf_thres, nms_thres, return_debug=False):
    """One image of utils/utils.py:235-273 in numpy with explicit fp32 steps.

    image_pred: [N, 5+C] fp32 in xywh.  Returns (rows [k,7] fp32, keep_idx [k]
    int64 = original row index of each kept box's head) or (None, empty).
    Ordering among equal scores is unspecified in the reference (unstable
    argsort); here ties are broken by original index."""
    p = np.asarray(image_pred, dtype=f32)
    xyxy = xywh2xyxy(p[:, :4])
    cand = np.nonzero(p[:, 4] >= f32(conf_thres))[0]
    if cand.size == 0:
        return (None, np.zeros(0, np.int64)) + (({},) if return_debug else ())
    cls = p[cand, 5:]
    cidx = cls.argmax(1)                      # first index on ties, like torch.max
    cconf = cls[np.arange(cand.size), cidx]
    score = p[cand, 4] * cconf
    order = np.argsort(-score, kind="stable")
    cand, cidx, cconf = cand[order], cidx[order], cconf[order]
    boxes = xyxy[cand]
    conf = p[cand, 4]
    n = cand.size
    alive = np.ones(n, bool)
    thr = f32(nms_thres)
    rows, keep = [], []
    for i in range(n):
        if not alive[i]:
            continue
        idx = np.nonzero(alive)[0]
        idx = idx[idx >= i]
        iou = bbox_iou_xyxy(boxes[i], boxes[idx])
        member = idx[(iou > thr) & (cidx[idx] == cidx[i])]
        w = conf[member]
        num = aten_sum_outer((w[:, None] * boxes[member]).astype(f32))
        den = aten_sum_inner(w)
        merged = (num / den).astype(f32)
        rows.append(np.concatenate([merged, [conf[i], cconf[i], f32(cidx[i])]]).astype(f32))
        keep.append(cand[i])
        alive[member] = False
    out = np.stack(rows).astype(f32)
    keep = np.asarray(keep, np.int64)
    if return_debug:
        return out, keep, {"order": cand}
    return out, keep


# --------------------------------------------------------------------------
# mergeDetections (core.py:366-423) + combineIfOverlapping (core.py:326-364)
# --------------------------------------------------------------------------


def _combine_pixelsets(b1, b2):
    """core.py:326-364, same cost structure (explicit pixel sets)."""
    x1, y1, w1, h1 = b1
    x2, y2, w2, h2 = b2
    pts1 = {(x, y) for x in range(x1, x1 + w1) for y in range(y1, y1 + h1)}
    pts2 = {(x, y) for x in range(x2, x2 + w2) for y in range(y2, y2 + h2)}
    if not (pts1 & pts2):
        return None
    left, right, top, bottom = 1000000, -1, 1000000, -1
    for px, py in pts1 | pts2:
        left = min(left, px); right = max(right, px)
        top = min(top, py); bottom = max(bottom, py)
    return left, top, right - left, bottom - top


def _combine_closed_form(b1, b2):
    """Same result as _combine_pixelsets without enumerating pixels
    (including the 1000000 / -1 loop sentinels)."""
    x1, y1, w1, h1 = b1
    x2, y2, w2, h2 = b2
    if w1 <= 0 or h1 <= 0 or w2 <= 0 or h2 <= 0:
        return None
    if not (max(x1, x2) < min(x1 + w1, x2 + w2) and max(y1, y2) < min(y1 + h1, y2 + h2)):
        return None
    left = min(1000000, x1, x2)
    top = min(1000000, y1, y2)
    right = max(-1, x1 + w1 - 1, x2 + w2 - 1)
    bottom = max(-1, y1 + h1 - 1, y2 + h2 - 1)
    return left, top, right - left, bottom - top


def _merge_generic(rows, combine):
    entries = [tuple(r) for r in np.asarray(rows, dtype=f32).astype(np.float64).tolist()]
    live = set(entries)
    gone = set()
    changed = True
    while changed:
        changed = False
        snap = list(live)
        for i in range(len(snap)):
            for j in range(i + 1, len(snap)):
                a, b = snap[i], snap[j]
                if not ((a[6] == 1 == b[6]) or (a[6] == 0 == b[6])):
                    continue
                if a in gone or b in gone:
                    continue
                box_a = (int(a[0]), int(a[1]), int(a[2] - a[0]), int(a[3] - a[1]))
                box_b = (int(b[0]), int(b[1]), int(b[2] - b[0]), int(b[3] - b[1]))
                hull = combine(box_a, box_b)
                if hull is None:
                    continue
                new = (hull[0], hull[1], hull[0] + hull[2], hull[1] + hull[3],
                       min(a[4], b[4]), min(a[5], b[5]), a[6])
                if new not in live:
                    live.add(new)
                    live.remove(a)
                    live.remove(b)
                    gone.add(a)
                    gone.add(b)
                    changed = True
    return np.asarray([list(e) for e in live], dtype=f32).reshape(-1, 7)


def merge_faithful(rows):
    """core.py:366-423 with the reference's cost structure (pixel sets, python
    set of tuples; row order = CPython set order).  CPU baseline uses this."""
    return _merge_generic(rows, _combine_pixelsets)


def merge_exact(rows):
    """Same values and row order, closed-form overlap test."""
    return _merge_generic(rows, _combine_closed_form)


# ---- CPython 3.12 hashing / set order, emulated with plain integers ---------
# (Objects/tupleobject.c tuplehash, Python/pyhash.c _Py_HashDouble,
#  Objects/setobject.c).  merge_emulated() must equal merge_exact() bit for bit;
#  it is the specification for oracle/nms_merge.c and the CUDA kernel.

_P61 = (1 << 61) - 1
_M64 = (1 << 64) - 1
_XXP1, _XXP2, _XXP5 = 11400714785074694791, 14029467366897019727, 2870177450012600261


def py_hash_number(v):
    """hash() of a finite python float or int, as a signed integer."""
    if isinstance(v, int) or float(v).is_integer():
        n = int(v)
        x = (abs(n) % _P61) * (1 if n >= 0 else -1)
        return -2 if x == -1 else x
    m, e = math.frexp(v)
    sign = 1
    if m < 0:
        sign, m = -1, -m
    x = 0
    while m:
        x = ((x << 28) & _P61) | (x >> 33)
        m *= 268435456.0
        e -= 28
        y = int(m)
        m -= y
        x += y
        if x >= _P61:
            x -= _P61
    e = e % 61 if e >= 0 else 60 - ((-1 - e) % 61)
    x = ((x << e) & _P61) | (x >> (61 - e))
    x *= sign
    return -2 if x == -1 else x


def py_hash_tuple(t):
    acc = _XXP5
    for item in t:
        lane = py_hash_number(item) & _M64
        acc = (acc + lane * _XXP2) & _M64
        acc = ((acc << 31) | (acc >> 33)) & _M64
        acc = (acc * _XXP1) & _M64
    acc = (acc + (len(t) ^ (_XXP5 ^ 3527539))) & _M64
    if acc == _M64:
        return 1546275796
    return acc - (1 << 64) if acc >> 63 else acc


class EmulatedSet:
    """Open-addressing table with CPython 3.12's probe sequence, dummy reuse
    and growth policy; iteration = ascending slot order."""
    _UNUSED, _DUMMY = 0, 1

    def __init__(self):
        self.mask = 7
        self.slots = [None] * 8          # None | "dummy" | (hash, key)
        self.fill = 0
        self.used = 0

    def _probe(self, h):
        mask = self.mask
        perturb = h & _M64
        i = perturb & mask
        while True:
            span = 10 if i + 9 <= mask else 1
            for s in range(i, i + span):
                yield s
            perturb >>= 5
            i = (i * 5 + 1 + perturb) & mask

    def find(self, key, h):
        for s in self._probe(h):
            e = self.slots[s]
            if e is None:
                return -1
            if e != "dummy" and e[0] == h and e[1] == key:
                return s

    def __contains__(self, key):
        return self.find(key, py_hash_tuple(key)) >= 0

    def add(self, key):
        h = py_hash_tuple(key)
        free = -1
        for s in self._probe(h):
            e = self.slots[s]
            if e is None:
                break
            if e == "dummy":
                free = s
            elif e[0] == h and e[1] == key:
                return
        if free >= 0:
            self.slots[free] = (h, key)
            self.used += 1
            return
        self.slots[s] = (h, key)
        self.fill += 1
        self.used += 1
        if self.fill * 5 >= self.mask * 3:
            self._resize(self.used * 2 if self.used > 50000 else self.used * 4)

    def _resize(self, minused):
        size = 8
        while size <= minused:
            size <<= 1
        old = [e for e in self.slots if e is not None and e != "dummy"]
        self.mask = size - 1
        self.slots = [None] * size
        for h, key in old:
            for s in self._probe(h):
                if self.slots[s] is None:
                    self.slots[s] = (h, key)
                    break
        self.fill = self.used = len(old)

    def remove(self, key):
        s = self.find(key, py_hash_tuple(key))
        assert s >= 0
        self.slots[s] = "dummy"
        self.used -= 1

    def items(self):
        return [e[1] for e in self.slots if e is not None and e != "dummy"]


def merge_emulated(rows):
    """mergeDetections with every CPython detail spelled out (no python set)."""
    entries = [tuple(r) for r in np.asarray(rows, dtype=f32).astype(np.float64).tolist()]
    live = EmulatedSet()
    for e in entries:
        live.add(e)
    gone = []
    changed = True
    while changed:
        changed = False
        snap = live.items()
        for i in range(len(snap)):
            for j in range(i + 1, len(snap)):
                a, b = snap[i], snap[j]
                if not ((a[6] == 1 == b[6]) or (a[6] == 0 == b[6])):
                    continue
                if any(a == g for g in gone) or any(b == g for g in gone):
                    continue
                box_a = (int(a[0]), int(a[1]), int(a[2] - a[0]), int(a[3] - a[1]))
                box_b = (int(b[0]), int(b[1]), int(b[2] - b[0]), int(b[3] - b[1]))
                hull = _combine_closed_form(box_a, box_b)
                if hull is None:
                    continue
                new = (hull[0], hull[1], hull[0] + hull[2], hull[1] + hull[3],
                       min(a[4], b[4]), min(a[5], b[5]), a[6])
                if new not in live:
                    live.add(new)
                    live.remove(a)
                    live.remove(b)
                    gone.append(a)
                    gone.append(b)
                    changed = True
    return np.asarray([list(e) for e in live.items()], dtype=f32).reshape(-1, 7)


# --------------------------------------------------------------------------
# evaluation statistics (utils/utils.py:69-190) - restated statement by statement
# --------------------------------------------------------------------------


def compute_ap_faithful(recall, precision):
    """utils/utils.py:123-149."""
    mrec = np.concatenate(([0.0], recall, [1.0]))
    mpre = np.concatenate(([0.0], precision, [0.0]))
    for i in range(mpre.size - 1, 0, -1):
        mpre[i - 1] = np.maximum(mpre[i - 1], mpre[i])
    i = np.where(mrec[1:] != mrec[:-1])[0]
    return np.sum((mrec[i + 1] - mrec[i]) * mpre[i + 1])


def ap_per_class_faithful(tp, conf, pred_cls, target_cls):
    """utils/utils.py:69-120."""
    i = np.argsort(-conf)
    tp, conf, pred_cls = tp[i], conf[i], pred_cls[i]
    unique_classes = np.unique(target_cls)
    ap, p, r = [], [], []
    for c in unique_classes:
        i = pred_cls == c
        n_gt = (target_cls == c).sum()
        n_p = i.sum()
        if n_p == 0 and n_gt == 0:
            continue
        elif n_p == 0 or n_gt == 0:
            ap.append(0)
            r.append(0)
            p.append(0)
        else:
            fpc = (1 - tp[i]).cumsum()
            tpc = (tp[i]).cumsum()
            recall_curve = tpc / (n_gt + 1e-16)
            r.append(recall_curve[-1])
            precision_curve = tpc / (tpc + fpc)
            p.append(precision_curve[-1])
            ap.append(compute_ap_faithful(recall_curve, precision_curve))
    p, r, ap = np.array(p), np.array(r), np.array(ap)
    f1 = 2 * p * r / (p + r + 1e-16)
    return p, r, ap, f1, unique_classes.astype("int32")


def get_batch_statistics_faithful(outputs, targets, iou_threshold):
    """utils/utils.py:152-190, with bbox_iou_xyxy (this file) in place of the reference's bbox_iou."""
    batch_metrics = []
    for sample_i in range(len(outputs)):
        if outputs[sample_i] is None:
            continue
        output = np.asarray(outputs[sample_i], dtype=f32)
        pred_boxes, pred_scores, pred_labels = output[:, :4], output[:, 4], output[:, -1]
        true_positives = np.zeros(pred_boxes.shape[0])
        annotations = targets[targets[:, 0] == sample_i][:, 1:]
        target_labels = annotations[:, 0] if len(annotations) else []
        if len(annotations):
            detected_boxes = []
            target_boxes = annotations[:, 1:].astype(f32)
            for pred_i, (pred_box, pred_label) in enumerate(zip(pred_boxes, pred_labels)):
                if len(detected_boxes) == len(annotations):
                    break
                if pred_label not in target_labels:
                    continue
                ious = bbox_iou_xyxy(pred_box, target_boxes)
                box_index = int(np.argmax(ious))
                if ious[box_index] >= iou_threshold and box_index not in detected_boxes:
                    true_positives[pred_i] = 1
                    detected_boxes += [box_index]
        batch_metrics.append([true_positives, pred_scores, pred_labels])
    return batch_metrics


# --------------------------------------------------------------------------
# consensus CAA filter: core.py:109-159 get256Img, 161-205 Net, 210-221 getClassPreds,
# 425-452 filterDetectionsByCAAModel
# --------------------------------------------------------------------------
CAA_CHANNELS = (3, 16, 32, 48, 64, 80, 96)


def caa_net_forward(state, x, dtype=torch.float64):
    """core.py:165-205 in eval mode: six (Conv2d 3x3 pad 1, BatchNorm2d, ReLU, MaxPool2d 2/2) stages on a
    [B,3,256,256] tensor, flatten in NCHW order, one Linear(96*4*4, 3).  ``state`` uses the reference's
    state_dict names (features.{4k}.weight/bias, features.{4k+1}.*, classifier.0.*).  Returns the LOGITS."""
    y = x.to(dtype)
    for k in range(6):
        c, b = "features.%d." % (4 * k), "features.%d." % (4 * k + 1)
        y = F.conv2d(y, state[c + "weight"].to(dtype), state[c + "bias"].to(dtype), padding=1)
        y = F.batch_norm(y, state[b + "running_mean"].to(dtype), state[b + "running_var"].to(dtype),
                         state[b + "weight"].to(dtype), state[b + "bias"].to(dtype), False, 0.0, 1e-5)
        y = F.max_pool2d(F.relu(y), 2, 2)
    y = y.reshape(y.shape[0], -1)
    return F.linear(y, state["classifier.0.weight"].to(dtype), state["classifier.0.bias"].to(dtype))


def get256img_box_faithful(bbox_coord, size=1536):
    """core.py:109-159: the crop window (x_left, x_right, y_top, y_bottom) the reference slices for a box
    (top-left x, top-left y, width, height).  Same branch order, including its quirks: the comparisons are strict,
    later ``if``s overwrite earlier ones, and a window that no branch sets stays (-1, -1, -1, -1)."""
    cx = int(bbox_coord[0] + (bbox_coord[2] / 2))
    cy = int(bbox_coord[1] + (bbox_coord[3] / 2))
    x_left_valid, x_right_valid = cx - 128 > 0, cx + 128 < size
    y_top_valid, y_bottom_valid = cy - 128 > 0, cy + 128 < size
    x_valid, y_valid = x_left_valid and x_right_valid, y_top_valid and y_bottom_valid
    xl, xr, yt, yb = -1, -1, -1, -1
    if x_valid and y_valid:
        xl, xr, yt, yb = cx - 128, cx + 128, cy - 128, cy + 128
    elif x_valid and not y_valid:
        if y_bottom_valid:
            xl, xr, yt, yb = cx - 128, cx + 128, 0, 256
        if y_top_valid:
            xl, xr, yt, yb = cx - 128, cx + 128, size - 256, size
    elif not x_valid and y_valid:
        if x_left_valid:
            xl, xr, yt, yb = size - 256, size, cy - 128, cy + 128
        if x_right_valid:
            xl, xr, yt, yb = 0, 256, cy - 128, cy + 128
    else:
        if x_right_valid and y_bottom_valid:
            xl, xr, yt, yb = 0, 256, 0, 256
        if x_left_valid and y_bottom_valid:
            xl, xr, yt, yb = size - 256, size, 0, 256
        if x_right_valid and y_top_valid:
            xl, xr, yt, yb = 0, 256, size - 256, size
        if x_left_valid and y_top_valid:
            xl, xr, yt, yb = size - 256, size, size - 256, size
    return xl, xr, yt, yb


def caa_normalise(crop_rgb_u8, mean, std):
    """core.py:214-217: ToTensor (/255, HWC -> CHW) + Normalize(mean, std) in fp32, as torchvision computes them."""
    t = torch.from_numpy(np.ascontiguousarray(crop_rgb_u8)).permute(2, 0, 1).to(torch.float32).div(255)
    m = torch.as_tensor(mean, dtype=torch.float32).view(-1, 1, 1)
    s = torch.as_tensor(std, dtype=torch.float32).view(-1, 1, 1)
    return t.sub(m).div(s)


def caa_filter_faithful(img_rgb_u8, detections, classes, state, mean, std, dtype=torch.float64, return_logits=False):
    """core.py:425-452 on an in-memory tile (``img_rgb_u8`` = cv2.imread + the BGR->RGB swap of core.py:214):
    detections [n,7] -> set of row tuples -> per box a 256^2 crop centred on it -> sigmoid(Net) -> CAA boxes with
    CAA_pred <= 0.5 are dropped.  Row order is the set's iteration order, as in the reference.  Returns an [m,7]
    float64-valued list of rows (and, on request, the CAA logit of every distinct row, in that order)."""
    rows = [tuple(r) for r in np.asarray(detections).tolist()]
    kept, logits = [], []
    for x1, y1, x2, y2, conf, cls_conf, cls_pred in set(rows):
        xl, xr, yt, yb = get256img_box_faithful((int(x1), int(y1), int(x2 - x1), int(y2 - y1)), img_rgb_u8.shape[0])
        crop = img_rgb_u8[yt:yb, xl:xr]
        with torch.no_grad():
            out = caa_net_forward(state, caa_normalise(crop, mean, std).unsqueeze(0), dtype)[0]
        caa_pred = float(torch.sigmoid(out[2]))
        logits.append(float(out[2]))
        if classes[int(cls_pred)] == "CAA" and caa_pred <= 0.5:
            continue
        kept.append((x1, y1, x2, y2, conf, cls_conf, cls_pred))
    return (kept, logits) if return_logits else kept


# --------------------------------------------------------------------------
# slide tiling: crop.py:13-25 (dzsave layout google, tile_size 1536, depth one) + the detector's resize.
# PARITY UNPINNED: pyvips / openslide are absent, the reference's tiler cannot be run here; this restates the published
# dzsave layout (tile (row, col) at (row*T, col*T), edge tiles padded white) without the JPEG round trip.
# --------------------------------------------------------------------------
def wsi_tiles_faithful(slide, tile_size=1536, background=255):
    """-> dict {(row, col): [T,T,3] uint8} in the reference's two steps: pad, cut."""
    h, w = slide.shape[:2]
    rows, cols = -(-h // tile_size), -(-w // tile_size)
    canvas = np.full((rows * tile_size, cols * tile_size, 3), background, dtype=np.uint8)
    canvas[:h, :w] = slide
    return {(r, c): canvas[r * tile_size:(r + 1) * tile_size, c * tile_size:(c + 1) * tile_size]
            for r in range(rows) for c in range(cols)}


def wsi_detector_inputs_faithful(slide, img_size=416, tile_size=1536):
    """Each padded tile through the nearest resize (utils/transforms.py:104-111) -> {(row, col): [S,S,3] uint8}."""
    idx = nearest_src_index(img_size, tile_size)
    return {k: t[idx][:, idx] for k, t in wsi_tiles_faithful(slide, tile_size).items()}


def divide_and_conquer_faithful(height, width, multiple=1536 * 16):
    """crop.py:63-92 with the branch structure of the reference: full-size crop unless the chunk touches the bottom
    and/or right edge, where the except-branches shrink it.  -> [(indice, left, top, w, h)]."""
    out, indice = [], 0
    tops, lefts = list(range(0, height, multiple)), list(range(0, width, multiple))
    for top in tops:
        for left in lefts:
            indice += 1
            if left + multiple <= width and top + multiple <= height:          # vips crop succeeds
                out.append((indice, left, top, multiple, multiple))
            elif top == tops[-1] and left == lefts[-1]:
                out.append((indice, left, top, width - left, height - top))
            else:
                if top == tops[-1]:
                    out.append((indice, left, top, multiple, height - top))
                if left == lefts[-1]:
                    out.append((indice, left, top, width - left, multiple))
    return out
