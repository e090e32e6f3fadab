#!/usr/bin/env python
"""Per-layer times of one forward (CUDA events between cfg blocks, ``ay_forward_profile``), next to the
layer's roofline bound, optionally under the convolution kernel's ablation switches.

    python tools/layer_times.py [--batch 256] [--net yolov3-custom] [--dbg 0,1,2,4,8] [--reps 3]

Columns: layer, type, Cin->Cout @H, k/s, best-of-reps microseconds per switch setting, and the bound
max(FLOP / bf16 peak, minimal bytes / HBM peak) in microseconds.  Run on the GPU box."""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--net", default="yolov3-custom")
    ap.add_argument("--dbg", default="0")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--tile", type=int, default=1536)
    ap.add_argument("--out", default="")
    ap.add_argument("--counters", action="store_true")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp16", "fp32"])
    args = ap.parse_args()
    import torch
    from amyloid_yolo_paper_b200 import _lib, synth
    from amyloid_yolo_paper_b200.config import ensure_cfg
    from amyloid_yolo_paper_b200.models import Darknet

    dev = torch.device("cuda", 0)
    model = Darknet(ensure_cfg(args.net))
    model.precision = args.precision
    synth.synthetic_state_dict(model, seed=0, conf_bias=-3.5)
    eng = model.engine(device=dev, max_batch=args.batch)
    B = args.batch
    tiles = torch.randint(0, 256, (B, args.tile, args.tile, 3), dtype=torch.uint8, device=dev)
    out = torch.empty((B, eng.rows, eng.cols), dtype=torch.float32, device=dev)
    n = len(model.module_defs)
    ms = (ctypes.c_float * n)()
    st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    peaks = {"bf16": 1385.1, "hbm": 6551.7}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        d = json.load(open(pk))
        peaks = {"bf16": float(d.get("bf16_tflops_sustained") or d["bf16_tflops"]), "hbm": float(d["hbm_gbs"])}
    settings = [int(v) for v in args.dbg.split(",")]
    cols = {}
    for dbg in settings:
        _lib.check(eng.lib.ay_debug_set(eng.ctx, dbg))
        best = [1e9] * n
        for r in range(args.reps + 1):
            _lib.check(eng.lib.ay_forward_profile(eng.ctx, ctypes.c_void_p(tiles.data_ptr()), _lib.AY_IN_U8_HWC, B,
                                                  args.tile, args.tile, ctypes.c_void_p(out.data_ptr()), st, ms, n))
            if r:
                best = [min(a, b) for a, b in zip(best, ms)]
        cols[dbg] = best
    # whole forward without the per-layer events (launch gaps included), 10 back-to-back calls
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        _lib.check(eng.lib.ay_forward(eng.ctx, ctypes.c_void_p(tiles.data_ptr()), _lib.AY_IN_U8_HWC, B, args.tile, args.tile,
                                      ctypes.c_void_p(out.data_ptr()), st))
    e1.record()
    torch.cuda.synchronize()
    whole_ms = e0.elapsed_time(e1) / 10
    cnt = (ctypes.c_ulonglong * (16 * n))()
    _lib.check(eng.lib.ay_debug_counters(eng.ctx, cnt, n))
    _lib.check(eng.lib.ay_debug_set(eng.ctx, 0))
    # shapes
    c, h = int(model.hyperparams["channels"]), int(model.hyperparams["height"])
    shapes = []
    chans = []
    for i, d in enumerate(model.module_defs):
        cc, hh, ww = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        eng.lib.ay_net_layer_shape(eng.ctx, i, ctypes.byref(cc), ctypes.byref(hh), ctypes.byref(ww))
        shapes.append((cc.value, hh.value))
    lines = []
    tot = {k: 0.0 for k in settings}
    tot_bound = 0.0
    prev_c = c
    for i, d in enumerate(model.module_defs):
        co, ho = shapes[i]
        t = d["type"]
        desc, bound = t, 0.0
        if t == "convolutional":
            k, s = int(d["size"]), int(d["stride"])
            cin = prev_c
            flop = 2.0 * B * ho * ho * co * cin * k * k
            byts = 2.0 * B * (ho * s * ho * s * cin + ho * ho * co)
            if i + 1 < n and model.module_defs[i + 1]["type"] == "shortcut":
                byts += 2.0 * B * ho * ho * co
            bound = max(flop / (peaks["bf16"] * 1e12), byts / (peaks["hbm"] * 1e9)) * 1e6
            desc = "conv %4d->%4d @%3d k%ds%d" % (cin, co, ho, k, s)
        prev_c = co
        if all(cols[k][i] < 0.002 for k in settings):
            continue
        for k in settings:
            tot[k] += cols[k][i] * 1e3
        tot_bound += bound
        lines.append("%3d %-28s " % (i, desc) + " ".join("%8.1f" % (cols[k][i] * 1e3) for k in settings) +
                     "   bound %7.1f" % bound)
    if args.counters:
        lines.append("cycle counters of CTA 0 (last setting): per stage: producer wait-empty / issue | MMA wait-full / issue+commit"
                     " ; per tile: MMA wait-acc, MMA total, epilogue warp 0 total; tiles, stages")
        for i, d in enumerate(model.module_defs):
            c = cnt[16 * i:16 * i + 16]
            if d["type"] != "convolutional" or c[7] == 0:
                continue
            st, tl = float(c[7]), float(c[6])
            lines.append("%3d  prod %6.0f %6.0f | mma %6.0f %6.0f ; acc-wait %7.0f total %7.0f epi %7.0f ; tiles %5d stages %6d" % (
                i, c[0] / st, c[1] / st, c[3] / st, c[4] / st, c[2] / tl, c[5] / tl, c[8] / max(1.0, float(c[9])), c[6], c[7]))
    hdr = "idx layer                        " + " ".join("dbg=%-4d" % k for k in settings)
    text = "\n".join([hdr] + lines + ["sum us                           " + " ".join("%8.0f" % tot[k] for k in settings) +
                                      "   bound %7.0f" % tot_bound,
                                      "whole forward, 10 calls back to back: %.0f us per call" % (whole_ms * 1e3),
                                      "batch %d, peaks %s" % (B, peaks)])
    print(text)
    if args.out:
        open(args.out, "w").write(text + "\n")


if __name__ == "__main__":
    main()
