#!/usr/bin/env python
"""Where the forward's time goes at SUSTAINED clocks: per-layer kernel spans, launch gaps,
prologues and tails of one forward taken in the middle of a seconds-long back-to-back loop.

    python tools/attribution.py [--batch 256] [--heat 2.5] [--reps 12] [--out file]

Three measurements, all after `--heat` seconds of back-to-back forwards (power cap engaged):
  1. whole forward, CUDA events, back to back (what bench.py's roofline uses);
  2. per-layer event times (ay_forward_profile) in the same hot state (mean over reps);
  3. per-CTA %globaltimer stamps (ay_forward_stamps: no events between launches, so gaps and the overlap of the
     chained launches are what a normal forward sees): span = last exit - first entry of a layer's CTAs,
     gap = first "dependency resolved" of the next conv - last exit of this one, prologue = entry -> end of prologue,
     tail = mean over CTAs of (last exit - own exit).
Run on the GPU box."""
import argparse
import ctypes
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--net", default="yolov3-custom")
    ap.add_argument("--heat", type=float, default=2.5)
    ap.add_argument("--reps", type=int, default=12)
    ap.add_argument("--tile", type=int, default=1536)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import numpy as np
    import torch
    from amyloid_yolo_paper_b200 import _lib, synth
    from amyloid_yolo_paper_b200.config import ensure_cfg
    from amyloid_yolo_paper_b200.models import Darknet

    dev = torch.device("cuda", 0)
    model = Darknet(ensure_cfg(args.net))
    synth.synthetic_state_dict(model, seed=0, conf_bias=-3.5)
    eng = model.engine(device=dev, max_batch=args.batch)
    B = args.batch
    tiles = torch.randint(0, 256, (B, args.tile, args.tile, 3), dtype=torch.uint8, device=dev)
    out = torch.empty((B, eng.rows, eng.cols), dtype=torch.float32, device=dev)
    n = len(model.module_defs)
    st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    tp, op = ctypes.c_void_p(tiles.data_ptr()), ctypes.c_void_p(out.data_ptr())

    def fwd():
        _lib.check(eng.lib.ay_forward(eng.ctx, tp, _lib.AY_IN_U8_HWC, B, args.tile, args.tile, op, st))

    clk = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits",
                            "-lms", "100"], stdout=subprocess.PIPE, text=True)
    # ---- heat
    t0 = time.time()
    while time.time() - t0 < args.heat:
        for _ in range(10):
            fwd()
        torch.cuda.synchronize()
    # ---- 1. whole forward, hot
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        fwd()
    e1.record()
    torch.cuda.synchronize()
    whole = e0.elapsed_time(e1) / 20 * 1e3
    # ---- 2./3. interleaved with plain forwards so the device stays hot
    ms = (ctypes.c_float * n)()
    words = eng.lib.ay_prof_words()
    stamps = (ctypes.c_ulonglong * (words * n))()
    lay = np.zeros(n)
    acc = {k: np.zeros(n) for k in ("span", "gap", "prol", "tail", "depw")}
    total_span = 0.0
    convs = [i for i, d in enumerate(model.module_defs) if d["type"] == "convolutional"]
    for r in range(args.reps):
        for _ in range(3):
            fwd()
        _lib.check(eng.lib.ay_forward_profile(eng.ctx, tp, _lib.AY_IN_U8_HWC, B, args.tile, args.tile, op, st, ms, n))
        lay += np.array(list(ms)) * 1e3
        for _ in range(3):
            fwd()
        _lib.check(eng.lib.ay_forward_stamps(eng.ctx, tp, _lib.AY_IN_U8_HWC, B, args.tile, args.tile, op, st, stamps, n))
        S = np.frombuffer(stamps, dtype=np.uint64).reshape(n, words).astype(np.float64)
        first_entry, last_exit = None, None
        prev_exit = None
        for i in convs:
            blk = S[i, 16:].reshape(4, -1)
            live = blk[0] > 0
            entry, prol, exit_, dep = blk[0][live], blk[1][live], blk[2][live], blk[3][live]
            acc["span"][i] += (exit_.max() - entry.min()) / 1e3
            acc["prol"][i] += (prol - entry).mean() / 1e3
            acc["tail"][i] += (exit_.max() - exit_).mean() / 1e3
            acc["depw"][i] += (dep - prol).mean() / 1e3
            if prev_exit is not None:
                acc["gap"][i] += (dep.min() - prev_exit) / 1e3
            prev_exit = exit_.max()
            first_entry = entry.min() if first_entry is None else first_entry
            last_exit = exit_.max()
        total_span += (last_exit - first_entry) / 1e3
    clk.terminate()
    clocks = [l.split(",") for l in clk.stdout.read().strip().splitlines()]
    sm = sorted(float(c[0]) for c in clocks[len(clocks) // 3:] if len(c) == 2)
    R = float(args.reps)
    lines = ["idx layer                       event_us  span_us   gap_us  prologue  wait_prev  tail_us"]
    prev_c = int(model.hyperparams["channels"])
    for i, d in enumerate(model.module_defs):
        cc, hh, ww = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        eng.lib.ay_net_layer_shape(eng.ctx, i, ctypes.byref(cc), ctypes.byref(hh), ctypes.byref(ww))
        if d["type"] == "convolutional":
            desc = "conv %4d->%4d @%3d k%ss%s" % (prev_c, cc.value, hh.value, d["size"], d["stride"])
            lines.append("%3d %-27s %8.1f %8.1f %8.1f %9.1f %10.1f %8.1f" % (
                i, desc, lay[i] / R, acc["span"][i] / R, acc["gap"][i] / R, acc["prol"][i] / R, acc["depw"][i] / R,
                acc["tail"][i] / R))
        prev_c = cc.value
    ev_sum = lay.sum() / R
    lines += [
        "",
        "whole forward, back to back, hot (CUDA events):       %9.1f us" % whole,
        "sum of per-layer event times, hot:                    %9.1f us  (includes %d event records)" % (ev_sum, n + 1),
        "first conv entry -> last conv exit (stamps, no events): %7.1f us" % (total_span / R),
        "  sum of conv spans:                                  %9.1f us" % (acc["span"].sum() / R),
        "  sum of gaps (prev last exit -> next dependency resolved): %5.1f us over %d launches" % (acc["gap"].sum() / R, len(convs) - 1),
        "  sum of mean prologues (entry -> barriers/TMEM/bias ready): %4.1f us (overlapped with the previous layer under PDL)" % (acc["prol"].sum() / R),
        "  sum of mean tails (CTA idle before the layer's last exit): %4.1f us" % (acc["tail"].sum() / R),
        "SM clock while measuring: median %.0f MHz, min %.0f, max %.0f (%d samples); pdl=%s" % (
            sm[len(sm) // 2] if sm else 0, sm[0] if sm else 0, sm[-1] if sm else 0, len(sm),
            "off" if os.environ.get("AY_NO_PDL") else "on"),
        "batch %d" % B,
    ]
    text = "\n".join(lines)
    print(text)
    if args.out:
        open(args.out, "w").write(text + "\n")


if __name__ == "__main__":
    main()
