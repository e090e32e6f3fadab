#!/usr/bin/env python
"""Benchmark of the tile-detection hot path (BASELINE.json metric: tiles/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config c3|c1|c2|c5]

Workloads (BASELINE.json `configs`; default c3, which is also C4 when N > 1):
  c3  yolov3-custom (Darknet-53, 2 classes), synthetic 1536x1536 uint8 WSI tiles nearest-resized to 416 inside the
      first convolution, bf16 tensor-core forward + fused YOLO decode + cluster-merging NMS (conf 0.8, nms 0.4) +
      rescale_boxes + mergeDetections, B = 256 tiles per GPU per step.  With N > 1 the tiles are sharded by rank through
      ``sweep.sweep`` and the detection lists are all-gathered over NCCL once per sweep (config C4).
  c1  the same network at B = 1 (the shape of the reference's only benchmark loop, validation.py:282-366).
  c2  yolov3-tiny (80 classes), fp32 [64,3,416,416] inputs, B = 64.
  c5  NMS stress: ay_nms alone on 128 tiles x 10 647 decoded rows, conf_thres 0.01 (every row a candidate).

One JSON line:
* ``value``    - tiles/s with the inputs already resident in HBM (CUDA events on the launching stream, max over ranks).
* ``e2e``      - the same metric through the reference-facing API with HOST buffers, host<->device copies inside the
                 timed region (c3: ``TileDetector.detect_stream`` over pinned uint8 [B,416,416,3] batches - the 416^2
                 pixels the resize samples, gathered by the loader; the fp32 [B,3,416,416] format is reported beside it).
* ``roofline`` - c1-c3: the conv stack (``ay_forward``) FLOPs / its CUDA-event time vs the measured bf16 peak;
                 c5: the IoU bit-mask kernel's algorithmic bytes / its CUDA-event time vs the measured HBM peak.
* ``cpu_baseline`` - the reference's own CPU implementation of the path on a bounded sample, all host threads:
                 the UNMODIFIED reference modules when ``oracle/_ref/reference_py.zip`` (or the checkout) is present
                 (kind "reference"), else the oracle port (kind "port").
``--impl reference`` times that CPU path as its own arm.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FLOP_PER_TILE = {"yolov3-custom": 65297143808, "yolov3-tiny": 5564961792}   # SURVEY.md §8(d), convs only
METRIC = "tiles/sec (416^2 YOLOv3 conv+decode+NMS)"
CONFIGS = {
    "c3": {"net": "yolov3-custom", "batch": 256, "conf_bias": "-6,-6,0.62", "cpu_tiles": 8, "ref_tiles": 4},
    "c1": {"net": "yolov3-custom", "batch": 1, "conf_bias": "-6,-6,0.62", "cpu_tiles": 1, "ref_tiles": 1},
    "c2": {"net": "yolov3-tiny", "batch": 64, "conf_bias": "-4,-0.38", "cpu_tiles": 16, "ref_tiles": 8},
    "c5": {"net": None, "batch": 128, "conf_bias": "", "cpu_tiles": 2, "ref_tiles": 1},
}


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", type=str, default="ours", choices=["ours", "reference"])
    p.add_argument("--config", type=str, default="c3", choices=sorted(CONFIGS))
    p.add_argument("--batch", type=int, default=0, help="tiles per GPU per step (default: the config's)")
    p.add_argument("--net", type=str, default="", choices=[""] + sorted(FLOP_PER_TILE))
    p.add_argument("--tile", type=int, default=1536, help="source tile edge (uint8 HWC)")
    p.add_argument("--conf_bias", type=str, default="",
                   help="objectness bias of the synthetic checkpoint per head: a few dozen candidates per tile at conf 0.8, "
                        "mostly small boxes as in the reference's plaque data")
    p.add_argument("--cpu-sample", type=int, default=0, help="tiles in the bounded CPU-baseline sample")
    p.add_argument("--precision", type=str, default="bf16", choices=["bf16", "fp32"],
                   help="fp32: the accuracy mode (1e-4 bar) - 3 x tf32 on the tensor cores (dtype f32x3), or with AY_F32_FFMA=1 "
                        "the CUDA-core FFMA kernel of round 1 (dtype f32-ffma)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true", help="device-resident value and roofline only (short runs under ncu)")
    args = p.parse_args()
    cfg = CONFIGS[args.config]
    args.net = args.net or cfg["net"]
    args.batch = args.batch or cfg["batch"]
    args.conf_bias = args.conf_bias or cfg["conf_bias"]
    return args


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return {"bf16": float(d.get("bf16_tflops_sustained") or d["bf16_tflops"]), "bf16_burst": float(d["bf16_tflops"]),
                "hbm": float(d["hbm_gbs"]), "source": "MEASURED_PEAKS.json"}
    return {"bf16": 1400.0, "bf16_burst": 1590.0, "hbm": 6650.0, "source": "fallback (B200_PROFILING.md)"}


def profiled_traffic(batch):
    """DRAM bytes of the conv stack per ay_forward, from the committed ncu capture (profiles/*_conv_traffic.json,
    tools/traffic_json.py), scaled per tile; None when no capture is committed."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_conv_traffic.json")))
    if not files:
        return None, None
    d = json.load(open(files[-1]))
    return float(d["bytes_per_tile"]) * batch, os.path.basename(files[-1])


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    FIELDS = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None
        self.spans = []          # wall-clock [begin, end] of the timed regions: only samples inside them are reported
        self._t0 = None

    def begin(self):
        self._t0 = time.time()

    def end(self):
        if self._t0 is not None:
            self.spans.append((self._t0, time.time()))
            self._t0 = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(prefix="clocks_", suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        import datetime
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = []                                             # (epoch seconds or None, sm MHz, set of active reasons)
        for line in open(self.path):
            parts = [v.strip() for v in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                mhz = float(parts[1])
                out["sm_max_mhz"] = float(parts[2])
            except ValueError:
                continue
            try:
                ts = datetime.datetime.strptime(parts[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
            except ValueError:
                ts = None
            rows.append((ts, mhz, {n for n, v in zip(names, parts[4:8]) if v.lower().startswith("active")}))
        os.unlink(self.path)
        # a sample counts when it was taken inside a timed region (a query takes a few ms: 20 ms of slack at the start)
        inside = [r for r in rows if r[0] is not None and any(a + 0.02 <= r[0] <= b for a, b in self.spans)]
        scope = "timed regions"
        if not inside:
            inside, scope = rows, "whole run (no sample fell inside a timed region)"
        sm = sorted(r[1] for r in inside)
        if sm:
            out["sm_mhz"] = sm[len(sm) // 2]
        out["samples"] = len(sm)
        out["reasons"] = sorted(set().union(*[r[2] for r in inside])) if inside else []
        out["scope"] = scope
        return out


# ----------------------------------------------------------------------------- inputs
def device_tiles(batch, size, device, seed, first_tile):
    """uint8 [batch,size,size,3] generated ON the device: smooth low-frequency
    field + noise, a function of (seed, tile id) only.  Data generation is not part of the path."""
    import torch
    import torch.nn.functional as F
    out = torch.empty((batch, size, size, 3), dtype=torch.uint8, device=device)
    chunk = 8
    for lo in range(0, batch, chunk):
        n = min(chunk, batch - lo)
        imgs = []
        for t in range(n):
            g = torch.Generator(device=device).manual_seed((seed << 20) + first_tile + lo + t)
            low = torch.rand((1, 3, 24, 24), device=device, generator=g)
            img = F.interpolate(low, size=(size, size), mode="bilinear", align_corners=False)
            img = 90.0 + 130.0 * img + 12.0 * torch.randn((1, 3, size, size), device=device, generator=g)
            imgs.append(img)
        img = torch.cat(imgs, 0).clamp_(0, 255).to(torch.uint8)
        out[lo:lo + n] = img.permute(0, 2, 3, 1)
    return out


def build_model(net, conf_bias):
    from amyloid_yolo_paper_b200 import synth
    from amyloid_yolo_paper_b200.config import ensure_cfg
    from amyloid_yolo_paper_b200.models import Darknet
    cfg = ensure_cfg(net)
    model = Darknet(cfg)
    bias = tuple(float(v) for v in str(conf_bias).split(","))
    sd = synth.synthetic_state_dict(model, seed=0, conf_bias=bias, wh_gain=0.15)
    return cfg, model, sd


# ----------------------------------------------------------------------------- CPU path (reference / oracle port)
class CpuPath:
    """The body of detect.py's loop on the host (detect.py:88-133, validation.py:333-346 without the CAA filter and
    file I/O): forward -> non_max_suppression -> rescale_boxes -> mergeDetections, tensors already in RAM.
    kind "reference": the unmodified reference modules (oracle/_ref archive or the mounted checkout);
    kind "port": the oracle's restatement (oracle/restate.py)."""

    def __init__(self, net, conf_bias):
        import torch
        from oracle import ref_import
        from oracle import restate as R
        self.R = R
        self.torch = torch
        self.cfg, _, self.sd = build_model(net, conf_bias) if net else (None, None, None)
        self.kind = "port"
        self.ref = None
        if ref_import.available():
            try:
                self.ref = ref_import.load()
                self.kind = "reference"
            except Exception as e:                       # an import problem on the box: fall back to the port, and say so
                print("reference modules unavailable (%s); timing the oracle port" % e, file=sys.stderr)
        if net and self.ref is not None:
            models = self.ref[0]
            self.model = models.Darknet(self.cfg)
            self.model.load_state_dict(self.sd)
            self.model.eval()
        elif net:
            self.blocks = R.parse_cfg(self.cfg)

    def step(self, x, conf_thres, nms_thres, orig):
        torch, R = self.torch, self.R
        n = 0
        if self.ref is not None:
            _, uu, core = self.ref
            with torch.no_grad():
                dec = self.model(x)
                dets = uu.non_max_suppression(dec, conf_thres, nms_thres)
            for d in dets:
                if d is None:
                    continue
                d = uu.rescale_boxes(d, x.shape[2], orig)
                n += len(core.mergeDetections(d))
            return n
        with torch.no_grad():
            dec = R.darknet_forward(self.blocks, self.sd, x)
        for d in R.nms_faithful(dec, conf_thres, nms_thres):
            if d is None:
                continue
            n += len(R.merge_faithful(R.rescale_boxes(d.numpy().copy(), x.shape[2], orig)))
        return n

    def nms_step(self, pred, conf_thres, nms_thres):
        if self.ref is not None:
            out = self.ref[1].non_max_suppression(pred.clone(), conf_thres, nms_thres)
        else:
            out = self.R.nms_faithful(pred.clone(), conf_thres, nms_thres)
        return sum(0 if d is None else len(d) for d in out)


def cpu_inputs(n_tiles, size=416, uniform=False):
    import torch
    g = torch.Generator().manual_seed(1234)
    if uniform:                                 # config C2: fp32 U[0,1) inputs
        return torch.rand((n_tiles, 3, size, size), generator=g)
    low = torch.rand((n_tiles, 3, 24, 24), generator=g)
    x = torch.nn.functional.interpolate(low, size=(size, size), mode="bilinear")
    return ((90.0 + 130.0 * x + 12.0 * torch.randn(x.shape, generator=g)).clamp(0, 255).floor() / 255.0).contiguous()


def run_cpu_baseline(args, steps, warmup, tiles_per_step):
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    if args.config == "c5":
        from amyloid_yolo_paper_b200 import synth
        path = CpuPath(None, None)
        pred = torch.from_numpy(synth.synthetic_predictions(tiles_per_step, seed=5))
        fn = lambda: path.nms_step(pred, 0.01, 0.4)                                    # noqa: E731
        what = "%d decoded tiles [10647,7], non_max_suppression(0.01, 0.4) only" % tiles_per_step
        warmup = 0                                                                      # one tile is ~6 s
    else:
        path = CpuPath(args.net, args.conf_bias)
        x = cpu_inputs(tiles_per_step, uniform=args.config == "c2")
        orig = (416, 416) if args.config == "c2" else (args.tile, args.tile)
        fn = lambda: path.step(x, 0.8, 0.4, orig)                                       # noqa: E731
        what = "%d fp32 [3,416,416] tiles, %s forward + NMS(0.8,0.4) + rescale + mergeDetections(pixel sets)" % (
            tiles_per_step, args.net)
        for _ in range(warmup):
            path.step(x[:1], 0.8, 0.4, orig)
    t0 = time.perf_counter()
    dets = 0
    for _ in range(steps):
        dets += fn()
    dt = time.perf_counter() - t0
    return {"tiles_per_s": steps * tiles_per_step / dt, "seconds": dt, "cores": cores, "tiles": steps * tiles_per_step,
            "detections": dets, "kind": path.kind, "what": what, "torch": torch.__version__}


def reference_arm(args):
    """--impl reference: the reference's own CPU implementation of the path, all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    per_step = args.cpu_sample or CONFIGS[args.config]["ref_tiles"]
    r = run_cpu_baseline(args, steps, warmup, per_step)
    sample = "%d steps x %s, %d warm-up step(s) of 1 tile, torch %s CPU, %d threads" % (
        steps, r["what"], 0 if args.config == "c5" else warmup, r["torch"], r["cores"])
    cfg = workload_config(args)
    cfg["batch_per_gpu"] = per_step                      # what this arm actually ran per step (a bounded sample)
    cfg["workload"] += " [reference arm: CPU, %d tile(s) per step]" % per_step
    line = {"impl": "reference", "metric": METRIC, "value": r["tiles_per_s"], "unit": "tiles/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": 0 if args.config == "c5" else warmup, "ms_per_step": 1000.0 * r["seconds"] / steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg,
            "cpu_baseline": {"value": r["tiles_per_s"], "unit": "tiles/s", "cores": r["cores"], "kind": r["kind"],
                             "sample": sample},
            "e2e": {"value": r["tiles_per_s"], "unit": "tiles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(args):
    B = args.batch
    if args.config == "c5":
        return {"workload": "C5: NMS stress, %d tiles x 10647 decoded rows x 7, conf_thres 0.01 (every row a candidate), "
                            "nms_thres 0.4, ay_nms only" % B,
                "batch_per_gpu": B, "rows": 10647, "classes": 2,
                "l2": "per step %.0f MB of IoU bit masks + %.0f MB of predictions (L2 is 126 MB)" % (B * 7.1, B * 0.298),
                "parallelism": "single GPU"}
    src = "fp32 [B,3,416,416] inputs" if args.config == "c2" else \
        "synthetic %dx%d uint8 tiles nearest-resized to 416" % (args.tile, args.tile)
    name = {"c3": "C3/C4", "c1": "C1", "c2": "C2"}[args.config]
    return {"workload": "%s: %s, %s, batch %d per GPU, conf_thres 0.8, nms_thres 0.4, merge_boxes" % (name, args.net, src, B),
            "net": args.net, "batch_per_gpu": B, "tile": args.tile, "img_size": 416,
            "checkpoint": "synthetic seed 0 (.pth key layout), conf_bias %s per head, wh_gain 0.15" % args.conf_bias,
            "l2": ("inputs larger than L2 (%.0f MB of tiles, >%.0f MB of activations per step)" % (
                B * args.tile * args.tile * 3 / 1e6, B * 76.7)) if B >= 8 else
                  "L2 flushed between timed steps (256 MB scratch write)",
            "parallelism": "tiles sharded by rank through sweep.sweep, detections all-gathered once per sweep"
                           if args.gpus > 1 else "single GPU"}


def emit(args, line, cpu_tiles=None):
    if not args.no_cpu_baseline:
        n = args.cpu_sample or cpu_tiles or CONFIGS[args.config]["cpu_tiles"]
        r = run_cpu_baseline(args, 1, 1, n)
        line["cpu_baseline"] = {"value": r["tiles_per_s"], "unit": "tiles/s", "cores": r["cores"], "kind": r["kind"],
                                "sample": "%s, %.1f s, torch %s CPU" % (r["what"], r["seconds"], r["torch"])}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- GPU arm: c5 (NMS only)
def nms_arm(args):
    import numpy as np
    import torch
    from amyloid_yolo_paper_b200 import _lib, synth
    from amyloid_yolo_paper_b200.utils.utils import non_max_suppression_device
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    B, K, W = args.batch, max(1, args.steps), max(3, args.warmup)
    base = synth.synthetic_predictions(min(B, 16), seed=5)
    P = np.concatenate([base] * (-(-B // len(base))), 0)[:B]
    N, cols = P.shape[1], P.shape[2]
    pred0 = torch.from_numpy(P).to(dev)
    pred = torch.empty_like(pred0)
    host = torch.from_numpy(P).pin_memory()
    lib, ctx = _lib.load(), _lib.context(0)
    stream = torch.cuda.current_stream(dev)
    st = ctypes.c_void_p(stream.cuda_stream)
    cap = N
    rows = torch.empty((B, cap, 7), dtype=torch.float32, device=dev)
    keep = torch.empty((B, cap), dtype=torch.int32, device=dev)
    counts = torch.empty((B,), dtype=torch.int32, device=dev)

    def step():
        pred.copy_(pred0)                   # ay_nms rewrites columns 0..3 in place (utils.py:244): fresh input every step
        _lib.check(lib.ay_nms(ctx, ctypes.c_void_p(pred.data_ptr()), B, N, cols - 5, 0.01, 0.4, ctypes.c_void_p(rows.data_ptr()),
                              ctypes.c_void_p(keep.data_ptr()), ctypes.c_void_p(counts.data_ptr()), cap, st))

    for _ in range(W):
        step()
    _lib.check(lib.ay_check(ctx, st))
    clocks = ClockSampler(0)
    clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    clocks.begin()
    e0.record(stream)
    for _ in range(K):
        step()
    e1.record(stream)
    torch.cuda.synchronize(dev)
    clocks.end()
    _lib.check(lib.ay_check(ctx, st))
    ms = e0.elapsed_time(e1)
    kept = int(counts.sum().item())
    # per-kernel times (events between the four kernels of one call)
    ms4 = (ctypes.c_float * 4)()
    parts = [0.0] * 4
    for _ in range(K):
        pred.copy_(pred0)
        _lib.check(lib.ay_nms_timed(ctx, ctypes.c_void_p(pred.data_ptr()), B, N, cols - 5, 0.01, 0.4,
                                    ctypes.c_void_p(rows.data_ptr()), ctypes.c_void_p(keep.data_ptr()),
                                    ctypes.c_void_p(counts.data_ptr()), cap, st, ms4))
        parts = [a + b for a, b in zip(parts, ms4)]
    parts = [v / K for v in parts]
    # e2e: pinned host predictions in, kept rows + keep indices + counts out
    e2e_s = d2h = None
    if not args.no_e2e:
        rows_h = torch.empty((B, cap, 7), dtype=torch.float32).pin_memory()
        keep_h = torch.empty((B, cap), dtype=torch.int32).pin_memory()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for _ in range(K):
            pred.copy_(host, non_blocking=True)
            r, k, c = non_max_suppression_device(pred, 0.01, 0.4, cap=cap)
            c_h = c.cpu()
            kmax = int(c_h.max())
            rows_h[:, :kmax].copy_(r[:, :kmax], non_blocking=True)
            keep_h[:, :kmax].copy_(k[:, :kmax], non_blocking=True)
            torch.cuda.synchronize(dev)
            d2h = B * 4 + B * kmax * 32
        e2e_s = time.perf_counter() - t0
    clk = clocks.stop()
    peaks = measured_peaks()
    mask_bytes = (N * N // 16)                            # upper triangle of the n x n same-class IoU bit matrix
    algo = B * (N * 20 + mask_bytes)                      # sorted boxes + classes in (20 B per candidate), bit rows out
    achieved = algo / (parts[1] * 1e-3) / 1e9
    line = {"metric": METRIC, "value": B * K / (ms * 1e-3), "unit": "tiles/s", "n_gpus": 1, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(args),
            "gpu_launches": 4 * K,
            "clocks": {"sm_mhz": clk["sm_mhz"], "sm_max_mhz": clk["sm_max_mhz"], "reasons": clk["reasons"], "samples": clk["samples"],
                       "scope": clk.get("scope")},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm"], "unit": "GB/s", "frac": achieved / peaks["hbm"],
                         "traffic": None, "kernel": "nms_mask_kernel (IoU > thres bit matrix, 64x64 blocks)",
                         "ms": parts[1], "algorithmic_bytes_per_launch": algo, "peak_source": peaks["source"] + " (hbm_gbs)",
                         "kernels_ms": {"prepare": parts[0], "mask": parts[1], "sweep": parts[2], "cluster": parts[3]}},
            "kept_per_tile": kept / B}
    if e2e_s is not None:
        line["e2e"] = {"value": B * K / e2e_s, "unit": "tiles/s", "h2d_bytes_per_step": int(host.numel() * 4),
                       "d2h_bytes_per_step": int(d2h), "input": "pinned fp32 [B,10647,7] decoded predictions",
                       "call": "non_max_suppression_device + read-back of counts, kept rows and keep indices"}
    emit(args, line)


# ----------------------------------------------------------------------------- GPU arm: c1 / c2 / c3
def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    if args.config == "c5":
        return nms_arm(args)

    import torch
    import torch.distributed as dist
    from amyloid_yolo_paper_b200 import _lib, sweep
    from amyloid_yolo_paper_b200.pipeline import TileDetector

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise _lib.EngineUnavailable("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    B, K, W = args.batch, max(1, args.steps), max(3, args.warmup)
    S = 416
    orig = (S, S) if args.config == "c2" else (args.tile, args.tile)
    cfg, model, sd = build_model(args.net, args.conf_bias)
    model.precision = args.precision
    cap = 128
    det = TileDetector(model, conf_thres=0.8, nms_thres=0.4, merge_boxes=True, max_batch=B, cap=cap, device=dev)
    float_in = args.config == "c2"
    if float_in:
        g = torch.Generator(device=dev).manual_seed(1000 + rank)
        tiles = torch.rand((B, 3, S, S), device=dev, generator=g)
        in_kind, sh, sw = _lib.AY_IN_F32_NCHW, S, S
    else:
        tiles = device_tiles(B, args.tile, dev, seed=0xA11CE, first_tile=rank * B)
        in_kind, sh, sw = _lib.AY_IN_U8_HWC, args.tile, args.tile
    torch.cuda.synchronize(dev)
    stream = torch.cuda.current_stream(dev)
    flush = torch.empty((256 << 20,), dtype=torch.uint8, device=dev) if B < 8 else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- value: device-resident inputs through sweep.sweep (the shipped C4 path: shard by rank, run, gather once)
    def resident(lo, hi):                     # this rank's tiles of a pass: the same B resident tiles every pass
        return tiles[:hi - lo]

    def run_sweep(steps):
        if B >= 8:
            rows, counts = sweep.sweep(det, resident, world * B * steps, B)
            return rows, counts
        for _ in range(steps):                # B < 8 (c1): one tile per step, L2 flushed between steps
            flush.zero_()
            r, c = det.run_device(tiles, orig)
        return r, c

    warm_rows, _ = run_sweep(W)
    det.check()
    if world > 1 and B >= 8 and not os.environ.get("AY_NO_WARM_GATHER"):
        # the K-step sweep's exchange is larger than the W-step one: set up its NCCL protocol and buffers outside the
        # timed region (sweep.warm_gather: zeros of the same shape; no detection work is moved out of the timing)
        sweep.warm_gather(world * B * K, warm_rows.shape[1], dev)
    launches_per_step = det.launches_per_batch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clocks = ClockSampler(local)
    clocks.start()               # (before the barrier: spawning the sampler takes tens of ms, different on every rank -
    barrier()                    #  ranks that entered the timed region early would wait for the late one in the gather)
    clocks.begin()
    if B >= 8:
        e0.record(stream)
        rows, counts = run_sweep(K)
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
    else:
        # single-tile latency: time every step on its own so that the L2 flush stays outside the sum
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        for a, b in evs:
            flush.zero_()
            a.record(stream)
            rows, counts = det.run_device(tiles, orig)
            b.record(stream)
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
    clocks.end()
    det.check()
    n_det = int(counts.clamp(min=0).sum().item()) // (world * K if B >= 8 else 1)

    # ---- conv stack alone (roofline numerator): events around ay_forward, same inputs, same stream
    fwd_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    clocks.begin()
    for a, b in fwd_ev:
        if flush is not None:
            flush.zero_()
        a.record(stream)
        _lib.check(det.lib.ay_forward(det.ctx, ctypes.c_void_p(tiles.data_ptr()), in_kind, B, sh, sw,
                                      ctypes.c_void_p(det.decoded.data_ptr()), ctypes.c_void_p(stream.cuda_stream)))
        b.record(stream)
    torch.cuda.synchronize(dev)
    clocks.end()
    fwd_mean = sum(a.elapsed_time(b) for a, b in fwd_ev) / K
    fwd_launches = det.engine.launch_count()

    # ---- gather check (N > 1): what arrives on rank 0 for rank r's tiles equals a single-GPU run of the same tile ids
    gather_verified = None
    if world > 1:
        def by_id(lo, hi):
            return device_tiles(hi - lo, args.tile, dev, seed=0xA11CE, first_tile=lo)
        all_rows, all_counts = sweep.sweep(det, by_id, world * B, B)
        ok = True
        if rank == 0:
            kk = all_rows.shape[1]
            col = torch.arange(kk, device=dev).view(1, -1, 1)
            zero = torch.zeros((), dtype=torch.int32, device=dev)
            for r in range(world):
                mine, mc = det.run_device(by_id(r * B, (r + 1) * B))
                got_r, got_c = all_rows[r * B:(r + 1) * B], all_counts[r * B:(r + 1) * B]
                ok = ok and torch.equal(got_c, mc) and int(mc.max()) <= kk
                valid = col < mc.view(-1, 1, 1)
                ok = ok and torch.equal(torch.where(valid, got_r.view(torch.int32), zero),
                                        torch.where(valid, mine[:, :kk].contiguous().view(torch.int32), zero))
                ok = ok and not bool(torch.where(valid, zero, got_r.view(torch.int32)).any())     # padding rows are zero
        flag = torch.tensor([1 if ok else 0], device=dev)
        dist.broadcast(flag, 0)
        gather_verified = bool(flag.item())
        barrier()

    # ---- e2e: host buffers through the reference-facing calls, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        if float_in:
            x_host = tiles.cpu().pin_memory()
            u8_host = None
        else:
            # what the loader hands over: the 416^2 source pixels the nearest resize samples (utils/transforms.py:104-111),
            # gathered on the host into pinned uint8 [B,416,416,3] (0.5 MB per tile); and, for comparison, the fp32
            # [B,3,416,416] tensor detect.py:94 hands to Darknet.forward (2 MB per tile)
            idx = torch.clamp(torch.floor(torch.arange(S, device=dev, dtype=torch.float32) * (float(args.tile) / S)).long(),
                              max=args.tile - 1)
            small = tiles[:, idx][:, :, idx].contiguous()
            u8_host = small.cpu().pin_memory()
            x_host = (small.permute(0, 3, 1, 2).float() / 255.0).cpu().pin_memory()

        def stream_rate(batch_host):
            for _ in range(2):
                det.detect(batch_host, orig)
            barrier()
            clocks.begin()
            t0 = time.perf_counter()
            n_out = 0
            per_tile = []
            for outs in det.detect_stream((batch_host for _ in range(K)), orig):
                n_out += len(outs)
                per_tile.extend(outs)
            if world > 1:                     # C4: the sweep's single exchange, on the detections of all K batches
                r_, c_ = sweep.pack_detections(per_tile, device=dev)
                sweep.allgather_detections(r_, c_, world * B * K)
            barrier()
            dt = time.perf_counter() - t0
            clocks.end()
            assert n_out == K * B
            return dt, det.d2h_bytes // K

        def sync_rate(batch_host):
            barrier()
            t0 = time.perf_counter()
            for _ in range(K):
                det.detect(batch_host, orig)
            barrier()
            return time.perf_counter() - t0

        main_host = x_host if float_in else u8_host
        e2e_s, d2h = stream_rate(main_host)
        e2e = {"unit": "tiles/s", "h2d_bytes_per_step": int(main_host.numel() * main_host.element_size()),
               "d2h_bytes_per_step": int(d2h),
               "input": "pinned fp32 [B,3,416,416]" if float_in else
                        "pinned uint8 [B,416,416,3]: the source pixels the nearest resize samples, gathered by the loader",
               "call": "TileDetector.detect_stream over the K host batches (one H2D / compute / D2H pipeline across steps)"}
        times = [e2e_s, sync_rate(main_host)]
        if not float_in:
            f32_s, _ = stream_rate(x_host)
            times.append(f32_s)
        if args.config == "c1":
            # the import-swap API of INTEGRATION.md §1, as detect.py:94-133 calls it: model(x) returns the decoded tensor on
            # the CPU, non_max_suppression / rescale_boxes / mergeDetections take and return CPU tensors
            from amyloid_yolo_paper_b200.core import mergeDetections
            from amyloid_yolo_paper_b200.utils.utils import non_max_suppression, rescale_boxes
            xs = x_host[:1]

            def swap_call():
                dec = model(xs)
                dets = non_max_suppression(dec, 0.8, 0.4)
                for d in dets:
                    if d is not None:
                        mergeDetections(rescale_boxes(d, S, orig))
            for _ in range(3):
                swap_call()
            barrier()
            t0 = time.perf_counter()
            for _ in range(K):
                swap_call()
            barrier()
            times.append(time.perf_counter() - t0)
        if world > 1:
            t = torch.tensor(times, dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            times = [float(v) for v in t.tolist()]
        e2e["value"] = world * B * K / times[0]
        e2e["per_call_sync"] = world * B * K / times[1]
        if not float_in:
            e2e["fp32_nchw_input"] = {"value": world * B * K / times[2], "h2d_bytes_per_step": int(x_host.numel() * 4)}
        if args.config == "c1":
            e2e["import_swap_api"] = {"value": K / times[-1], "ms_per_tile": 1000.0 * times[-1] / K,
                                      "call": "Darknet.forward (CPU tensor in, decoded CPU tensor out) + non_max_suppression + "
                                              "rescale_boxes + mergeDetections"}
    clk = clocks.stop()          # samples taken inside the timed regions (value, conv stack, e2e streams)

    per_rank = None
    if world > 1:
        t = torch.tensor([ms, fwd_mean], dtype=torch.float64, device=dev)
        every = torch.empty((world, 2), dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(every, t)
        per_rank = {"ms_per_step": [round(float(v) / K, 3) for v in every[:, 0].tolist()],
                    "conv_stack_ms": [round(float(v), 3) for v in every[:, 1].tolist()]}
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, fwd_mean = (float(v) for v in t.tolist())

    if rank == 0:
        peaks = measured_peaks()
        flop = FLOP_PER_TILE[args.net]
        achieved = B * flop / (fwd_mean * 1e-3) / 1e12
        traffic, traffic_src = profiled_traffic(B) if args.config == "c3" else (None, None)
        line = {
            "metric": METRIC, "value": world * B * K / (ms * 1e-3), "unit": "tiles/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if args.precision == "bf16" else ("f32-ffma" if os.environ.get("AY_F32_FFMA") else "f32x3"),
            "data": "synthetic", "config": workload_config(args),
            "gpu_launches": int(launches_per_step * K),
            "clocks": {"sm_mhz": clk["sm_mhz"], "sm_max_mhz": clk["sm_max_mhz"], "reasons": clk["reasons"],
                       "samples": clk["samples"], "scope": clk.get("scope")},
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peaks["bf16"], "unit": "TFLOP/s",
                         "frac": achieved / peaks["bf16"], "traffic": traffic, "traffic_unit": "DRAM bytes per ay_forward",
                         "traffic_source": traffic_src,
                         "kernel": "conv stack (ay_forward: %d launches, tcgen05 implicit GEMM)" % fwd_launches,
                         "ms": fwd_mean, "flop_per_tile": flop, "peak_source": peaks["source"] + " (sustained bf16)",
                         "frac_of_burst": achieved / peaks["bf16_burst"]},
            "detections_per_step": n_det,
        }
        if args.precision == "fp32":
            # 3-pass tf32 at half the bf16 rate: the tensor ceiling of this mode is a sixth of the bf16 peak
            line["roofline"]["peak"] = peaks["bf16"] / 6.0
            line["roofline"]["frac"] = achieved / (peaks["bf16"] / 6.0)
            line["roofline"]["peak_source"] += " / 6 (tf32 is half rate, three passes)"
            line["roofline"]["kernel"] = "conv stack of the fp32 accuracy mode (ay_forward: %d launches)" % fwd_launches
            line["config"]["workload"] += " [precision fp32: accuracy mode]"
        if e2e is not None:
            line["e2e"] = e2e
        if gather_verified is not None:
            line["gather_verified"] = gather_verified
        if per_rank is not None:
            line["per_rank"] = per_rank          # max over ranks is what `value` and the roofline use
        emit(args, line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
