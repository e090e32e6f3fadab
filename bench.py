#!/usr/bin/env python
"""Benchmark of the tile-detection hot path (BASELINE.json metric: tiles/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B]

Workload (config C3/C4 of BASELINE.json): yolov3-custom (Darknet-53, 2 classes), synthetic
1536x1536 uint8 WSI tiles, nearest-resized to 416 inside the first convolution, bf16 tensor-core
forward + fused YOLO decode + cluster-merging NMS (conf 0.8, nms 0.4) + rescale_boxes +
mergeDetections.  One step = one batch of B tiles per GPU through that whole chain.

* ``value``  - tiles/s with the uint8 tiles already resident in HBM (device-timed, CUDA events,
               max over ranks).  Per-GPU work is fixed as N grows ("weak"); with N > 1 each step
               ends with the NCCL all-gather of the step's detection lists (config C4).
* ``e2e``    - the same metric through the reference-facing API with HOST buffers
               (``TileDetector.detect`` on a pinned fp32 [B,3,416,416] batch = the tensor
               ``Darknet.forward`` receives in detect.py:94): H2D + chain + D2H of the detections
               inside the timed region.
* ``roofline`` - conv stack (``ay_forward``: 74 tcgen05 implicit-GEMM launches + first-layer and
               gather kernels) FLOPs / its CUDA-event time vs the measured bf16 peak.
* ``cpu_baseline`` - the oracle port of the reference path (torch CPU forward + python NMS loop +
               pixel-set mergeDetections) on a bounded sample, all host threads.

``--impl reference`` times that CPU path as the reference arm (the reference is pure Python and
is not present on the GPU box).
"""
import argparse
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


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", type=str, default="ours", choices=["ours", "reference"])
    p.add_argument("--batch", type=int, default=256, help="tiles per GPU per step")
    p.add_argument("--net", type=str, default="yolov3-custom", choices=sorted(FLOP_PER_TILE))
    p.add_argument("--tile", type=int, default=1536, help="source tile edge (uint8 HWC)")
    p.add_argument("--conf_bias", type=str, default="-6,-6,0.62",
                   help="objectness bias of the synthetic checkpoint per head (grids 13,26,52): a few dozen candidates per "
                        "tile at conf 0.8, mostly small boxes as in the reference's plaque data")
    p.add_argument("--cpu-sample", type=int, default=8, help="tiles in the bounded CPU-baseline sample (about 2 s each)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--layers", action="store_true", help="also print a per-layer timing table to stderr")
    return p.parse_args()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return {"bf16": float(d.get("bf16_tflops_sustained") or d["bf16_tflops"]), "bf16_burst": float(d["bf16_tflops"]),
                "hbm": float(d["hbm_gbs"]), "source": "MEASURED_PEAKS.json (sustained bf16)"}
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
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(prefix="clocks_", suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
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
        sm, names = [], ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        seen = set()
        for line in open(self.path):
            parts = [v.strip() for v in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                out["sm_max_mhz"] = float(parts[1])
            except ValueError:
                continue
            for name, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    seen.add(name)
        os.unlink(self.path)
        if sm:
            sm.sort()
            out["sm_mhz"] = sm[len(sm) // 2]
        out["samples"] = len(sm)
        out["reasons"] = sorted(seen)
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


# ----------------------------------------------------------------------------- CPU path (oracle port)
def cpu_path_step(blocks, sd, x, conf_thres, nms_thres, orig):
    """The body of detect.py's loop as the reference runs it on the host (oracle port):
    forward (models.py:237-255) -> non_max_suppression (utils.py:235-273) -> rescale_boxes
    (utils.py:36-50) -> mergeDetections with pixel-set enumeration (core.py:326-423)."""
    import torch
    from oracle import restate as R
    with torch.no_grad():
        dec = R.darknet_forward(blocks, sd, x)
    dets = R.nms_faithful(dec, conf_thres, nms_thres)
    n = 0
    for d in dets:
        if d is None:
            continue
        rows = R.rescale_boxes(d.numpy().copy(), x.shape[2], orig)
        n += len(R.merge_faithful(rows))
    return n


def cpu_inputs(net, conf_bias, n_tiles, tile):
    import torch
    from amyloid_yolo_paper_b200 import synth
    from oracle import restate as R
    cfg, model, sd = build_model(net, conf_bias)
    blocks = R.parse_cfg(cfg)
    size = int(blocks[0]["height"])
    g = torch.Generator().manual_seed(1234)
    low = torch.rand((n_tiles, 3, 24, 24), generator=g)
    x = torch.nn.functional.interpolate(low, size=(size, size), mode="bilinear")
    x = ((90.0 + 130.0 * x + 12.0 * torch.randn(x.shape, generator=g)).clamp(0, 255).floor() / 255.0).contiguous()
    return blocks, sd, x


def run_cpu_baseline(args, steps, warmup):
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    blocks, sd, x = cpu_inputs(args.net, args.conf_bias, args.cpu_sample, args.tile)
    orig = (args.tile, args.tile)
    for _ in range(warmup):
        cpu_path_step(blocks, sd, x[:1], 0.8, 0.4, orig)
    t0 = time.perf_counter()
    dets = 0
    for _ in range(steps):
        dets += cpu_path_step(blocks, sd, x, 0.8, 0.4, orig)
    dt = time.perf_counter() - t0
    return {"tiles_per_s": steps * x.shape[0] / dt, "seconds": dt, "cores": cores, "tiles": steps * x.shape[0],
            "detections": dets}


def reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port; the reference
    itself is pure Python on torch and is not on the GPU box), all host threads, bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    steps, warmup = max(1, args.steps), max(0, min(args.warmup, 3))
    args.cpu_sample = min(args.cpu_sample, 4)        # about 8 s of host work per step
    r = run_cpu_baseline(args, steps, min(warmup, 1))
    sample = "%d steps x %d fp32 [3,416,416] tiles, forward+NMS(0.8,0.4)+rescale+mergeDetections(pixel sets), torch %s" % (
        steps, args.cpu_sample, torch.__version__)
    line = {"impl": "reference", "metric": METRIC, "value": r["tiles_per_s"], "unit": "tiles/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": 1000.0 * r["seconds"] / steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, cpu=True),
            "cpu_baseline": {"value": r["tiles_per_s"], "unit": "tiles/s", "cores": r["cores"], "kind": "port",
                             "sample": sample},
            "e2e": {"value": r["tiles_per_s"], "unit": "tiles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(args, cpu=False):
    return {"workload": "C3/C4: %s, synthetic %dx%d uint8 tiles nearest-resized to 416, batch %d per GPU, "
                        "conf_thres 0.8, nms_thres 0.4, merge_boxes" % (args.net, args.tile, args.tile, args.batch),
            "net": args.net, "batch_per_gpu": args.batch, "tile": args.tile, "img_size": 416,
            "checkpoint": "synthetic seed 0 (.pth key layout), conf_bias %s per head, wh_gain 0.15" % args.conf_bias,
            "l2": "inputs larger than L2 (%.0f MB of tiles, >%.0f MB of activations per step)" % (
                args.batch * args.tile * args.tile * 3 / 1e6, args.batch * 76.7),
            "parallelism": "tiles sharded by rank, detections all-gathered" if args.gpus > 1 else "single GPU"}


# ----------------------------------------------------------------------------- GPU arm
def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import torch.distributed as dist
    from amyloid_yolo_paper_b200 import _lib
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
    cfg, model, sd = build_model(args.net, args.conf_bias)
    cap = 128
    det = TileDetector(model, conf_thres=0.8, nms_thres=0.4, merge_boxes=True, max_batch=B, cap=cap, device=dev)
    tiles = device_tiles(B, args.tile, dev, seed=0xA11CE, first_tile=rank * B)
    torch.cuda.synchronize(dev)
    stream = torch.cuda.current_stream(dev)

    gather_rows = gather_counts = None
    if world > 1:
        gather_rows = torch.empty((world * B, cap, 7), dtype=torch.float32, device=dev)
        gather_counts = torch.empty((world * B,), dtype=torch.int32, device=dev)

    def step():
        rows, counts = det.run_device(tiles, (args.tile, args.tile))
        if world > 1:      # config C4: only the per-tile detection lists cross NVLink
            dist.all_gather_into_tensor(gather_counts, counts)
            dist.all_gather_into_tensor(gather_rows, rows)
        return rows, counts

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(W):
        step()
    det.check()
    launches_per_step = det.launches_per_batch

    # ---- value: device-resident inputs, CUDA events on the launching stream
    fwd_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    e0.record(stream)
    for k in range(K):
        rows, counts = step()
    e1.record(stream)
    barrier()
    det.check()
    ms = e0.elapsed_time(e1)
    n_det = int(counts.clamp(min=0).sum().item())

    # ---- conv stack alone (roofline numerator): events around ay_forward, same inputs, same stream
    import ctypes
    fwd_ms = []
    for k in range(K):
        a, b = fwd_ev[k]
        a.record(stream)
        _lib.check(det.lib.ay_forward(det.ctx, ctypes.c_void_p(tiles.data_ptr()), _lib.AY_IN_U8_HWC, B, args.tile,
                                      args.tile, ctypes.c_void_p(det.decoded.data_ptr()),
                                      ctypes.c_void_p(stream.cuda_stream)))
        b.record(stream)
    torch.cuda.synchronize(dev)
    fwd_ms = [a.elapsed_time(b) for a, b in fwd_ev]
    fwd_mean = sum(fwd_ms) / len(fwd_ms)
    fwd_launches = det.engine.launch_count()

    # ---- e2e: host buffers through the reference-facing call, copies inside the timed region
    # the same tiles as the device-resident arm, preprocessed as detect.py's loader does (ToTensor + nearest Resize)
    x_host = torch.empty((B, 3, 416, 416), dtype=torch.float32).pin_memory()
    idx = torch.clamp(torch.floor(torch.arange(416, device=dev, dtype=torch.float32) * (float(args.tile) / 416.0)).long(),
                      max=args.tile - 1)
    x_host.copy_((tiles[:, idx][:, :, idx].permute(0, 3, 1, 2).float() / 255.0))
    for _ in range(2):
        det.detect(x_host, (args.tile, args.tile))
    # (a) one synchronous detect() per step, as a single-batch caller sees it
    barrier()
    t0 = time.perf_counter()
    for k in range(K):
        det.detect(x_host, (args.tile, args.tile))
    barrier()
    e2e_sync_s = time.perf_counter() - t0
    # (b) the streaming call a batch loop makes (detect.py:88 iterates a DataLoader): K host batches through
    # detect_stream; every step's input is copied H2D and its detections D2H inside the timed region, the copies of
    # step k+1 / k-1 overlap the compute of step k
    barrier()
    t0 = time.perf_counter()
    d2h = 0
    n_out = 0
    for outs in det.detect_stream((x_host for _ in range(K)), (args.tile, args.tile)):
        n_out += len(outs)
    d2h = det.d2h_bytes                       # bytes actually copied device -> host (counts + padded detection rows)
    barrier()
    e2e_s = time.perf_counter() - t0
    assert n_out == K * B
    clk = clocks.stop()          # sampled over the three timed regions (value, conv stack, e2e): all of them are under load

    if world > 1:
        t = torch.tensor([ms, e2e_s, fwd_mean, e2e_sync_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_s, fwd_mean, e2e_sync_s = (float(v) for v in t.tolist())

    if rank == 0:
        peaks = measured_peaks()
        flop = FLOP_PER_TILE[args.net]
        achieved = B * flop / (fwd_mean * 1e-3) / 1e12
        traffic, traffic_src = profiled_traffic(B)
        line = {
            "metric": METRIC, "value": world * B * K / (ms * 1e-3), "unit": "tiles/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic", "config": workload_config(args),
            "e2e": {"value": world * B * K / e2e_s, "unit": "tiles/s",
                    "h2d_bytes_per_step": int(x_host.numel() * 4), "d2h_bytes_per_step": int(d2h // K),
                    "input": "pinned fp32 [B,3,416,416] (the tensor detect.py:94 hands to Darknet.forward)",
                    "call": "TileDetector.detect_stream over the K host batches (one pipeline across steps)",
                    "per_call_sync": world * B * K / e2e_sync_s},
            "gpu_launches": int(launches_per_step * K),
            "clocks": {"sm_mhz": clk["sm_mhz"], "sm_max_mhz": clk["sm_max_mhz"], "reasons": clk["reasons"],
                       "samples": clk["samples"]},
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peaks["bf16"], "unit": "TFLOP/s",
                         "frac": achieved / peaks["bf16"], "traffic": traffic, "traffic_unit": "DRAM bytes per ay_forward",
                         "traffic_source": traffic_src,
                         "kernel": "conv stack (ay_forward: %d launches, tcgen05 implicit GEMM)" % fwd_launches,
                         "ms": fwd_mean, "flop_per_tile": flop, "peak_source": peaks["source"],
                         "frac_of_burst": achieved / peaks["bf16_burst"]},
            "detections_last_step": n_det,
        }
        if not args.no_cpu_baseline:
            r = run_cpu_baseline(args, 1, 1)
            line["cpu_baseline"] = {"value": r["tiles_per_s"], "unit": "tiles/s", "cores": r["cores"], "kind": "port",
                                    "sample": "%d fp32 [3,416,416] tiles, forward+NMS+rescale+mergeDetections "
                                              "(oracle port of the reference path), %.1f s" % (r["tiles"], r["seconds"])}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
