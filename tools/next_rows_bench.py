#!/usr/bin/env python
"""Throughput of the SURVEY 8(f) rows built on the engine (not BASELINE configs; recorded under profiles/):
  N2  consensus CAA filter: crops/s of caa.Net (bf16 and fp32 mode, batch 64, crops cut and normalised by ay_crop_normalize),
      the whole filterDetectionsByCAAModel call for 32 boxes of a tile, and the reference's own class (core.Net, torch CPU,
      all host threads) on a bounded sample when the reference modules are available.
  N4  crop.detect_slide over a synthetic slide resident in HBM against TileDetector.detect on explicitly cut tiles.
python tools/next_rows_bench.py [--out FILE]"""
import argparse
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(fn, reps, dev):
    fn()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize(dev)
    return (time.perf_counter() - t0) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    from amyloid_yolo_paper_b200 import caa, crop, synth
    from amyloid_yolo_paper_b200.config import ensure_cfg
    from amyloid_yolo_paper_b200.models import Darknet
    from amyloid_yolo_paper_b200.pipeline import TileDetector
    from oracle.make_golden_caa import NORM, caa_tile, seeded_caa_state          # seeded weights / synthetic tile only
    dev = torch.device("cuda", 0)
    lines = ["# %s, torch %s" % (torch.cuda.get_device_name(0), torch.__version__)]
    tile = torch.from_numpy(caa_tile(11)).to(dev)
    state = seeded_caa_state(5)
    rng = np.random.default_rng(0)
    wins = [caa.crop_window((int(rng.integers(0, 1500)), int(rng.integers(0, 1500)), 30, 30)) for _ in range(64)]
    for prec in ("bf16", "fp32"):
        net = caa.Net(precision=prec)
        net.load_state_dict(state)
        eng = net.engine(dev, max_batch=64)

        def step():
            x = caa.normalised_crops(tile, wins, NORM["mean"], NORM["std"])
            eng.forward_nchw(x)
        dt = timed(step, 20, dev)
        lines.append("N2 caa.Net %-4s: %8.0f crops/s (64 crops per batch: crop + normalise + 6 conv stages + classifier, %.2f ms)"
                     % (prec, 64 / dt, 1e3 * dt))
    net = caa.Net()
    net.load_state_dict(state)
    dets = torch.tensor([[x, y, x + 40.0, y + 30.0, 0.9, 0.8, float(i & 1)] for i, (x, y) in
                         enumerate(rng.integers(0, 1400, (32, 2)).tolist())])
    dt = timed(lambda: caa.filterDetectionsByCAAModel(tile, dets, ["Cored", "CAA"], model=net, norm=NORM, device=dev), 20, dev)
    lines.append("N2 filterDetectionsByCAAModel, 32 boxes of one tile (fp32 mode, tile already on the GPU): %.2f ms per tile" % (1e3 * dt))
    try:
        from oracle import ref_import
        _, _, core = ref_import.load()
        ref = core.Net()
        ref.load_state_dict(state)
        ref.eval()
        x = caa.normalised_crops(tile, wins[:16], NORM["mean"], NORM["std"]).cpu()
        with torch.no_grad():
            ref(x[:1])
            t0 = time.perf_counter()
            for i in range(16):
                ref(x[i:i + 1])                        # one model(...) call per box, as core.py:443-444
            dt = (time.perf_counter() - t0) / 16
        lines.append("N2 reference core.Net on the host cores (%d threads, one call per crop as core.py:443): %.1f crops/s"
                     % (torch.get_num_threads(), 1 / dt))
    except Exception as e:                               # noqa: BLE001
        lines.append("N2 reference core.Net: unavailable here (%s)" % str(e).splitlines()[0])
    # N4
    m = Darknet(ensure_cfg("yolov3-custom"))
    synth.synthetic_state_dict(m, seed=1, conf_bias=-2.5, wh_gain=0.15)
    det = TileDetector(m, conf_thres=0.8, nms_thres=0.4, merge_boxes=True, max_batch=64, cap=256, device=dev)
    tiles = synth.synthetic_tiles(64, size=1536, seed=3)
    slide = torch.from_numpy(np.ascontiguousarray(tiles.reshape(8, 8, 1536, 1536, 3).transpose(0, 2, 1, 3, 4)
                                                  .reshape(8 * 1536, 8 * 1536, 3))).to(dev)
    tiles_d = torch.from_numpy(tiles).to(dev)
    dt_s = timed(lambda: list(crop.detect_slide(det, slide)), 5, dev)
    dt_t = timed(lambda: det.detect(tiles_d), 5, dev)
    lines.append("N4 crop.detect_slide, 12288^2 slide in HBM (64 tiles, one gather + one engine batch): %.0f tiles/s (%.2f ms)"
                 % (64 / dt_s, 1e3 * dt_s))
    lines.append("N4 TileDetector.detect on the same 64 tiles cut beforehand ([64,1536,1536,3] in HBM):    %.0f tiles/s (%.2f ms)"
                 % (64 / dt_t, 1e3 * dt_t))
    text = "\n".join(lines)
    print(text)
    if args.out:
        open(args.out, "w").write(text + "\n")


if __name__ == "__main__":
    main()
