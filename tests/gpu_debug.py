"""First-contact script for a GPU box: runs each stage, prints per-layer errors
against the CPU oracle, never raises before the end.  Not a pytest file.

    python tests/gpu_debug.py [mini|custom|tiny|nms|merge ...]
"""
import sys
import tempfile
import time
import traceback

import numpy as np
import torch

from helpers import ROOT, rel_max, write_mini_cfg  # noqa: F401  (sets sys.path)

from amyloid_yolo_paper_b200 import synth
from amyloid_yolo_paper_b200.config import ensure_cfg
from amyloid_yolo_paper_b200.models import Darknet
from oracle import restate as R


def layerwise(cfg, batch, src, seed=0, conf_bias=-3.0, precision="bf16"):
    m = Darknet(cfg)
    m.precision = precision
    sd = synth.synthetic_state_dict(m, seed=seed, conf_bias=conf_bias)
    blocks = R.parse_cfg(cfg)
    S = int(blocks[0]["height"])
    tiles = synth.synthetic_tiles(batch, size=src, seed=7)
    x = torch.stack([R.preprocess_u8(t, S) for t in tiles])
    with torch.no_grad():
        ref32 = R.darknet_forward(blocks, sd, x)
        emu, emu_layers = R.darknet_forward(blocks, sd, x, emulate_bf16=(precision == "bf16"), return_layers=True)
    eng = m.engine(device=torch.device("cuda", 0), max_batch=batch)
    results = {}
    for mode in ("ref", "tc"):
        eng.use_reference_conv(mode == "ref")
        t0 = time.time()
        out_f32 = eng.forward_nchw(x)
        eng.check()
        out_u8 = eng.forward_u8(torch.from_numpy(tiles))
        eng.check()
        dt = time.time() - t0
        results[mode] = out_u8.cpu()
        print("[%s] forward ok in %.3fs launches=%d  f32-vs-u8 input max diff %.3e" %
              (mode, dt, eng.launch_count(), float((out_f32 - out_u8).abs().max())))
        print("[%s] decoded vs emulation rel_max %.3e   vs fp32 oracle %.3e" %
              (mode, rel_max(results[mode], emu), rel_max(results[mode], ref32)))
        worst = []
        for i, d in enumerate(blocks[1:]):
            if d["type"] == "yolo":
                continue
            try:
                got = eng.layer_output(i, batch).cpu()
            except Exception:
                continue
            err = rel_max(got, emu_layers[i])
            worst.append((err, i, d["type"]))
            if err > 2e-2 or not np.isfinite(err):
                print("   layer %3d %-13s shape %s rel_max %.3e  <-- suspicious" % (i, d["type"], tuple(got.shape), err))
        worst.sort(reverse=True)
        print("[%s] worst layers:" % mode, [(round(e, 5), i, t) for e, i, t in worst[:5]])
    print("tc vs ref decoded rel_max %.3e" % rel_max(results["tc"], results["ref"]))
    eng.use_reference_conv(False)
    return m, eng


def run_nms():
    from amyloid_yolo_paper_b200.utils.utils import non_max_suppression
    P = synth.synthetic_predictions(3, clusters=(200, 400))
    for thr in (0.8, 0.3, 0.01):
        pred = torch.from_numpy(P.copy())
        t0 = time.time()
        out, keep = non_max_suppression(pred, thr, 0.4, return_keep=True)
        dt = time.time() - t0
        ok_all = True
        for b in range(P.shape[0]):
            rows, kidx = R.nms_exact(P[b], thr, 0.4)
            got = out[b].numpy()
            ok = got.shape == rows.shape and np.array_equal(got.view(np.int32), rows.view(np.int32)) and \
                np.array_equal(keep[b].numpy(), kidx)
            if not ok:
                ok_all = False
                n = min(len(got), len(rows))
                bad = np.nonzero((got[:n].view(np.int32) != rows[:n].view(np.int32)).any(1))[0]
                print("   nms thr %.2f tile %d MISMATCH shapes %s %s first bad rows %s" % (thr, b, got.shape, rows.shape, bad[:5]))
                if len(bad):
                    print("      got", got[bad[0]], "want", rows[bad[0]])
        xy_ok = np.array_equal(pred[..., :4].numpy().view(np.int32), R.xywh2xyxy(P[..., :4]).view(np.int32))
        print("nms thr %.2f: bit-exact=%s in-place xyxy=%s (%.3fs, %d kept in tile 0)" % (thr, ok_all, xy_ok, dt, len(out[0])))


def run_merge():
    from amyloid_yolo_paper_b200.core import mergeDetections
    bad = 0
    merges = 0
    for s in range(200):
        rng = np.random.default_rng(s)
        rows = synth.synthetic_detections(int(rng.integers(1, 40)), seed=s, overlap=0.6)
        if s % 5 == 0:
            rows[:, 6] = rng.integers(0, 3, len(rows))
        want = R.merge_exact(rows)
        got = mergeDetections(torch.from_numpy(rows.copy())).numpy()
        merges += len(rows) - len(want)
        if got.shape != want.shape or not np.array_equal(got.view(np.int32), want.view(np.int32)):
            bad += 1
            if bad < 4:
                print("   merge seed %d mismatch: got %s want %s" % (s, got.shape, want.shape))
    print("merge: %d/200 scenes mismatch (%d merges total)" % (bad, merges))


def main(argv):
    what = argv or ["mini", "nms", "merge", "custom"]
    print("device:", torch.cuda.get_device_name(0))
    for w in what:
        print("=" * 20, w)
        try:
            if w == "mini":
                with tempfile.TemporaryDirectory() as d:
                    layerwise(write_mini_cfg(d), 3, 96)
            elif w == "custom":
                layerwise(ensure_cfg("yolov3-custom"), 2, 512)
            elif w == "custom16":
                layerwise(ensure_cfg("yolov3-custom"), 2, 512, precision="fp16")
            elif w == "tiny":
                layerwise(ensure_cfg("yolov3-tiny"), 2, 512)
            elif w == "nms":
                run_nms()
            elif w == "merge":
                run_merge()
        except Exception:
            traceback.print_exc()
    print("done")


if __name__ == "__main__":
    main(sys.argv[1:])
