"""TEST INFRASTRUCTURE ONLY — writes tests/golden/* from the UNMODIFIED reference.

Run in the build container (the reference is mounted at /root/reference):

    python -m oracle.make_golden

Every fixture is an input/output pair of a reference function executed here
(CPython 3.12.3, torch 2.11.0 CPU), so that the GPU box — which has no copy of the
reference — can still check the oracle (``-m "not gpu"``) and the CUDA path
(``-m gpu``) against the real thing:

  nms_ref.npz          utils.utils.non_max_suppression    (utils/utils.py:235-273)
  merge_ref.npz        core.mergeDetections               (core.py:366-423)
  rescale_ref.npz      utils.utils.rescale_boxes          (utils/utils.py:36-50)
  iou_ref.npz          utils.utils.bbox_iou + core.IOU    (utils/utils.py:202-232, unit_test.py:137-150)
  forward_mini_ref.npz models.Darknet.forward             (models.py:237-255) on tests/helpers.MINI_CFG
  forward_heads_ref.npz  the three YOLOLayer decodes of yolov3-custom on planted raw head tensors (models.py:127-172)
  preprocess_ref.npz   ToTensor + Resize arithmetic       (utils/transforms.py:89-111)
  cfg_blocks.json      utils.parse_config.parse_model_config of the reference's cfg files
  state_checksums.json float64 sums of the seeded synthetic checkpoints (guards RNG drift)
"""
import json
import os
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from amyloid_yolo_paper_b200 import synth  # noqa: E402
from oracle import ref_import  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")

NMS_CASES = [  # (seed, batch, classes, conf_thres, nms_thres)
    (11, 3, 2, 0.30, 0.4), (12, 2, 2, 0.01, 0.4), (13, 2, 80, 0.25, 0.5), (14, 2, 2, 0.80, 0.4), (15, 1, 2, 0.999, 0.4),
]


def small_predictions(seed, batch, classes):
    """240-row tiles (grids 4 and 8 on a 128 px image) with planted clusters."""
    return synth.synthetic_predictions(batch, num_classes=classes, seed=seed, img=128, grids=(4, 8), clusters=(8, 16),
                                       anchors={4: ((40, 30), (50, 60), (90, 80)), 8: ((10, 20), (20, 15), (20, 40))})


def merge_scene(seed):
    rng = np.random.default_rng(1000 + seed)
    rows = synth.synthetic_detections(int(rng.integers(1, 14)), seed=seed, extent=160.0, overlap=0.65, wh=(4, 36),
                                      jitter=8)
    if seed % 5 == 0:
        rows[:, 6] = rng.integers(0, 3, len(rows))       # class 2 never merges
    if seed % 7 == 0 and len(rows) > 1:
        rows[1] = rows[0]                                 # exact duplicate row collapses in the set
    return rows


def state_checksum(sd):
    return float(sum(v.double().sum() for v in sd.values()))


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    models, uu, core = ref_import.load()
    torch.set_num_threads(1)

    # ---------------- NMS
    out = {}
    for ci, (seed, batch, classes, ct, nt) in enumerate(NMS_CASES):
        P = small_predictions(seed, batch, classes)
        pred = torch.from_numpy(P.copy())
        res = uu.non_max_suppression(pred, conf_thres=ct, nms_thres=nt)
        out["in_%d" % ci] = P
        out["xyxy_%d" % ci] = pred[..., :4].numpy()          # the in-place mutation
        out["par_%d" % ci] = np.array([ct, nt, classes], dtype=np.float64)
        for b, r in enumerate(res):
            out["out_%d_%d" % (ci, b)] = np.zeros((0, 7), np.float32) if r is None else r.numpy()
    np.savez_compressed(os.path.join(GOLDEN, "nms_ref.npz"), **out)

    # ---------------- merge
    out = {}
    n_scenes = 60
    for s in range(n_scenes):
        rows = merge_scene(s)
        got = core.mergeDetections(torch.from_numpy(rows.copy()))
        out["in_%d" % s] = rows
        out["out_%d" % s] = got.numpy().astype(np.float32).reshape(-1, 7)
    # the SURVEY D6 chain scene (order dependence of chains >= 3)
    chain = np.array([[10, 10, 60, 60, .9, .8, 1], [50, 50, 100, 100, .95, .7, 1], [90, 90, 150, 150, .85, .9, 1],
                      [200, 200, 240, 240, .9, .9, 0], [220, 220, 260, 260, .8, .6, 0],
                      [300, 300, 340, 340, .9, .9, 2], [310, 310, 350, 350, .9, .9, 2]], dtype=np.float32)
    out["in_%d" % n_scenes] = chain
    out["out_%d" % n_scenes] = core.mergeDetections(torch.from_numpy(chain.copy())).numpy().astype(np.float32)
    out["count"] = np.array([n_scenes + 1])
    np.savez_compressed(os.path.join(GOLDEN, "merge_ref.npz"), **out)

    # ---------------- rescale
    out = {}
    rng = np.random.default_rng(3)
    for ci, shape in enumerate([(1536, 1536), (1000, 1536), (1536, 700), (416, 416), (333, 517)]):
        boxes = rng.uniform(-20, 436, (9, 7)).astype(np.float32)
        got = uu.rescale_boxes(torch.from_numpy(boxes.copy()), 416, shape)
        out["in_%d" % ci] = boxes
        out["shape_%d" % ci] = np.array(shape)
        out["out_%d" % ci] = got.numpy()
    np.savez_compressed(os.path.join(GOLDEN, "rescale_ref.npz"), **out)

    # ---------------- IoU (tensor function + the scalar known answers of unit_test.py:137-150)
    rng = np.random.default_rng(4)
    b = rng.uniform(0, 416, (64, 4)).astype(np.float32)
    b[:, 2:] = b[:, :2] + rng.uniform(1, 200, (64, 2)).astype(np.float32)
    b[5] = b[0]
    iou = uu.bbox_iou(torch.from_numpy(b[:1]), torch.from_numpy(b)).numpy()
    kat_boxes = np.array([[100, 100, 200, 200], [100, 100, 200, 200], [201, 201, 300, 300], [150, 150, 200, 200]],
                         dtype=np.float32)
    kat = np.array([core.IOU(list(kat_boxes[0]), list(kat_boxes[j])) for j in (1, 2, 3)], dtype=np.float64)
    np.savez_compressed(os.path.join(GOLDEN, "iou_ref.npz"), boxes=b, iou=iou, kat_boxes=kat_boxes, kat=kat)

    # ---------------- forward on the mini network (whole Darknet.forward incl. decode)
    from helpers import write_mini_cfg
    checks = {}
    with tempfile.TemporaryDirectory() as d:
        cfg = write_mini_cfg(d)
        ref = models.Darknet(cfg, img_size=64)
        from amyloid_yolo_paper_b200.models import Darknet as OurDarknet
        ours = OurDarknet(cfg, img_size=64)
        sd = synth.synthetic_state_dict(ours, seed=0, conf_bias=-1.0)
        ref.load_state_dict(sd)
        ref.eval()
        tiles = synth.synthetic_tiles(2, size=96, seed=7)
        from oracle import restate as R
        x = torch.stack([R.preprocess_u8(t, 64) for t in tiles])
        with torch.no_grad():
            y = ref(x)
        checks["mini_seed0"] = state_checksum(sd)
        np.savez_compressed(os.path.join(GOLDEN, "forward_mini_ref.npz"), tiles=tiles, x=x.numpy(), y=y.numpy())

    # ---------------- the three decode layers of yolov3-custom on planted raw head tensors
    cfg = os.path.join(ref_import.REFERENCE_ROOT, "config", "yolov3-custom.cfg")
    ref = models.Darknet(cfg)
    out = {}
    g = torch.Generator().manual_seed(9)
    for li, yl in enumerate(ref.yolo_layers):
        G = (13, 26, 52)[li]
        raw = torch.randn((1, 21, G, G), generator=g) * 2.0
        with torch.no_grad():
            dec, _ = yl(raw, None, 416)
        out["raw_%d" % li] = raw.numpy()
        out["dec_%d" % li] = dec.numpy()
    np.savez_compressed(os.path.join(GOLDEN, "forward_heads_ref.npz"), **out)

    # ---------------- preprocessing arithmetic (ToTensor: /255; Resize: F.interpolate nearest)
    import torch.nn.functional as F
    tile = synth.synthetic_tiles_fast(1, size=1536, seed=21)[0]
    t = torch.from_numpy(tile).permute(2, 0, 1).float().div(255)          # transforms.py:96 via torchvision to_tensor
    r = F.interpolate(t.unsqueeze(0), size=416, mode="nearest").squeeze(0)  # transforms.py:110
    idx = (F.interpolate(torch.arange(1536, dtype=torch.float32).view(1, 1, 1, -1).expand(1, 1, 2, -1), size=(2, 416),
                         mode="nearest")[0, 0, 0]).to(torch.int64).numpy()
    np.savez_compressed(os.path.join(GOLDEN, "preprocess_ref.npz"), corner=tile[:64, :64].copy(), idx_1536_416=idx,
                        sample_rows=r[:, ::52, :].numpy(), sample_src=tile[idx[::52]][:, idx].copy())

    # ---------------- cfg blocks + checkpoint checksums
    blocks = {}
    for name in ("yolov3-custom", "yolov3-tiny", "yolov3"):
        sys.modules.update(ref_import._cache["ref_modules"])
        try:
            blocks[name] = sys.modules["utils.parse_config"].parse_model_config(
                os.path.join(ref_import.REFERENCE_ROOT, "config", name + ".cfg"))
        finally:
            for k in ref_import._cache["ref_modules"]:
                sys.modules.pop(k, None)
    json.dump(blocks, open(os.path.join(GOLDEN, "cfg_blocks.json"), "w"))
    json.dump(checks, open(os.path.join(GOLDEN, "state_checksums.json"), "w"))
    total = sum(os.path.getsize(os.path.join(GOLDEN, f)) for f in os.listdir(GOLDEN))
    print("golden fixtures written to %s (%d files, %.1f KB)" % (GOLDEN, len(os.listdir(GOLDEN)), total / 1024))


if __name__ == "__main__":
    main()
