"""TEST INFRASTRUCTURE ONLY — writes tests/golden/eval_ref.npz from the UNMODIFIED reference
(utils/utils.py:69-190: ap_per_class / compute_ap / get_batch_statistics).   python -m oracle.make_golden_eval"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_import  # noqa: E402


def eval_case(seed):
    """Random ranked predictions against random labels: (tp, conf, pred_cls, target_cls)."""
    rng = np.random.default_rng(seed)
    n, classes = int(rng.integers(5, 400)), int(rng.integers(1, 4))
    tp = (rng.random(n) < 0.6).astype(np.float64)
    conf = rng.random(n).astype(np.float32)
    pred_cls = rng.integers(0, classes + (seed % 3 == 0), n).astype(np.float32)
    target_cls = rng.integers(0, classes, int(rng.integers(1, 300))).astype(np.float64).tolist()
    return tp, conf, pred_cls, target_cls


def stats_case(seed):
    """Detections of a batch of 4 images (one empty) in 416-pixel coordinates against jittered copies as labels."""
    rng = np.random.default_rng(100 + seed)
    outputs, targets = [], []
    for s in range(4):
        k = int(rng.integers(0, 12)) if s != 2 else 0
        if k == 0:
            outputs.append(None)
        else:
            xy = rng.uniform(20, 380, (k, 2))
            wh = rng.uniform(10, 60, (k, 2))
            rows = np.concatenate([xy - wh / 2, xy + wh / 2, rng.uniform(0.8, 1, (k, 2)), rng.integers(0, 2, (k, 1))], 1)
            rows = rows[np.argsort(-rows[:, 4])].astype(np.float32)
            outputs.append(torch.from_numpy(rows))
        n_t = int(rng.integers(0, 8))
        for _ in range(n_t):
            if outputs[-1] is not None and rng.random() < 0.7:
                src = outputs[-1][int(rng.integers(0, len(outputs[-1])))].numpy()
                box = src[:4] + rng.normal(0, 3, 4)
                cls = src[6] if rng.random() < 0.8 else 1 - src[6]
            else:
                c, w = rng.uniform(30, 380, 2), rng.uniform(10, 60, 2)
                box, cls = np.concatenate([c - w / 2, c + w / 2]), float(rng.integers(0, 2))
            targets.append([s, cls, *box])
    targets = torch.tensor(targets, dtype=torch.float32).reshape(-1, 6)
    return outputs, targets


def main():
    _, uu, _ = ref_import.load()
    out = {}
    for s in range(12):
        tp, conf, pred_cls, target_cls = eval_case(s)
        p, r, ap, f1, cls = uu.ap_per_class(tp.copy(), conf.copy(), pred_cls.copy(), list(target_cls))
        out.update({"ap_p_%d" % s: p, "ap_r_%d" % s: r, "ap_ap_%d" % s: ap, "ap_f1_%d" % s: f1, "ap_cls_%d" % s: cls})
    for s in range(8):
        outputs, targets = stats_case(s)
        for thr in (0.5, 0.75):
            m = uu.get_batch_statistics([o.clone() if o is not None else None for o in outputs], targets.clone(), thr)
            out["st_n_%d_%d" % (s, int(thr * 100))] = np.array([len(m)])
            for i, (tp, sc, lb) in enumerate(m):
                out["st_tp_%d_%d_%d" % (s, int(thr * 100), i)] = tp
                out["st_sc_%d_%d_%d" % (s, int(thr * 100), i)] = sc.numpy()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "eval_ref.npz"), **out)
    print("wrote eval_ref.npz with %d arrays" % len(out))


if __name__ == "__main__":
    main()
