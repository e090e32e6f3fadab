"""The oracle (oracle/restate.py) against the golden vectors produced by the real
reference (oracle/make_golden.py) and against the reference's own known answers.
CPU only; these pin the checker that the GPU parity tests rely on."""
import os
import tempfile

import numpy as np
import pytest
import torch

from helpers import write_mini_cfg
from oracle import restate as R


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.int32)


# ----------------------------------------------------------------- IoU / rescale
def test_iou_known_answers_of_reference_unit_test(golden):
    """unit_test.py:137-150: identical -> 1, disjoint -> 0, quarter overlap in [0.25,0.26] (+1 convention)."""
    g = golden("iou_ref.npz")
    kb = g["kat_boxes"]
    got = R.bbox_iou_xyxy(kb[0], kb[1:])
    assert got[0] == 1.0 and got[1] == 0.0 and 0.25 <= got[2] <= 0.26
    assert np.allclose(got, g["kat"], rtol=0, atol=1e-6)
    assert abs(float(got[2]) - 51.0 ** 2 / 101.0 ** 2) < 1e-6


def test_iou_bitwise_vs_reference(golden):
    g = golden("iou_ref.npz")
    got = R.bbox_iou_xyxy(g["boxes"][0], g["boxes"])
    assert np.array_equal(bits(got), bits(g["iou"]))


def test_rescale_bitwise_vs_reference(golden):
    g = golden("rescale_ref.npz")
    for ci in range(5):
        got = R.rescale_boxes(g["in_%d" % ci], 416, tuple(int(v) for v in g["shape_%d" % ci]))
        assert np.array_equal(bits(got[:, :4]), bits(g["out_%d" % ci][:, :4])), ci
        assert np.array_equal(bits(got[:, 4:]), bits(g["in_%d" % ci][:, 4:]))


# ----------------------------------------------------------------- ATen sum order
def test_aten_sum_order_matches_this_torch():
    """The cluster merge of the reference's NMS goes through torch.sum (utils/utils.py:267);
    the restated cascade order (and its lane width) must be the one this host's ATen uses."""
    rng = np.random.default_rng(0)
    for k in list(range(1, 70)) + [100, 127, 128, 129, 255, 256, 257, 300, 513, 1000, 4100]:
        m = rng.uniform(0, 400, (k, 4)).astype(np.float32)
        v = rng.uniform(0, 1, k).astype(np.float32)
        assert np.array_equal(bits(R.aten_sum_outer(m)), bits(torch.from_numpy(m).sum(0).numpy())), k
        assert bits(np.array([R.aten_sum_inner(v)]))[0] == bits(torch.from_numpy(v).sum().numpy().reshape(1))[0], k


# ----------------------------------------------------------------- NMS
def test_nms_exact_bitwise_vs_reference_golden(golden):
    g = golden("nms_ref.npz")
    ci = 0
    checked = 0
    while "in_%d" % ci in g:
        P = g["in_%d" % ci]
        ct, nt, _ = g["par_%d" % ci]
        assert np.array_equal(bits(R.xywh2xyxy(P[..., :4])), bits(g["xyxy_%d" % ci]))
        for b in range(P.shape[0]):
            want = g["out_%d_%d" % (ci, b)]
            rows, keep = R.nms_exact(P[b], ct, nt)
            if want.shape[0] == 0:
                assert rows is None
                continue
            assert rows.shape == want.shape, (ci, b)
            assert np.array_equal(bits(rows), bits(want)), (ci, b)
            # keep indices point at the head of each cluster: conf / class columns are untouched copies
            xy = R.xywh2xyxy(P[b, :, :4])
            assert np.array_equal(bits(P[b, keep, 4]), bits(want[:, 4]))
            assert len(np.unique(keep)) == len(keep)
            checked += 1
        ci += 1
    assert ci == 5 and checked >= 8


def test_nms_faithful_equals_exact():
    from amyloid_yolo_paper_b200 import synth
    P = synth.synthetic_predictions(2, seed=3, img=128, grids=(4, 8), clusters=(8, 16),
                                    anchors={4: ((40, 30), (50, 60), (90, 80)), 8: ((10, 20), (20, 15), (20, 40))})
    res = R.nms_faithful(torch.from_numpy(P.copy()), 0.2, 0.4)
    for b in range(2):
        rows, _ = R.nms_exact(P[b], 0.2, 0.4)
        assert np.array_equal(bits(res[b].numpy()), bits(rows))


def test_nms_empty_and_single():
    P = np.zeros((1, 5, 7), np.float32)
    assert R.nms_exact(P[0], 0.5, 0.4)[0] is None
    assert R.nms_faithful(torch.from_numpy(P.copy()), 0.5, 0.4) == [None]
    P[0, 2] = [50, 60, 10, 20, 0.9, 0.2, 0.7]
    rows, keep = R.nms_exact(P[0], 0.5, 0.4)
    assert keep.tolist() == [2] and rows.shape == (1, 7) and rows[0, 6] == 1.0
    assert np.array_equal(bits(rows[0, :4]), bits(np.array([45, 50, 55, 70], np.float32)))


# ----------------------------------------------------------------- merge
def test_merge_exact_and_emulated_vs_reference_golden(golden):
    g = golden("merge_ref.npz")
    n = int(g["count"][0])
    merged = 0
    for s in range(n):
        rows, want = g["in_%d" % s], g["out_%d" % s]
        a = R.merge_exact(rows)
        e = R.merge_emulated(rows)
        assert a.shape == want.shape and np.array_equal(bits(a), bits(want)), s
        assert e.shape == want.shape and np.array_equal(bits(e), bits(want)), s      # values AND row order
        merged += len(rows) - len(want)
    assert merged > 40


def test_merge_chain_is_order_dependent_and_matches_reference(golden):
    """a 3-chain shrinks by one pixel per re-merge; class 2 never merges."""
    g = golden("merge_ref.npz")
    s = int(g["count"][0]) - 1
    want = g["out_%d" % s]
    got = R.merge_emulated(g["in_%d" % s])
    assert np.array_equal(bits(got), bits(want))
    assert (want[:, 6] == 2).sum() == 2 and (want[:, 6] == 1).sum() == 1 and (want[:, 6] == 0).sum() == 1


def test_merge_faithful_equals_closed_form():
    from amyloid_yolo_paper_b200 import synth
    for s in range(12):
        rows = synth.synthetic_detections(6, seed=s, extent=80.0, overlap=0.7, wh=(3, 20), jitter=5)
        assert np.array_equal(bits(R.merge_faithful(rows)), bits(R.merge_exact(rows))), s


def test_python_hash_and_set_emulation():
    rng = np.random.default_rng(1)
    for _ in range(300):
        t = tuple(float(np.float32(v)) if rng.random() < 0.6 else float(int(v)) for v in rng.uniform(-50, 1600, 7))
        assert R.py_hash_tuple(t) == hash(t)
    for trial in range(60):
        real, emu = set(), R.EmulatedSet()
        keys = [tuple(float(np.float32(v)) for v in rng.uniform(0, 1500, 7)) for _ in range(int(rng.integers(1, 90)))]
        for k in keys:
            real.add(k); emu.add(k)
        for k in keys[::3]:
            real.remove(k); emu.remove(k)
        for k in keys[1::5]:
            real.add(k); emu.add(k)
        assert list(real) == emu.items(), trial


# ----------------------------------------------------------------- forward / decode / preprocess
def test_forward_mini_vs_reference_golden(golden):
    from amyloid_yolo_paper_b200 import synth
    from amyloid_yolo_paper_b200.models import Darknet
    g = golden("forward_mini_ref.npz")
    with tempfile.TemporaryDirectory() as d:
        cfg = write_mini_cfg(d)
        m = Darknet(cfg, img_size=64)
        sd = synth.synthetic_state_dict(m, seed=0, conf_bias=-1.0)
        chk = golden("state_checksums.json")["mini_seed0"]
        assert abs(float(sum(v.double().sum() for v in sd.values())) - chk) < 1e-6 * max(1.0, abs(chk))
        x = torch.stack([R.preprocess_u8(t, 64) for t in g["tiles"]])
        assert np.array_equal(bits(x.numpy()), bits(g["x"]))
        with torch.no_grad():
            y = R.darknet_forward(R.parse_cfg(cfg), sd, x)
    want = torch.from_numpy(g["y"])
    assert y.shape == want.shape
    # same torch ops on the same host: conv accumulation order may differ between module and functional calls
    assert float((y - want).abs().max()) <= 1e-4 * float(want.abs().max())


def test_decode_bitwise_vs_reference_golden(golden):
    from amyloid_yolo_paper_b200 import synth
    g = golden("forward_heads_ref.npz")
    for li, G in enumerate((13, 26, 52)):
        dec = R.decode_head(torch.from_numpy(g["raw_%d" % li]), synth.ANCHORS_V3[G], 2, 416)
        assert np.array_equal(bits(dec.numpy()), bits(g["dec_%d" % li])), G


def test_preprocess_vs_reference_golden(golden):
    g = golden("preprocess_ref.npz")
    idx = R.nearest_src_index(416, 1536)
    assert np.array_equal(idx, g["idx_1536_416"])
    assert np.array_equal(idx, (np.arange(416) * 48) // 13)                # SURVEY probe D2
    src = g["sample_src"]                                                  # [8,416,3] uint8 gathered pixels
    want = g["sample_rows"]                                                # [3,8,416] fp32
    got = (torch.from_numpy(src).permute(2, 0, 1).float() / 255).numpy()
    assert np.array_equal(bits(got), bits(want))


def test_evaluation_statistics_vs_reference_golden(golden):
    """ap_per_class / compute_ap (product, host numpy - as in the reference) and the oracle's restatement against the
    reference's outputs (utils/utils.py:69-149), and the oracle's get_batch_statistics against utils/utils.py:152-190."""
    from amyloid_yolo_paper_b200.utils.utils import ap_per_class
    from oracle.make_golden_eval import eval_case, stats_case
    g = golden("eval_ref.npz")
    for s in range(12):
        tp, conf, pred_cls, target_cls = eval_case(s)
        for fn in (ap_per_class, R.ap_per_class_faithful):
            p, r, ap, f1, cls = fn(tp.copy(), conf.copy(), pred_cls.copy(), np.asarray(target_cls))
            assert np.array_equal(cls, g["ap_cls_%d" % s])
            for name, v in (("p", p), ("r", r), ("ap", ap), ("f1", f1)):
                assert np.array_equal(v, g["ap_%s_%d" % (name, s)]), (fn.__name__, s, name)
    for s in range(8):
        outputs, targets = stats_case(s)
        for thr in (0.5, 0.75):
            m = R.get_batch_statistics_faithful([o.numpy() if o is not None else None for o in outputs], targets.numpy(), thr)
            assert len(m) == int(g["st_n_%d_%d" % (s, int(thr * 100))][0])
            for i, (tp, sc, lb) in enumerate(m):
                assert np.array_equal(tp, g["st_tp_%d_%d_%d" % (s, int(thr * 100), i)])


def test_caa_filter_pieces_vs_reference_golden(golden):
    """SURVEY §8f N2.  Against outputs of the unmodified reference (oracle/make_golden_caa.py: core.get256Img, core.Net,
    torchvision's ToTensor + Normalize): the oracle's crop windows, its normalisation and its float64 restatement of the
    network; and the product's host-side crop rule (caa.crop_window) against the oracle's branch-for-branch restatement
    on every centre position near the comparisons."""
    from amyloid_yolo_paper_b200 import caa
    from oracle.make_golden_caa import BOXES, NORM, caa_tile, seeded_caa_state
    g = golden("caa_ref.npz")
    tile, state = caa_tile(11), seeded_caa_state(5)
    xs = []
    for box, win, ns in zip(BOXES, g["windows"], g["norm_sum"]):
        assert R.get256img_box_faithful(box) == tuple(int(v) for v in win), box
        assert caa.crop_window(box) == tuple(int(v) for v in win), box
        xl, xr, yt, yb = win
        assert np.array_equal(caa.get256Img(tile, box), tile[yt:yb, xl:xr])
        x = R.caa_normalise(tile[yt:yb, xl:xr], NORM["mean"], NORM["std"])
        assert float(x.double().sum()) == ns[0] and float(x[:, 17, 201].double().sum()) == ns[1]     # same fp32 bits
        xs.append(x)
    with torch.no_grad():
        logits = R.caa_net_forward(state, torch.stack(xs)).numpy()
    assert np.abs(logits - g["logits"]).max() <= 1e-4 * np.abs(g["logits"]).max()                     # fp32 reference vs float64
    # the product's per-axis rule == the reference's nine branches, for every centre within 3 pixels of a comparison
    edge = [c for c in range(0, 1537) if min(abs(c - 128), abs(c - 1408), c, 1536 - c) <= 3] + [700]
    for cx in edge:
        for cy in edge:
            box = (cx - 10, cy - 7, 20, 14)
            assert caa.crop_window(box) == R.get256img_box_faithful(box), box
    # parameter names of the product's Net are the reference's (the pickled model's state_dict loads)
    net = caa.Net()
    assert set(net.state_dict()) == set(state)
    net.load_state_dict(state)
    w, b = caa.fold_stage(state, 2)
    y_ref = torch.nn.functional.batch_norm(
        torch.nn.functional.conv2d(torch.ones(1, 32, 5, 5, dtype=torch.float64), state["features.8.weight"].double(),
                                   state["features.8.bias"].double(), padding=1),
        state["features.9.running_mean"].double(), state["features.9.running_var"].double(),
        state["features.9.weight"].double(), state["features.9.bias"].double(), False, 0.0, 1e-5)
    y_fold = torch.nn.functional.conv2d(torch.ones(1, 32, 5, 5, dtype=torch.float64), torch.from_numpy(w).double(),
                                        torch.from_numpy(b).double(), padding=1)
    assert float((y_ref - y_fold).abs().max()) < 1e-5
