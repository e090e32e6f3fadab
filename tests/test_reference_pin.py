"""Container-only: the oracle against the UNMODIFIED reference imported from /root/reference
(more cases than the committed golden vectors).  Skipped where the reference is not mounted."""
import numpy as np
import pytest
import torch

from amyloid_yolo_paper_b200 import synth
from amyloid_yolo_paper_b200.config import ensure_cfg
from oracle import ref_import, restate as R

pytestmark = pytest.mark.reference


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.int32)


@pytest.fixture(scope="module")
def ref():
    return ref_import.load()


def test_nms_exact_vs_reference(ref):
    _, uu, _ = ref
    anchors = {4: ((40, 30), (50, 60), (90, 80)), 8: ((10, 20), (20, 15), (20, 40))}
    for seed in range(6):
        classes = 2 if seed % 3 else 5
        P = synth.synthetic_predictions(2, num_classes=classes, seed=100 + seed, img=128, grids=(4, 8),
                                        clusters=(6, 20), anchors=anchors)
        for ct in (0.05, 0.5):
            res = uu.non_max_suppression(torch.from_numpy(P.copy()), ct, 0.4)
            for b in range(2):
                rows, _ = R.nms_exact(P[b], ct, 0.4)
                assert (rows is None) == (res[b] is None)
                if rows is not None:
                    assert np.array_equal(bits(rows), bits(res[b].numpy())), (seed, ct, b)


def test_merge_emulated_vs_reference(ref):
    _, _, core = ref
    for s in range(80):
        rng = np.random.default_rng(s)
        rows = synth.synthetic_detections(int(rng.integers(1, 12)), seed=500 + s, extent=120.0, overlap=0.7, wh=(3, 30),
                                          jitter=6)
        want = core.mergeDetections(torch.from_numpy(rows.copy())).numpy().reshape(-1, 7)
        assert np.array_equal(bits(R.merge_emulated(rows)), bits(want)), s


def test_forward_vs_reference_tiny(ref):
    models, _, _ = ref
    from amyloid_yolo_paper_b200.models import Darknet
    cfg = ensure_cfg("yolov3-tiny")
    ours = Darknet(cfg)
    sd = synth.synthetic_state_dict(ours, seed=1, conf_bias=-2.0)
    m = models.Darknet(cfg)
    m.load_state_dict(sd)
    m.eval()
    x = torch.rand(1, 3, 416, 416, generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        want = m(x)
        got = R.darknet_forward(R.parse_cfg(cfg), sd, x)
    assert got.shape == want.shape == (1, 2535, 85)
    assert float((got - want).abs().max()) <= 1e-4 * float(want.abs().max())


def test_state_dict_keys_equal_reference(ref):
    models, _, _ = ref
    from amyloid_yolo_paper_b200.models import Darknet
    for name in ("yolov3-custom", "yolov3-tiny"):
        cfg = ensure_cfg(name)
        a, b = Darknet(cfg).state_dict(), models.Darknet(cfg).state_dict()
        assert list(a.keys()) == list(b.keys())
        assert all(a[k].shape == b[k].shape and a[k].dtype == b[k].dtype for k in a)
