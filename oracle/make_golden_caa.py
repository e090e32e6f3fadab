"""TEST INFRASTRUCTURE ONLY — writes tests/golden/caa_ref.npz from the UNMODIFIED reference
(core.py:109-159 get256Img, 161-205 Net) and torchvision's ToTensor + Normalize (core.py:435).
    python -m oracle.make_golden_caa

The reference's trained filter (pickles/model_all_fold_3_thresholding_2_l2.pkl, an LFS pointer) and its
normalisation file are not in the checkout, so the network is the reference's ``Net`` class with a seeded
state_dict and the normalisation constants are NORM below; the inputs are regenerated from their seeds by the tests
(``seeded_caa_state``, ``caa_tile``, BOXES), only the reference's outputs are stored."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

NORM = {"mean": [0.77, 0.59, 0.72], "std": [0.13, 0.17, 0.12]}      # stained-tissue-like channel statistics
CLASSES = ["Cored", "CAA"]

# (top-left x, top-left y, width, height): interior, the four edges, the four corners, and centres exactly on the
# strict comparisons of get256Img (128 and 1408)
BOXES = [(700, 640, 60, 50), (300, 900, 35, 80), (1100, 400, 120, 90), (640, 10, 50, 40), (30, 700, 40, 60),
         (800, 1480, 60, 50), (1490, 600, 40, 44), (5, 8, 30, 30), (1480, 12, 50, 40), (10, 1470, 60, 60),
         (1470, 1490, 60, 40), (108, 500, 40, 40), (500, 108, 40, 40), (1388, 500, 40, 40), (500, 1388, 40, 40),
         (109, 1387, 40, 40), (0, 0, 1536, 1536), (900, 900, 3, 3)]


def seeded_caa_state(seed):
    """A state_dict for core.Net (core.py:165-198) with He-scaled convolutions and non-trivial BatchNorm statistics;
    the classifier is scaled and biased so that the CAA logit of textured crops straddles zero."""
    g = torch.Generator().manual_seed(seed)
    chans = (3, 16, 32, 48, 64, 80, 96)
    sd = {}
    for k in range(6):
        cin, cout = chans[k], chans[k + 1]
        c, b = "features.%d." % (4 * k), "features.%d." % (4 * k + 1)
        sd[c + "weight"] = torch.randn((cout, cin, 3, 3), generator=g) * (2.0 / (9 * cin)) ** 0.5
        sd[c + "bias"] = torch.randn((cout,), generator=g) * 0.05
        sd[b + "weight"] = 0.8 + 0.4 * torch.rand((cout,), generator=g)
        sd[b + "bias"] = torch.randn((cout,), generator=g) * 0.1
        sd[b + "running_mean"] = torch.randn((cout,), generator=g) * 0.1
        sd[b + "running_var"] = 0.5 + torch.rand((cout,), generator=g)
        sd[b + "num_batches_tracked"] = torch.tensor(100)
    sd["classifier.0.weight"] = torch.randn((3, 96 * 4 * 4), generator=g) * (4.0 / (96 * 16) ** 0.5)
    sd["classifier.0.bias"] = torch.tensor([0.3, -0.2, -3.0])
    return sd


def caa_tile(seed, size=1536):
    """A synthetic RGB uint8 tile: pinkish background, blocky low-frequency texture, darker round plaques."""
    rng = np.random.default_rng(seed)
    base = np.array([205, 160, 190], dtype=np.float32)
    coarse = rng.normal(0, 18, (size // 32, size // 32, 3)).astype(np.float32)
    img = base + np.kron(coarse, np.ones((32, 32, 1), dtype=np.float32))
    yy, xx = np.mgrid[0:size, 0:size].astype(np.float32)
    for _ in range(60):
        cx, cy, r = rng.uniform(0, size), rng.uniform(0, size), rng.uniform(10, 70)
        m = np.exp(-((xx - cx) ** 2 + (yy - cy) ** 2) / (2 * r * r))
        img -= m[..., None] * rng.uniform(40, 120, 3).astype(np.float32)
    img += rng.normal(0, 6, img.shape).astype(np.float32)
    return np.clip(img, 0, 255).astype(np.uint8)


def main():
    from oracle import ref_import
    from torchvision import transforms
    _, _, core = ref_import.load()
    tf = transforms.Compose([transforms.ToTensor(), transforms.Normalize(NORM["mean"], NORM["std"])])
    net = core.Net()
    net.load_state_dict(seeded_caa_state(5))
    net.eval()
    tile = caa_tile(11)
    windows, logits, norm_sum = [], [], []
    for box in BOXES:
        crop = core.get256Img(tile, box)                               # core.py:109-159, unmodified
        assert crop.shape == (256, 256, 3), (box, crop.shape)
        ys = np.flatnonzero((tile[:, :, 0] == crop[0, 0, 0]).any(axis=1))   # recover the window by search
        found = None
        for y in ys:
            for x in np.flatnonzero(tile[y, :, 0] == crop[0, 0, 0]):
                if y + 256 <= tile.shape[0] and x + 256 <= tile.shape[1] and np.array_equal(tile[y:y + 256, x:x + 256], crop):
                    found = (x, x + 256, y, y + 256)
                    break
            if found:
                break
        assert found, box
        windows.append(found)
        from PIL import Image
        x = tf(Image.fromarray(crop.astype("uint8"), "RGB"))           # core.py:215-217
        with torch.no_grad():
            out = net(x.view(1, 3, 256, 256).contiguous())
        logits.append(out[0].numpy())
        norm_sum.append([float(x.double().sum()), float(x[:, 17, 201].double().sum())])
    out = {"windows": np.asarray(windows, dtype=np.int32), "logits": np.asarray(logits, dtype=np.float32),
           "norm_sum": np.asarray(norm_sum, dtype=np.float64)}
    path = os.path.join(ROOT, "tests", "golden", "caa_ref.npz")
    np.savez_compressed(path, **out)
    print("wrote %s: %d boxes, CAA logits %s" % (path, len(BOXES), np.round(out["logits"][:, 2], 2).tolist()))


if __name__ == "__main__":
    main()
