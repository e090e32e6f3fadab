"""Inference-side image loading (reference ``utils/datasets.py:40-62`` + ``utils/transforms.py:80-111``).

The reference's ``ImageFolder`` decodes an image with PIL, pads it to a square (``PadSquare``, centred, zeros),
converts ALL of its pixels to float (``ToTensor``) and then keeps ``img_size**2`` of them (``Resize`` is a nearest
neighbour gather: ``out[y, x] = in[idx[y], idx[x]]`` with ``idx[i] = floor(i * float(src) / float(dst))`` in fp32).
For a 1536 x 1536 tile that is 7 M conversions to keep 173 k pixels, and 2 MB of fp32 per tile on the way to the GPU.

Here the gather happens first, on the uint8 image, with the same index table: what crosses PCIe is the pinned uint8
``[B, S, S, 3]`` block of exactly the pixels the network reads (0.5 MB per tile at S = 416); ``/ 255`` is folded into the
first convolution's weights on the device.  Decoding stays on the host with PIL, spread over ``n_cpu`` threads (PIL
releases the GIL while it decodes).  The result is bit-identical to the reference chain (tests/test_host_cpu.py).
"""
import glob
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch


def nearest_index(dst, src):
    """Source index of ``F.interpolate(mode="nearest")`` (ATen ``nearest_neighbor_compute_source_index``):
    ``min(floor(i * float32(src / dst)), src - 1)``, all in fp32."""
    scale = np.float32(src) / np.float32(dst)
    return np.minimum(np.floor(np.arange(dst, dtype=np.float32) * scale).astype(np.int64), src - 1)


def load_image(path):
    """``np.array(Image.open(path).convert('RGB'), dtype=np.uint8)`` (utils/datasets.py:47-50)."""
    from PIL import Image, ImageFile
    ImageFile.LOAD_TRUNCATED_IMAGES = True                    # utils/datasets.py:10
    return np.array(Image.open(path).convert("RGB"), dtype=np.uint8)


def gather_resized(img, size, out=None):
    """PadSquare + nearest Resize of an ``[h, w, 3]`` uint8 image, as uint8 ``[size, size, 3]``.

    Equals ``Resize(size)(ToTensor(PadSquare(img))) * 255`` of the reference element for element: the padded square of
    edge ``s = max(h, w)`` (image centred, zeros around, ``transforms.py:80-86``) is never built; rows / columns of the
    index table that fall into the padding produce zeros."""
    h, w = img.shape[:2]
    s = max(h, w)
    top, left = (s - h) // 2, (s - w) // 2
    idx = nearest_index(size, s)
    ry, rx = idx - top, idx - left
    oky, okx = (ry >= 0) & (ry < h), (rx >= 0) & (rx < w)
    if out is None:
        out = np.empty((size, size, 3), dtype=np.uint8)
    g = img[np.clip(ry, 0, h - 1)][:, np.clip(rx, 0, w - 1)]
    if not (oky.all() and okx.all()):
        g = g * (oky[:, None, None] & okx[None, :, None])
    out[...] = g
    return out


class ImageFolder(torch.utils.data.Dataset):
    """Drop-in for the reference's ``ImageFolder`` (same ``files`` order: ``sorted(glob("%s/*.*"))``).  ``__getitem__``
    returns ``(path, uint8 [S, S, 3] pre-gathered pixels, (h, w))`` instead of the reference's fp32 ``[3, S, S]``."""

    def __init__(self, folder_path, img_size=416, transform=None):
        self.files = sorted(glob.glob("%s/*.*" % folder_path))
        self.img_size = int(img_size)
        self.transform = transform                          # accepted for signature compatibility; unused

    def __getitem__(self, index):
        path = self.files[index % len(self.files)]
        img = load_image(path)
        return path, torch.from_numpy(gather_resized(img, self.img_size)), tuple(img.shape[:2])

    def __len__(self):
        return len(self.files)


class TileLoader:
    """Batches of an :class:`ImageFolder` for ``TileDetector.detect_stream``: yields
    ``(paths, pinned uint8 [b, S, S, 3], (h, w))`` with ``b <= batch_size``.  Images of one batch share their original
    shape (``rescale_boxes`` takes one shape per call); a batch closes early when the shape changes.  ``n_cpu`` worker
    threads decode and gather ahead of the consumer; two pinned buffers rotate, so a batch stays valid until the batch
    after the next one is requested."""

    def __init__(self, folder_path, img_size=416, batch_size=1, n_cpu=0, pin=True, depth=3):
        self.dataset = ImageFolder(folder_path, img_size)
        self.batch_size = max(1, int(batch_size))
        self.n_cpu = max(1, int(n_cpu))
        self.pin = bool(pin) and torch.cuda.is_available()
        self.depth = max(2, int(depth))                      # batches in flight: the consumer keeps up to depth - 1 alive

    def __len__(self):
        return len(self.dataset)

    def _decode(self, path, slot):
        img = load_image(path)
        gather_resized(img, self.dataset.img_size, out=slot)
        return tuple(img.shape[:2])

    def __iter__(self):
        S, B = self.dataset.img_size, self.batch_size
        bufs = []
        for _ in range(self.depth):
            t = torch.empty((B, S, S, 3), dtype=torch.uint8)
            bufs.append(t.pin_memory() if self.pin else t)
        files = self.dataset.files
        turn = 0
        with ThreadPoolExecutor(max_workers=self.n_cpu) as pool:
            pos = 0
            while pos < len(files):
                buf = bufs[turn % self.depth]
                turn += 1
                view = buf.numpy()
                chunk = files[pos:pos + B]
                shapes = list(pool.map(self._decode, chunk, [view[i] for i in range(len(chunk))]))
                n = 1
                while n < len(chunk) and shapes[n] == shapes[0]:
                    n += 1
                # images after a shape change are decoded again with the next batch (rare: slide tiles share one size)
                yield chunk[:n], buf[:n], shapes[0]
                pos += n
