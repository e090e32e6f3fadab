"""Slide tiling: ``crop.py`` of the reference on an in-memory slide.

The reference reads an Aperio ``.svs`` with pyvips/openslide, halves 40x slides, and lets ``dzsave`` (layout "google",
tile_size 1536, depth "one", JPEG Q=90) write ``<save_dir>/<slide>/0/<row>/<col>.jpg`` (``crop.py:13-25,27-61``); slides
that crash dzsave are first cut into 24 576-pixel chunks (``divideAndConquerLargeWSI``, ``crop.py:63-92``).  The detector
later re-reads every JPEG and resizes it to 416^2.  pyvips / openslide are not available here, so slide DECODING is out of
reach (and with it any pinning against the reference's own tiles: parity of this module is unpinned); what this module
provides is everything after the decode, for a slide or slide region held as an ``[H, W, 3]`` uint8 array:

* the tile grid, tile names and the divide-and-conquer chunk list with the reference's arithmetic;
* ``save_and_tile``: the same directory layout written with PIL (JPEG Q=90, edge tiles padded white as dzsave does);
* ``detector_inputs`` / ``detect_slide``: with the slide in HBM, tiling and the nearest resize collapse into ONE gather
  kernel (``ay_wsi_gather``) that reads the 416^2 sampled pixels of every tile straight from the slide - no tile files, no
  JPEG round trip, 7 % of the slide's bytes read - feeding ``TileDetector.run_device``.
"""
import ctypes
import os

import numpy as np
import torch

from . import _lib

TILE = 1536
CHUNK = TILE * 16            # crop.py:71 `multiple`
BACKGROUND = 255             # dzsave pads edge tiles with its background, white by default


def tile_grid(height, width, tile_size=TILE):
    """(rows, cols) of the google layout: every started tile is written (edge tiles padded)."""
    return -(-int(height) // tile_size), -(-int(width) // tile_size)


def tile_name(row, col, level=0, suffix=".jpg"):
    """Path of tile (row, col) below the slide's directory: dzsave layout google is ``<z>/<y>/<x><suffix>``."""
    return "%d/%d/%d%s" % (level, row, col, suffix)


def divide_and_conquer_chunks(height, width, multiple=CHUNK):
    """``crop.py:63-92``: (indice, left, top, chunk_width, chunk_height) of every chunk, indices counting row-major from
    1 as the reference's ``indice`` does; chunks on the right / bottom edge are clipped to the slide (the reference
    reaches the same sizes through its except-branches after the full-size crop fails)."""
    out, indice = [], 0
    for top in range(0, height, multiple):
        for left in range(0, width, multiple):
            indice += 1
            out.append((indice, left, top, min(multiple, width - left), min(multiple, height - top)))
    return out


def chunk_basename(indice, slide_name):
    """``crop.py:103-104``: the name a chunk is tiled under (merged again by ``merge1536Subdirectories``)."""
    return "%d_%s" % (indice, slide_name)


def halve(slide):
    """Stand-in for ``vips_img.resize(0.5)`` of 40x slides (``crop.py:45-48``): 2x2 box average with round-half-up.
    (libvips' resize uses a lanczos3 kernel after its own box shrink; without pyvips that filter is not restated.)"""
    a = np.asarray(slide)
    h, w = a.shape[0] // 2 * 2, a.shape[1] // 2 * 2
    a = a[:h, :w].astype(np.uint16)
    return ((a[0::2, 0::2] + a[0::2, 1::2] + a[1::2, 0::2] + a[1::2, 1::2] + 2) // 4).astype(np.uint8)


def padded_tile(slide, row, col, tile_size=TILE):
    """Tile (row, col) of a host slide as dzsave cuts it: ``tile_size``^2, white outside the slide."""
    a = np.asarray(slide)
    out = np.full((tile_size, tile_size, 3), BACKGROUND, dtype=np.uint8)
    part = a[row * tile_size:(row + 1) * tile_size, col * tile_size:(col + 1) * tile_size]
    out[:part.shape[0], :part.shape[1]] = part
    return out


def save_and_tile(slide, output_dir, slide_name, tile_size=TILE, quality=90):
    """``crop.py:13-25`` for a host array: writes ``<output_dir>/<slide_name>/0/<row>/<col>.jpg`` (``.svs`` stripped from
    the name as the reference does) and returns the list of files."""
    from PIL import Image
    base = os.path.join(output_dir, slide_name.split(".svs")[0])
    rows, cols = tile_grid(slide.shape[0], slide.shape[1], tile_size)
    files = []
    for r in range(rows):
        os.makedirs(os.path.join(base, "0", str(r)), exist_ok=True)
        for c in range(cols):
            path = os.path.join(base, tile_name(r, c))
            Image.fromarray(padded_tile(slide, r, c, tile_size), "RGB").save(path, quality=quality)
            files.append(path)
    return files


# ----------------------------------------------------------------------------- the device path
def detector_inputs(slide, row0, col0, n_rows, n_cols, img_size=416, tile_size=TILE, out=None):
    """slide: CUDA uint8 [H,W,3] (rows may be strided).  -> uint8 [n_rows*n_cols, img_size, img_size, 3] on the same
    device: for every tile of the block (row0.., col0..), row-major, the pixels the nearest resize would sample from the
    padded 1536^2 tile (``ay_wsi_gather``).  There is no CPU path."""
    if not (torch.is_tensor(slide) and slide.is_cuda and slide.dtype == torch.uint8 and slide.dim() == 3
            and slide.shape[2] == 3 and slide.stride(2) == 1 and slide.stride(1) == 3):
        raise ValueError("expected a CUDA uint8 slide [H,W,3] with packed pixels")
    dev = slide.device
    n = n_rows * n_cols
    if out is None:
        out = torch.empty((n, img_size, img_size, 3), dtype=torch.uint8, device=dev)
    lib, ctx = _lib.load(), _lib.context(dev.index)
    with torch.cuda.device(dev):
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(lib.ay_wsi_gather(ctx, ctypes.c_void_p(slide.data_ptr()), slide.shape[0], slide.shape[1],
                                     slide.stride(0), tile_size, img_size, row0, col0, n_rows, n_cols,
                                     ctypes.c_void_p(out.data_ptr()), BACKGROUND, st))
    return out[:n]


def detect_slide(detector, slide, tile_size=TILE):
    """All tiles of a slide resident on the detector's GPU through ``detector`` (a ``pipeline.TileDetector``), at most
    ``detector.max_batch`` tiles per engine batch, in row-major tile order.  Yields ``(row, col, detections)`` with
    ``detections`` as ``TileDetector.detect`` returns them (``[k,7]`` CPU tensor in TILE pixel coordinates, or None) -
    what ``detect.py`` would produce for ``<slide>/0/<row>/<col>.jpg``."""
    rows, cols = tile_grid(slide.shape[0], slide.shape[1], tile_size)
    order = [(r, c) for r in range(rows) for c in range(cols)]
    B = detector.max_batch
    staged = torch.empty((B, detector.img_size, detector.img_size, 3), dtype=torch.uint8, device=slide.device)
    for lo in range(0, len(order), B):
        block = order[lo:lo + B]
        k = 0
        while k < len(block):                      # runs of consecutive tiles within one tile row: one gather launch each
            r, c = block[k]
            run = 1
            while k + run < len(block) and block[k + run] == (r, c + run):
                run += 1
            detector_inputs(slide, r, c, 1, run, detector.img_size, tile_size, out=staged[k:k + run])
            k += run
        out = []
        rows_d, counts_d = detector.run_device(staged[:len(block)], (tile_size, tile_size))
        detector._collect(rows_d, counts_d, out)
        for (r, c), det in zip(block, out):
            yield r, c, det
