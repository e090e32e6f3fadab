"""Tile-detection pipeline: the body of ``detect.py``'s batch loop as one device-side chain.

Reference flow per batch (``detect.py:88-133``, ``validation.py:333-346``)::

    ImageFolder -> ToTensor -> Resize(416)          utils/transforms.py:89-111
    model(input_imgs)                               models.py:237-255   (returns a CPU tensor)
    non_max_suppression(det, conf_thres, nms_thres) utils/utils.py:235-273  (python loop, CPU)
    rescale_boxes(det, img_size, (H, W))            utils/utils.py:36-50
    mergeDetections(det)   [--merge_boxes True]     core.py:366-423

Here the decoded tensor never leaves the GPU: ``ay_forward`` (uint8 tiles in, the
resize fused into the first convolution) -> ``ay_nms`` -> ``ay_rescale_merge`` run
back to back on one stream, and only ``counts [B]`` plus the first ``kmax`` rows per
tile are copied back.  Results equal what the reference's functions return on the
same decoded input (bit-exact, ``tests/test_gpu_parity.py``).
"""
import ctypes

import torch

from . import _lib


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


class TileDetector:
    """Runs forward + NMS + rescale (+ merge) for batches of square uint8 tiles.

    ``model`` is a :class:`amyloid_yolo_paper_b200.models.Darknet` whose weights are
    already loaded.  ``cap`` bounds the detections kept per tile (the reference has
    no bound; exceeding it raises ``EngineError(AY_ECAPACITY)`` at the next check
    instead of truncating silently).
    """

    def __init__(self, model, conf_thres=0.8, nms_thres=0.4, merge_boxes=True, max_batch=64, cap=256,
                 device=None):
        if not torch.cuda.is_available():
            raise _lib.EngineUnavailable("no CUDA device: the B200 pipeline has no CPU fallback")
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.model = model
        self.conf_thres = float(conf_thres)
        self.nms_thres = float(nms_thres)
        self.merge_boxes = bool(merge_boxes)
        self.max_batch = int(max_batch)
        self.engine = model.engine(device=self.device, max_batch=self.max_batch)
        self.img_size = self.engine.img_size
        self.rows, self.cols = self.engine.rows, self.engine.cols
        self.cap = int(min(cap, self.rows))
        self.lib = self.engine.lib
        self.ctx = self.engine.ctx          # NMS / merge scratch lives in the same context as the network
        B, dev = self.max_batch, self.device
        self.decoded = torch.empty((B, self.rows, self.cols), dtype=torch.float32, device=dev)
        self.nms_rows = torch.empty((B, self.cap, 7), dtype=torch.float32, device=dev)
        self.nms_keep = torch.empty((B, self.cap), dtype=torch.int32, device=dev)
        self.nms_counts = torch.empty((B,), dtype=torch.int32, device=dev)
        # two result slots: the host path reads slot i while the device fills slot i+1
        self._out_rows = [torch.empty((B, self.cap, 7), dtype=torch.float32, device=dev) for _ in range(2)]
        self._out_counts = [torch.empty((B,), dtype=torch.int32, device=dev) for _ in range(2)]
        self.out_rows, self.out_counts = self._out_rows[0], self._out_counts[0]
        self.launches_per_batch = 0

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # ------------------------------------------------------------------ device side
    def run_device(self, tiles, original_shape=None, slot=0):
        """tiles: CUDA uint8 [B,H,H,3] (B <= max_batch) or CUDA fp32 [B,3,S,S].
        Enqueues the whole chain on the current stream; returns views
        ``(rows [B,cap,7], counts [B])`` of internal buffers (valid until the next call)."""
        B = tiles.shape[0]
        if B < 1 or B > self.max_batch:
            raise ValueError("batch %d outside 1..%d" % (B, self.max_batch))
        if not tiles.is_cuda or not tiles.is_contiguous():
            raise ValueError("run_device expects a contiguous CUDA tensor")
        if tiles.dtype == torch.uint8:
            if tiles.dim() != 4 or tiles.shape[3] != 3 or tiles.shape[1] != tiles.shape[2]:
                raise ValueError("expected uint8 tiles [B,H,H,3], got %s" % (tuple(tiles.shape),))
            kind, sh, sw = _lib.AY_IN_U8_HWC, tiles.shape[1], tiles.shape[2]
        elif tiles.dtype == torch.float32:
            kind, sh, sw = _lib.AY_IN_F32_NCHW, tiles.shape[2], tiles.shape[3]
        else:
            raise ValueError("tiles must be uint8 HWC or float32 NCHW")
        if original_shape is None:
            original_shape = (sh, sw)
        st = self._stream()
        lib, ctx = self.lib, self.ctx
        with torch.cuda.device(self.device):
            _lib.check(lib.ay_forward(ctx, _ptr(tiles), kind, B, sh, sw, _ptr(self.decoded), st))
            _lib.check(lib.ay_nms(ctx, _ptr(self.decoded), B, self.rows, self.cols - 5, self.conf_thres,
                                  self.nms_thres, _ptr(self.nms_rows), _ptr(self.nms_keep), _ptr(self.nms_counts),
                                  self.cap, st))
            _lib.check(lib.ay_rescale_merge(ctx, _ptr(self.nms_rows), _ptr(self.nms_counts), B, self.cap,
                                            self.img_size, int(original_shape[0]), int(original_shape[1]),
                                            1 if self.merge_boxes else 0, _ptr(self._out_rows[slot]),
                                            _ptr(self._out_counts[slot]), st))
        # forward kernels + 4 NMS kernels + 1 rescale/merge kernel
        self.launches_per_batch = self.engine.launch_count() + 4 + 1
        return self._out_rows[slot][:B], self._out_counts[slot][:B]

    def check(self):
        with torch.cuda.device(self.device):
            _lib.check(self.lib.ay_check(self.ctx, self._stream()))

    # ------------------------------------------------------------------ host side
    def detect(self, tiles, original_shape=None):
        """tiles: host (ideally pinned) or device uint8 [B,H,H,3] / fp32 [B,3,S,S], any B.
        -> list (len B) of ``[k,7]`` fp32 CPU tensors (x1,y1,x2,y2,conf,cls_conf,cls_pred in
        original-tile pixels) or ``None`` where nothing was detected — the per-image value
        ``detect.py:127-133`` ends up with."""
        out = []
        if not tiles.is_cuda:
            return self._detect_host(tiles, original_shape)
        for lo in range(0, tiles.shape[0], self.max_batch):
            chunk = tiles[lo:lo + self.max_batch]
            rows, counts = self.run_device(chunk.contiguous(), original_shape)
            self._collect(rows, counts, out)
        return out

    def _collect(self, rows, counts, out):
        counts_h = counts.cpu()                 # synchronises the compute stream
        self.check()
        kmax = int(counts_h.max()) if counts_h.numel() else 0
        rows_h = rows[:, :max(kmax, 1)].cpu()
        self.d2h_bytes = getattr(self, "d2h_bytes", 0) + counts_h.numel() * 4 + rows_h.numel() * 4
        for b, k in enumerate(counts_h.tolist()):
            out.append(rows_h[b, :k].clone() if k > 0 else None)

    def _detect_host(self, tiles, original_shape):
        """Host input.  Sub-batches are uploaded on a copy stream into two device buffers while the previous one
        computes (a small first sub-batch keeps the exposed part of the H2D short); results come back on a third
        stream from two result slots, so the compute stream never waits for the host between sub-batches."""
        n = tiles.shape[0]
        main = torch.cuda.current_stream(self.device)
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._d2h_stream = torch.cuda.Stream(device=self.device)
            self._stage = [None, None]
            self._counts_h = [torch.empty((self.max_batch,), dtype=torch.int32).pin_memory() for _ in range(2)]
            self._rows_h = [torch.empty((self.max_batch, self.cap, 7), dtype=torch.float32).pin_memory() for _ in range(2)]
        big = self.max_batch if self.max_batch < 128 else self.max_batch // 2
        plan = getattr(self, "sub_batches", None)        # explicit sub-batch sizes of one max_batch (tuning / tests)
        spans, lo = [], 0
        while lo < n:
            left = n - lo
            if plan:
                k, acc = 0, 0
                while k < len(plan) - 1 and acc + plan[k] <= lo % self.max_batch:
                    acc += plan[k]
                    k += 1
                size = min(int(plan[k]), big)
            elif self.max_batch >= 128 and lo % self.max_batch == 0 and left > big // 2:
                size = big // 4                      # first sub-batch of a batch: short, its upload is exposed
            elif self.max_batch >= 128 and lo % self.max_batch == big // 4:
                size = big - big // 4
            else:
                size = big
            spans.append((lo, min(n, lo + size)))
            lo = spans[-1][1]
        shape = (big,) + tuple(tiles.shape[1:])
        for i in range(2):
            if self._stage[i] is None or self._stage[i].shape != shape or self._stage[i].dtype != tiles.dtype:
                self._stage[i] = torch.empty(shape, dtype=tiles.dtype, device=self.device)
        copied = [torch.cuda.Event() for _ in spans]
        computed = [torch.cuda.Event() for _ in spans]
        fetched = [None] * len(spans)

        def upload(i):
            lo, hi = spans[i]
            with torch.cuda.stream(self._copy_stream):
                if i >= 2:
                    self._copy_stream.wait_event(computed[i - 2])    # the compute that read this buffer has finished
                self._stage[i % 2][:hi - lo].copy_(tiles[lo:hi], non_blocking=True)
                copied[i].record(self._copy_stream)

        def fetch(i):            # D2H of slot i % 2 on its own stream, after the compute of sub-batch i
            lo, hi = spans[i]
            with torch.cuda.stream(self._d2h_stream):
                self._d2h_stream.wait_event(computed[i])
                self._counts_h[i % 2][:hi - lo].copy_(self._out_counts[i % 2][:hi - lo], non_blocking=True)
                self._rows_h[i % 2][:hi - lo].copy_(self._out_rows[i % 2][:hi - lo], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self._d2h_stream)
                fetched[i] = ev

        def collect(i, out):
            lo, hi = spans[i]
            fetched[i].synchronize()
            counts = self._counts_h[i % 2][:hi - lo].tolist()
            if min(counts) < 0:
                self.check()                         # raises the capacity error with its message
            rows_h = self._rows_h[i % 2]
            for b, k in enumerate(counts):
                out.append(rows_h[b, :k].clone() if k > 0 else None)

        out = []
        self.d2h_bytes = sum((hi - lo) * (4 + self.cap * 28) for lo, hi in spans)    # counts + padded rows copied back
        upload(0)
        for i, (lo, hi) in enumerate(spans):
            if i + 1 < len(spans):
                upload(i + 1)
            main.wait_event(copied[i])
            if i >= 2:
                main.wait_event(fetched[i - 2])      # result slot i % 2 has been read back
            self.run_device(self._stage[i % 2][:hi - lo], original_shape, slot=i % 2)
            computed[i].record(main)
            fetch(i)
            if i >= 1:
                collect(i - 1, out)
        collect(len(spans) - 1, out)
        self.check()
        return out
