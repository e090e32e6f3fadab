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
                 device=None, img_size=None):
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
        # network input size: what detect.py passes as --img_size to Darknet(...) and to Resize/rescale_boxes
        # (detect.py:59,73,129); the engine is planned for that size, any multiple of 32
        self.img_size = int(img_size if img_size is not None else getattr(model, "img_size", 0) or
                            model.hyperparams["height"])
        self.engine = None
        self._bind_engine()
        self.cap = int(min(cap, self.rows))
        B, dev = self.max_batch, self.device
        self.decoded = torch.empty((B, self.rows, self.cols), dtype=torch.float32, device=dev)
        self.nms_rows = torch.empty((B, self.cap, 7), dtype=torch.float32, device=dev)
        self.nms_keep = torch.empty((B, self.cap), dtype=torch.int32, device=dev)
        self.nms_counts = torch.empty((B,), dtype=torch.int32, device=dev)
        # candidate lists filled by the head convolutions' epilogue (conf >= conf_thres, utils/utils.py:248): NMS starts
        # from these instead of re-reading the dense decoded tensor.  A tile with more candidates than cand_cap is
        # handled by a dense scan of that tile (never truncated).
        self.cand_cap = int(min(self.rows, max(1024, 4 * self.cap)))
        self.cand_idx = torch.empty((B, self.cand_cap), dtype=torch.int32, device=dev)
        self.cand_count = torch.zeros((B,), dtype=torch.int32, device=dev)
        self.fuse_threshold = True          # False: ay_forward + ay_nms (dense scan, in-place xyxy) as in round 1
        # two result slots: the host path reads slot i while the device fills slot i+1
        self._out_rows = [torch.empty((B, self.cap, 7), dtype=torch.float32, device=dev) for _ in range(2)]
        self._out_counts = [torch.empty((B,), dtype=torch.int32, device=dev) for _ in range(2)]
        self.out_rows, self.out_counts = self._out_rows[0], self._out_counts[0]
        self.launches_per_batch = 0

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _bind_engine(self):
        """(Re)fetch the model's engine.  ``Darknet.engine()`` replaces its engine when the weights, the
        precision or the batch capacity change; this detector keeps a reference to the engine object it
        runs on (never a bare ``ay_ctx`` pointer), so a replaced engine stays alive until this call has
        switched over and nothing ever runs on a destroyed context."""
        eng = self.model.engine(device=self.device, max_batch=self.max_batch, img_size=self.img_size)
        if eng is not self.engine:
            if self.engine is not None and (eng.rows, eng.cols) != (self.rows, self.cols):
                raise RuntimeError("the model's output shape changed under a live TileDetector")
            self.engine = eng
            self.rows, self.cols = eng.rows, eng.cols
            self.lib = eng.lib
        return eng

    @property
    def ctx(self):
        return self.engine.ctx              # NMS / merge scratch lives in the same context as the network

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
        self._bind_engine()
        lib, ctx = self.lib, self.ctx
        with torch.cuda.device(self.device):
            if self.fuse_threshold and self.engine.precision != "fp32":
                _lib.check(lib.ay_forward_candidates(ctx, _ptr(tiles), kind, B, sh, sw, _ptr(self.decoded), self.conf_thres,
                                                     _ptr(self.cand_idx), _ptr(self.cand_count), self.cand_cap, st))
                _lib.check(lib.ay_nms_candidates(ctx, _ptr(self.decoded), _ptr(self.cand_idx), _ptr(self.cand_count),
                                                 self.cand_cap, B, self.rows, self.cols - 5, self.conf_thres, self.nms_thres,
                                                 _ptr(self.nms_rows), _ptr(self.nms_keep), _ptr(self.nms_counts), self.cap, st))
            else:
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
        """Host input, one call: the stream below over a single batch."""
        return next(self.detect_stream([tiles], original_shape))

    def _plan_spans(self, n, stream_head=True):
        """Sub-batches of one host batch of n tiles: (lo, hi) spans.  At the head of a stream a short first sub-batch of
        every max_batch keeps the exposed part of the first upload short; later batches of a stream upload under the
        previous batch's compute and use full-size sub-batches only."""
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
                size = min(int(plan[k]), self.max_batch)
            elif not stream_head:
                size = self.max_batch                # steady state of a stream: whole batches, fewest waves wasted
            elif stream_head and self.max_batch >= 128 and lo % self.max_batch == 0 and left > big // 2:
                size = big // 4
            elif stream_head and self.max_batch >= 128 and lo % self.max_batch == big // 4:
                size = big - big // 4
            else:
                size = big
            spans.append((lo, min(n, lo + size)))
            lo = spans[-1][1]
        return spans

    def detect_stream(self, batches, original_shape=None):
        """``batches``: iterable of HOST tensors (ideally pinned; uint8 [B,H,H,3] or fp32 [B,3,S,S], any B each) - what
        ``detect.py:88-95`` pulls from its DataLoader - or of ``(tensor, (h, w))`` pairs when the original image shape
        differs per batch (``utils.datasets.TileLoader``).  Yields, per batch and in order, the list ``detect()`` returns.

        One software pipeline runs across the whole iterable: sub-batches are uploaded on a copy stream into two device
        buffers while the previous one computes, results come back on a third stream from two result slots and are
        unpacked on the host one sub-batch behind the compute - so the upload of batch k+1 and the read-back of batch k-1
        overlap the compute of batch k, and only the first upload and the last read-back of the stream are exposed."""
        main = torch.cuda.current_stream(self.device)
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._d2h_stream = torch.cuda.Stream(device=self.device)
            self._stage = [None, None]
            self._counts_h = [torch.empty((self.max_batch,), dtype=torch.int32).pin_memory() for _ in range(2)]
            self._rows_h = [torch.empty((self.max_batch, self.cap, 7), dtype=torch.float32).pin_memory() for _ in range(2)]
        self.d2h_bytes = 0

        def work():                                   # (batch index, tiles, lo, hi, last span of its batch)
            for bi, item in enumerate(batches):
                tiles, shape = item if isinstance(item, tuple) else (item, original_shape)
                if tiles.is_cuda:
                    raise ValueError("detect_stream takes host batches; use detect() for device tensors")
                spans = self._plan_spans(tiles.shape[0], stream_head=(bi == 0))
                for k, (lo, hi) in enumerate(spans):
                    yield bi, tiles, lo, hi, k == len(spans) - 1, shape
                if not spans:
                    yield bi, tiles, 0, 0, True, shape

        copied, computed, fetched = {}, {}, {}

        def upload(i, item):
            _, tiles, lo, hi, _, _ = item
            shape = (self.max_batch,) + tuple(tiles.shape[1:])
            buf = self._stage[i % 2]
            if buf is None or buf.shape != shape or buf.dtype != tiles.dtype:
                torch.cuda.synchronize(self.device)          # (first use or a new tile shape: rare)
                buf = self._stage[i % 2] = torch.empty(shape, dtype=tiles.dtype, device=self.device)
            with torch.cuda.stream(self._copy_stream):
                if i - 2 in computed:
                    self._copy_stream.wait_event(computed[i - 2])    # the compute that read this buffer has finished
                if hi > lo:
                    buf[:hi - lo].copy_(tiles[lo:hi], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self._copy_stream)
                copied[i] = ev

        def fetch(i, n):         # D2H of slot i % 2 on its own stream, after the compute of span i
            with torch.cuda.stream(self._d2h_stream):
                self._d2h_stream.wait_event(computed[i])
                self._counts_h[i % 2][:n].copy_(self._out_counts[i % 2][:n], non_blocking=True)
                self._rows_h[i % 2][:n].copy_(self._out_rows[i % 2][:n], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self._d2h_stream)
                fetched[i] = ev
            self.d2h_bytes += n * (4 + self.cap * 28)        # counts + padded rows copied back

        def collect(i, n, out):
            fetched[i].synchronize()
            counts = self._counts_h[i % 2][:n].tolist()
            if counts and min(counts) < 0:
                self.check()                         # raises the capacity error with its message
            rows_h = self._rows_h[i % 2]
            for b, k in enumerate(counts):
                out.append(rows_h[b, :k].clone() if k > 0 else None)
            for d in (copied, computed, fetched):
                d.pop(i - 2, None)

        it = work()
        cur = next(it, None)
        if cur is not None:
            upload(0, cur)
        i, prev, out = 0, None, []
        while cur is not None:
            nxt = next(it, None)
            if nxt is not None:
                upload(i + 1, nxt)
            n = cur[3] - cur[2]
            main.wait_event(copied[i])
            if i - 2 in fetched:
                main.wait_event(fetched[i - 2])      # result slot i % 2 has been read back
            if n > 0:
                self.run_device(self._stage[i % 2][:n], cur[5], slot=i % 2)
            ev = torch.cuda.Event()
            ev.record(main)
            computed[i] = ev
            fetch(i, n)
            if prev is not None:
                collect(i - 1, prev[3] - prev[2], out)
                if prev[4]:
                    yield out
                    out = []
            prev, cur, i = cur, nxt, i + 1
        if prev is not None:
            collect(i - 1, prev[3] - prev[2], out)
            self.check()
            yield out
