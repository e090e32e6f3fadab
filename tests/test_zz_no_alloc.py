"""ay_nms_static: the allocation-free variant of ay_nms (scratch in a file-scope ``static __device__`` array,
csrc/nms.cu).  It has no counterpart in the reference; it exists for callers that cannot call cudaMalloc.  Called
through the C ABI and checked bit for bit against ay_nms on the same decoded rows, on two calls in a row with
DIFFERENT inputs (the array keeps the first call's contents, the second call must not see them).

Collected last on purpose (test_zz_*).
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _decoded(batch, n_rows, n_classes, seed, frac):
    """Decoded YOLO rows (xywh, conf, class scores) with clustered boxes, so that NMS has clusters to merge; about
    `frac` of the rows pass conf 0.5."""
    g = np.random.default_rng(seed)
    centres = g.uniform(40, 376, size=(batch, 24, 2)).astype(np.float32)
    pick = g.integers(0, 24, size=(batch, n_rows))
    xy = np.take_along_axis(centres, pick[..., None].repeat(2, -1), axis=1) + g.normal(0, 6, size=(batch, n_rows, 2))
    wh = g.uniform(12, 90, size=(batch, n_rows, 2))
    conf = np.where(g.random((batch, n_rows)) < frac, g.uniform(0.5, 1.0, (batch, n_rows)), g.uniform(0, 0.5, (batch, n_rows)))
    cls = g.random((batch, n_rows, n_classes))
    return np.concatenate([xy, wh, conf[..., None], cls], axis=-1).astype(np.float32)


def _outputs(B, cap, dev):
    return (torch.full((B, cap, 7), float("nan"), dtype=torch.float32, device=dev),
            torch.full((B, cap), -7, dtype=torch.int32, device=dev), torch.full((B,), -9, dtype=torch.int32, device=dev))


def _run(lib, ctx, pred, n_classes, conf, thr, cap, st, static, outs=None):
    from amyloid_yolo_paper_b200 import _lib
    B, N, _ = pred.shape
    rows, keep, counts = outs if outs is not None else _outputs(B, cap, pred.device)
    P = lambda t: ctypes.c_void_p(t.data_ptr())                     # noqa: E731
    if static:
        _lib.check(lib.ay_nms_static(P(pred), B, N, n_classes, conf, thr, P(rows), P(keep), P(counts), cap, st))
    else:
        _lib.check(lib.ay_nms(ctx, P(pred), B, N, n_classes, conf, thr, P(rows), P(keep), P(counts), cap, st))
    torch.cuda.synchronize(pred.device)
    return rows, keep, counts


def _same(a, b):
    (ra, ka, ca), (rb, kb, cb) = a, b
    assert torch.equal(ca, cb), (ca.tolist(), cb.tolist())
    for t in range(ca.numel()):
        k = int(ca[t])
        assert k >= 0
        assert torch.equal(ka[t, :k], kb[t, :k])
        assert torch.equal(ra[t, :k].view(torch.int32), rb[t, :k].view(torch.int32))      # bit-exact, as every NMS test
    return int(ca.sum())


def test_ay_nms_static_matches_ay_nms_twice_in_a_row():
    from amyloid_yolo_paper_b200 import _lib
    dev = torch.device("cuda", 0)
    lib, ctx = _lib.load(), _lib.context(0)
    room = int(lib.ay_nms_static_capacity())
    assert 16 << 20 <= room < 32 << 20
    with torch.cuda.device(dev):
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        total = 0
        # warm-up on a small problem: the first use loads the module (which is when the driver backs the static array)
        # and the kernels, so that the memory check below sees only what a call itself takes
        w_in = torch.from_numpy(_decoded(1, 300, 2, seed=1, frac=0.3)).to(dev)
        _same(_run(lib, ctx, w_in.clone(), 2, 0.5, 0.4, 300, st, True), _run(lib, ctx, w_in.clone(), 2, 0.5, 0.4, 300, st, False))
        # call 1: the full yolov3 416^2 shape at the static capacity (2 tiles); call 2: a smaller, different problem
        # whose carve-up overlaps call 1's leftovers; call 3: call 1's shape again with other data
        for seed, (B, N, C, frac) in enumerate([(2, 10647, 2, 0.08), (5, 2535, 80, 0.3), (2, 10647, 2, 0.02)]):
            host = _decoded(B, N, C, seed=100 + seed, frac=frac)
            a_in = torch.from_numpy(host).to(dev)
            b_in = a_in.clone()
            outs = _outputs(B, N, dev)
            torch.cuda.synchronize(dev)
            free0 = torch.cuda.mem_get_info(dev)[0]
            got = _run(lib, ctx, a_in, C, 0.5, 0.4, N, st, static=True, outs=outs)
            free1 = torch.cuda.mem_get_info(dev)[0]
            want = _run(lib, ctx, b_in, C, 0.5, 0.4, N, st, static=False)
            total += _same(got, want)
            assert torch.equal(a_in.view(torch.int32), b_in.view(torch.int32))            # both rewrote xywh -> xyxy in place
            assert free1 >= free0 - (1 << 20), "ay_nms_static took %d bytes of device memory" % (free0 - free1)
        assert total > 100
        # overflow of `cap`: no context to ay_check, the tile reports -1
        host = _decoded(1, 2535, 2, seed=7, frac=0.5)
        _, _, counts = _run(lib, ctx, torch.from_numpy(host).to(dev), 2, 0.5, 0.4, 4, st, static=True)
        assert counts.tolist() == [-1]
        # and the next call is unaffected by it
        a_in = torch.from_numpy(host).to(dev)
        b_in = a_in.clone()
        _same(_run(lib, ctx, a_in, 2, 0.5, 0.4, 2535, st, True), _run(lib, ctx, b_in, 2, 0.5, 0.4, 2535, st, False))
        lib.ay_check(ctx, st)                                                             # clear the context's flag if any


def test_ay_nms_static_rejects_what_does_not_fit():
    from amyloid_yolo_paper_b200 import _lib
    dev = torch.device("cuda", 0)
    lib = _lib.load()
    with torch.cuda.device(dev):
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        pred = torch.zeros((3, 10647, 7), dtype=torch.float32, device=dev)
        out = torch.zeros((3, 8, 7), dtype=torch.float32, device=dev)
        keep = torch.zeros((3, 8), dtype=torch.int32, device=dev)
        cnt = torch.full((3,), 5, dtype=torch.int32, device=dev)
        P = lambda t: ctypes.c_void_p(t.data_ptr())                 # noqa: E731
        rc = lib.ay_nms_static(P(pred), 3, 10647, 2, 0.5, 0.4, P(out), P(keep), P(cnt), 8, st)
        assert rc == -5 and b"static array holds" in lib.ay_last_error()            # AY_ECAPACITY, nothing launched
        rc = lib.ay_nms_static(P(pred), 1 << 20, 16384, 2, 0.5, 0.4, P(out), P(keep), P(cnt), 8, st)
        assert rc == -5
        rc = lib.ay_nms_static(P(pred), 1, 20000, 2, 0.5, 0.4, P(out), P(keep), P(cnt), 8, st)
        assert rc == -1                                                             # AY_EINVAL: beyond the on-chip sort
        rc = lib.ay_nms_static(None, 1, 100, 2, 0.5, 0.4, P(out), P(keep), P(cnt), 8, st)
        assert rc == -1
        torch.cuda.synchronize(dev)
        assert cnt.tolist() == [5, 5, 5]
