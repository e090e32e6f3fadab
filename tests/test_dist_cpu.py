"""World-size-2 gloo test of the sweep's only exchange step (detection allgather)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from amyloid_yolo_paper_b200.sweep import (allgather_detections, pack_detections, shard_range, unpack_detections)


def _tile_detections(t):
    """Deterministic fake detections of global tile t (k differs per tile, some tiles empty)."""
    k = (t * 7) % 5
    if k == 0:
        return None
    g = torch.Generator().manual_seed(t)
    return torch.rand((k, 7), generator=g)


def _worker(rank, world, port, n_tiles, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from amyloid_yolo_paper_b200.sweep import warm_gather
        warm_gather(n_tiles, 5, "cpu")                  # the untimed exchange of the sweep's shape (bench warm-up): no side effects
        lo, hi = shard_range(n_tiles, rank, world)
        rows, counts = pack_detections([_tile_detections(t) for t in range(lo, hi)])
        all_rows, all_counts = allgather_detections(rows, counts, n_tiles)
        torch.save((all_rows, all_counts), os.path.join(out_dir, "r%d.pt" % rank))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_allgather_detections_gloo_world2(tmp_path):
    for n_tiles in (9, 4):
        mp.spawn(_worker, args=(2, _free_port(), n_tiles, str(tmp_path)), nprocs=2, join=True)
        want = [_tile_detections(t) for t in range(n_tiles)]
        for r in range(2):
            rows, counts = torch.load(os.path.join(str(tmp_path), "r%d.pt" % r))
            got = unpack_detections(rows, counts)
            assert len(got) == n_tiles
            for a, b in zip(got, want):
                assert (a is None) == (b is None)
                if a is not None:
                    assert torch.equal(a, b)


def test_pack_unpack_roundtrip():
    per_tile = [_tile_detections(t) for t in range(12)]
    rows, counts = pack_detections(per_tile)
    assert rows.shape[0] == 12 and rows.shape[2] == 7 and counts.dtype == torch.int32
    back = unpack_detections(rows, counts)
    for a, b in zip(back, per_tile):
        assert (a is None) == (b is None) and (a is None or torch.equal(a, b))
    rows, counts = pack_detections([None, None])
    assert counts.tolist() == [0, 0] and unpack_detections(rows, counts) == [None, None]


class _FakeDetector:
    """Stands in for TileDetector on the CPU: 'detects' from the tile's first byte (= its global id)."""
    device = torch.device("cpu")
    cap = 8

    def detect_stream(self, batches, original_shape=None):
        for tiles in batches:
            yield [_tile_detections(int(t[0, 0, 0])) for t in tiles]


def _sweep_worker(rank, world, port, n_tiles, out_dir):
    from amyloid_yolo_paper_b200.sweep import sweep
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        def source(lo, hi):                       # host tiles whose first byte is the global tile id
            t = torch.zeros((hi - lo, 2, 2, 3), dtype=torch.uint8)
            t[:, 0, 0, 0] = torch.arange(lo, hi, dtype=torch.uint8)
            return t
        rows, counts = sweep(_FakeDetector(), source, n_tiles, batch=3)
        torch.save((rows, counts), os.path.join(out_dir, "s%d.pt" % rank))
    finally:
        dist.destroy_process_group()


def test_sweep_shards_and_gathers_gloo_world2(tmp_path):
    """sweep(): contiguous shards, ragged last batch, uneven shard sizes (11 tiles over 2 ranks), gather in tile order."""
    n_tiles = 11
    mp.spawn(_sweep_worker, args=(2, _free_port(), n_tiles, str(tmp_path)), nprocs=2, join=True)
    want = [_tile_detections(t) for t in range(n_tiles)]
    for r in range(2):
        rows, counts = torch.load(os.path.join(str(tmp_path), "s%d.pt" % r))
        got = unpack_detections(rows, counts)
        assert len(got) == n_tiles
        for a, b in zip(got, want):
            assert (a is None) == (b is None) and (a is None or torch.equal(a, b))
