"""Whole-slide sweep: tiles sharded over the GPUs of one box, detections gathered.

The reference walks a slide's tiles one by one in a single process
(``validation.py:80-135``, ``prospective.py:52-93``).  Every tile is independent
through forward, NMS, rescale and merge, so the sweep shards by tile: rank ``r`` of
``R`` owns the contiguous block ``shard_range(n_tiles, r, R)``; the only exchange is
the final gather of per-tile detection lists —

    all_gather(counts int32 [tiles_per_rank])  +  all_gather(rows fp32 [tiles_per_rank, kmax, 7])

over NCCL (NVLink/NVSwitch) on the GPU box, gloo in the CPU tests.  No activation or
weight traffic; each rank holds a full weight replica.
"""
import itertools
import os
import sys
import time

import torch
import torch.distributed as dist


def shard_range(n_tiles, rank, world):
    """Contiguous block of tile ids owned by ``rank`` (sizes differ by at most one)."""
    base, extra = divmod(int(n_tiles), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def pack_detections(per_tile, kmax=None, device=None):
    """list of ``[k,7]`` tensors / None -> (rows [T,kmax,7] fp32 zero padded, counts [T] int32).
    Vectorised (one concatenation and one scatter): a sweep packs tens of thousands of tiles."""
    ks = [0 if d is None else int(d.shape[0]) for d in per_tile]
    counts = torch.tensor(ks, dtype=torch.int32)
    kmax = int(kmax if kmax is not None else (max(ks) if ks else 0))
    rows = torch.zeros((len(per_tile), max(kmax, 1), 7), dtype=torch.float32)
    # (detect_stream hands over fp32 CPU tensors: no per-tile conversion then - this loop runs once per tile of the sweep)
    live = [d if (d.dtype == torch.float32 and not d.is_cuda) else d.to(torch.float32).cpu()
            for d in per_tile if d is not None and d.shape[0]]
    if live:
        flat = torch.cat(live, 0)
        c64 = counts.to(torch.int64)
        tile = torch.repeat_interleave(torch.arange(len(per_tile)), c64)
        start = torch.cumsum(c64, 0) - c64
        pos = torch.arange(flat.shape[0]) - torch.repeat_interleave(start, c64)
        rows[tile, pos] = flat
    if device is not None:
        rows, counts = rows.to(device), counts.to(device)
    return rows, counts


def unpack_detections(rows, counts):
    rows, counts = rows.cpu(), counts.cpu().tolist()
    return [rows[t, :k].clone() if k > 0 else None for t, k in enumerate(counts)]


def allgather_detections(rows, counts, n_tiles, group=None):
    """Gather every rank's ``(rows [t_r,k_r,7], counts [t_r])`` (tensors on the backend's
    device: CUDA for NCCL, CPU for gloo) into the full sweep on every rank.

    -> (rows [n_tiles, kmax, 7], counts [n_tiles]) in global tile order."""
    if not (dist.is_available() and dist.is_initialized()):
        return rows, counts
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lo, hi = shard_range(n_tiles, rank, world)
    if counts.shape[0] != hi - lo:
        raise ValueError("rank %d owns %d tiles but passed %d" % (rank, hi - lo, counts.shape[0]))
    dev = rows.device
    per_rank = -(-int(n_tiles) // world)                     # ceil: all_gather needs equal shapes
    # agree on the row padding: one tiny all_reduce(MAX)
    kmax = torch.tensor([rows.shape[1] if rows.shape[0] else 1], dtype=torch.int32, device=dev)
    dist.all_reduce(kmax, op=dist.ReduceOp.MAX, group=group)
    k = int(kmax.item())
    send_counts = torch.zeros((per_rank,), dtype=torch.int32, device=dev)
    send_counts[:hi - lo] = counts
    send_rows = torch.zeros((per_rank, k, 7), dtype=torch.float32, device=dev)
    if hi > lo:
        send_rows[:hi - lo, :rows.shape[1]] = rows
    all_counts = torch.empty((world * per_rank,), dtype=torch.int32, device=dev)
    all_rows = torch.empty((world * per_rank, k, 7), dtype=torch.float32, device=dev)
    dist.all_gather_into_tensor(all_counts, send_counts, group=group)
    dist.all_gather_into_tensor(all_rows, send_rows, group=group)
    if per_rank * world == int(n_tiles):                     # even shards: nothing to drop
        return all_rows, all_counts
    # drop the padding tiles of the ranks that own one tile fewer
    sel = []
    for r in range(world):
        rlo, rhi = shard_range(n_tiles, r, world)
        sel.extend(range(r * per_rank, r * per_rank + (rhi - rlo)))
    sel = torch.tensor(sel, dtype=torch.long, device=dev)
    return all_rows.index_select(0, sel), all_counts.index_select(0, sel)


def warm_gather(n_tiles, kmax, device, group=None):
    """One exchange of the shape a sweep over ``n_tiles`` tiles with lists of up to ``kmax`` rows will make, on zeros.
    NCCL sets up the connections of each protocol the first time a message size selects it, and the caching allocator
    has to get the gather buffers from the driver once: tens of milliseconds at 8 ranks, which a production sweep of
    10^5 tiles never notices but a 10-step measurement does.  Call it once before timing anything."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) < 2:
        return
    lo, hi = shard_range(n_tiles, dist.get_rank(group), dist.get_world_size(group))
    rows = torch.zeros((hi - lo, max(int(kmax), 1), 7), dtype=torch.float32, device=device)
    counts = torch.zeros((hi - lo,), dtype=torch.int32, device=device)
    allgather_detections(rows, counts, n_tiles, group)
    if torch.device(device).type == "cuda":
        torch.cuda.synchronize(device)


def _phase_timer(device):
    """AY_SWEEP_TIMING=1: wall-clock of the sweep's phases on stderr (synchronises at every phase boundary)."""
    if not os.environ.get("AY_SWEEP_TIMING"):
        return lambda name: None
    state = {"t": time.perf_counter()}

    def mark(name):
        if torch.device(device).type == "cuda":
            torch.cuda.synchronize(device)
        now = time.perf_counter()
        rank = dist.get_rank() if dist.is_initialized() else 0
        if rank == 0:
            sys.stderr.write("[sweep] %-10s %8.2f ms\n" % (name, 1e3 * (now - state["t"])))
        state["t"] = now
    return mark


def sweep(detector, tile_source, n_tiles, batch, group=None, gather=True):
    """Run ``detector`` (a :class:`TileDetector`) over this rank's shard of ``n_tiles``.

    ``tile_source(lo, hi)`` returns the tiles ``[lo, hi)`` as a uint8 ``[hi-lo,H,H,3]`` tensor
    (host or device).  -> (rows, counts) of the whole sweep if ``gather`` else of the shard.

    Device tiles stay on the device end to end: every batch's result rows are appended to one shard buffer on the
    compute stream (no host round trip per batch), the row padding is trimmed to the shard's longest list with a
    single read-back at the end, and the gather moves ``[tiles, kmax, 7]`` once per sweep (config C4)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    lo, hi = shard_range(n_tiles, rank, world)
    nccl = dist.is_initialized() and dist.get_backend(group) == "nccl"
    mark = _phase_timer(detector.device)
    batches = (tile_source(b0, min(hi, b0 + batch)) for b0 in range(lo, hi, batch))
    first = next(batches, None)
    if first is not None and first.is_cuda:
        cap = detector.cap
        rows = torch.zeros((hi - lo, cap, 7), dtype=torch.float32, device=detector.device)
        counts = torch.zeros((hi - lo,), dtype=torch.int32, device=detector.device)
        at = 0
        for tiles in itertools.chain([first], batches):
            n = tiles.shape[0]
            r, c = detector.run_device(tiles)
            rows[at:at + n].copy_(r)
            counts[at:at + n].copy_(c)
            at += n
        detector.check()                                   # capacity overflows / device faults surface here
        mark("compute")
        kmax = max(int(counts.max().item()), 1) if hi > lo else 1
        rows = rows[:, :kmax].contiguous()
        # rows beyond a tile's count hold whatever the previous occupant of the slot left: zero them so that the
        # gathered block is a function of the detections only
        valid = torch.arange(kmax, device=rows.device).view(1, -1) < counts.view(-1, 1)
        rows = torch.where(valid.unsqueeze(-1), rows, torch.zeros((), dtype=rows.dtype, device=rows.device))
        if not nccl and dist.is_initialized():
            rows, counts = rows.cpu(), counts.cpu()
    else:
        per_tile = []
        if first is not None:
            # host tiles: one software pipeline over the whole shard (upload of batch k+1 under the compute of batch k)
            for outs in detector.detect_stream(itertools.chain([first], batches)):
                per_tile.extend(outs)
        rows, counts = pack_detections(per_tile, device=detector.device if nccl else None)
    mark("pack")
    if gather and world > 1:
        out = allgather_detections(rows, counts, n_tiles, group)
        mark("gather")
        return out
    return rows, counts
