"""Whole-view deploy: tile 1008x3456 wire planes into overlapping 512x512 crops, run the net, stitch back.

Reference semantics (deploy/run_ubresnet_wholeview.py:174-188,218-242,264-272): planes are dense (rows, cols)
float32 arrays; a crop is the numpy slice [r0:r0+512, c0:c0+512] (:238-239); results are written back at the same
offsets with overwrite semantics (larcv Image2D::overlay kOverWrite, Image2D.cxx:362-367).  The reference's crop
boxes come from `larcv.UBSplitDetector`, which is not in the tree; the tile plan here is builder-defined
(SURVEY.md 8c): unit-pixel meta, evenly spread origins, row-major tile order, later tiles overwrite earlier ones.
Everything between the host planes and the stitched class maps stays on the device; crops (or whole events) are
the unit of multi-GPU sharding -- no collective is needed when events are sharded.
"""
import math

import torch

from .. import ops


def tile_starts(extent, crop, ntiles=None):
    if extent <= crop:
        return [0]
    if ntiles is None:
        ntiles = -(-extent // crop)
        while ntiles * crop - extent < (ntiles - 1) * 16:      # keep >= 16 px of overlap per seam
            ntiles += 1
    span = extent - crop
    return [int(math.floor(i * span / (ntiles - 1) + 0.5)) for i in range(ntiles)]


def tile_plan(rows=1008, cols=3456, crop=512):
    return [(r, c) for r in tile_starts(rows, crop) for c in tile_starts(cols, crop)]


def _origins(plan, device):
    return torch.tensor(plan, dtype=torch.int32, device=device).contiguous()


def gather(plane, plan, crop=512):
    """plane (rows, cols) fp32 CUDA -> (ntiles, 1, crop, crop)."""
    return ops.tile_gather(plane.contiguous(), _origins(plan, plane.device), crop)


def stitch(tiles, plan, rows, cols):
    """tiles (ntiles, C, crop, crop) fp32 CUDA -> (C, rows, cols); highest-index covering tile wins."""
    return ops.stitch_scatter(tiles.contiguous(), _origins(plan, tiles.device), rows, cols)


@torch.no_grad()
def run_event(model, planes, crop=512, batch=None, rank=0, world=1):
    """planes (P, rows, cols) fp32 (CPU or CUDA).  Returns (P, C, rows, cols) log-prob maps on the model's device.
    The crops of ALL planes of the event go through the network together (`batch` = crops per forward call; None = the
    whole event, 42 crops for 3 planes of 1008x3456).  With world > 1 the caller shards EVENTS round-robin
    (event i -> rank i % world); this function is per event."""
    dev = next(model.parameters()).device
    planes = planes.to(dev, non_blocking=True).float()
    nplanes, rows, cols = planes.shape
    plan = tile_plan(rows, cols, crop)
    crops = torch.cat([gather(planes[p], plan, crop) for p in range(nplanes)], 0)
    per_plane = crops.shape[0] // nplanes
    batch = batch or crops.shape[0]
    res = [model(crops[i:i + batch]) for i in range(0, crops.shape[0], batch)]
    res = res[0] if len(res) == 1 else torch.cat(res, 0)
    return torch.stack([stitch(res[p * per_plane:(p + 1) * per_plane], plan, rows, cols) for p in range(nplanes)], 0)


@torch.no_grad()
def run_events(model, events, crop=512, batch=None):
    """Deploy loop over many events with the host<->device copies off the critical path.

    `events`: iterable of (P, rows, cols) fp32 HOST tensors of one shape (pinned memory makes the copies
    asynchronous).  Yields, in order, one uint8 class map (P, rows, cols) per event in pinned host memory (argmax over
    the stitched log-prob maps, what deploy/run_ubresnet_wholeview.py:264-272 writes out).  The planes of event i+1 are
    uploaded and the class map of event i-1 is downloaded on a copy stream while event i runs, through two persistent
    device input buffers and two device / pinned-host output buffers (no allocation inside the loop).  A yielded map
    stays valid until the next one is requested (copy it to keep it)."""
    dev = next(model.parameters()).device
    main = torch.cuda.current_stream(dev)
    copy = torch.cuda.Stream(device=dev)
    it = iter(events)
    try:
        first = next(it)
    except StopIteration:
        return
    nplanes, rows, cols = first.shape
    d_in = [torch.empty((nplanes, rows, cols), dtype=torch.float32, device=dev) for _ in range(2)]
    d_cls = [torch.empty((nplanes, rows, cols), dtype=torch.uint8, device=dev) for _ in range(2)]
    h_cls = [torch.empty((nplanes, rows, cols), dtype=torch.uint8).pin_memory() for _ in range(2)]
    in_free = [None, None]             # main-stream event: the last forward that read d_in[j] has consumed it
    d2h = [None, None]                 # copy-stream event: h_cls[j] / d_cls[j] hold a finished download

    def upload(planes, j):
        if tuple(planes.shape) != (nplanes, rows, cols):
            raise RuntimeError("run_events: every event must have the shape of the first one")
        with torch.cuda.stream(copy):
            if in_free[j] is not None:
                copy.wait_event(in_free[j])
            d_in[j].copy_(planes, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy)
        return ev

    up = upload(first, 0)
    pending = None                     # buffer index of the previous event
    i = 0
    while up is not None:
        j = i & 1
        main.wait_event(up)
        try:
            nxt = next(it)
        except StopIteration:
            nxt = None
        up = upload(nxt, j ^ 1) if nxt is not None else None            # overlaps with this event's forward pass
        maps = run_event(model, d_in[j], crop=crop, batch=batch)
        in_free[j] = torch.cuda.Event()
        in_free[j].record(main)
        if d2h[j] is not None:
            main.wait_event(d2h[j])                                     # the download two events ago has left d_cls[j]
        d_cls[j].copy_(maps.argmax(1))
        done = torch.cuda.Event()
        done.record(main)
        with torch.cuda.stream(copy):
            copy.wait_event(done)
            h_cls[j].copy_(d_cls[j], non_blocking=True)
            d2h[j] = torch.cuda.Event()
            d2h[j].record(copy)
        if pending is not None:
            d2h[pending].synchronize()
            yield h_cls[pending]
        pending = j
        i += 1
    d2h[pending].synchronize()
    yield h_cls[pending]
