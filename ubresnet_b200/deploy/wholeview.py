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
        while ntiles * crop - extent < (ntiles - 1) * min(16, crop // 4):      # keep >= 16 px of overlap per seam
            ntiles += 1
    span = extent - crop
    return [int(math.floor(i * span / (ntiles - 1) + 0.5)) for i in range(ntiles)]


def tile_plan(rows=1008, cols=3456, crop=512):
    return [(r, c) for r in tile_starts(rows, crop) for c in tile_starts(cols, crop)]


_ORIGINS = {}


def _origins(plan, device):
    """int32 [ntiles][2] device table of a tile plan, cached per (plan, device): building it with torch.tensor(...,
    device=cuda) is a SYNCHRONOUS host-to-device copy, six of which per event serialised the deploy loop with the host"""
    key = (tuple(map(tuple, plan)), str(device))
    t = _ORIGINS.get(key)
    if t is None:
        if len(_ORIGINS) > 256:
            _ORIGINS.clear()
        t = _ORIGINS[key] = torch.tensor(plan, dtype=torch.int32).to(device).contiguous()
    return t


def gather(plane, plan, crop=512):
    """plane (rows, cols) fp32 CUDA -> (ntiles, 1, crop, crop); crop may be (crop_h, crop_w)."""
    return ops.tile_gather(plane.contiguous(), _origins(plan, plane.device), crop)


def stitch(tiles, plan, rows, cols):
    """tiles (ntiles, C, crop, crop) fp32 CUDA -> (C, rows, cols); highest-index covering tile wins."""
    return ops.stitch_scatter(tiles.contiguous(), _origins(plan, tiles.device), rows, cols)


@torch.no_grad()
def run_event(model, planes, crop=512, batch=None):
    """planes (P, rows, cols) fp32 (CPU or CUDA).  Returns (P, C, rows, cols) log-prob maps on the model's device.
    The crops of ALL planes of the event go through the network together (`batch` = crops per forward call; None = the
    whole event, 42 crops for 3 planes of 1008x3456).  Multi-GPU: run_events(rank=, world=) shards events,
    run_event_sharded shards the crops of one event."""
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


def stitch_classes(class_tiles, plan, rows, cols, out=None):
    """class_tiles (ntiles, crop, crop) uint8 CUDA -> (rows, cols) uint8; highest-index covering tile wins."""
    return ops.stitch_scatter_u8(class_tiles.contiguous(), _origins(plan, class_tiles.device), rows, cols, out=out)


@torch.no_grad()
def run_event_classes(model, planes, crop=512, batch=None, out=None):
    """run_event reduced to what the deploy script keeps (run_ubresnet_wholeview.py:264-272 writes the per-pixel
    winner): (P, rows, cols) uint8 class maps.  The argmax is taken per crop (1 byte per pixel) and the TILES are
    stitched -- bit-identical to argmax(run_event(...), 1) because stitching copies whole pixels -- so the
    (P, C, rows, cols) fp32 map (125 MB per event) is never materialised."""
    dev = next(model.parameters()).device
    planes = planes.to(dev, non_blocking=True).float()
    nplanes, rows, cols = planes.shape
    plan = tile_plan(rows, cols, crop)
    crops = torch.cat([gather(planes[p], plan, crop) for p in range(nplanes)], 0)
    per_plane = len(plan)
    batch = batch or crops.shape[0]
    cls = torch.empty((crops.shape[0], crop, crop), dtype=torch.uint8, device=dev)
    for i in range(0, crops.shape[0], batch):
        ops.argmax_classes(model(crops[i:i + batch]), out=cls[i:i + batch])
    if out is None:
        out = torch.empty((nplanes, rows, cols), dtype=torch.uint8, device=dev)
    for p in range(nplanes):
        stitch_classes(cls[p * per_plane:(p + 1) * per_plane], plan, rows, cols, out=out[p])
    return out


@torch.no_grad()
def run_event_classes_3d(model, planes, box=(512, 832), covered_z=310, batch=None):
    """Whole-view deploy on the 3D-consistent crop sets of deploy/run_ubresnet_wholeview.py:33-47 (512 x 832 boxes, one
    per plane and z slab, `crop_geometry.split_detector`): planes (3, rows, cols) in U, V, Y order -> (3, rows, cols)
    uint8 class maps.  Pixels no box covers (U/V wires beyond the planes' 2400) stay 0."""
    from . import crop_geometry
    dev = next(model.parameters()).device
    planes = planes.to(dev, non_blocking=True).float()
    nplanes, rows, cols = planes.shape
    plans = crop_geometry.plane_plans(crop_geometry.split_detector(rows, cols, box[0], box[1], covered_z))[:nplanes]
    crops = torch.cat([gather(planes[p], plans[p], box) for p in range(nplanes)], 0)
    per_plane = len(plans[0])
    batch = batch or crops.shape[0]
    cls = torch.empty((crops.shape[0], box[0], box[1]), dtype=torch.uint8, device=dev)
    for i in range(0, crops.shape[0], batch):
        ops.argmax_classes(model(crops[i:i + batch]), out=cls[i:i + batch])
    out = torch.empty((nplanes, rows, cols), dtype=torch.uint8, device=dev)
    for p in range(nplanes):
        stitch_classes(cls[p * per_plane:(p + 1) * per_plane], plans[p], rows, cols, out=out[p])
    return out


# ---------------------------------------------------------------------------------------------------------------
# multi-GPU deploy (kernel id C2).  Two shardings, both over torch.distributed:
#   events : event i -> rank i % world, no collective at all (throughput mode)           -> run_events(rank=, world=)
#   crops  : the 42 crops of ONE event are split into `world` contiguous blocks; every rank runs its block and the
#            1-byte class tiles are all-gathered (NCCL over NVLink: 11 MB per event) so the stitch owner -- every rank,
#            the stitch is 20 us -- holds the whole event (single-event latency mode)    -> run_event_sharded
# ---------------------------------------------------------------------------------------------------------------
def crop_shard(ncrops, rank, world):
    """(first, count, per_rank): rank's contiguous block of the event's crops; per_rank = ceil(ncrops / world) is the
    padded block size of the all-gather (the last ranks may hold fewer, or zero, real crops)."""
    per = -(-ncrops // world)
    first = min(rank * per, ncrops)
    return first, max(0, min(per, ncrops - first)), per


def gather_class_tiles(local_tiles, ncrops, group=None):
    """local_tiles (per_rank, crop, crop) uint8 (rows beyond this rank's real crops are padding) -> (ncrops, crop, crop)
    in global crop order on every rank.  One all-gather of 1-byte tiles; works on CUDA (NCCL) and CPU (gloo) tensors."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    per = local_tiles.shape[0]
    full = torch.empty((world * per,) + tuple(local_tiles.shape[1:]), dtype=local_tiles.dtype, device=local_tiles.device)
    dist.all_gather_into_tensor(full, local_tiles.contiguous(), group=group)
    return full[:ncrops]


@torch.no_grad()
def run_event_sharded(model, planes, crop=512, group=None, batch=None):
    """ONE event, its crops sharded over the ranks of `group` (every rank passes the same `planes`; host tensors are
    uploaded by each rank; every rank must hold the same model state -- a loaded checkpoint, or
    parallel.broadcast_module after training, whose BatchNorm running statistics are per rank).  Returns the
    (P, rows, cols) uint8 class maps on every rank -- bit-identical to
    run_event_classes on one GPU (the per-crop forward does not depend on which other crops share the batch in eval
    mode; `tools/dp_check.py wholeview` and bench.py check it)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return run_event_classes(model, planes, crop=crop, batch=batch)
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    dev = next(model.parameters()).device
    planes = planes.to(dev, non_blocking=True).float()
    nplanes, rows, cols = planes.shape
    plan = tile_plan(rows, cols, crop)
    per_plane = len(plan)
    ncrops = nplanes * per_plane
    first, count, per = crop_shard(ncrops, rank, world)
    local = torch.zeros((per, crop, crop), dtype=torch.uint8, device=dev)
    if count:
        parts = []
        for p in range(first // per_plane, (first + count - 1) // per_plane + 1):       # planes this block touches
            lo = max(first, p * per_plane) - p * per_plane
            hi = min(first + count, (p + 1) * per_plane) - p * per_plane
            parts.append(gather(planes[p], plan[lo:hi], crop))
        crops = parts[0] if len(parts) == 1 else torch.cat(parts, 0)
        b = batch or count
        for i in range(0, count, b):
            ops.argmax_classes(model(crops[i:i + b]), out=local[i:min(i + b, count)])
    tiles = gather_class_tiles(local, ncrops, group)
    out = torch.empty((nplanes, rows, cols), dtype=torch.uint8, device=dev)
    for p in range(nplanes):
        stitch_classes(tiles[p * per_plane:(p + 1) * per_plane], plan, rows, cols, out=out[p])
    return out


@torch.no_grad()
def run_events(model, events, crop=512, batch=None, rank=0, world=1):
    """Deploy loop over many events with the host<->device copies off the critical path.

    `events`: iterable of (P, rows, cols) fp32 HOST tensors of one shape (pinned memory makes the copies
    asynchronous).  Yields, in order, one uint8 class map (P, rows, cols) per event in pinned host memory (argmax over
    the stitched log-prob maps, what deploy/run_ubresnet_wholeview.py:264-272 writes out).  The planes of event i+1 are
    uploaded and the class map of event i-1 is downloaded on a copy stream while event i runs, through two persistent
    device input buffers and two device / pinned-host output buffers (no allocation inside the loop).  A yielded map
    stays valid until the next one is requested (copy it to keep it).

    Multi-GPU: with world > 1 this rank processes events rank, rank + world, ... (round-robin event sharding, no
    collective -- SURVEY 8e) and yields only those, in order; event i of the input is the (i // world)-th map of rank
    i % world."""
    if world > 1:
        events = (e for i, e in enumerate(events) if i % world == rank)
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
        if d2h[j] is not None:
            main.wait_event(d2h[j])                                     # the download two events ago has left d_cls[j]
        run_event_classes(model, d_in[j], crop=crop, batch=batch, out=d_cls[j])
        in_free[j] = torch.cuda.Event()
        in_free[j].record(main)
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
