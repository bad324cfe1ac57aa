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
def run_event(model, planes, crop=512, batch=14, rank=0, world=1):
    """planes (P, rows, cols) fp32 (CPU or CUDA).  Returns (P, C, rows, cols) log-prob maps on the model's device.
    With world > 1 the caller shards EVENTS round-robin (event i -> rank i % world); this function is per event."""
    dev = next(model.parameters()).device
    planes = planes.to(dev, non_blocking=True).float()
    nplanes, rows, cols = planes.shape
    plan = tile_plan(rows, cols, crop)
    outs = []
    for p in range(nplanes):
        crops = gather(planes[p], plan, crop)
        res = []
        for i in range(0, crops.shape[0], batch):
            res.append(model(crops[i:i + batch]))
        outs.append(stitch(torch.cat(res, 0), plan, rows, cols))
    return torch.stack(outs, 0)
