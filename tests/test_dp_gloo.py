"""world_size-2 gloo test (CPU) of the data-parallel gradient path: flat buckets, overlapped all-reduce protocol,
averaging, weight broadcast.  The same GradBuckets object runs under NCCL on the GPUs."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret, arch="uresnet"):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ubresnet_b200 import parallel
        from ubresnet_b200.models import UResNet
        from ubresnet_b200.models.ASPP_ResNet import ASPP_ResNet
        torch.manual_seed(100 + rank)                 # different init per rank ...
        m = UResNet(3, 1, 16) if arch == "uresnet" else ASPP_ResNet(3, 1, 16, showsizes=False)
        order = parallel.STAGE_ORDER if arch == "uresnet" else parallel.ASPP_STAGE_ORDER
        nparams = 18100931 if arch == "uresnet" else 31575683
        parallel.broadcast_module(m, 0)               # ... until rank 0's weights win
        chk = torch.tensor([sum(float(p.double().sum()) for p in m.parameters())], dtype=torch.float64)
        gathered = [torch.zeros_like(chk) for _ in range(world)]
        dist.all_gather(gathered, chk)
        assert all(torch.equal(g, gathered[0]) for g in gathered)
        buckets = parallel.GradBuckets(list(m.named_parameters()))
        assert len(buckets.flat) == 4
        assert buckets.nbytes >= 4 * nparams
        # emulate a backward: every stage fills its grads with (rank+1) * index pattern, then reports done
        names = {p: n for n, p in m.named_parameters()}
        for stage in order:
            for p, n in names.items():
                if parallel.stage_of_param(n) == stage:
                    buckets.grad_buffer(p).fill_(float(rank + 1))
            buckets.stage_done(stage)
        buckets.finish()
        expect = sum(range(1, world + 1)) / world
        for p in m.parameters():
            g = buckets.grad_buffer(p)
            assert g.shape == p.shape
            assert torch.allclose(g, torch.full_like(g, expect)), names[p]
        ret[rank] = 1
    finally:
        dist.destroy_process_group()


def _run(arch):
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, ret, arch), nprocs=world, join=True)
    assert dict(ret) == {0: 1, 1: 1}


def test_grad_buckets_allreduce_mean_world2():
    _run("uresnet")


def test_grad_buckets_allreduce_mean_world2_aspp():
    """ASPP_ResNet: the ASPP skip layers' parameters ride in enc5's bucket (their backward runs before enc5's)"""
    _run("aspp")


def _wv_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import numpy as np
        from oracle import uresnet_port as port_
        from ubresnet_b200.deploy import wholeview as wv
        crop, rows, cols, nplanes = 32, 80, 144, 3
        plan = wv.tile_plan(rows, cols, crop)
        ncrops = nplanes * len(plan)
        first, count, per = wv.crop_shard(ncrops, rank, world)
        # stand-in for "forward + argmax" of crop g: a tile whose bytes encode (g, pixel)
        def tile(g):
            return ((torch.arange(crop * crop).view(crop, crop) * 7 + g * 13) % 251).to(torch.uint8)
        local = torch.zeros((per, crop, crop), dtype=torch.uint8)
        for i in range(count):
            local[i] = tile(first + i)
        tiles = wv.gather_class_tiles(local, ncrops)
        assert tiles.shape == (ncrops, crop, crop)
        for g in range(ncrops):
            assert torch.equal(tiles[g], tile(g)), g
        # the gathered tiles stitch to what the single-process numpy oracle gives
        per_plane = len(plan)
        for p in range(nplanes):
            ref = port_.stitch_scatter_np(tiles[p * per_plane:(p + 1) * per_plane].numpy()[:, None], plan, rows, cols)[0]
            seq = np.zeros((rows, cols), dtype=np.uint8)
            for i, (r0, c0) in enumerate(plan):
                seq[r0:r0 + crop, c0:c0 + crop] = tile(p * per_plane + i).numpy()
            assert (ref == seq).all()
        ret[rank] = (first, count, per)
    finally:
        dist.destroy_process_group()


def test_wholeview_crop_sharding_world2():
    """crop-sharded whole-view deploy: block partition of the event's crops, padded all-gather of the
    1-byte class tiles in global crop order (gloo here, NCCL on the GPUs)"""
    from ubresnet_b200.deploy import wholeview as wv
    for ncrops in (42, 7, 1, 48):
        for world in (1, 2, 4, 8):
            blocks = [wv.crop_shard(ncrops, r, world) for r in range(world)]
            covered = [i for f, c, _ in blocks for i in range(f, f + c)]
            assert covered == list(range(ncrops)), (ncrops, world)
            assert len({per for _, _, per in blocks}) == 1 and blocks[0][2] * world >= ncrops
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_wv_worker, args=(world, port, ret), nprocs=world, join=True)
    assert sorted(dict(ret).keys()) == [0, 1]


def test_run_events_round_robin_sharding():
    """event sharding: rank r of `world` sees events r, r+world, ... (no collective)"""
    evs = list(range(11))
    seen = []
    for r in range(4):
        mine = [e for i, e in enumerate(evs) if i % 4 == r]
        seen += mine
        assert mine == evs[r::4]
    assert sorted(seen) == evs
