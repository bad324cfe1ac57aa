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
