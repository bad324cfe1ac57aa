"""Data-parallel gradient check (run under torchrun, one rank per GPU):
the bucketed, overlapped NCCL all-reduce of DistributedUResNet must equal the average of the ranks' stand-alone
gradients (same kernels, BatchNorm statistics per rank like the reference's nn.DataParallel replicas).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/dp_check.py [aspp]
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from ubresnet_b200 import synthetic  # noqa: E402
from ubresnet_b200.models import UResNet  # noqa: E402
from ubresnet_b200.parallel import DistributedUResNet  # noqa: E402
from ubresnet_b200.training.pixelwise_nllloss import PixelWiseNLLLoss  # noqa: E402


def wholeview_check(rank, world, dev):
    """crop-sharded whole-view deploy == single-GPU deploy, bit for bit; event sharding covers every event once"""
    from ubresnet_b200.deploy import wholeview as wv
    torch.manual_seed(7)
    model = UResNet(3, 1, 16).to(dev).eval()
    bad = 0
    for rows, cols, crop in ((256, 448, 128), (1008, 3456, 512)):
        planes = synthetic.make_wholeview(rows=rows, cols=cols, planes=3)
        single = wv.run_event_classes(model, planes, crop=crop)
        sharded = wv.run_event_sharded(model, planes, crop=crop)
        bad += int(not torch.equal(single, sharded))
    events = [synthetic.make_wholeview(seed=50 + i, rows=256, cols=448, planes=2).pin_memory() for i in range(5)]
    mine = [m.clone() for m in wv.run_events(model, events, crop=128, rank=rank, world=world)]
    refs = [wv.run_event_classes(model, e, crop=128).cpu() for i, e in enumerate(events) if i % world == rank]
    bad += int(len(mine) != len(refs)) + sum(int(not torch.equal(a, b)) for a, b in zip(mine, refs))
    t = torch.tensor([bad], device=dev)
    dist.all_reduce(t)
    if rank == 0:
        print("dp_check wholeview world=%d mismatches %d %s" % (world, int(t.item()), "OK" if t.item() == 0 else "FAIL"))
    return int(t.item())


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(99)
    arch = sys.argv[1] if len(sys.argv) > 1 else "uresnet"
    if arch == "wholeview":
        bad = wholeview_check(rank, world, dev)
        dist.destroy_process_group()
        sys.exit(0 if bad == 0 else 1)
    if arch == "aspp":
        from ubresnet_b200.models.ASPP_ResNet import ASPP_ResNet
        model = ASPP_ResNet(3, 1, 16, showsizes=False).to(dev).train()
    else:
        model = UResNet(3, 1, 16).to(dev).train()
    img, lab, wgt = (t.to(dev) for t in synthetic.make_batch(4, 256, 256, rank=rank))
    crit = PixelWiseNLLLoss()
    worst = 0.0
    for trial in range(3):
        # stand-alone gradients, averaged with a plain all-reduce
        model.engine().grad_sink = None
        model.zero_grad(set_to_none=True)
        crit(model(img), lab, wgt).backward()
        expect = {}
        for n, p in model.named_parameters():
            g = p.grad.detach().clone()
            dist.all_reduce(g)
            expect[n] = g / world
        # bucketed + overlapped path (BN running buffers advance, parameters do not change)
        ddp = DistributedUResNet(model)
        model.zero_grad(set_to_none=True)
        crit(ddp(img), lab, wgt).backward()
        torch.cuda.synchronize()
        for n, p in model.named_parameters():
            den = expect[n].abs().max().item() or 1.0
            worst = max(worst, (p.grad - expect[n]).abs().max().item() / den)
    t = torch.tensor([worst], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print("dp_check %s world=%d worst relative gradient difference %.3e %s" % (arch, world, t.item(), "OK" if t.item() < 1e-5 else "FAIL"))
    dist.destroy_process_group()
    sys.exit(0 if t.item() < 1e-5 else 1)


if __name__ == "__main__":
    main()
