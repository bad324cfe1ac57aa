"""Run-to-run determinism of a training step (forward + loss + backward) under allocator perturbation.

Every reduction in the kernels has a fixed order and the side streams only change WHEN a kernel runs, so repeated steps
on the same weights and batch must give bit-identical losses and gradients; a lifetime bug (a side-stream kernel reading
memory the allocator already handed out again) shows up here as a difference.

  python tools/determinism_check.py [uresnet|aspp] [reps]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ubresnet_b200 import synthetic  # noqa: E402
from ubresnet_b200.models import UResNet  # noqa: E402
from ubresnet_b200.models.ASPP_ResNet import ASPP_ResNet  # noqa: E402
from ubresnet_b200.training.pixelwise_nllloss import PixelWiseNLLLoss  # noqa: E402


def main():
    arch = sys.argv[1] if len(sys.argv) > 1 else "uresnet"
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    dev = torch.device("cuda", 0)
    torch.manual_seed(5)
    model = (ASPP_ResNet(3, 1, 16, showsizes=False) if arch == "aspp" else UResNet(3, 1, 16)).to(dev).train()
    img, lab, wgt = (t.to(dev) for t in synthetic.make_batch(4, 256, 256))
    crit = PixelWiseNLLLoss()
    g = torch.Generator().manual_seed(1)
    ref = None
    worst = 0.0
    junk = []
    for r in range(reps):
        # perturb the caching allocator: hold / release blocks of random sizes between steps
        junk = [torch.empty(int(torch.randint(1, 64, (1,), generator=g)) << 20, dtype=torch.uint8, device=dev)
                for _ in range(int(torch.randint(0, 6, (1,), generator=g)))]
        if r % 3 == 2:
            torch.cuda.empty_cache()
        model.zero_grad(set_to_none=True)
        loss = crit(model(img), lab, wgt)
        loss.backward()
        torch.cuda.synchronize()
        cur = [loss.detach().clone()] + [p.grad.detach().clone() for p in model.parameters()]
        if ref is None:
            ref = cur
        else:
            for a, b in zip(ref, cur):
                if not torch.equal(a, b):
                    worst = max(worst, (a - b).abs().max().item() / (a.abs().max().item() or 1.0))
    print("determinism %s: %d repeated steps, worst relative difference %.3e %s" % (arch, reps, worst, "OK" if worst == 0 else "FAIL"))
    sys.exit(0 if worst == 0 else 1)


if __name__ == "__main__":
    main()
