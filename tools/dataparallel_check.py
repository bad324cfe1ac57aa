"""`nn.DataParallel(model)` (the reference's only multi-GPU code, training/train_ubresnet2018_wlarcv2.py:99,103) over this
repo's UResNet on 2+ GPUs of one process: eval outputs equal the single-GPU forward bit for bit, training gradients equal
the sum of the per-replica gradients computed by hand.   usage: python tools/dataparallel_check.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ubresnet_b200 import synthetic  # noqa: E402
from ubresnet_b200.models import UResNet  # noqa: E402
from ubresnet_b200.training.pixelwise_nllloss import PixelWiseNLLLoss  # noqa: E402


def main():
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        print("dataparallel_check: needs >= 2 GPUs, found %d" % ngpu)
        return 0
    torch.manual_seed(3)
    model = UResNet(3, 1, 16).cuda(0)
    dp = torch.nn.DataParallel(model, device_ids=[0, 1])
    img, lab, wgt = (t.cuda(0) for t in synthetic.make_batch(4, 128, 128))
    crit = PixelWiseNLLLoss()
    bad = 0
    # eval: batch-independent, so the scattered forward must equal the single-GPU one exactly
    model.eval()
    with torch.no_grad():
        ref = model(img)
        out = dp(img)
    ok = torch.equal(ref, out)
    print("eval: DataParallel output == single-GPU output:", ok)
    bad += int(not ok)
    # train: replicas see 2 crops each (BatchNorm statistics per replica, like the reference); the gathered log-probs
    # go through the loss on device 0 and the gradients are reduce-added onto the original parameters
    model.train()
    model.zero_grad(set_to_none=True)
    loss = crit(dp(img), lab, wgt)
    loss.backward()
    got = {n: p.grad.detach().clone() for n, p in model.named_parameters()}
    model.zero_grad(set_to_none=True)
    exp = None
    for h in range(2):                       # by hand: each half through the model on GPU 0, loss normalised over all 4
        sl = slice(2 * h, 2 * h + 2)
        lp = model(img[sl])
        # PixelWiseNLLLoss = mean over the pixels of ITS input: half the pixels -> weight 1/2 in the full-batch mean
        (0.5 * crit(lp, lab[sl], wgt[sl])).backward()
    exp = {n: p.grad.detach().clone() for n, p in model.named_parameters()}
    worst = 0.0
    for n in got:
        den = exp[n].abs().max().item() or 1.0
        worst = max(worst, (got[n] - exp[n]).abs().max().item() / den)
    print("train: worst relative gradient difference DataParallel vs by-hand replicas: %.3e" % worst)
    bad += int(not worst < 1e-5)
    print("dataparallel_check", "OK" if bad == 0 else "FAIL")
    return bad


if __name__ == "__main__":
    sys.exit(main())
