"""per C-ABI-call timing of one training step (batch 32, 512x512, side stream off): where the step goes, by entry point AND
by resolution of the tensor the call works on.  usage: python tools/train_probe.py [--calls] [--aspp]"""
import collections
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("UBR_WGRAD_STREAM", "0")
os.environ.setdefault("UBR_BRANCH_MAX_RES", "0")
from ubresnet_b200 import _lib, synthetic                                   # noqa: E402
from ubresnet_b200.models import UResNet                                    # noqa: E402
from ubresnet_b200.models.ASPP_ResNet import ASPP_ResNet                    # noqa: E402
from ubresnet_b200.training.pixelwise_nllloss import PixelWiseNLLLoss       # noqa: E402

torch.manual_seed(1)
m = (ASPP_ResNet(3, 1, 16, showsizes=False) if "--aspp" in sys.argv else UResNet(3, 1, 16)).cuda().train()
img, lab, wgt = [t.cuda() for t in synthetic.make_batch(32, 512, 512)]
crit = PixelWiseNLLLoss()


def step():
    for p in m.parameters():
        p.grad = None
    loss = crit(m(img), lab, wgt)
    loss.backward()


for _ in range(3):
    step()
torch.cuda.synchronize()
prof = _lib.enable_profiling()
prof.records.clear()
step()
torch.cuda.synchronize()
rows = []
for name, e0, e1, kern, summ in prof.records:
    views = [a for a in summ if isinstance(a, tuple) and len(a) == 4 and a[1] > 0]
    res = max((v[1] for v in views), default=0)
    rows.append((name, e0.elapsed_time(e1), kern, res, views[:2]))
_lib.disable_profiling()
tot = sum(r[1] for r in rows)
print("train step: %d calls, %.3f ms serialised" % (len(rows), tot))
by = collections.OrderedDict()
for name, ms, kern, res, _ in rows:
    e = by.setdefault((res, name), [0, 0.0])
    e[0] += 1
    e[1] += ms
levels = sorted({k[0] for k in by}, reverse=True)
for lv in levels:
    sub = [(k[1], v) for k, v in by.items() if k[0] == lv]
    print("-- resolution %4d: %7.3f ms (%4.1f %%), %d calls" % (lv, sum(v[1] for _, v in sub), 100 * sum(v[1] for _, v in sub) / tot,
                                                                 sum(v[0] for _, v in sub)))
    for name, v in sorted(sub, key=lambda kv: -kv[1][1]):
        print("     %-30s n=%3d %7.3f ms" % (name, v[0], v[1]))
if "--calls" in sys.argv:
    for name, ms, kern, res, views in rows:
        print("%-30s %7.3f ms  %-28s %s" % (name, ms, kern[:28], views))
