"""per C-ABI-call timing of one eval forward (batch 32, 512x512): where the inference time goes"""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ubresnet_b200 import _lib, synthetic
from ubresnet_b200.models import UResNet
torch.manual_seed(1)
m = UResNet(3, 1, 16).cuda().eval()
x = synthetic.make_batch(32, 512, 512)[0].cuda()
with torch.no_grad():
    for _ in range(3):
        m(x)
    torch.cuda.synchronize()
    prof = _lib.enable_profiling()
    prof.records.clear()
    m(x)
    torch.cuda.synchronize()
    rows = [(name, e0.elapsed_time(e1), kern, summ) for name, e0, e1, kern, summ in prof.records]
    _lib.disable_profiling()
tot = sum(r[1] for r in rows)
print("eval forward: %d calls, %.3f ms serialised" % (len(rows), tot))
for name, ms, kern, summ in rows:
    views = [a for a in summ if isinstance(a, tuple)]
    print("%-28s %7.3f ms  %-28s %s" % (name, ms, kern[:28], views[:2]))
