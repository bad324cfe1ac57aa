"""Timing of the halo-resident conv kernel's epilogue modes on one layer shape (default: the 16->16 3x3 layer at 512^2,
batch 32): plain forward, forward + fused BatchNorm statistics/finalize, data gradient, accumulating data gradient, data
gradient + fused BatchNorm-backward reduction.  usage: python tools/halo_probe.py [cin cout k res] [--once]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ubresnet_b200 import ops  # noqa: E402
from ubresnet_b200.ops import Act  # noqa: E402

DEV = "cuda:0"
args = [a for a in sys.argv[1:] if not a.startswith("--")]
cin, cout, k, res = (int(a) for a in args) if len(args) == 4 else (16, 16, 3, 512)
once = "--once" in sys.argv
B = 32


def timeit(fn, warm=3, reps=10):
    if once:
        fn()
        torch.cuda.synchronize()
        return 0.0
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


x = Act(torch.randn(B, res, res, cin, device=DEV).to(torch.bfloat16))
dy = Act(torch.randn(B, res, res, cout, device=DEV).to(torch.bfloat16))
w = torch.randn(cout, cin, k, k, device=DEV) * 0.05
wf, wd = ops.pack_weights(w, False, True), ops.pack_weights(w, False, False)
y = Act.empty(B, res, res, cout, DEV)
dx = Act.zeros(B, res, res, cin, DEV)
ws = ops.bn_workspace(64, DEV)
gamma, beta = torch.ones(cout, device=DEV), torch.zeros(cout, device=DEV)
gi, bi = torch.ones(cin, device=DEV), torch.zeros(cin, device=DEV)
ypre = Act(torch.randn(B, res, res, cin, device=DEV).to(torch.bfloat16))
st = ops.bn_train_stats(ypre, gi, bi, None, None, ws)
byt = 2.0 * B * res * res * (cin + cout)
rows = [
    ("fwd", lambda: ops.conv2d_fwd(x, wf, None, cout, k, 1, y=y), byt),
    ("fwd + bn stats", lambda: ops.conv2d_fwd_bn(x, wf, None, cout, k, 1, gamma, beta, None, None, ws, y=y), byt),
    ("dgrad", lambda: ops.conv2d_dgrad(dy, wd, cin, k, 1, dx=dx), byt),
    ("dgrad acc", lambda: ops.conv2d_dgrad_acc(dy, wd, cin, k, 1, dx), byt + 2.0 * B * res * res * cin),
    ("dgrad + bn reduce", lambda: ops.conv2d_dgrad_bn_bwd(dy, wd, cin, k, 1, ypre, st, True, ws), byt + 6.0 * B * res * res * cin),
    ("bn_act", lambda: ops.bn_act(ypre, st, True, out=dx), 4.0 * B * res * res * cin),
]
print("%dx%d conv %d->%d at %d^2, batch %d" % (k, k, cin, cout, res, B))
for name, fn, nbytes in rows:
    t = timeit(fn)
    print("%-20s %7.3f ms  %6.0f GB/s (all bytes the call moves)" % (name, t, nbytes / max(t, 1e-9) / 1e6))
