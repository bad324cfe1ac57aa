"""Per-layer timing of the conv kernels at the BASELINE shapes (batch 32, 512x512 crop): fwd / dgrad / wgrad.

  python tools/layer_bench.py [--batch 32] [--only 512]       (UBR_NO_HALO=1 / UBR_NO_HALO_WGRAD=1 for the old kernels)
Prints ms, algorithmic GB/s (read x + write y, bf16) and TFLOP/s per layer; CUDA events, 3 warm-up + 5 timed.
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ubresnet_b200 import ops  # noqa: E402
from ubresnet_b200.ops import Act  # noqa: E402

DEV = "cuda:0"

# (name, cin, cout, k, stride, input resolution)
LAYERS = [
    ("dec1.res1.conv1", 32, 16, 3, 1, 512), ("dec1.conv3x3", 16, 16, 3, 1, 512), ("dec1.bypass", 32, 16, 1, 1, 512),
    ("conv10", 16, 16, 7, 1, 512), ("conv11", 16, 3, 7, 1, 512),
    ("enc1.res1.conv1", 16, 32, 3, 1, 256), ("enc1.conv3x3", 32, 32, 3, 1, 256), ("dec2.res1.conv1", 64, 32, 3, 1, 256),
    ("dec2.bypass", 64, 32, 1, 1, 256), ("enc2.res1.conv1", 32, 64, 3, 2, 256),
    ("enc2.conv3x3", 64, 64, 3, 1, 128), ("dec3.res1.conv1", 128, 64, 3, 1, 128), ("enc3.res1.conv1", 64, 128, 3, 2, 128),
    ("enc3.conv3x3", 128, 128, 3, 1, 64), ("dec4.res1.conv1", 256, 128, 3, 1, 64),
    ("enc4.conv3x3", 256, 256, 3, 1, 32), ("dec5.res1.conv1", 512, 256, 3, 1, 32),
    ("enc5.conv3x3", 512, 512, 3, 1, 16),
]


def timeit(fn, warm=3, reps=5):
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--only", type=int, default=0)
    ap.add_argument("--name", default="", help="substring filter on the layer name")
    args = ap.parse_args()
    B = args.batch
    # a big scratch write between runs is unnecessary: every tensor here is >> L2 at the high resolutions
    print("%-18s %4s %4s k s  res |  fwd ms   GB/s   TF/s | dgrad ms   GB/s | wgrad ms   GB/s" % ("layer", "cin", "cout"))
    tot = [0.0, 0.0, 0.0]
    for name, cin, cout, k, s, res in LAYERS:
        if (args.only and res != args.only) or (args.name and args.name not in name):
            continue
        ro = res // s
        cop = (cout + 15) // 16 * 16
        x = Act(torch.randn(B, res, res, cin, device=DEV).to(torch.bfloat16))
        dy = Act(torch.randn(B, ro, ro, cop, device=DEV).to(torch.bfloat16))
        w = torch.randn(cout, cin, k, k, device=DEV) * 0.05
        wf = ops.pack_weights(w, False, True)
        wd = ops.pack_weights(w, False, False)
        y = Act.empty(B, ro, ro, cop, DEV)
        dx = Act.empty(B, res, res, cin, DEV)
        dw = torch.empty_like(w)
        ws = torch.empty(ops.wgrad_workspace_bytes(B, ro, ro, cop, cin, k * k) // 4 + 64, dtype=torch.float32, device=DEV)
        if cout % 16 == 0:
            t_f = timeit(lambda: ops.conv2d_fwd(x, wf, None, cout, k, s, y=y))
        else:
            lg = torch.empty(B * ro * ro, 4, dtype=torch.float32, device=DEV)
            t_f = timeit(lambda: ops.conv2d_fwd(x, wf, None, cout, k, s, y_f32=lg))
        t_d = timeit(lambda: ops.conv2d_dgrad(dy, wd, cin, k, s, dx=dx))
        t_w = timeit(lambda: ops.conv2d_wgrad(x, dy, cout, k, s, dw, False, ws))
        byt = 2.0 * B * (res * res * cin + ro * ro * cop)
        fl = 2.0 * B * ro * ro * cout * cin * k * k
        print("%-18s %4d %4d %d %d %4d | %7.3f %6.0f %6.1f | %7.3f %6.0f | %7.3f %6.0f" %
              (name, cin, cout, k, s, res, t_f, byt / t_f / 1e6, fl / t_f / 1e9, t_d, byt / t_d / 1e6, t_w, byt / t_w / 1e6),
              flush=True)
        tot[0] += t_f; tot[1] += t_d; tot[2] += t_w
        del x, dy, y, dx, ws
    # ConvTranspose2d(k4,s2,p1) layers: (name, cin, cout, input res)
    for name, cin, cout, res in [("dec1.deconv", 32, 16, 256), ("dec2.deconv", 64, 32, 128), ("dec3.deconv", 128, 64, 64),
                                 ("dec4.deconv", 256, 128, 32), ("dec5.deconv", 512, 256, 16)]:
        if (args.only and res != args.only) or (args.name and args.name not in name):
            continue
        x = Act(torch.randn(B, res, res, cin, device=DEV).to(torch.bfloat16))
        dy = Act(torch.randn(B, 2 * res, 2 * res, cout, device=DEV).to(torch.bfloat16))
        w = torch.randn(cin, cout, 4, 4, device=DEV) * 0.05
        wf = ops.pack_weights(w, True, True)
        wd = ops.pack_weights(w, True, False)
        y = Act.empty(B, 2 * res, 2 * res, cout, DEV)
        dx = Act.empty(B, res, res, cin, DEV)
        dw = torch.empty_like(w)
        ws = torch.empty(ops.wgrad_workspace_bytes(B, res, res, cin, cout, 16) // 4 + 64, dtype=torch.float32, device=DEV)
        w3 = ops.pack_weights_subpixel(w, True)
        t_f = timeit(lambda: ops.deconv_fwd_subpixel(x, w3, cout, y) or ops.deconv_fwd(x, wf, cout, y=y))
        t_d = timeit(lambda: ops.deconv_dgrad(dy, wd, cin, dx=dx))
        t_w = timeit(lambda: ops.deconv_wgrad(x, dy, dw, False, ws))
        byt = 2.0 * B * (res * res * cin + 4 * res * res * cout)
        fl = 2.0 * B * res * res * cout * cin * 16
        print("%-18s %4d %4d 4 T %4d | %7.3f %6.0f %6.1f | %7.3f %6.0f | %7.3f %6.0f" %
              (name, cin, cout, res, t_f, byt / t_f / 1e6, fl / t_f / 1e9, t_d, byt / t_d / 1e6, t_w, byt / t_w / 1e6), flush=True)
        del x, dy, y, dx, ws
    if not args.name or args.name in "conv1.stem":
        # stem: fp32 1-channel input -> 16 bf16 channels (7x7); bytes = read x (4 B/px) + write y or read dy (32 B/px)
        res = 512
        x = torch.randn(B, 1, res, res, device=DEV) * 30
        w = torch.randn(16, 1, 7, 7, device=DEV) * 0.05
        b = torch.zeros(16, device=DEV)
        y = Act(torch.empty(B, res, res, 16, device=DEV, dtype=torch.bfloat16))
        dy = Act(torch.randn(B, res, res, 16, device=DEV).to(torch.bfloat16))
        dw, db = torch.empty_like(w), torch.empty_like(b)
        ws = torch.empty(ops.stem_wgrad_workspace_bytes(B, res, res, 16) // 4, dtype=torch.float32, device=DEV)
        t_f = timeit(lambda: ops.stem_conv_fwd(x, w, b, y=y))
        t_w = timeit(lambda: ops.stem_conv_wgrad(x, dy, dw, db, False, ws))
        byt = 36.0 * B * res * res
        print("%-18s %4d %4d 7 1 %4d | %7.3f %6.0f %6.1f |       -      - | %7.3f %6.0f" %
              ("conv1.stem", 1, 16, res, t_f, byt / t_f / 1e6, 2.0 * B * res * res * 16 * 49 / t_f / 1e9, t_w,
               byt / t_w / 1e6), flush=True)
    print("sum (one of each listed layer): fwd %.3f dgrad %.3f wgrad %.3f ms" % tuple(tot))


if __name__ == "__main__":
    main()
