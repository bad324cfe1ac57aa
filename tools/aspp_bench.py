"""ASPP_ResNet (BASELINE.json configs[4]) training-step and eval-forward throughput on one B200 (bf16 storage).
usage: python tools/aspp_bench.py [batch] [size] [steps]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from ubresnet_b200 import _lib, synthetic  # noqa: E402
from ubresnet_b200.models import ASPP_ResNet  # noqa: E402
from ubresnet_b200.training.pixelwise_nllloss import PixelWiseNLLLoss  # noqa: E402

GF_TRAIN, GF_FWD = 310.5, 103.65           # algorithmic conv GFLOP per 512x512 crop


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    size = int(sys.argv[2]) if len(sys.argv) > 2 else 512
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 8
    dev = torch.device("cuda:0")
    lib = _lib.load()
    torch.manual_seed(5)
    model = ASPP_ResNet(3, 1, 16, showsizes=False).to(dev).train()
    crit = PixelWiseNLLLoss()
    opt = torch.optim.SGD(model.parameters(), lr=1e-4, momentum=0.9, weight_decay=1e-3, foreach=True)
    img, lab, wgt = (t.to(dev) for t in synthetic.make_batch(batch, size, size))

    def train():
        opt.zero_grad(set_to_none=True)
        crit(model(img), lab, wgt).backward()
        opt.step()

    def timed(fn, n, warm=3):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = lib.ubr_launch_count()
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n, (lib.ubr_launch_count() - n0) // n

    ms_t, l_t = timed(train, steps)
    model.eval()
    with torch.no_grad():
        ms_i, l_i = timed(lambda: model(img), steps)
    scale = (size / 512.0) ** 2
    print(json.dumps({"model": "ASPP_ResNet(3,1,16)", "batch": batch, "size": size,
                      "train_crops_per_sec": batch / (ms_t * 1e-3), "train_ms_per_step": ms_t, "train_launches": l_t,
                      "train_conv_tflops": GF_TRAIN * scale * batch / ms_t,
                      "infer_crops_per_sec": batch / (ms_i * 1e-3), "infer_ms_per_step": ms_i, "infer_launches": l_i,
                      "infer_conv_tflops": GF_FWD * scale * batch / ms_i,
                      "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}))


if __name__ == "__main__":
    main()
