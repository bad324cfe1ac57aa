#!/usr/bin/env python
"""bench.py -- 512x512 crops/sec of the ub_uresnet training step (BASELINE.json metric) on N B200s.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU under torchrun)
  python bench.py --impl reference --gpus N --steps K ...   # the reference algorithm on the host CPU cores

A "step" = forward + PixelWiseNLLLoss + backward + SGD update (+ gradient all-reduce when N > 1) over one batch of
32 synthetic 512x512 single-channel crops per GPU (configs[1] of BASELINE.json; weak scaling).
`value`  : whole-job crops/s, inputs resident in HBM when the timed region starts.
`e2e`    : same step through the public nn.Module API with the batch copied host->device from pinned memory and
           the loss read back device->host inside the timed region, every step.
`roofline`: the dominant kernel class (tcgen05 implicit-GEMM conv), algorithmic conv FLOPs / its CUDA-event time.
`cpu_baseline`: the oracle port (CPU fp32 restatement of the reference) timed on the box's host cores (N=1 only).
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "train_crops_per_sec_512x512"
UNIT = "crops/s"
# algorithmic conv FLOPs per crop @512x512, UResNet(3,1,16) (SURVEY.md 8d / BASELINE.md 3)
GF_FWD = 66.606
GF_STEM = 0.411
GF_TRAIN = 199.41


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as fh:
            p = json.load(fh)
        return dict(hbm_gbs=p.get("hbm_gbs", 6650.0), tf_burst=p.get("bf16_tflops", 1590.0),
                    tf_sustained=p.get("bf16_tflops_sustained", 1400.0), source="measured")
    return dict(hbm_gbs=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


class ClockSampler(threading.Thread):
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.reasons = set()
        self._stop_evt = threading.Event()

    def run(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                self.samples.append((float(f[0]), float(f[1])))
                for nm, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(nm)
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=5)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": sorted(self.reasons)}
        sm = sorted(s[0] for s in self.samples)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(s[1] for s in self.samples), "reasons": sorted(self.reasons)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port (reference algorithm) on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_train_rate(batch, size, steps, warmup, threads=None):
    from oracle import uresnet_port as port
    from ubresnet_b200 import synthetic
    if threads:
        torch.set_num_threads(threads)
    sd = port.init_state_dict(3, 1, 16, seed=1)
    params = port.clone_state_dict(sd, requires_grad=True)
    leaves = [v for v in params.values() if v.requires_grad]
    opt = torch.optim.SGD(leaves, lr=1e-3, momentum=0.9, weight_decay=1e-3)
    img, lab, wgt = synthetic.make_batch(batch, size, size)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        logp = port.forward(params, img, training=True)
        loss = port.pixelwise_nll(logp, lab, wgt)
        loss.backward()
        opt.step()
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
    mean = sum(times) / len(times)
    return batch / mean, mean * 1e3, torch.get_num_threads()


def run_reference(args, rank):
    if rank != 0:
        return
    batch = args.cpu_batch
    rate, ms, threads = cpu_train_rate(batch, args.size, max(1, args.steps), max(0, min(args.warmup, 1)))
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "ub_uresnet(inplanes=16,1ch,3cls) train step fwd+weighted-NLL+bwd+SGD, 512x512 crops",
                   "per_gpu_batch": args.batch, "l2": "working set >> 126 MB L2 (no flush needed)"},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "oracle/uresnet_port.py (CPU fp32 restatement of the reference), batch %d of %dx%d "
                                   "crops per step, %d timed steps" % (batch, args.size, args.size, max(1, args.steps))},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# CUDA arm
# ------------------------------------------------------------------------------------------------
def run_own(args, rank, world, local_rank):
    import torch.distributed as dist
    from ubresnet_b200 import _lib, synthetic
    from ubresnet_b200.models import UResNet
    from ubresnet_b200.parallel import DistributedUResNet
    from ubresnet_b200.training.pixelwise_nllloss import PixelWiseNLLLoss

    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; the ubresnet_b200 path has no CPU fallback")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    lib = _lib.load()
    _lib.check(lib.ubr_device_check(), "device_check")

    torch.manual_seed(1234)
    model = UResNet(num_classes=3, input_channels=1, inplanes=16).to(dev).train()
    net = DistributedUResNet(model) if world > 1 else model
    crit = PixelWiseNLLLoss()
    opt = torch.optim.SGD(model.parameters(), lr=1e-4, momentum=0.9, weight_decay=1e-3, foreach=True)

    B, S = args.batch, args.size
    nbuf = 2
    host = []
    for i in range(nbuf):
        img, lab, wgt = synthetic.make_batch(B, S, S, rank=rank, step=i)
        host.append((img.pin_memory(), lab.pin_memory(), wgt.pin_memory()))
    resident = [(a.to(dev), b.to(dev), c.to(dev)) for a, b, c in host]
    h2d_bytes = sum(t.numel() * t.element_size() for t in host[0])

    def step_resident(i):
        img, lab, wgt = resident[i % nbuf]
        opt.zero_grad(set_to_none=True)
        loss = crit(net(img), lab, wgt)
        loss.backward()
        opt.step()
        return loss

    def step_e2e(i):
        img, lab, wgt = host[i % nbuf]
        img = img.to(dev, non_blocking=True)
        lab = lab.to(dev, non_blocking=True)
        wgt = wgt.to(dev, non_blocking=True)
        opt.zero_grad(set_to_none=True)
        loss = crit(net(img), lab, wgt)
        loss.backward()
        opt.step()
        return float(loss.item())            # device -> host read of the step's result

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = lib.ubr_launch_count()
        e0.record()
        for i in range(steps):
            fn(warmup + i)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        launches = lib.ubr_launch_count() - n0
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    ms, launches = timed(step_resident, args.steps, args.warmup)
    clocks = sampler.stop() if sampler else None
    ms_e2e, _ = timed(step_e2e, args.steps, 1)
    value = world * B * args.steps / (ms * 1e-3)
    e2e = world * B * args.steps / (ms_e2e * 1e-3)

    # ---- live per-kernel-class timing (CUDA events around every C-ABI call, separate short pass)
    roof = None
    kernels = None
    # every rank runs the extra steps (they contain the gradient all-reduce); only rank 0 records events
    nprof = 2
    prof = None
    if rank == 0:
        prof = _lib.enable_profiling()
        prof.records.clear()
    for i in range(nprof):
        step_resident(i)
    barrier()
    if rank == 0:
        summ = prof.summary()
        _lib.disable_profiling()
        tot = sum(t for _, t in summ.values())
        gemm_fns = ("ubr_conv2d_fwd", "ubr_conv2d_dgrad", "ubr_deconv4s2_fwd", "ubr_deconv4s2_dgrad")
        wg_fns = ("ubr_conv2d_wgrad", "ubr_deconv4s2_wgrad")
        t_fd = sum(summ[f][1] for f in gemm_fns if f in summ) / nprof
        t_wg = sum(summ[f][1] for f in wg_fns if f in summ) / nprof
        t_gemm = t_fd + t_wg
        n_gemm = sum(summ[f][0] for f in gemm_fns + wg_fns if f in summ) / nprof
        peaks = load_peaks()
        # algorithmic conv FLOPs of one step: fwd + dgrad + wgrad of every tensor-core conv (stem runs on CUDA cores)
        gf_gemm = 3.0 * (GF_FWD - GF_STEM) * B
        gf_wg = (GF_FWD - GF_STEM) * B
        ach = gf_gemm / t_gemm                               # GFLOP / ms = TFLOP/s
        roof = {"bound": "tensor",
                "kernel": "tcgen05 conv kernels: conv_halo_kernel / conv_gemm_kernel (fwd, dgrad) + wgrad_halo_kernel / "
                          "wgrad_gemm_kernel, %d C-ABI conv calls per step" % n_gemm,
                "achieved": ach, "peak": peaks["tf_sustained"], "unit": "TFLOP/s", "frac": ach / peaks["tf_sustained"],
                "traffic": None, "peak_source": peaks["source"] + " (bf16 sustained)",
                "share_of_step": t_gemm * nprof / tot if tot else None,
                "note": "41% of these FLOPs sit in C<=32 layers (AI 72-144 FLOP/B < ridge 213) and are HBM-bound; "
                        "see profiles/ per-layer tables for their GB/s against the measured HBM peak",
                "fwd_dgrad_tflops": 2.0 * (GF_FWD - GF_STEM) * B / t_fd if t_fd else None}
        kernels = {k: {"calls_per_step": c / nprof, "ms_per_step": t / nprof} for k, (c, t) in
                   sorted(summ.items(), key=lambda kv: -kv[1][1])}
        kernels["_wgrad_gemm_tflops"] = gf_wg / t_wg if t_wg else None

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        rate, cms, threads = cpu_train_rate(args.cpu_batch, S, 2, 1)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": "oracle/uresnet_port.py train step, batch %d of %dx%d crops, best-effort mean of 2 steps after 1 "
                         "warm-up (%.0f ms/step)" % (args.cpu_batch, S, S, cms)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "ub_uresnet(inplanes=16,1ch,3cls) train step fwd+weighted-NLL+bwd+SGD, 512x512 crops",
                       "per_gpu_batch": B, "global_batch": B * world, "crop": S,
                       "parallelism": "dp%d" % world if world > 1 else "single",
                       "l2": "per-step working set (>8 GB) >> 126 MB L2; no flush needed"},
            "clocks": clocks,
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes * world, "d2h_bytes_per_step": 4 * world,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches),
            "utilisation": {"conv_tflops_whole_step": GF_TRAIN * B * world * args.steps / (ms * 1e-3) / 1e3 / world,
                            "note": "algorithmic conv FLOPs (199.41 GF/crop) / step time, per GPU"},
        }
        if roof:
            line["roofline"] = roof
        if cpu:
            line["cpu_baseline"] = cpu
        if kernels and args.verbose:
            line["kernels"] = kernels
        print(json.dumps(line), flush=True)
        if kernels and args.kernel_log:
            with open(args.kernel_log, "w") as fh:
                json.dump(kernels, fh, indent=1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--batch", type=int, default=32, help="crops per GPU per step")
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--cpu-batch", type=int, default=4, help="crops per CPU step (bounded sample)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--kernel-log", default="")
    args = ap.parse_args()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_own(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
