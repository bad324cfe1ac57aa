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
def host_cores():
    """cores this process may use (torchrun exports OMP_NUM_THREADS=1; the CPU arm runs alone on rank 0 and takes
    every core of the box)"""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_train_rate(batch, size, steps, warmup, threads=None):
    from oracle import uresnet_port as port
    from ubresnet_b200 import synthetic
    torch.set_num_threads(threads or host_cores())
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


def library_train_rate(batch, size, steps, warmup, mode, dev):
    """The oracle port (plain torch.nn.functional ops == what the reference's nn.Modules call) on the GPU through
    PyTorch/cuDNN: the "library" comparison of SURVEY.md 8d cfg-2.  mode: "fp32" (TF32 off), "tf32", "bf16" (autocast +
    channels_last).  Baseline leg only -- never part of the product path."""
    from oracle import uresnet_port as port
    from ubresnet_b200 import synthetic
    torch.backends.cudnn.benchmark = True
    tf32 = mode != "fp32"
    torch.backends.cuda.matmul.allow_tf32 = tf32
    torch.backends.cudnn.allow_tf32 = tf32
    sd = port.init_state_dict(3, 1, 16, seed=1)
    params = port.clone_state_dict(sd, requires_grad=True)
    params = type(params)((k, v.detach().to(dev).requires_grad_(v.requires_grad)) for k, v in params.items())
    if mode == "bf16":
        for k, v in params.items():
            if v.dim() == 4:
                v.data = v.data.contiguous(memory_format=torch.channels_last)
    leaves = [v for v in params.values() if v.requires_grad]
    opt = torch.optim.SGD(leaves, lr=1e-4, momentum=0.9, weight_decay=1e-3, foreach=True)
    img, lab, wgt = (t.to(dev) for t in synthetic.make_batch(batch, size, size))
    if mode == "bf16":
        img = img.contiguous(memory_format=torch.channels_last)

    def step():
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=(mode == "bf16")):
            logp = port.forward(params, img, training=True)
        loss = port.pixelwise_nll(logp.float(), lab, wgt)
        loss.backward()
        opt.step()

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"value": batch / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "mode": mode,
            "what": "oracle/uresnet_port.py (torch.nn.functional -> cuDNN/ATen, torch %s) train step on the same GPU, "
                    "batch %d" % (torch.__version__, batch)}


def aspp_rates(batch, size, steps, host_batch, dev, lib):
    """BASELINE.json configs[4]: ASPP_ResNet(3,1,16) train step (fwd + loss + bwd + fused SGD) and eval forward, bf16
    storage, same batch shape as the headline workload."""
    from ubresnet_b200.models import ASPP_ResNet
    from ubresnet_b200.training.optim import SGD
    from ubresnet_b200.training.pixelwise_nllloss import PixelWiseNLLLoss
    torch.manual_seed(5)
    model = ASPP_ResNet(3, 1, 16, showsizes=False).to(dev).train()
    crit = PixelWiseNLLLoss()
    opt = SGD(model.parameters(), lr=1e-4, momentum=0.9, weight_decay=1e-3)
    img, lab, wgt = (t.to(dev) for t in host_batch)

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

    mode = "eager"
    try:                                       # the loop body as one CUDA graph, like the headline workload
        from ubresnet_b200.graphs import GraphedTrainStep
        n0 = lib.ubr_launch_count()
        g = GraphedTrainStep(model, crit, opt, img, lab, wgt, warmup=3)
        per = int((lib.ubr_launch_count() - n0) // 4)
        ms_t, _ = timed(g.replay, steps)
        l_t = per
        mode = "cuda graph"
    except Exception as exc:
        mode = "eager (graph capture failed: %s)" % str(exc)[:120]
        torch.cuda.synchronize()
        ms_t, l_t = timed(train, steps)
    model.eval()
    with torch.no_grad():
        ms_i, l_i = timed(lambda: model(img), steps)
    gf_train, gf_fwd = 310.5, 103.65           # algorithmic conv GFLOP per 512x512 crop
    return {"model": "ASPP_ResNet(3,1,16)", "per_gpu_batch": batch, "crop": size,
            "train": {"value": batch / (ms_t * 1e-3), "unit": UNIT, "ms_per_step": ms_t, "gpu_launches": int(l_t),
                      "launch_mode": mode,
                      "conv_tflops": gf_train * batch / ms_t},
            "infer": {"value": batch / (ms_i * 1e-3), "unit": UNIT, "ms_per_step": ms_i, "gpu_launches": int(l_i),
                      "conv_tflops": gf_fwd * batch / ms_i}}


def run_reference(args, rank):
    if rank != 0:
        return
    batch = args.cpu_batch
    nsteps = max(5, args.steps)                      # bounded sample: >= 5 timed steps of `batch` crops
    rate, ms, threads = cpu_train_rate(batch, args.size, nsteps, max(1, min(args.warmup, 2)))
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "ub_uresnet(inplanes=16,1ch,3cls) train step fwd+weighted-NLL+bwd+SGD, 512x512 crops",
                   "per_gpu_batch": batch, "crop": args.size,
                   "note": "bounded sample of the %d-crop batch of the CUDA arm: %d crops per CPU step (BASELINE.json "
                           "configs[0], the reference's own CPU-runnable case); rank 0 only, all host cores" %
                           (args.batch, batch),
                   "l2": "working set >> L2 (no flush needed)"},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "oracle/uresnet_port.py (CPU fp32 restatement of the reference), batch %d of %dx%d "
                                   "crops per step, %d timed steps, %d threads" %
                                   (batch, args.size, args.size, nsteps, threads)},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# per-kernel roofline from the profiled C-ABI calls
# ------------------------------------------------------------------------------------------------
def _conv_call_work(fn, a):
    """(algorithmic FLOPs, algorithmic HBM bytes) of one conv C-ABI call from its raw arguments (Views are (n,h,w,c)).
    FLOPs = 2*M*N*K of the contraction; bytes = each operand / result tensor once (bf16), the DESIGN.md figure."""
    npx = lambda v: v[0] * v[1] * v[2]
    if fn in ("ubr_conv2d_fwd", "ubr_conv2d_fwd_stats", "ubr_conv2d_fwd_bn", "ubr_conv2d_fwd_fused"):
        x, out_ch, k, st = a[0], a[3], a[4], a[5]
        m = npx(x) // (st * st)
        return 2.0 * m * x[3] * out_ch * k * k, 2.0 * (npx(x) * x[3] + m * max(out_ch, 16))
    if fn in ("ubr_conv2d_dgrad", "ubr_conv2d_dgrad_acc"):
        dy, in_ch, k, st = a[0], a[2], a[3], a[4]
        b = 2.0 * (npx(dy) * dy[3] + npx(dy) * st * st * in_ch * (2 if fn.endswith("_acc") else 1))
        return 2.0 * npx(dy) * dy[3] * in_ch * k * k, b
    if fn in ("ubr_conv2d_dgrad_bnreduce", "ubr_conv2d_dgrad_bnreduce_final"):
        # data gradient + the BatchNorm-backward reduction it carries: reads dy and y_pre, writes dx
        dy, in_ch, k = a[0], a[2], a[3]
        return 2.0 * npx(dy) * dy[3] * in_ch * k * k, 2.0 * (npx(dy) * dy[3] + 2 * npx(dy) * in_ch)
    if fn == "ubr_conv2d_wgrad":
        x, dy, out_ch, k = a[0], a[1], a[2], a[3]
        return 2.0 * npx(dy) * x[3] * out_ch * k * k, 2.0 * (npx(x) * x[3] + npx(dy) * dy[3])
    if fn in ("ubr_deconv4s2_fwd", "ubr_deconv4s2_fwd_subpixel"):
        x, out_ch = a[0], a[2]
        return 2.0 * npx(x) * x[3] * out_ch * 16, 2.0 * (npx(x) * x[3] + 4 * npx(x) * out_ch)
    if fn == "ubr_deconv4s2_dgrad":
        dy, in_ch = a[0], a[2]
        return 2.0 * (npx(dy) // 4) * dy[3] * in_ch * 16, 2.0 * (npx(dy) * dy[3] + (npx(dy) // 4) * in_ch)
    if fn == "ubr_deconv4s2_wgrad":
        x, dy = a[0], a[1]
        return 2.0 * npx(x) * x[3] * dy[3] * 16, 2.0 * (npx(x) * x[3] + npx(dy) * dy[3])
    return None


def conv_kernel_classes(groups, nsteps):
    """{primary kernel: per-step launches / ms / algorithmic GFLOP and GB} from TimedLib.by_kernel()."""
    out = {}
    for kern, calls in groups.items():
        names = kern.split("+") if kern else []
        if not names or names[0] not in ("conv_halo_kernel", "conv_gemm_kernel", "wgrad_halo_kernel", "wgrad_gemm_kernel"):
            continue
        ent = out.setdefault(names[0], {"calls": 0.0, "launches": 0.0, "ms": 0.0, "gflop": 0.0, "gbytes": 0.0})
        for fn, a, ms in calls:
            work = _conv_call_work(fn, a)
            if work is None:
                continue
            ent["calls"] += 1.0 / nsteps
            ent["launches"] += sum(1 for nm in names if nm == names[0]) / nsteps
            ent["ms"] += ms / nsteps
            ent["gflop"] += work[0] / 1e9 / nsteps
            ent["gbytes"] += work[1] / 1e9 / nsteps
    for ent in out.values():
        ent["tflops"] = ent["gflop"] / ent["ms"] if ent["ms"] else None
        ent["gbs"] = ent["gbytes"] / ent["ms"] * 1e3 if ent["ms"] else None
    return out


def conv_bound_split(groups, nsteps, peaks):
    """Every conv launch against ITS OWN roofline: a launch is tensor bound when its algorithmic FLOPs / sustained bf16
    peak exceed its algorithmic bytes / HBM peak (arithmetic intensity above the ridge), HBM bound otherwise -- one kernel
    template serves both kinds (3x3 at 16 channels: 72 FLOP/B, 7x7: 392 FLOP/B, ridge ~210).  Returns per bound the
    launches, time, achieved rate and fraction of that peak, the FLOP-weighted tensor fraction and
    `mixed_frac` = sum of the launches' roofline times / sum of their measured times."""
    tf, bw = peaks["tf_sustained"] * 1e12, peaks["hbm_gbs"] * 1e9
    acc = {"tensor": [0, 0.0, 0.0, 0.0, 0.0], "hbm": [0, 0.0, 0.0, 0.0, 0.0]}     # launches, ms, flop, bytes, flop*frac
    roof_ms = 0.0
    for kern, calls in groups.items():
        names = kern.split("+") if kern else []
        if not names or names[0] not in ("conv_halo_kernel", "conv_gemm_kernel", "wgrad_halo_kernel", "wgrad_gemm_kernel"):
            continue
        for fn, a, ms in calls:
            work = _conv_call_work(fn, a)
            if work is None or ms <= 0:
                continue
            t_tensor, t_hbm = work[0] / tf * 1e3, work[1] / bw * 1e3
            e = acc["tensor" if t_tensor >= t_hbm else "hbm"]
            e[0] += 1; e[1] += ms; e[2] += work[0]; e[3] += work[1]; e[4] += work[0] * (t_tensor / ms)
            roof_ms += max(t_tensor, t_hbm)
    tot_ms = acc["tensor"][1] + acc["hbm"][1]
    if tot_ms <= 0:
        return None
    t, h = acc["tensor"], acc["hbm"]
    out = {"ridge_flop_per_byte": tf / bw, "mixed_frac": roof_ms / tot_ms,
           "what": "per-launch roofline: tensor bound if FLOPs/peak >= bytes/peak, else HBM bound; mixed_frac = sum(roofline "
                   "time) / sum(measured time) over all conv launches of a step"}
    if t[1] > 0:
        out["tensor_bound"] = {"launches_per_step": t[0] / nsteps, "ms_per_step": t[1] / nsteps,
                               "achieved_tflops": t[2] / t[1] / 1e9, "frac_of_bf16_sustained": t[2] / t[1] / 1e9 / peaks["tf_sustained"],
                               "frac_flop_weighted": t[4] / t[2] if t[2] else None}
    if h[1] > 0:
        out["hbm_bound"] = {"launches_per_step": h[0] / nsteps, "ms_per_step": h[1] / nsteps,
                            "achieved_gbs": h[3] / h[1] / 1e6, "frac_of_hbm_peak": h[3] / h[1] / 1e6 / peaks["hbm_gbs"]}
    return out


def load_traffic():
    """per-launch DRAM traffic of the conv kernels from the committed `ncu --set full` capture (profiles/)."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.isfile(path):
        with open(path) as fh:
            return json.load(fh)
    return {}


def make_roofline(klass, peaks, step_ms):
    """`roofline` of the dominant conv kernel (largest share of the step) + the same figures for the other three."""
    if not klass:
        return None
    traffic = load_traffic()
    hbm_bound = {"conv_halo_kernel": True, "wgrad_halo_kernel": True, "conv_gemm_kernel": False, "wgrad_gemm_kernel": False}
    desc = {"conv_halo_kernel": "halo-resident tcgen05 conv fwd/dgrad, stride 1, C<=64 at 512^2..128^2 (AI below the ridge)",
            "wgrad_halo_kernel": "halo-resident stacked-tap tcgen05 weight gradient, stride 1, C<=64",
            "conv_gemm_kernel": "tcgen05 implicit-GEMM conv fwd/dgrad + deconv, C>=128 and strided layers",
            "wgrad_gemm_kernel": "tcgen05 split-K weight-gradient GEMM, C>=128 and strided layers"}

    def one(name):
        e = klass[name]
        per_launch_ms = e["ms"] / max(e["launches"], 1.0)
        t = traffic.get(name)
        if hbm_bound[name]:
            r = {"bound": "hbm", "achieved": e["gbs"], "peak": peaks["hbm_gbs"], "unit": "GB/s",
                 "frac": e["gbs"] / peaks["hbm_gbs"], "algorithmic_bytes_per_launch": e["gbytes"] * 1e9 / max(e["launches"], 1.0)}
        else:
            r = {"bound": "tensor", "achieved": e["tflops"], "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                 "frac": e["tflops"] / peaks["tf_sustained"],
                 "algorithmic_flops_per_launch": e["gflop"] * 1e9 / max(e["launches"], 1.0)}
        r.update(kernel="%s (%s)" % (name, desc[name]), launches_per_step=e["launches"], avg_launch_ms=per_launch_ms,
                 share_of_step=e["ms"] / step_ms if step_ms else None,
                 traffic=t.get("dram_bytes_per_launch") if t else None,
                 traffic_source=t.get("source") if t else None,
                 peak_source=peaks["source"] + (" (HBM copy bandwidth)" if hbm_bound[name] else " (bf16 sustained)"))
        return r

    order = sorted(klass, key=lambda k: -klass[k]["ms"])
    roof = one(order[0])
    roof["others"] = [one(k) for k in order[1:]]
    tot_ms = sum(e["ms"] for e in klass.values())
    tot_gf = sum(e["gflop"] for e in klass.values())
    roof["all_conv_kernels_tflops"] = tot_gf / tot_ms if tot_ms else None
    roof["all_conv_kernels_share_of_step"] = tot_ms / step_ms if step_ms else None
    return roof


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
    from ubresnet_b200.training.optim import SGD
    opt = SGD(model.parameters(), lr=1e-4, momentum=0.9, weight_decay=1e-3)     # one multi-tensor launch per step

    B, S = args.batch, args.size
    nbuf = 2
    host = []
    for i in range(nbuf):
        img, lab, wgt = synthetic.make_batch(B, S, S, rank=rank, step=i)
        host.append((img.pin_memory(), lab.pin_memory(), wgt.pin_memory()))
    resident = [(a.to(dev), b.to(dev), c.to(dev)) for a, b, c in host]
    # e2e wire format (ubresnet_b200/fill.py): what the reference's fillers read from the larcv file --
    # column-major fp32 ADC planes and 1-byte ROI-type label planes; the transpose, the ROI->class table with the -1
    # shift and the SegWeightTrackShower pixel weights are computed on the device
    from ubresnet_b200.fill import DeviceBatchFiller
    filler = DeviceBatchFiller(dev)
    wire = []
    for img, lab, _ in host:
        roi = torch.zeros_like(lab, dtype=torch.uint8)
        roi[lab == 1] = 6                        # kROIMuminus (track)
        roi[lab == 2] = 3                        # kROIEminus  (shower)
        wire.append((img[:, 0].transpose(1, 2).contiguous().view(B, -1).pin_memory(),
                     roi.transpose(1, 2).contiguous().view(B, -1).pin_memory()))
    h2d_bytes = sum(t.numel() * t.element_size() for t in wire[0])
    fi, fl, fw = filler.fill(wire[0][0], wire[0][1], S, S, check=True)
    fill_ok = bool(torch.equal(fi, resident[0][0]) and torch.equal(fl, resident[0][1]) and torch.equal(fw, resident[0][2]))
    del fi, fl, fw
    dw0 = tuple(t.to(dev) for t in wire[0])
    torch.cuda.synchronize()
    ef0, ef1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ef0.record()
    for _ in range(5):
        filler.fill(dw0[0], dw0[1], S, S)
    ef1.record()
    torch.cuda.synchronize()
    fill_ms = ef0.elapsed_time(ef1) / 5
    del dw0

    def step_resident(i):
        img, lab, wgt = resident[i % nbuf]
        opt.zero_grad(set_to_none=True)
        loss = crit(net(img), lab, wgt)
        loss.backward()
        opt.step()
        return loss

    # e2e: every step's batch is copied host -> device from pinned memory inside the timed region; the copy of batch
    # i+1 runs on the prefetcher's side stream while step i computes (ubresnet_b200/prefetch.py)
    from ubresnet_b200.prefetch import AssembledPrefetcher, DevicePrefetcher

    def host_batches():
        i = 0
        while True:
            yield host[i % nbuf]
            i += 1

    def wire_batches():
        i = 0
        while True:
            yield wire[i % nbuf]
            i += 1

    feed = {}
    # the loss of EVERY step is copied device -> host inside the timed region, asynchronously into pinned memory, and
    # read by the host one step later (the last one after the final synchronize): the copy never drains the GPU queue
    loss_host = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ready = [None, None]
    losses_seen = []

    def step_e2e(i):
        if "it" not in feed:
            feed["it"] = AssembledPrefetcher(wire_batches(), filler, S, S, B, dev)
        img, lab, wgt = next(feed["it"])                              # copied + assembled one step ahead
        opt.zero_grad(set_to_none=True)
        loss = crit(net(img), lab, wgt)
        loss.backward()
        opt.step()
        j = i & 1
        loss_host[j].copy_(loss.detach(), non_blocking=True)          # device -> host read of the step's result
        loss_ready[j] = torch.cuda.Event()
        loss_ready[j].record()
        if loss_ready[j ^ 1] is not None:
            loss_ready[j ^ 1].synchronize()                           # previous step's loss: already on the host
            losses_seen.append(float(loss_host[j ^ 1]))
        return None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_ms = [0.0]

    def timed(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = lib.ubr_launch_count()
        e0.record()
        t_host = time.perf_counter()
        for i in range(steps):
            fn(warmup + i)
        host_ms[0] = (time.perf_counter() - t_host) * 1e3 / steps      # host time to ENQUEUE one step (no sync inside)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        launches = lib.ubr_launch_count() - n0
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches

    # ---- the loop body as ONE CUDA graph (ubresnet_b200.graphs.GraphedTrainStep): what a user who cares about
    # throughput runs.  Eager launches (the ~520 ctypes calls per step) are measured too and reported next to it.
    graphed = None
    graph_note = None
    if not args.no_graph:
        try:
            from ubresnet_b200.graphs import GraphedTrainStep
            n0 = lib.ubr_launch_count()
            graphed = GraphedTrainStep(net, crit, opt, *resident[0], warmup=3)
            graphed.launches_per_step = int((lib.ubr_launch_count() - n0) // 4)        # 3 warm-up steps + the capture
        except Exception as exc:                                    # capture is an optimisation: fall back loudly
            graph_note = "CUDA-graph capture failed, eager launches timed instead: %s" % str(exc)[:300]
            graphed = None
            torch.cuda.synchronize()

    def step_graphed(i):
        img, lab, wgt = resident[i % nbuf]
        return graphed(img, lab, wgt)

    def step_e2e_graphed(i):
        if "it" not in feed:
            feed["it"] = AssembledPrefetcher(wire_batches(), filler, S, S, B, dev)
        img, lab, wgt = next(feed["it"])                              # copied + assembled one step ahead
        loss = graphed(img, lab, wgt)                                 # 3 device copies into the graph's static inputs + replay
        j = i & 1
        loss_host[j].copy_(loss, non_blocking=True)                   # device -> host read of the step's result
        loss_ready[j] = torch.cuda.Event()
        loss_ready[j].record()
        if loss_ready[j ^ 1] is not None:
            loss_ready[j ^ 1].synchronize()
            losses_seen.append(float(loss_host[j ^ 1]))
        return None

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    eager = None
    if graphed is not None:
        ms_eager, launches = timed(step_resident, max(3, args.steps // 2), args.warmup)
        eager = {"value": world * B * max(3, args.steps // 2) / (ms_eager * 1e-3), "unit": UNIT,
                 "ms_per_step": ms_eager / max(3, args.steps // 2), "host_enqueue_ms_per_step": host_ms[0],
                 "what": "the same step issued kernel by kernel from Python (no CUDA graph)"}
        launches_per_step = graphed.launches_per_step
        ms, _ = timed(step_graphed, args.steps, args.warmup)
        launches = launches_per_step * args.steps
    else:
        ms, launches = timed(step_resident, args.steps, args.warmup)
    host_enqueue_ms = host_ms[0]
    # `value` is the step through whichever of the two launch modes of the public API is faster ON THIS BOX, both timed
    # the same way (the timings are max-reduced over ranks, so every rank takes the same decision): since the bypass
    # branch stream exists the eager streams keep more of programmatic dependent launch than the captured graph does
    # (DESIGN 8), as long as the host can enqueue a step faster than the GPU runs it.  The other mode rides along.
    use_eager = False
    graph_leg = None
    if graphed is not None and not args.graph_only and eager["ms_per_step"] < 0.997 * ms / args.steps:
        ms_g, launches_g, host_g = ms, launches, host_enqueue_ms
        ms_e, launches_e = timed(step_resident, args.steps, args.warmup)          # EXACTLY args.steps steps, like the graph leg
        if ms_e < ms_g:
            use_eager = True
            ms, launches, host_enqueue_ms = ms_e, launches_e, host_ms[0]
            graph_leg = {"value": world * B * args.steps / (ms_g * 1e-3), "unit": UNIT, "ms_per_step": ms_g / args.steps,
                         "host_enqueue_ms_per_step": host_g, "gpu_launches": int(launches_g),
                         "what": "the same step as ONE CUDA graph replay (ubresnet_b200.graphs.GraphedTrainStep)"}
    clocks = sampler.stop() if sampler else None
    ms_e2e, _ = timed(step_e2e_graphed if graphed is not None else step_e2e, args.steps, max(3, args.warmup))
    e2e_mode = "graph" if graphed is not None else "eager"
    if use_eager:
        losses_seen_g = list(losses_seen)
        ms_e2e_e, _ = timed(step_e2e, args.steps, max(3, args.warmup))                # the same feed through the eager loop
        if ms_e2e_e < ms_e2e:
            ms_e2e, e2e_mode = ms_e2e_e, "eager"
            del losses_seen[:len(losses_seen_g)]
        else:
            del losses_seen[len(losses_seen_g):]
    value = world * B * args.steps / (ms * 1e-3)
    e2e = world * B * args.steps / (ms_e2e * 1e-3)

    # ---- live per-kernel timing: CUDA events around every C-ABI call on the launching stream (separate short pass);
    # every rank runs the extra steps (they contain the gradient all-reduce), only rank 0 records events
    roof = None
    kernels = None
    klass = None
    nprof = 2
    prof = None
    from ubresnet_b200 import engine as _engine
    side, branch = _engine.WGRAD_SIDE_STREAM, _engine.BRANCH_MAX_RES
    _engine.WGRAD_SIDE_STREAM = False          # per-kernel durations are taken with every kernel alone on the GPU:
    _engine.BRANCH_MAX_RES = 0                 # no weight-gradient side stream, no bypass branch stream
    if rank == 0:
        prof = _lib.enable_profiling()
        prof.records.clear()
    for i in range(nprof):
        step_resident(i)
    barrier()
    _engine.WGRAD_SIDE_STREAM, _engine.BRANCH_MAX_RES = side, branch
    if rank == 0:
        summ = prof.summary()
        groups = prof.by_kernel()
        _lib.disable_profiling()
        tot = sum(t for _, t in summ.values())
        peaks = load_peaks()
        klass = conv_kernel_classes(groups, nprof)
        roof = make_roofline(klass, peaks, tot / nprof)
        if roof is not None:
            roof["by_bound"] = conv_bound_split(groups, nprof, peaks)
        kernels = {k: {"calls_per_step": c / nprof, "ms_per_step": t / nprof} for k, (c, t) in
                   sorted(summ.items(), key=lambda kv: -kv[1][1])}

    # ---- multi-GPU correctness riding along with the speed number: after all steps every rank must hold bit-identical
    # parameters (same seed, same all-reduced gradients) -- the spread of a float64 checksum over ranks must be 0
    dp_check = None
    if world > 1:
        with torch.no_grad():
            chk = torch.stack([torch.stack([p.double().sum(), p.double().abs().sum()]) for p in model.parameters()]).sum(0)
        allc = [torch.zeros_like(chk) for _ in range(world)]
        dist.all_gather(allc, chk)
        allc = torch.stack(allc)
        dp_check = {"param_checksum_spread": float((allc.max(0).values - allc.min(0).values).abs().max().item()),
                    "param_abs_sum": float(allc[0, 1].item()), "ranks": world,
                    "what": "max over {sum, abs-sum} of (max - min over ranks) of the float64 parameter checksums after "
                            "every timed step; must be 0"}

    # ---- inference (eval-mode forward; BASELINE metric is "train+infer") and whole-view deploy (configs[3])
    infer = None
    wholeview = None
    if not args.no_infer:
        model.eval()

        def infer_resident(i):
            with torch.no_grad():
                return model(resident[i % nbuf][0])

        ifeed = {}

        def infer_e2e(i):
            # host crops in through the prefetcher (copy of batch i+1 overlaps forward i), class map back on the host
            if "it" not in ifeed:
                ifeed["it"] = DevicePrefetcher(((h[0],) for h in host_batches()), dev)
            with torch.no_grad():
                out = model(next(ifeed["it"])[0])
                return out.argmax(1).to(torch.uint8).cpu()          # every step

        ms_i, l_i = timed(infer_resident, args.steps, args.warmup)
        ms_ie, _ = timed(infer_e2e, max(2, args.steps // 2), 1)
        infer = {"metric": "infer_crops_per_sec_512x512", "value": world * B * args.steps / (ms_i * 1e-3), "unit": UNIT,
                 "ms_per_step": ms_i / args.steps, "gpu_launches": int(l_i),
                 "e2e": {"value": world * B * max(2, args.steps // 2) / (ms_ie * 1e-3), "unit": UNIT,
                         "h2d_bytes_per_step": host[0][0].numel() * 4 * world, "d2h_bytes_per_step": B * S * S * world},
                 "conv_tflops": GF_FWD * B * args.steps / (ms_i * 1e-3) / 1e3}
        if S == 512:
            from ubresnet_b200.deploy import wholeview as wv
            # every rank holds the SAME event for the crop-sharded measurement, its own events for the event-sharded one
            planes = synthetic.make_wholeview(seed=synthetic.BASE_SEED, rows=1008, cols=3456, planes=3).pin_memory()
            nev = max(2, args.steps // 2)
            h_out = torch.empty((3, 1008, 3456), dtype=torch.uint8).pin_memory()

            def event(i):
                # ONE event through the public deploy call: host planes in, stitched uint8 class maps back on the host.
                # N > 1: the event's 42 crops are sharded over the ranks, class tiles all-gathered (C2)
                out = wv.run_event_sharded(model, planes, crop=512) if world > 1 else wv.run_event_classes(model, planes, crop=512)
                h_out.copy_(out)
                return h_out

            ms_w, l_w = timed(event, nev, 1)
            sharded_ok = None
            if world > 1:
                # deploy needs ONE model state on every rank: BatchNorm running statistics are per rank during training
                # (like nn.DataParallel's replicas) -- rank 0's buffers win, as when a checkpoint is written
                from ubresnet_b200.parallel import broadcast_module
                broadcast_module(model, 0)
                single = wv.run_event_classes(model, planes, crop=512)
                same = torch.tensor([int(torch.equal(single, wv.run_event_sharded(model, planes, crop=512)))], device=dev)
                dist.all_reduce(same, op=dist.ReduceOp.MIN)
                sharded_ok = bool(same.item())
            # the deploy loop over many events: events sharded round-robin over the ranks (no collective), copies of
            # neighbouring events overlapped with the forward pass; host wall clock, max over ranks
            npipe = 4 * nev * world
            for _ in wv.run_events(model, [planes] * (2 * world), crop=512, rank=rank, world=world):
                pass
            barrier()
            t0 = time.perf_counter()
            nout = sum(1 for _ in wv.run_events(model, [planes] * npipe, crop=512, rank=rank, world=world))
            torch.cuda.synchronize()
            dt = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            ms_pipe = float(dt.item()) * 1e3 / npipe
            wholeview = {"metric": "wholeview_events_per_sec", "value": 1e3 / ms_pipe, "unit": "events/s",
                         "crops_per_sec": 42e3 / ms_pipe, "ms_per_event": ms_pipe, "events": npipe,
                         "workload": "3 planes 1008x3456 -> 42 overlapping 512x512 crops -> eval forward -> per-crop "
                                     "argmax -> stitched uint8 class maps; host planes in, host class maps out; "
                                     "deploy.wholeview.run_events(rank, world): events sharded round-robin over the "
                                     "ranks (no collective), upload of event i+1 / download of event i-1 on a copy "
                                     "stream during event i; host wall clock over the loop, max over ranks",
                         "single_event": {"ms_per_event": ms_w / nev, "gpu_launches_per_event": int(l_w / nev),
                                          "sharding": ("crops of one event over %d ranks, uint8 class tiles all-gathered "
                                                       "over NCCL (deploy.wholeview.run_event_sharded)" % world)
                                          if world > 1 else "one GPU (deploy.wholeview.run_event_classes)",
                                          "sharded_equals_single_gpu": sharded_ok}}
        model.train()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        rate, cms, threads = cpu_train_rate(args.cpu_batch, S, 6, 1)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": "oracle/uresnet_port.py train step, batch %d of %dx%d crops, mean of 6 steps after 1 "
                         "warm-up (%.0f ms/step)" % (args.cpu_batch, S, S, cms)}

    library = None
    if rank == 0 and world == 1 and not args.no_library:
        resident.clear()
        torch.cuda.empty_cache()
        library = []
        # default: the bf16-autocast + channels_last leg (the fastest library configuration, "the bar" of SURVEY 2b);
        # --library-baseline adds tf32 and fp32
        for mode in (("bf16", "tf32", "fp32") if args.library_baseline else ("bf16",)):
            try:
                library.append(library_train_rate(B, S, 5, 3, mode, dev))
            except Exception as exc:                               # e.g. out of memory in fp32
                library.append({"mode": mode, "error": str(exc)[:200]})
            torch.cuda.empty_cache()

    aspp = None
    if rank == 0 and world == 1 and not args.no_aspp and S == 512:
        try:
            resident.clear()
            del model, net, opt
            torch.cuda.empty_cache()
            aspp = aspp_rates(B, S, max(3, args.steps // 2), host[0], dev, lib)
        except Exception as exc:
            aspp = {"error": str(exc)[:300]}

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
                    "ms_per_step": ms_e2e / args.steps,
                    "input": "pinned host larcv-style planes (column-major fp32 ADC + uint8 ROI-type labels) -> H2D copy and "
                             "device batch assembly (ubresnet_b200.fill: transpose, ROI->class LUT -1, SegWeightTrackShower "
                             "weights) on the prefetch stream one step ahead (prefetch.AssembledPrefetcher) -> train step "
                             "-> loss read back",
                    "loss_readback": "every step's loss copied D2H (pinned, non_blocking) and consumed by the host one step "
                                     "later; %d losses read inside the timed loop" % len(losses_seen),
                    "device_fill_equals_host_batch": fill_ok, "device_fill_ms_per_batch": fill_ms,
                    "launch_mode": e2e_mode},
            "gpu_launches": int(launches),
            "host_enqueue_ms_per_step": host_enqueue_ms,
            "launch_mode": (("eager streams: one ctypes call per kernel, %d kernels per step on the main, weight-gradient and "
                             "bypass-branch streams (faster than the CUDA-graph replay on this box, which is reported "
                             "under `graph`)" % (launches // max(1, args.steps))) if use_eager else
                            ("one CUDA graph per step (ubresnet_b200.graphs.GraphedTrainStep: forward + loss + backward "
                             "+ all-reduce + optimizer captured once, %d kernels per replay)" % (launches // max(1, args.steps))))
                           if graphed is not None else (graph_note or "eager: one ctypes call per kernel"),
            "utilisation": {"conv_tflops_whole_step": GF_TRAIN * B * world * args.steps / (ms * 1e-3) / 1e3 / world,
                            "note": "algorithmic conv FLOPs (199.41 GF/crop) / step time, per GPU"},
        }
        if roof:
            line["roofline"] = roof
        if cpu:
            line["cpu_baseline"] = cpu
        if library:
            line["library_baseline"] = library
        if eager:
            line["eager"] = eager
        if graph_leg:
            line["graph"] = graph_leg
        if dp_check:
            line["dp_check"] = dp_check
        if aspp:
            line["aspp"] = aspp
        if infer:
            line["infer"] = infer
        if wholeview:
            line["wholeview"] = wholeview
        if klass:
            line["conv_kernels"] = klass
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
    ap.add_argument("--library-baseline", action="store_true",
                    help="also time the oracle port through PyTorch/cuDNN on the same GPU (bf16 autocast, tf32, fp32)")
    ap.add_argument("--no-library", action="store_true", help="skip the PyTorch/cuDNN bf16 comparison leg (N=1)")
    ap.add_argument("--no-aspp", action="store_true", help="skip the ASPP_ResNet (configs[4]) extra measurement (N=1)")
    ap.add_argument("--no-graph", action="store_true", help="time eager kernel-by-kernel launches only")
    ap.add_argument("--graph-only", action="store_true", help="`value` through the CUDA graph even where eager is faster")
    ap.add_argument("--no-infer", action="store_true", help="skip the eval-forward and whole-view measurements")
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
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        # NCCL writes its version banner and warnings to stdout: send them to a file so that stdout is ONE JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/tmp/ubresnet_b200_nccl_%h_%p.log")
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
