"""Whole-step CUDA graph of the training loop body.

The reference's hot loop (training/train_ubresnet2018_wlarcv2.py:319-344) is `zero_grad -> forward -> criterion ->
backward -> optimizer.step`.  Through this library that is ~520 kernel launches per step, each issued from Python
through ctypes (~16 ms of host time per step against ~21.6 ms of GPU time at batch 32 on a B200): the GPU stays fed only
while the host thread runs undisturbed, and with eight ranks on one host it does not.  `GraphedTrainStep` captures the
loop body ONCE into a CUDA graph -- forward, loss, the hand-written backward with its weight-gradient side stream, the
data-parallel bucket all-reduces (NCCL is capturable) and the fused optimizer step -- and replays it with one
`cudaGraphLaunch` per step.  Capture is legal because nothing in the step touches the host: tensor maps are encoded at
capture time for buffers of the graph's private memory pool (stable addresses), the optimizer / weight-pack tables are
uploaded from pinned memory, statistics partial-row counts are host integers known at launch time.

    step = GraphedTrainStep(model_or_ddp, criterion, optimizer, img, lab, wgt)     # device tensors of the batch shape
    loss = step(img, lab, wgt)          # copies into the static inputs, replays; `loss` is a device scalar (static)

The `warmup` eager steps before capture are real optimizer steps (the momentum buffers and the allocator have to
exist); run them on data you would train on anyway.
"""
import torch


class GraphedTrainStep:
    def __init__(self, net, criterion, optimizer, img, lab, wgt, warmup=3):
        if not img.is_cuda:
            raise RuntimeError("GraphedTrainStep: pass CUDA tensors of the batch shape (there is no CPU path)")
        self.net, self.criterion, self.optimizer = net, criterion, optimizer
        self.static = (img.clone(), lab.clone(), wgt.clone())
        dev = img.device
        cur = torch.cuda.current_stream(dev)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side):                   # torch's recipe: warm up off the default stream
            for _ in range(max(1, warmup)):
                self._body()
        cur.wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = self._body()
        self.replays = 0
        mod = net.module if hasattr(net, "module") else net
        self._touched = list(mod.parameters()) + [b for b in mod.buffers() if b.is_floating_point()]
        self.launches_per_step = None

    def _body(self):
        img, lab, wgt = self.static
        self.optimizer.zero_grad(set_to_none=True)
        loss = self.criterion(self.net(img), lab, wgt)
        loss.backward()
        self.optimizer.step()
        return loss.detach()

    @property
    def inputs(self):
        """the static input tensors (image, label, weight): fill them in place (e.g. `DeviceBatchFiller.fill(out=...)`)
        and call `replay()` to skip the copy of `__call__`"""
        return self.static

    def replay(self):
        self.graph.replay()
        self.replays += 1
        # the replayed kernels rewrite parameters, BatchNorm buffers and optimizer state through raw pointers: advance
        # the version counters so that everything keyed on them (the engine's bf16 operand and BatchNorm-fold caches,
        # used by a later eager / eval forward) notices
        torch.autograd.graph.increment_version(self._touched)
        return self.loss

    def __call__(self, img, lab, wgt):
        s = self.static
        s[0].copy_(img, non_blocking=True)
        s[1].copy_(lab, non_blocking=True)
        s[2].copy_(wgt, non_blocking=True)
        return self.replay()
