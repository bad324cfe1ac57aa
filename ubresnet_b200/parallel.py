"""Data-parallel training: one process per GPU, NCCL gradient all-reduce over NVLink overlapped with backward.

The reference's only multi-GPU code is single-process `nn.DataParallel` (training/train_ubresnet2018_wlarcv2.py:99,103):
scatter the batch, replicate the module every step, reduce-add gradients onto device 0; BatchNorm statistics
stay per replica.  The B200 design keeps those semantics (per-rank BN statistics, averaged gradients) with
one process per GPU:

  * gradients are written by the wgrad kernels straight into a few flat fp32 BUCKETS laid out in
    backward-completion order (head+dec1-3 | dec4-5 | enc5 | enc4-1+stem);
  * the moment the last stage of a bucket finishes, its all-reduce is enqueued on a side stream, so the
    transfer of dec/enc5 (75 % of the 72 MB) overlaps with the remaining backward kernels;
  * `finish()` joins the side stream and scales by 1/world.

The path has exactly one exchange step (the gradient all-reduce); activations never leave the GPU.
"""
import os

import torch
import torch.distributed as dist

# UBR_DP_OVERLAP=0: launch every bucket's all-reduce at the end of backward instead of as soon as the bucket is complete
DP_OVERLAP = os.environ.get("UBR_DP_OVERLAP", "1") != "0"

STAGE_ORDER = ["head", "dec1", "dec2", "dec3", "dec4", "dec5", "enc5", "enc4", "enc3", "enc2", "enc1", "stem"]
ASPP_STAGE_ORDER = STAGE_ORDER[:6] + ["aspp"] + STAGE_ORDER[6:]          # backward order of engine_aspp.ASPPEngine
# "aspp" (ASPP_ResNet only): the three ASPP skip layers run their backward between dec5 and enc5, so they share
# enc5's bucket
BUCKET_OF_STAGE = {"head": 0, "dec1": 0, "dec2": 0, "dec3": 0, "dec4": 1, "dec5": 1, "aspp": 2, "enc5": 2, "enc4": 3,
                   "enc3": 3, "enc2": 3, "enc1": 3, "stem": 3}
LAST_STAGE_OF_BUCKET = {0: "dec3", 1: "dec5", 2: "enc5", 3: "stem"}


def stage_of_param(name):
    """state_dict-style parameter name -> backward stage."""
    head = name.split(".")[0]
    if head in ("conv10", "bn10", "conv11"):
        return "head"
    if head in ("conv1", "bn1"):
        return "stem"
    if head.startswith("enc_layer"):
        return "enc" + head[len("enc_layer"):]
    if head.startswith("dec_layer"):
        return "dec" + head[len("dec_layer"):]
    if head.startswith("ASPP_layer_enc") or head.startswith("ASPP_combine_enc"):
        return "aspp"
    raise KeyError(name)


class GradBuckets:
    """Flat fp32 gradient buckets + overlapped all-reduce.  Works on CUDA (NCCL, side stream) and on CPU
    tensors (gloo; used by the world_size-2 CPU tests)."""

    def __init__(self, named_params, process_group=None, average=True):
        self.group = process_group
        self.average = average
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        named_params = [(n, p) for n, p in named_params if p.requires_grad]
        device = named_params[0][1].device
        sizes = {}
        members = {}
        for n, p in named_params:
            b = BUCKET_OF_STAGE[stage_of_param(n)]
            members.setdefault(b, []).append((n, p))
            sizes[b] = sizes.get(b, 0) + (p.numel() + 3) // 4 * 4      # keep every view 16-byte aligned
        self.flat = {b: torch.zeros(sizes[b], dtype=torch.float32, device=device) for b in sizes}
        self.views = {}
        for b, lst in members.items():
            off = 0
            for n, p in lst:
                self.views[p] = self.flat[b][off:off + p.numel()].view(p.shape)
                off += (p.numel() + 3) // 4 * 4
        self.is_cuda = device.type == "cuda"
        self.comm_stream = torch.cuda.Stream(device=device) if self.is_cuda else None
        self._pending = []
        self._deferred = []
        self.nbytes = sum(t.numel() * 4 for t in self.flat.values())

    # ---- Engine sink protocol
    def grad_buffer(self, param):
        # a FRESH tensor object over the bucket storage: autograd's AccumulateGrad adopts a gradient it holds the only
        # reference to, but clones one that something else (our `views` table) also references -- 165 copy kernels/step
        return self.views[param].view(param.shape)

    def aliases(self, param):
        """True when `param.grad` still lives in this bucket's storage (zero_grad(set_to_none=False), or a second
        backward before the optimizer step): the next backward would overwrite it and AccumulateGrad would then add the
        new gradient onto itself."""
        g = param.grad
        return g is not None and g.data_ptr() == self.views[param].data_ptr()

    def wants(self, stage):
        """True when `stage` closes a bucket (its all-reduce is launched by stage_done)."""
        b = BUCKET_OF_STAGE[stage]
        return LAST_STAGE_OF_BUCKET[b] == stage and b in self.flat

    def stage_done(self, stage, wait_events=()):
        """`wait_events`: CUDA events of other streams (the engine's weight-gradient side stream) that the bucket's
        all-reduce must also wait for."""
        b = BUCKET_OF_STAGE[stage]
        if LAST_STAGE_OF_BUCKET[b] != stage or b not in self.flat:
            return
        if not DP_OVERLAP:
            self._deferred.append(b)
            return
        self._launch(b, wait_events)

    def _launch(self, b, wait_events=()):
        flat = self.flat[b]
        if self.world == 1:
            return
        if self.is_cuda:
            ready = torch.cuda.Event()
            ready.record()                              # on the compute (current) stream
            self.comm_stream.wait_event(ready)
            for ev in wait_events:
                self.comm_stream.wait_event(ev)
            with torch.cuda.stream(self.comm_stream):
                work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            self._pending.append((b, work))
        else:
            work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            self._pending.append((b, work))

    def finish(self):
        for b in self._deferred:          # UBR_DP_OVERLAP=0: backward (incl. the weight-gradient stream) is complete here
            self._launch(b)
        self._deferred = []
        for b, work in self._pending:
            if self.is_cuda:
                with torch.cuda.stream(self.comm_stream):
                    work.wait()
                    if self.average:
                        from . import ops
                        ops.scale_f32_(self.flat[b], 1.0 / self.world)
            else:
                work.wait()
                if self.average:
                    self.flat[b].mul_(1.0 / self.world)
        if self.is_cuda and self._pending:
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        self._pending = []


def broadcast_module(module, src=0, process_group=None):
    """Identical weights and BN buffers on every rank (rank `src` wins, as DataParallel's replica 0)."""
    if not dist.is_initialized() or dist.get_world_size(process_group) == 1:
        return
    with torch.no_grad():
        ts = list(module.parameters()) + list(module.buffers())
        for t in ts:
            dist.broadcast(t.data, src=src, group=process_group)
        # `.data` writes do not advance the version counters the engine's bf16 operand caches are keyed on
        torch.autograd.graph.increment_version(ts)


class DistributedUResNet(torch.nn.Module):
    """`model = DistributedUResNet(UResNet(...).cuda())` -- takes nn.DataParallel's place in the training
    scripts (checkpoints keep the `module.` prefix that deploy/ubresnet_funcs.py:52-64 strips).  Wraps `UResNet`
    and `ASPP_ResNet` alike."""

    """Gradients are written straight into the flat all-reduce buckets and handed to autograd as views of them, so
    `param.grad` must be released before the next backward: use `optimizer.zero_grad(set_to_none=True)` (the default
    since torch 2.0).  Gradient accumulation over micro-batches / `zero_grad(set_to_none=False)` is detected and
    handled by a scratch copy (engine.backward), at the cost of one extra pass over the gradients."""

    def __init__(self, module, process_group=None):
        super().__init__()
        self.module = module
        broadcast_module(module, 0, process_group)
        self.buckets = GradBuckets(list(module.named_parameters()), process_group)
        module.engine().grad_sink = self.buckets

    def forward(self, x):
        return self.module(x)
