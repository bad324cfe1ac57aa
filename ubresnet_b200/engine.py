"""Whole-network kernel schedule for the U-ResNet (forward, backward) on top of the C-ABI ops.

The reference runs 58 convolutions, 52 BatchNorms, 62 ReLUs, 5 concats, ... as separate torch.nn library calls
under autograd (models/ub_uresnet.py:88-147, models/common_layers.py:39-58,127-132).  Here the whole network is
ONE autograd node: `Engine.forward` launches the sm_100a kernels in order and records the few tensors the
hand-written backward needs; `Engine.backward` replays the network in reverse (dgrad / wgrad GEMMs, fused
BN+ReLU backward, block-tail backward) and hands PyTorch one gradient per parameter.

Data layout: activations NHWC bf16; `torch.cat([deconv, skip], 1)` is zero-copy -- encoder outputs and deconv
outputs are written directly into channel slices of the decoder's concat buffer.  Parameters stay fp32 in the
reference layouts (state_dict compatibility); bf16 GEMM operand copies are re-derived when a parameter changes.
"""
import os

import torch

from . import ops
from .ops import Act

# Weight gradients have no consumer inside the backward pass, so they run on a side stream concurrently with the
# dgrad -> BN-backward chain of the main stream (deep 16^2..64^2 layers launch far fewer CTAs than the GPU has SMs).
WGRAD_SIDE_STREAM = os.environ.get("UBR_WGRAD_STREAM", "1") != "0"
# The 1x1 bypass conv of a projecting block (forward) and its data gradient (backward) depend only on the block input /
# the residual gradient: they run on a branch stream beside the conv1 -> bn1 -> conv2 chain (measured: -0.24 ms per
# batch-32 step, all of it from the 256^2 / 512^2 blocks).  UBR_BRANCH_MAX_RES: largest output width that branches (0 = off).
BRANCH_MAX_RES = int(os.environ.get("UBR_BRANCH_MAX_RES", "512"))


def model_params(m):
    """`list(m.parameters())`, also for an `nn.DataParallel` replica: replicate() leaves a replica's `_parameters` empty
    and keeps the broadcast copies (non-leaf tensors) as plain attributes + `_former_parameters`"""
    ps = list(m.parameters())
    if ps:
        return ps
    out = []
    for mod in m.modules():
        out.extend(v for v in getattr(mod, "_former_parameters", {}).values() if v is not None)
    return out


class _FoldedConv:
    """Inference operands of conv -> BatchNorm(eval): W' = W * gamma/sqrt(var+eps) per output channel (packed bf16),
    b' = beta - mean*scale (+ conv bias*scale), fp32.  Re-derived when any of the six tensors changes."""
    __slots__ = ("w", "b", "_key")

    def __init__(self):
        self.w = self.b = self._key = None

    def refresh(self, conv, bn, packed=True):
        """packed=False keeps W' as fp32 OIHW (the stem kernel reads fp32 weights)"""
        srcs = [conv.weight, conv.bias, bn.weight, bn.bias, bn.running_mean, bn.running_var]
        key = tuple((t.data_ptr(), t._version) if t is not None else None for t in srcs) + (bn.eps,)
        if key == self._key:
            return self
        with torch.no_grad():
            scale = torch.rsqrt(bn.running_var.float() + bn.eps)
            if bn.weight is not None:
                scale = scale * bn.weight.float()
            shift = -bn.running_mean.float() * scale
            if bn.bias is not None:
                shift = shift + bn.bias.float()
            if conv.bias is not None:
                shift = shift + conv.bias.float() * scale
            wf = (conv.weight.float() * scale.view(-1, 1, 1, 1)).contiguous()
            self.w = ops.pack_weights(wf, False, True) if packed else wf
            self.b = shift.contiguous()
        self._key = key
        return self


FOLD_EVAL_BN = os.environ.get("UBR_FOLD_EVAL_BN", "1") != "0"
CHAIN_EVAL_BLOCKS = os.environ.get("UBR_NO_CHAIN", "0") in ("", "0")    # eval: conv->ReLU->conv chains in one kernel


class _ConvW:
    """bf16 packed operands of one conv / deconv parameter, refreshed when the parameter changes."""
    __slots__ = ("param", "transposed", "fwd", "dgrad", "fwd3", "dgrad3", "_key", "_dgrad_stale")

    def __init__(self, param, transposed=False):
        self.param = param
        self.transposed = transposed
        self.fwd = None
        self.dgrad = None
        self.fwd3 = None            # ConvTranspose only: 3x3 sub-pixel forward operand (ops.pack_weights_subpixel)
        self.dgrad3 = None          # ConvTranspose only: 3x3 sub-pixel data-gradient operand (4*out_ch <= 64)
        self._key = None
        self._dgrad_stale = False

    def refresh(self, need_dgrad):
        key = (self.param.data_ptr(), self.param._version)
        if key != self._key:
            self.fwd = ops.pack_weights(self.param, self.transposed, True)
            self.fwd3 = ops.pack_weights_subpixel(self.param, True) if self.sub_fwd() else None
            self.dgrad = None
            self.dgrad3 = None
            self._key = key
        if need_dgrad and (self.dgrad is None or self._dgrad_stale):
            self.dgrad = ops.pack_weights(self.param, self.transposed, False)
            self.dgrad3 = ops.pack_weights_subpixel(self.param, False) if self.sub_bwd() else None
            self._dgrad_stale = False

    # sub-pixel (3x3 on the input grid) formulations of the ConvTranspose run on the halo-resident kernels, which
    # take <= 64 input channels: forward needs in_ch <= 64, backward needs 4*out_ch <= 64 (and in_ch <= 64 for wgrad)
    def sub_fwd(self):
        return self.transposed and self.param.shape[0] <= 64

    def sub_bwd(self):
        return self.transposed and 4 * self.param.shape[1] <= 64 and self.param.shape[0] <= 64

    # ---- batched refresh (one launch for the whole network, Engine._refresh_packs)
    def dims(self):
        w = self.param
        if self.transposed:
            in_ch, out_ch, kh, kw = w.shape
        else:
            out_ch, in_ch, kh, kw = w.shape
        return out_ch, in_ch, kh, kw

    def jobs(self, need_dgrad):
        """[(n_is_out, dst)] that must be (re)packed; allocates destination buffers on first use."""
        key = (self.param.data_ptr(), self.param._version)
        changed = key != self._key
        out_ch, in_ch, kh, kw = self.dims()
        pad = lambda v: (v + 15) // 16 * 16
        res = []
        lay = 1 if self.transposed else 0
        if changed or self.fwd is None:
            if self.fwd is None or self.fwd.device != self.param.device:
                self.fwd = torch.empty((kh * kw + 1, pad(out_ch), pad(in_ch)), dtype=torch.bfloat16, device=self.param.device)
            res.append((lay, 1, kh, kw, self.fwd))
            if self.sub_fwd():
                if self.fwd3 is None or self.fwd3.device != self.param.device:
                    self.fwd3 = torch.empty((10, pad(4 * out_ch), pad(in_ch)), dtype=torch.bfloat16, device=self.param.device)
                res.append((2, 1, 3, 3, self.fwd3))
        if need_dgrad and (changed or self.dgrad is None or self._dgrad_stale):
            if self.dgrad is None or self.dgrad.device != self.param.device:
                self.dgrad = torch.empty((kh * kw + 1, pad(in_ch), pad(out_ch)), dtype=torch.bfloat16, device=self.param.device)
            res.append((lay, 0, kh, kw, self.dgrad))
            if self.sub_bwd():
                if self.dgrad3 is None or self.dgrad3.device != self.param.device:
                    self.dgrad3 = torch.empty((10, pad(in_ch), pad(4 * out_ch)), dtype=torch.bfloat16,
                                              device=self.param.device)
                res.append((2, 0, 3, 3, self.dgrad3))
            self._dgrad_stale = False
        elif changed:
            self._dgrad_stale = True
        self._key = key
        return res


class _BlockRec:
    __slots__ = ("x", "y1", "st1", "a1", "y2", "st2", "yb", "stb", "out")


class Engine:
    def __init__(self, model):
        self.model = model
        self._packs = {}
        self._folds = {}           # id(conv) -> _FoldedConv (eval-mode conv+BN operands)
        self._ws = None
        self._ws_key = None
        self._w_on = False
        self.grad_sink = None      # optional data-parallel gradient sink (ubresnet_b200.parallel.GradBuckets)
        self._wstream = None       # side stream of the weight-gradient kernels (backward only)
        self._bstream = None       # branch stream of the bypass convs (see BRANCH_MAX_RES)
        self._bn_ws_b = None       # its BatchNorm workspace
        self.keep_logits = False   # parity aid: keep the raw conv11 output of the last forward (what a forward hook on
        self.last_logits = None    # the reference's conv11 sees, ub_uresnet.py:141) as (N,C,H,W) fp32

    # ------------------------------------------------------------------ helpers
    def _pack(self, conv, need_dgrad, transposed=False):
        ent = self._packs.get(id(conv))
        if ent is None or ent.param is not conv.weight:
            ent = _ConvW(conv.weight, transposed)
            self._packs[id(conv)] = ent
        ent.refresh(need_dgrad)
        return ent

    def _refresh_packs(self, need_dgrad):
        """Re-derive the bf16 GEMM operands of every conv whose fp32 parameter changed (optimizer step,
        load_state_dict) with ONE kernel launch; the device job table is cached while pointers are stable."""
        m = self.model
        convs = getattr(self, "_conv_list", None)
        if convs is None:                       # the module tree is fixed after construction: walk it once
            convs = []
            for mod in m.modules():
                if mod is m.conv1:
                    continue
                if isinstance(mod, torch.nn.ConvTranspose2d):
                    convs.append((mod, True))
                elif isinstance(mod, torch.nn.Conv2d):
                    convs.append((mod, False))
            self._conv_list = convs
        jobs = []
        for conv, transposed in convs:
            ent = self._packs.get(id(conv))
            if ent is None or ent.param is not conv.weight:
                ent = _ConvW(conv.weight, transposed)
                self._packs[id(conv)] = ent
            for lay, n_is_out, kh, kw, dst in ent.jobs(need_dgrad):
                out_ch, in_ch = ent.dims()[:2]
                jobs.append((ent.param.data_ptr(), dst.data_ptr(), lay, n_is_out, out_ch, in_ch, kh, kw,
                             dst.shape[1], dst.shape[2]))
        if not jobs:
            return
        key = tuple(jobs)
        cache = getattr(self, "_job_cache", None)
        if cache is None or cache[0] != key:
            host = torch.tensor(jobs, dtype=torch.int64).pin_memory()       # pinned + async: legal under graph capture
            table = host.to(m.conv1.weight.device, non_blocking=True)
            self._job_cache = cache = (key, table, host)
        ops.pack_weights_batch(cache[1], len(jobs))

    def _workspaces(self, n, h, w, device):
        """(bn workspace, wgrad workspace, stem workspace) sized for the largest layer."""
        m = self.model
        key = (n, h, w, str(device), m.inplanes)
        if self._ws_key == key:
            return self._ws
        p = m.inplanes
        cmax = 32 * p
        bn_ws = ops.bn_workspace(max(cmax, 16), device)
        need = 0
        shapes = self._wgrad_shapes(n, h, w)
        for s_ in shapes:
            need = max(need, ops.wgrad_workspace_bytes(*s_))
        wg_ws = torch.empty(need // 4 + 64, dtype=torch.float32, device=device)
        stem_ws = torch.empty(ops.stem_wgrad_workspace_bytes(n, h, w, p) // 4 + 16, dtype=torch.float32, device=device)
        self._ws = (bn_ws, wg_ws, stem_ws)
        self._bn_ws_b = ops.bn_workspace(max(cmax, 16), device)
        self._ws_key = key
        return self._ws

    # ------------------------------------------------------------------ bypass branch stream
    def _branch(self, out_w, device):
        """True if a bypass conv producing `out_w`-wide maps should run on the branch stream (forked here)."""
        if BRANCH_MAX_RES <= 0 or out_w > BRANCH_MAX_RES or self._bn_ws_b is None:
            return False
        if self._bstream is None:
            self._bstream = torch.cuda.Stream(device=device)
        ev = torch.cuda.Event()
        ev.record()
        self._bstream.wait_event(ev)
        return True

    def _branch_join(self):
        ev = torch.cuda.Event()
        ev.record(self._bstream)
        torch.cuda.current_stream().wait_event(ev)

    def _wgrad_shapes(self, n, h, w):
        """(grid_n, grid_h, grid_w, m_ch, n_ch, taps) for every weight-gradient GEMM in the net."""
        m = self.model
        p = m.inplanes
        pad16 = lambda v: (v + 15) // 16 * 16
        shapes = [(n, h, w, pad16(m.nkernels), p, 49), (n, h, w, 16, pad16(m.nkernels), 49)]
        for i in range(1, 6):                      # encoder stage i works at H/2^i
            ci, co = p * 2 ** (i - 1), p * 2 ** i
            hi, wi = h // 2 ** i, w // 2 ** i
            shapes += [(n, hi, wi, co, ci, 9), (n, hi, wi, co, co, 9), (n, hi, wi, co, ci, 1)]
        for i in (5, 4, 3, 2, 1):                  # decoder stage i: deconv input at H/2^i, blocks at H/2^(i-1)
            ci, cd = p * 2 ** i, p * 2 ** (i - 1)
            hin, win = h // 2 ** i, w // 2 ** i
            shapes += [(n, hin, win, ci, cd, 16), (n, 2 * hin, 2 * win, cd, 2 * cd, 9), (n, 2 * hin, 2 * win, cd, cd, 9),
                       (n, 2 * hin, 2 * win, cd, 2 * cd, 1)]
            if 4 * cd <= 64:
                shapes.append((n, hin, win, 4 * cd, ci, 9))       # sub-pixel ConvTranspose weight gradient
        return shapes

    @staticmethod
    def _bn(y, bn, training, bn_ws):
        if training:
            st = ops.bn_train_stats(y, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn_ws,
                                    momentum=bn.momentum if bn.momentum is not None else 0.1, eps=bn.eps)
        else:
            st = ops.bn_eval_affine(bn.weight, bn.bias, bn.running_mean, bn.running_var, eps=bn.eps)
            st.count = float(y.npix)
        return st

    def _conv_bn(self, x, wpk, bias, out_ch, k, stride, bn, training, bn_ws, y=None, state=None):
        """conv -> BatchNorm statistics/affine.  In training the statistics come out of the conv epilogue where the
        kernel supports it (ops.conv2d_fwd_bn), otherwise from the stand-alone bn_stats pass."""
        if training:
            return ops.conv2d_fwd_bn(x, wpk, bias, out_ch, k, stride, bn.weight, bn.bias, bn.running_mean, bn.running_var,
                                     bn_ws, y=y, momentum=bn.momentum if bn.momentum is not None else 0.1, eps=bn.eps,
                                     state=state)
        y = ops.conv2d_fwd(x, wpk, bias, out_ch, k, stride)
        return y, self._bn(y, bn, False, bn_ws)

    def _grad_targets(self, sink):
        """{param: fp32 tensor the weight-gradient kernels write}.  With a data-parallel sink the targets are views of
        its flat all-reduce buckets; frozen parameters (not in any bucket) get scratch.  A `param.grad` that still
        aliases bucket storage (zero_grad(set_to_none=False), gradient accumulation over micro-batches) is moved out
        first -- otherwise this backward would overwrite it and AccumulateGrad would add the new gradient onto itself."""
        m = self.model
        if sink is None:
            return {prm: torch.empty_like(prm, dtype=torch.float32) for prm in model_params(m)}
        grads = {}
        with torch.no_grad():
            for prm in model_params(m):
                if prm.requires_grad and prm in sink.views:
                    if sink.aliases(prm):
                        prm.grad = prm.grad.clone()
                    grads[prm] = sink.grad_buffer(prm)
                else:
                    grads[prm] = torch.empty_like(prm, dtype=torch.float32)
        return grads

    # ------------------------------------------------------------------ weight gradients on a side stream
    def _wgrad_begin(self, device):
        if WGRAD_SIDE_STREAM and self._wstream is None:
            self._wstream = torch.cuda.Stream(device=device)
        self._w_on = WGRAD_SIDE_STREAM
        self._w_keep = []          # inputs of side-stream kernels queued in the current stage (kept alive, see below)
        self._w_prev = None        # (side-stream event, inputs) of the previous stage

    def _wgrad(self, fn, *acts):
        """Run `fn` (weight-gradient launches) on the side stream once the main stream has produced its inputs.
        `acts` are inputs the main stream would otherwise release before the side kernels ran: they are kept alive
        until the main stream has waited for the side stream (`_wgrad_stage`), which keeps the caching allocator's
        reuse pattern deterministic (tensor.record_stream would make it depend on kernel timing)."""
        if not self._w_on:
            fn()
            return
        ev = torch.cuda.Event()
        ev.record()
        self._wstream.wait_event(ev)
        with torch.cuda.stream(self._wstream):
            fn()
        self._w_keep.extend(acts)

    def _wgrad_stage(self):
        """Stage boundary: the main stream waits for the side stream's work of the PREVIOUS stage (already finished or
        nearly so) and releases that stage's kept inputs."""
        if not self._w_on:
            return
        ev = torch.cuda.Event()
        ev.record(self._wstream)
        if self._w_prev is not None:
            torch.cuda.current_stream().wait_event(self._w_prev[0])
        self._w_prev = (ev, self._w_keep)
        self._w_keep = []

    def _wgrad_join(self):
        """main stream waits for every weight gradient queued so far (before a gradient bucket is reduced / returned)"""
        if self._w_on:
            torch.cuda.current_stream().wait_stream(self._wstream)
            self._w_prev = None
            self._w_keep = []

    # ------------------------------------------------------------------ stem / head (shared with engine_aspp)
    def _stem_fwd(self, x, x0_out, training, need_grad, bn_ws):
        """conv1 -> bn1 -> relu1 -> pool1 (ub_uresnet.py:94-97); x0 lands in `x0_out` (the dec1 skip slice)."""
        m = self.model
        x0 = None
        if FOLD_EVAL_BN and not training and not need_grad:
            f0 = self._folded(m.conv1, m.bn1, packed=False)
            x0 = ops.stem_conv_fwd_relu(x, f0.w, f0.b, y=x0_out)
            y0 = st0 = None
        if x0 is None:
            y0 = ops.stem_conv_fwd(x, m.conv1.weight, m.conv1.bias)
            st0 = self._bn(y0, m.bn1, training, bn_ws)
            x0 = ops.bn_act(y0, st0, True, out=x0_out)
        if need_grad:
            h0, pool_idx = ops.maxpool_fwd_idx(x0)
        else:
            h0, pool_idx = ops.maxpool_fwd(x0), None
        return y0, st0, x0, h0, pool_idx

    def _head_fwd(self, cur, training, need_grad, bn_ws):
        """conv10 -> bn10 -> relu10 -> conv11 -> log-softmax (ub_uresnet.py:137-143); bumps num_batches_tracked."""
        m = self.model
        n, h, w = cur.n, cur.h, cur.w
        if FOLD_EVAL_BN and not training and not need_grad:
            f10 = self._folded(m.conv10, m.bn10)
            y10 = st10 = None
            a10 = ops.conv2d_fwd_fused(cur, f10.w, f10.b, m.nkernels, 7, 1, True)
        else:
            w10 = self._pack(m.conv10, need_grad)
            y10, st10 = self._conv_bn(cur, w10.fwd, m.conv10.bias, m.nkernels, 7, 1, m.bn10, training, bn_ws)
            a10 = ops.bn_act(y10, st10, True)
        w11 = self._pack(m.conv11, need_grad)
        ncls = m.conv11.out_channels
        ld = 4 if ncls <= 4 else 16
        logits = torch.empty((n * h * w, ld), dtype=torch.float32, device=cur.buf.device)
        ops.conv2d_fwd(a10, w11.fwd, m.conv11.bias, ncls, 7, 1, y_f32=logits)
        logp = ops.logsoftmax_fwd(logits, n, ncls, h, w)
        if self.keep_logits:
            self.last_logits = logits.view(n, h, w, ld)[..., :ncls].permute(0, 3, 1, 2).contiguous()
        if training:
            self._bump_batches_tracked()
        return y10, st10, a10, logp

    def _bump_batches_tracked(self):
        """num_batches_tracked += 1 for every BatchNorm2d of the model: one launch over a device pointer table"""
        m = self.model
        bns = [mod.num_batches_tracked for mod in m.modules()
               if isinstance(mod, torch.nn.BatchNorm2d) and mod.num_batches_tracked is not None]
        if not bns:
            return
        key = tuple(t.data_ptr() for t in bns)
        cache = getattr(self, "_nbt_cache", None)
        if cache is None or cache[0] != key:
            host = torch.tensor(key, dtype=torch.int64).pin_memory()
            cache = self._nbt_cache = (key, host.to(bns[0].device, non_blocking=True), host)
        ops._lib.check(ops._lib.load().ubr_increment_i64(cache[1].data_ptr(), len(bns), ops._stream()), "increment_i64")
        torch.autograd.graph.increment_version(bns)

    def _head_bwd(self, saved, dlogp, grads, bn_ws, wg_ws):
        """backward of the head; returns the gradient wrt the last decoder output."""
        m = self.model
        p = m.inplanes
        ncls = m.conv11.out_channels
        dl = ops.logsoftmax_bwd(saved["logp"], dlogp.float())
        self._wgrad(lambda: ops.conv2d_wgrad(saved["a10"], dl, ncls, 7, 1, grads[m.conv11.weight], False, wg_ws), dl)
        _bias_grad(dl, grads[m.conv11.bias], bn_ws)
        w11 = self._pack(m.conv11, True)
        dy10 = ops.conv2d_dgrad_bn_bwd(dl, w11.dgrad, m.nkernels, 7, 1, saved["y10"], saved["st10"], True, bn_ws,
                                       grads[m.bn10.weight], grads[m.bn10.bias])
        del dl
        self._wgrad(lambda: ops.conv2d_wgrad(saved["d1"], dy10, m.nkernels, 7, 1, grads[m.conv10.weight], False, wg_ws),
                    dy10)
        _bias_grad(dy10, grads[m.conv10.bias], bn_ws)
        w10 = self._pack(m.conv10, True)
        return ops.conv2d_dgrad(dy10, w10.dgrad, p, 7, 1)

    def _stem_bwd(self, saved, dcur, dx0, grads, bn_ws, stem_ws):
        """x0 gets the pool gradient plus the dec1 skip gradient `dx0` (accumulated in place), then bn1 / conv1."""
        m = self.model
        ops.maxpool_bwd_idx(saved["pool_idx"], dcur, dx0, True)
        dy0 = ops.bn_act_bwd(dx0, saved["y0"], saved["st0"], True, bn_ws, grads[m.bn1.weight], grads[m.bn1.bias])
        ops.stem_conv_wgrad(saved["x"], dy0, grads[m.conv1.weight], grads[m.conv1.bias], False, stem_ws)

    # ------------------------------------------------------------------ ConvTranspose (common_layers.py:125-128)
    @staticmethod
    def _deconv_fwd(x, wd, out_ch, y):
        if wd.fwd3 is not None and ops.deconv_fwd_subpixel(x, wd.fwd3, out_ch, y):
            return y
        return ops.deconv_fwd(x, wd.fwd, out_ch, y=y)

    def _deconv_bwd(self, x, dy, wd, dw, wg_ws):
        """weight gradient into `dw` (IOHW fp32); returns the data gradient.  High-resolution stages (4*out_ch <= 64,
        W % 128 == 0) use the sub-pixel formulation on the halo-resident kernels: one pixel-unshuffle pass, then a 3x3
        conv (data gradient) and a 3x3 weight gradient on the input grid."""
        in_ch, out_ch = wd.param.shape[0], wd.param.shape[1]
        if wd.dgrad3 is not None and x.w % 128 == 0:
            dy4 = ops.pixel_unshuffle2(dy)
            dw3 = torch.empty((4 * out_ch, in_ch, 3, 3), dtype=torch.float32, device=dw.device)

            def wg():
                ops.conv2d_wgrad(x, dy4, 4 * out_ch, 3, 1, dw3, False, wg_ws)
                ops.deconv_wgrad_gather(dw3, dw)
            self._wgrad(wg, dy4, dw3)
            return ops.conv2d_fwd(dy4, wd.dgrad3, None, in_ch, 3, 1)
        self._wgrad(lambda: ops.deconv_wgrad(x, dy, dw, False, wg_ws), dy)
        return ops.deconv_dgrad(dy, wd.dgrad, in_ch)

    # ------------------------------------------------------------------ forward pieces
    def _folded(self, conv, bn, packed=True):
        ent = self._folds.get(id(conv))
        if ent is None:
            ent = self._folds[id(conv)] = _FoldedConv()
        return ent.refresh(conv, bn, packed)

    def _block_fwd_folded(self, blk, x, out):
        """eval-mode BasicBlock (common_layers.py:39-58) in at most three launches: BatchNorm folded into the conv operands,
        ReLU / residual add / ReLU in the conv epilogues -- no stand-alone pointwise pass."""
        cout = blk.conv1.out_channels
        s = blk.stride
        f1 = self._folded(blk.conv1, blk.bn1)
        f2 = self._folded(blk.conv2, blk.bn2)
        if blk.bypass is not None:
            fb = self._folded(blk.bypass, blk.bnpass)
            res = ops.conv2d_fwd_fused(x, fb.w, fb.b, cout, 1, s, False)
        else:
            res = x
        if s == 1 and CHAIN_EVAL_BLOCKS:
            # conv1 -> ReLU -> conv2 with the intermediate activation kept in shared memory (512^2 .. 128^2 layers)
            y = ops.block_chain_fwd(x, f1.w, f1.b, f2.w, f2.b, cout, res, y=out)
            if y is not None:
                return y
        a1 = ops.conv2d_fwd_fused(x, f1.w, f1.b, cout, 3, s, True)
        return ops.conv2d_fwd_fused(a1, f2.w, f2.b, cout, 3, 1, True, res, True, y=out)

    def _block_fwd(self, blk, x, out, training, need_grad, bn_ws):
        if FOLD_EVAL_BN and not training and not need_grad:
            return self._block_fwd_folded(blk, x, out), None
        rec = _BlockRec()
        cout = blk.conv1.out_channels
        s = blk.stride
        w1 = self._pack(blk.conv1, need_grad)
        w2 = self._pack(blk.conv2, need_grad)
        branch = False
        if blk.bypass is not None:
            wb = self._pack(blk.bypass, need_grad)
            dev = x.buf.device
            if training and self._branch(x.w // s, dev):
                # outputs are allocated by the main stream (which also frees them); only the launches move
                branch = True
                yb, stb = Act.empty(x.n, x.h // s, x.w // s, cout, dev), ops.BNState(cout, dev)
                with torch.cuda.stream(self._bstream):
                    self._conv_bn(x, wb.fwd, None, cout, 1, s, blk.bnpass, training, self._bn_ws_b, y=yb, state=stb)
        y1, st1 = self._conv_bn(x, w1.fwd, None, cout, 3, s, blk.bn1, training, bn_ws)
        a1 = ops.bn_act(y1, st1, True)
        y2, st2 = self._conv_bn(a1, w2.fwd, None, cout, 3, 1, blk.bn2, training, bn_ws)
        if blk.bypass is not None:
            if branch:
                self._branch_join()
            else:
                yb, stb = self._conv_bn(x, wb.fwd, None, cout, 1, s, blk.bnpass, training, bn_ws)
            out = ops.block_tail(y2, st2, yb, stb, out)
        else:
            yb, stb = None, None
            out = ops.block_tail(y2, st2, x, None, out)
        if need_grad:
            rec.x, rec.y1, rec.st1, rec.a1, rec.y2, rec.st2, rec.yb, rec.stb, rec.out = x, y1, st1, a1, y2, st2, yb, stb, out
        return out, rec

    def _double_fwd(self, dbl, x, out, training, need_grad, bn_ws):
        mid, r1 = self._block_fwd(dbl.res1, x, None, training, need_grad, bn_ws)
        o, r2 = self._block_fwd(dbl.res2, mid, out, training, need_grad, bn_ws)
        return o, (r1, r2)

    def forward(self, x):
        m = self.model
        ops.require_cuda(x, "UResNet input")
        if x.dim() != 4 or x.shape[1] != m.conv1.in_channels:
            raise RuntimeError("UResNet: expected input (B,%d,H,W), got %s" % (m.conv1.in_channels, tuple(x.shape)))
        if m.conv1.in_channels != 1:
            raise RuntimeError("ubresnet_b200: the bf16 tensor-core schedule takes input_channels=1 (wire-plane crops); "
                               "other channel counts run in the fp32 mode (UResNet(..., precision='fp32'))")
        if m.inplanes % 16 != 0:
            raise RuntimeError("ubresnet_b200: inplanes must be a multiple of 16 (tensor-core tile width)")
        if x.shape[2] % 32 or x.shape[3] % 32:
            raise RuntimeError("UResNet: H and W must be multiples of 32 (5 stride-2 stages), got %dx%d" % tuple(x.shape[2:]))
        if m.conv1.weight.device != x.device:
            raise RuntimeError("UResNet: parameters on %s but input on %s" % (m.conv1.weight.device, x.device))
        params = model_params(m)                         # ONE walk of the module tree per call (host time before
        trainable = torch.is_grad_enabled() and any(p.requires_grad for p in params)                  # the first launch)
        if x.requires_grad and torch.is_grad_enabled():
            raise RuntimeError("ubresnet_b200: gradients with respect to the input image are not implemented (the "
                               "reference scripts never request them); detach the input")
        if trainable and not m.training:
            raise RuntimeError("ubresnet_b200: autograd through an eval()-mode forward is not implemented; call "
                               "model.train() for training or wrap inference in torch.no_grad()")
        need_grad = trainable and m.training
        with torch.cuda.device(x.device):
            if need_grad:
                return _UResNetFn.apply(self, x, *params)
            with torch.no_grad():
                logp, _ = self._forward_impl(x.detach(), m.training, False)
            return logp

    def _forward_impl(self, x, training, need_grad):
        m = self.model
        x = x.contiguous().float()
        n, _, h, w = x.shape
        dev = x.device
        p = m.inplanes
        bn_ws, _, _ = self._workspaces(n, h, w, dev)
        saved = {}
        # concat buffers: cat[i] feeds dec_layer{i}; channels [0,Cd) deconv output, [Cd,2Cd) skip x_{i-1}
        cat = {}
        for i in (1, 2, 3, 4, 5):
            cd = p * 2 ** (i - 1)
            hh, ww = (h, w) if i == 1 else (h // 2 ** (i - 1), w // 2 ** (i - 1))
            cat[i] = torch.empty((n, hh, ww, 2 * cd), dtype=torch.bfloat16, device=dev)
        # ---- stem (ub_uresnet.py:94-97)
        ev0 = None
        if BRANCH_MAX_RES > 0 and self._bn_ws_b is not None and os.environ.get("UBR_PACK_BRANCH", "1") != "0":
            ev0 = torch.cuda.Event()
            ev0.record()
        y0, st0, x0, h0, pool_idx = self._stem_fwd(x, Act(cat[1], p, p), training, need_grad, bn_ws)
        # the stem reads the fp32 conv1 parameter directly: the (host-side) bookkeeping of the bf16 operand refresh
        # runs while its kernels are already executing -- and the re-pack launch itself goes to the branch stream, ordered
        # after everything BEFORE the stem (the optimizer step), so that it overlaps the stem's HBM-bound kernels
        if ev0 is not None:
            if self._bstream is None:
                self._bstream = torch.cuda.Stream(device=dev)
            self._bstream.wait_event(ev0)
            with torch.cuda.stream(self._bstream):
                self._refresh_packs(need_grad)
            self._branch_join()
        else:
            self._refresh_packs(need_grad)
        # ---- encoder (:103-107)
        recs = {}
        cur = h0
        for i in range(1, 6):
            co = p * 2 ** i
            out = Act(cat[i + 1], co, co) if i < 5 else None
            cur, recs["enc%d" % i] = self._double_fwd(getattr(m, "enc_layer%d" % i), cur, out, training, need_grad, bn_ws)
        # ---- decoder (:116-133; common_layers.py:127-132)
        dec_in = {}
        for i in (5, 4, 3, 2, 1):
            layer = getattr(m, "dec_layer%d" % i)
            cd = p * 2 ** (i - 1)
            wd = self._pack(layer.deconv, need_grad, transposed=True)
            self._deconv_fwd(cur, wd, cd, Act(cat[i], 0, cd))
            dec_in[i] = cur
            cur, recs["dec%d" % i] = self._double_fwd(layer.res, Act(cat[i]), None, training, need_grad, bn_ws)
        # ---- head (:137-143)
        y10, st10, a10, logp = self._head_fwd(cur, training, need_grad, bn_ws)
        if need_grad:
            saved.update(x=x, y0=y0, st0=st0, x0=x0, h0=h0, pool_idx=pool_idx, recs=recs, dec_in=dec_in, d1=cur, y10=y10, st10=st10,
                         a10=a10, logp=logp, shape=(n, h, w))
        return logp, saved

    # ------------------------------------------------------------------ backward pieces
    def _block_bwd(self, blk, rec, dout, grads, bn_ws, wg_ws, extra=None, need_dx=True):
        """Returns grad wrt the block input (optionally + `extra`, a same-shape Act)."""
        cout = blk.conv1.out_channels
        cin = blk.conv1.in_channels
        s = blk.stride
        g = grads
        has_bp = blk.bypass is not None
        dy2, dres = ops.block_tail_bwd(
            dout, rec.out, rec.y2, rec.st2, rec.yb if has_bp else None, rec.stb, blk.bnpass.weight if has_bp else None,
            bn_ws, g[blk.bn2.weight], g[blk.bn2.bias], g[blk.bnpass.weight] if has_bp else None,
            g[blk.bnpass.bias] if has_bp else None)
        w2 = self._pack(blk.conv2, True)
        branch = False
        if has_bp and need_dx:
            wb = self._pack(blk.bypass, True)
            if self._branch(rec.y2.w, dout.buf.device):
                # residual-branch data gradient beside the conv2 -> bn1 -> conv1 chain of the main stream
                branch = True
                base = Act.empty(rec.x.n, rec.x.h, rec.x.w, cin, dout.buf.device)
                with torch.cuda.stream(self._bstream):
                    ops.conv2d_dgrad(dres, wb.dgrad, cin, 1, s, dx=base)
        self._wgrad(lambda: ops.conv2d_wgrad(rec.a1, dy2, cout, 3, 1, g[blk.conv2.weight], False, wg_ws), dy2)
        # conv2 data gradient -> bn1/relu1 backward (the reduction pass rides in the conv epilogue where it can)
        dy1 = ops.conv2d_dgrad_bn_bwd(dy2, w2.dgrad, cout, 3, 1, rec.y1, rec.st1, True, bn_ws, g[blk.bn1.weight],
                                      g[blk.bn1.bias])
        del dy2
        def wg1():
            ops.conv2d_wgrad(rec.x, dy1, cout, 3, s, g[blk.conv1.weight], False, wg_ws)
            if has_bp:
                ops.conv2d_wgrad(rec.x, dres, cout, 1, s, g[blk.bypass.weight], False, wg_ws)
        self._wgrad(wg1, dy1, dres)
        if not need_dx:
            return None
        w1 = self._pack(blk.conv1, True)
        # grad wrt the block input = dgrad(conv1) + residual-branch grad (+ skip grad): the residual term is
        # produced first and the 3x3 data gradient accumulates onto it inside its epilogue where a fused path exists
        if branch:
            self._branch_join()
        elif has_bp:
            base = ops.conv2d_dgrad(dres, wb.dgrad, cin, 1, s)
        else:
            base = dres
        if ops.conv2d_dgrad_acc(dy1, w1.dgrad, cin, 3, s, base):
            return ops.add(base, extra, None, out=base) if extra is not None else base
        dx1 = ops.conv2d_dgrad(dy1, w1.dgrad, cin, 3, s)
        return ops.add(dx1, base, extra, out=dx1)

    def _double_bwd(self, dbl, recs, dout, grads, bn_ws, wg_ws, extra=None):
        r1, r2 = recs
        dmid = self._block_bwd(dbl.res2, r2, dout, grads, bn_ws, wg_ws)
        return self._block_bwd(dbl.res1, r1, dmid, grads, bn_ws, wg_ws, extra=extra)

    def backward(self, saved, dlogp):
        """dlogp: (B,C,H,W) fp32 grad wrt the log-probabilities.  Returns {param: fp32 grad}."""
        m = self.model
        n, h, w = saved["shape"]
        dev = dlogp.device
        p = m.inplanes
        bn_ws, wg_ws, stem_ws = self._workspaces(n, h, w, dev)
        sink = self.grad_sink
        grads = self._grad_targets(sink)
        self._wgrad_begin(dev)

        def done(name):
            self._wgrad_stage()
            if sink is not None and sink.wants(name):
                # the bucket's all-reduce waits (on the communication stream) for the weight gradients queued on the
                # side stream up to here; the main stream itself does not stop
                side = [self._w_prev[0]] if (self._w_on and self._w_prev is not None) else []
                sink.stage_done(name, side)
            elif sink is not None:
                sink.stage_done(name)
        # ---- head
        dcur = self._head_bwd(saved, dlogp, grads, bn_ws, wg_ws)
        done("head")
        # ---- decoder, shallow to deep
        recs = saved["recs"]
        dskip = {}
        for i in (1, 2, 3, 4, 5):
            layer = getattr(m, "dec_layer%d" % i)
            cd = p * 2 ** (i - 1)
            dcat = self._double_bwd(layer.res, recs["dec%d" % i], dcur, grads, bn_ws, wg_ws)
            ddec = dcat.slice(0, cd)
            dskip[i - 1] = dcat.slice(cd, cd)                      # grad wrt x_{i-1}
            xin = saved["dec_in"][i]
            wd = self._pack(layer.deconv, True, transposed=True)
            dcur = self._deconv_bwd(xin, ddec, wd, grads[layer.deconv.weight], wg_ws)
            done("dec%d" % i)
        # ---- encoder, deep to shallow: grad(x_i) = (from enc_{i+1} | dec5 deconv) + skip grad
        for i in (5, 4, 3, 2, 1):
            extra = dskip[i - 1] if i > 1 else None
            dcur = self._double_bwd(getattr(m, "enc_layer%d" % i), recs["enc%d" % i], dcur, grads, bn_ws, wg_ws,
                                    extra=extra)
            done("enc%d" % i)
        # ---- stem: x0 gets the pool gradient plus the dec1 skip gradient (accumulated in place)
        self._stem_bwd(saved, dcur, dskip[0], grads, bn_ws, stem_ws)
        done("stem")
        self._wgrad_join()
        if sink is not None:
            sink.finish()
        return grads


def _bias_grad(a, out, ws):
    """out[:] = per-channel sum over pixels of the first out.numel() channels of `a` (the bias gradient of a biased
    conv), written by the library into the gradient tensor: two launches, no torch op"""
    ops._lib.check(ops._lib.load().ubr_channel_sum(a.view(), out.numel(), ws.data_ptr(), out.data_ptr(), 0, ops._stream()),
                   "channel_sum")


class _UResNetFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, engine, x, *params):
        logp, saved = engine._forward_impl(x.detach(), True, True)
        ctx.engine = engine
        ctx.saved = saved
        ctx.nparams = len(params)
        ctx.param_list = params            # model.parameters() order, as passed by Engine.forward
        return logp

    @staticmethod
    def backward(ctx, dlogp):
        engine = ctx.engine
        saved = ctx.saved
        ctx.saved = None
        with torch.cuda.device(dlogp.device):
            grads = engine.backward(saved, dlogp.contiguous())
        out = [grads[p] if p.requires_grad else None for p in ctx.param_list]
        return (None, None) + tuple(out)


# ------------------------------------------------------------------------------------------------
# stand-alone sub-module execution (BasicBlock / DoubleResNet / ConvTransposeLayer called directly).
# Inference-style (no autograd through the block): the reference scripts never call blocks on their own.
# ------------------------------------------------------------------------------------------------
def run_submodule(module, *inputs):
    from .models.common_layers import BasicBlock, ConvTransposeLayer, DoubleResNet
    x = inputs[0]
    ops.require_cuda(x, type(module).__name__ + " input")
    if torch.is_grad_enabled() and module.training and any(p.requires_grad for p in module.parameters()):
        raise RuntimeError("ubresnet_b200: autograd through a stand-alone %s is not supported; call the whole "
                           "UResNet (one fused autograd node)" % type(module).__name__)
    eng = Engine(module)
    with torch.no_grad(), torch.cuda.device(x.device):
        bn_ws = ops.bn_workspace(4096, x.device)
        xa = Act.from_nchw(x.float())
        if isinstance(module, BasicBlock):
            out, _ = eng._block_fwd(module, xa, None, module.training, False, bn_ws)
        elif isinstance(module, DoubleResNet):
            out, _ = eng._double_fwd(module, xa, None, module.training, False, bn_ws)
        elif isinstance(module, ConvTransposeLayer):
            skip = inputs[1]
            cd = module.deconv.out_channels
            n, _, hh, ww = skip.shape
            catbuf = torch.empty((n, hh, ww, cd + skip.shape[1]), dtype=torch.bfloat16, device=x.device)
            Act(catbuf, cd, skip.shape[1]).torch().copy_(skip.permute(0, 2, 3, 1))
            wd = eng._pack(module.deconv, False, transposed=True)
            eng._deconv_fwd(xa, wd, cd, Act(catbuf, 0, cd))
            out, _ = eng._double_fwd(module.res, Act(catbuf), None, module.training, False, bn_ws)
        else:
            raise RuntimeError("unsupported sub-module %s" % type(module).__name__)
        return out.nchw_float()
