"""Kernel schedule of ASPP_ResNet (reference models/ASPP_ResNet.py:416-523) on top of the U-ResNet engine.

Same building blocks as `engine.Engine` (stem, BasicBlock / DoubleResNet, ConvTranspose + zero-copy concat, head);
what is new here is the ASPP skip path on enc3 / enc4 / enc5:

    b1..b4 = relu(bn(conv_{1x1 | 3x3 d1 | 3x3 d3 | 3x3 d5}(e)))    16 channels each, biased    (:221-239)
    b5     = maxpool3x3/s1(e)                                                                  (:242)
    aspp   = relu(bn(conv1x1(cat[b1..b5])))                         biased, Cin channels       (:281-283)
    skip   = cat[aspp, e]                                                                      (:459,471,483)

All concatenations are zero-copy: the branch activations and the pooled input are written straight into channel
slices of the (64 + Cin)-channel ASPP buffer, and `aspp` / `e` into slices of the decoder's concat buffer
[deconv | aspp | e].  The dilated branches run on the tcgen05 implicit-GEMM kernel with dilation-scaled taps
(`ubr_conv2d_dilated_{fwd,dgrad,wgrad}`).
"""
import torch

from . import engine as _engine
from . import ops
from .engine import Engine, _UResNetFn, _bias_grad  # noqa: F401
from .ops import Act

_BRANCHES = (("B1", 1, 1), ("B2", 3, 1), ("B3", 3, 3), ("B4", 3, 5))      # (prefix, ksize, dilation)


class _AsppRec:
    __slots__ = ("e", "ys", "sts", "acat", "idx", "yc", "stc")


class ASPPEngine(Engine):

    # ------------------------------------------------------------------ workspace sizing
    def _wgrad_shapes(self, n, h, w):
        m = self.model
        p = m.inplanes
        pad16 = lambda v: (v + 15) // 16 * 16
        shapes = [(n, h, w, pad16(m.nkernels), p, 49), (n, h, w, 16, p, 49)]
        for i in range(1, 6):
            ci, co = p * 2 ** (i - 1), p * 2 ** i
            hi, wi = h // 2 ** i, w // 2 ** i
            shapes += [(n, hi, wi, co, ci, 9), (n, hi, wi, co, co, 9), (n, hi, wi, co, ci, 1)]
        for lvl in (3, 4, 5):                      # ASPP on enc{lvl}: Cin = p * 2^lvl at H / 2^lvl
            ci = p * 2 ** lvl
            hi, wi = h // 2 ** lvl, w // 2 ** lvl
            shapes += [(n, hi, wi, 16, ci, 1), (n, hi, wi, 16, ci, 9), (n, hi, wi, ci, ci + 64, 1)]
        for i in (5, 4, 3, 2, 1):
            layer = getattr(m, "dec_layer%d" % i)
            ci, cd = layer.deconv.in_channels, layer.deconv.out_channels
            cr_in, cr = layer.res.res1.conv1.in_channels, layer.res.res1.conv1.out_channels
            hin, win = h // 2 ** i, w // 2 ** i
            shapes += [(n, hin, win, ci, cd, 16), (n, 2 * hin, 2 * win, cr, cr_in, 9), (n, 2 * hin, 2 * win, cr, cr, 9),
                       (n, 2 * hin, 2 * win, cr, cr_in, 1)]
            if 4 * cd <= 64:
                shapes.append((n, hin, win, 4 * cd, ci, 9))
        return shapes

    # ------------------------------------------------------------------ ASPP forward / backward
    def _aspp_fwd(self, layer, post, e, out, training, need_grad, bn_ws):
        """e: Act (Cin channels).  Writes relu(bn(conv1x1(cat[b1..b4, maxpool(e)]))) into `out`."""
        rec = _AsppRec()
        cin = e.c
        acat = torch.empty((e.n, e.h, e.w, 64 + cin), dtype=torch.bfloat16, device=e.buf.device)
        if _engine.FOLD_EVAL_BN and not training and not need_grad:
            # inference: BatchNorm folded into every branch conv and into ASPP_conv, ReLU in the conv epilogues, branch
            # outputs written straight into their channel slices of the concat buffer
            for j, (pre, k, dil) in enumerate(_BRANCHES):
                f = self._folded(getattr(layer, pre + "_conv"), getattr(layer, pre + "_bn"))
                dst = Act(acat, 16 * j, 16)
                if dil == 1:
                    ops.conv2d_fwd_fused(e, f.w, f.b, 16, k, 1, True, y=dst)
                else:
                    ops.conv2d_dilated_fwd(e, f.w, f.b, 16, dil, y=dst, relu=True)
            ops.maxpool3s1_fwd(e, y=Act(acat, 64, cin), want_idx=False)
            fc = self._folded(post.ASPP_conv, post.ASPP_bn)
            ops.conv2d_fwd_fused(Act(acat), fc.w, fc.b, cin, 1, 1, True, y=out)
            return rec
        ys, sts = [], []
        for j, (pre, k, dil) in enumerate(_BRANCHES):
            conv = getattr(layer, pre + "_conv")
            bn = getattr(layer, pre + "_bn")
            wpk = self._pack(conv, need_grad)
            if dil == 1:
                y, st = self._conv_bn(e, wpk.fwd, conv.bias, 16, k, 1, bn, training, bn_ws)
            else:
                y = ops.conv2d_dilated_fwd(e, wpk.fwd, conv.bias, 16, dil)
                st = self._bn(y, bn, training, bn_ws)
            ops.bn_act(y, st, True, out=Act(acat, 16 * j, 16))
            ys.append(y)
            sts.append(st)
        _, idx = ops.maxpool3s1_fwd(e, y=Act(acat, 64, cin), want_idx=need_grad)
        wc = self._pack(post.ASPP_conv, need_grad)
        yc, stc = self._conv_bn(Act(acat), wc.fwd, post.ASPP_conv.bias, cin, 1, 1, post.ASPP_bn, training, bn_ws)
        ops.bn_act(yc, stc, True, out=out)
        if need_grad:
            rec.e, rec.ys, rec.sts, rec.acat, rec.idx, rec.yc, rec.stc = e, ys, sts, Act(acat), idx, yc, stc
        return rec

    def _aspp_bwd(self, layer, post, rec, d_aspp, acc, grads, bn_ws, wg_ws):
        """d_aspp: grad wrt the ASPP_post output.  Adds the gradient wrt the ASPP input e into `acc` (in place)."""
        g = grads
        cin = rec.e.c
        dyc = ops.bn_act_bwd(d_aspp, rec.yc, rec.stc, True, bn_ws, g[post.ASPP_bn.weight], g[post.ASPP_bn.bias])
        self._wgrad(lambda d=dyc: ops.conv2d_wgrad(rec.acat, d, cin, 1, 1, g[post.ASPP_conv.weight], False, wg_ws), dyc)
        _bias_grad(dyc, g[post.ASPP_conv.bias], bn_ws)
        wc = self._pack(post.ASPP_conv, True)
        dacat = ops.conv2d_dgrad(dyc, wc.dgrad, 64 + cin, 1, 1)
        del dyc
        ops.maxpool3s1_bwd_idx(rec.idx, dacat.slice(64, cin), acc, True)
        parts = []
        for j, (pre, k, dil) in enumerate(_BRANCHES):
            conv = getattr(layer, pre + "_conv")
            bn = getattr(layer, pre + "_bn")
            dyb = ops.bn_act_bwd(dacat.slice(16 * j, 16), rec.ys[j], rec.sts[j], True, bn_ws, g[bn.weight], g[bn.bias])
            wpk = self._pack(conv, True)
            if k == 1:
                self._wgrad(lambda d=dyb, c=conv: ops.conv2d_wgrad(rec.e, d, 16, 1, 1, g[c.weight], False, wg_ws), dyb)
                parts.append(ops.conv2d_dgrad(dyb, wpk.dgrad, cin, 1, 1))
            else:
                self._wgrad(lambda d=dyb, c=conv, dl=dil: ops.conv2d_dilated_wgrad(rec.e, d, 16, dl, g[c.weight], False,
                                                                                 wg_ws), dyb)
                parts.append(ops.conv2d_dilated_dgrad(dyb, wpk.dgrad, cin, dil))
            _bias_grad(dyb, g[conv.bias], bn_ws)
            if len(parts) == 2:
                ops.add(acc, parts[0], parts[1], out=acc)
                parts = []
        return acc

    # ------------------------------------------------------------------ whole network
    def _forward_impl(self, x, training, need_grad):
        m = self.model
        x = x.contiguous().float()
        n, _, h, w = x.shape
        dev = x.device
        p = m.inplanes
        if m.conv10.in_channels != p or m.conv11.in_channels != m.nkernels:
            raise RuntimeError("ASPP_ResNet: the reference head (conv11 in_channels == inplanes, nkernels == 16, "
                               "models/ASPP_ResNet.py:376-386) is only consistent at inplanes=16")
        bn_ws, _, _ = self._workspaces(n, h, w, dev)
        new = lambda hh, ww, c: torch.empty((n, hh, ww, c), dtype=torch.bfloat16, device=dev)
        # decoder concat buffers [deconv | skip]; skip of dec5 / dec4 is itself [aspp | e]
        cat = {1: new(h, w, 2 * p), 2: new(h // 2, w // 2, 4 * p), 3: new(h // 4, w // 4, 8 * p),
               4: new(h // 8, w // 8, 24 * p), 5: new(h // 16, w // 16, 48 * p)}
        e5skip = new(h // 32, w // 32, 64 * p)
        y0, st0, x0, h0, pool_idx = self._stem_fwd(x, Act(cat[1], p, p), training, need_grad, bn_ws)
        self._refresh_packs(need_grad)          # after the stem launches (they do not use packed operands)
        # ---- encoder: e_i lands in the slice of the buffer that consumes it as a skip
        e_out = {1: Act(cat[2], 2 * p, 2 * p), 2: Act(cat[3], 4 * p, 4 * p), 3: Act(cat[4], 16 * p, 8 * p),
                 4: Act(cat[5], 32 * p, 16 * p), 5: Act(e5skip, 32 * p, 32 * p)}
        a_out = {3: Act(cat[4], 8 * p, 8 * p), 4: Act(cat[5], 16 * p, 16 * p), 5: Act(e5skip, 0, 32 * p)}
        recs = {}
        cur = h0
        for i in range(1, 6):
            cur, recs["enc%d" % i] = self._double_fwd(getattr(m, "enc_layer%d" % i), cur, e_out[i], training, need_grad,
                                                      bn_ws)
        # ---- ASPP on e3, e4, e5 (:448-483)
        for lvl in (3, 4, 5):
            recs["aspp%d" % lvl] = self._aspp_fwd(getattr(m, "ASPP_layer_enc%d" % lvl), getattr(m, "ASPP_combine_enc%d" % lvl),
                                                  e_out[lvl], a_out[lvl], training, need_grad, bn_ws)
        # ---- decoder (:488-503)
        dec_in = {}
        cur = Act(e5skip)
        for i in (5, 4, 3, 2, 1):
            layer = getattr(m, "dec_layer%d" % i)
            cd = layer.deconv.out_channels
            wd = self._pack(layer.deconv, need_grad, transposed=True)
            self._deconv_fwd(cur, wd, cd, Act(cat[i], 0, cd))
            dec_in[i] = cur
            cur, recs["dec%d" % i] = self._double_fwd(layer.res, Act(cat[i]), None, training, need_grad, bn_ws)
        y10, st10, a10, logp = self._head_fwd(cur, training, need_grad, bn_ws)
        saved = {}
        if need_grad:
            saved.update(x=x, y0=y0, st0=st0, x0=x0, h0=h0, pool_idx=pool_idx, recs=recs, dec_in=dec_in, d1=cur, y10=y10,
                         st10=st10, a10=a10, logp=logp, shape=(n, h, w))
        return logp, saved

    def backward(self, saved, dlogp):
        m = self.model
        n, h, w = saved["shape"]
        p = m.inplanes
        bn_ws, wg_ws, stem_ws = self._workspaces(n, h, w, dlogp.device)
        sink = self.grad_sink
        grads = self._grad_targets(sink)     # data-parallel: gradients land in the flat all-reduce buckets (parallel.py)
        self._wgrad_begin(dlogp.device)     # every weight gradient (ASPP layers included) goes to the side stream

        def done(name):
            self._wgrad_stage()
            if sink is not None:
                side = [self._w_prev[0]] if (self._w_on and self._w_prev is not None) else []
                sink.stage_done(name, side)
        recs = saved["recs"]
        dcur = self._head_bwd(saved, dlogp, grads, bn_ws, wg_ws)
        done("head")
        # ---- decoder, shallow to deep; dcat splits into [ddeconv | dskip]
        dskip = {}
        for i in (1, 2, 3, 4, 5):
            layer = getattr(m, "dec_layer%d" % i)
            cd = layer.deconv.out_channels
            dcat = self._double_bwd(layer.res, recs["dec%d" % i], dcur, grads, bn_ws, wg_ws)
            ddec = dcat.slice(0, cd)
            dskip[i] = dcat.slice(cd, dcat.c - cd)
            wd = self._pack(layer.deconv, True, transposed=True)
            dcur = self._deconv_bwd(saved["dec_in"][i], ddec, wd, grads[layer.deconv.weight], wg_ws)
            done("dec%d" % i)
        # dcur = grad wrt cat[aspp5, e5]; dskip[5] = grad wrt cat[aspp4, e4]; dskip[4] = grad wrt cat[aspp3, e3]
        d_as = {5: dcur.slice(0, 32 * p), 4: dskip[5].slice(0, 16 * p), 3: dskip[4].slice(0, 8 * p)}
        d_e = {5: dcur.slice(32 * p, 32 * p), 4: dskip[5].slice(16 * p, 16 * p), 3: dskip[4].slice(8 * p, 8 * p),
               2: dskip[3], 1: dskip[2]}
        for lvl in (5, 4, 3):                      # ASPP input gradients accumulate onto the direct skip gradients
            self._aspp_bwd(getattr(m, "ASPP_layer_enc%d" % lvl), getattr(m, "ASPP_combine_enc%d" % lvl),
                           recs["aspp%d" % lvl], d_as[lvl], d_e[lvl], grads, bn_ws, wg_ws)
        # ---- encoder, deep to shallow: grad(e_i) = (from enc_{i+1}) + skip (+ ASPP) gradient
        dcur = d_e[5]
        for i in (5, 4, 3, 2, 1):
            extra = d_e[i - 1] if i > 1 else None
            dcur = self._double_bwd(getattr(m, "enc_layer%d" % i), recs["enc%d" % i], dcur, grads, bn_ws, wg_ws, extra=extra)
            done("enc%d" % i)
        self._stem_bwd(saved, dcur, dskip[1], grads, bn_ws, stem_ws)
        done("stem")
        self._wgrad_join()
        if sink is not None:
            sink.finish()
        return grads


def run_aspp_submodule(module, x):
    """Stand-alone ASPP / ASPP_post call (inference-style; the reference scripts never call them on their own)."""
    from .models.ASPP_ResNet import ASPP, ASPP_post
    ops.require_cuda(x, type(module).__name__ + " input")
    if torch.is_grad_enabled() and module.training and any(p.requires_grad for p in module.parameters()):
        raise RuntimeError("ubresnet_b200: autograd through a stand-alone %s is not supported; call the whole "
                           "ASPP_ResNet (one fused autograd node)" % type(module).__name__)
    eng = ASPPEngine(module)
    with torch.no_grad(), torch.cuda.device(x.device):
        bn_ws = ops.bn_workspace(4096, x.device)
        xa = Act.from_nchw(x.float())
        if isinstance(module, ASPP):
            n, cin, h, w = x.shape
            acat = torch.empty((n, h, w, 64 + cin), dtype=torch.bfloat16, device=x.device)
            for j, (pre, k, dil) in enumerate(_BRANCHES):
                conv, bn = getattr(module, pre + "_conv"), getattr(module, pre + "_bn")
                wpk = eng._pack(conv, False)
                if dil == 1:
                    y, st = eng._conv_bn(xa, wpk.fwd, conv.bias, 16, k, 1, bn, module.training, bn_ws)
                else:
                    y = ops.conv2d_dilated_fwd(xa, wpk.fwd, conv.bias, 16, dil)
                    st = eng._bn(y, bn, module.training, bn_ws)
                ops.bn_act(y, st, True, out=Act(acat, 16 * j, 16))
            ops.maxpool3s1_fwd(xa, y=Act(acat, 64, cin))
            return Act(acat).nchw_float()
        if isinstance(module, ASPP_post):
            wc = eng._pack(module.ASPP_conv, False)
            yc, stc = eng._conv_bn(xa, wc.fwd, module.ASPP_conv.bias, module.outplanes, 1, 1, module.ASPP_bn,
                                   module.training, bn_ws)
            return ops.bn_act(yc, stc, True).nchw_float()
        raise RuntimeError("unsupported sub-module %s" % type(module).__name__)
