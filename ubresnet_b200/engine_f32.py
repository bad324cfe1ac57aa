"""fp32 precision mode of the U-ResNet schedule (UBR_PRECISION=fp32 / UResNet(..., precision="fp32")).

The reference is fp32 end to end (models/ub_uresnet.py:41-70, models/common_layers.py:13-58); BASELINE.json asks for
a mode that matches it to 1e-4.  Same network walk as `engine.Engine` (one autograd node, zero-copy concat buffers,
hand-written backward), but every tensor is fp32 NHWC and every kernel is one of the `ubr_f32_*` entry points
(csrc/fp32_path.cu): fp32 FMA on the CUDA cores, float64 reductions.  It is the parity mode -- the bf16 tensor-core
schedule is the fast one.  Any `input_channels` is accepted here (the bf16 stem kernels are single-channel).
"""
import torch

from . import _lib
from ._lib import View
from .ops import require_cuda, _ptr, _stream, _bump, logsoftmax_fwd

_NULL = View(0, 0, 0, 0, 0, 0)


class ActF:
    """channel slice [c0, c0+c) of an NHWC fp32 buffer (N,H,W,LD)"""
    __slots__ = ("buf", "c0", "c")

    def __init__(self, buf, c0=0, c=None):
        assert buf.dtype == torch.float32 and buf.dim() == 4 and buf.is_contiguous()
        self.buf, self.c0 = buf, c0
        self.c = buf.shape[3] - c0 if c is None else c

    @staticmethod
    def empty(n, h, w, c, device):
        return ActF(torch.empty((n, h, w, c), dtype=torch.float32, device=device))

    n = property(lambda s: s.buf.shape[0])
    h = property(lambda s: s.buf.shape[1])
    w = property(lambda s: s.buf.shape[2])
    npix = property(lambda s: s.buf.shape[0] * s.buf.shape[1] * s.buf.shape[2])

    def slice(self, c0, c):
        return ActF(self.buf, self.c0 + c0, c)

    def view(self):
        b = self.buf
        return View(b.data_ptr() + 4 * self.c0, b.shape[0], b.shape[1], b.shape[2], self.c, b.shape[3])

    def like(self):
        return ActF.empty(self.n, self.h, self.w, self.c, self.buf.device)


class _St:
    __slots__ = ("scale", "shift", "mean", "invstd", "count")


def _ck(rc, what):
    _lib.check(rc, what)


class EngineF32:
    def __init__(self, model):
        self.model = model
        self.keep_logits = False
        self.last_logits = None
        self.grad_sink = None
        self._ws = None

    # ------------------------------------------------------------------ primitives
    def _workspace(self, device):
        """float64 scratch: the largest weight gradient (enc5/dec5 3x3, 32p x 32p x 9) + statistics"""
        need = max(4096, (32 * self.model.inplanes) ** 2 * 16 + 4096)
        if self._ws is None or self._ws.numel() < need or self._ws.device != device:
            self._ws = torch.empty(need, dtype=torch.float64, device=device)
        return self._ws

    def _conv(self, x, conv, y=None):
        k, s, p, d = conv.kernel_size[0], conv.stride[0], conv.padding[0], conv.dilation[0]
        oh = (x.h + 2 * p - d * (k - 1) - 1) // s + 1
        ow = (x.w + 2 * p - d * (k - 1) - 1) // s + 1
        if y is None:
            y = ActF.empty(x.n, oh, ow, conv.out_channels, x.buf.device)
        _ck(self.lib.ubr_f32_conv2d_fwd(x.view(), conv.weight.data_ptr(), _ptr(conv.bias), conv.out_channels, k, s, p, d,
                                        y.view(), _stream()), "f32_conv2d_fwd")
        return y

    def _conv_bwd(self, x, conv, dy, grads, ws, need_dx=True, dx=None, accumulate=False):
        k, s, p, d = conv.kernel_size[0], conv.stride[0], conv.padding[0], conv.dilation[0]
        _ck(self.lib.ubr_f32_conv2d_wgrad(x.view(), dy.view(), k, s, p, d, grads[conv.weight].data_ptr(), 0, ws.data_ptr(),
                                          ws.numel() * 8, _stream()), "f32_conv2d_wgrad")
        if conv.bias is not None:
            _ck(self.lib.ubr_f32_channel_sum(dy.view(), ws.data_ptr(), grads[conv.bias].data_ptr(), 0, _stream()),
                "f32_channel_sum")
        if not need_dx:
            return None
        if dx is None:
            dx = x.like()
        _ck(self.lib.ubr_f32_conv2d_dgrad(dy.view(), conv.weight.data_ptr(), conv.in_channels, k, s, p, d, dx.view(),
                                          int(accumulate), _stream()), "f32_conv2d_dgrad")
        return dx

    def _bn(self, y, bn, training, ws):
        dev = y.buf.device
        st = _St()
        buf = torch.empty((4, y.c), dtype=torch.float32, device=dev)
        st.scale, st.shift, st.mean, st.invstd = buf[0], buf[1], buf[2], buf[3]
        st.count = float(y.npix)
        if training:
            _ck(self.lib.ubr_f32_bn_stats(y.view(), ws.data_ptr(), _stream()), "f32_bn_stats")
            mom = bn.momentum if bn.momentum is not None else 0.1
            _ck(self.lib.ubr_f32_bn_finalize(ws.data_ptr(), y.c, st.count, _ptr(bn.weight), _ptr(bn.bias),
                                             _ptr(bn.running_mean), _ptr(bn.running_var), mom, bn.eps, st.scale.data_ptr(),
                                             st.shift.data_ptr(), st.mean.data_ptr(), st.invstd.data_ptr(), _stream()),
                "f32_bn_finalize")
            _bump(bn.running_mean, bn.running_var)
        else:
            with torch.no_grad():
                inv = torch.rsqrt(bn.running_var.double() + bn.eps)
                sc = inv * (bn.weight.double() if bn.weight is not None else 1.0)
                st.scale.copy_(sc)
                st.shift.copy_((bn.bias.double() if bn.bias is not None else 0.0) - bn.running_mean.double() * sc)
                st.mean.copy_(bn.running_mean)
                st.invstd.copy_(inv)
        return st

    def _bn_act(self, y, st, relu, out=None):
        out = out or y.like()
        _ck(self.lib.ubr_f32_bn_act(y.view(), st.scale.data_ptr(), st.shift.data_ptr(), int(relu), out.view(), _stream()),
            "f32_bn_act")
        return out

    def _bn_act_bwd(self, dout, y, st, relu, bn, grads, ws):
        dy = y.like()
        _ck(self.lib.ubr_f32_bn_act_bwd(dout.view(), y.view(), st.scale.data_ptr(), st.shift.data_ptr(), st.mean.data_ptr(),
                                        st.invstd.data_ptr(), int(relu), ws.data_ptr(), st.count, dy.view(),
                                        grads[bn.weight].data_ptr(), grads[bn.bias].data_ptr(), 0, _stream()),
            "f32_bn_act_bwd")
        return dy

    def _add(self, a, b, c=None, out=None):
        out = out or a.like()
        _ck(self.lib.ubr_f32_add(a.view(), b.view() if b is not None else _NULL, c.view() if c is not None else _NULL,
                                 out.view(), _stream()), "f32_add")
        return out

    # ------------------------------------------------------------------ residual block (common_layers.py:39-58)
    def _block_fwd(self, blk, x, out, training, ws):
        y1 = self._conv(x, blk.conv1)
        st1 = self._bn(y1, blk.bn1, training, ws)
        a1 = self._bn_act(y1, st1, True)
        y2 = self._conv(a1, blk.conv2)
        st2 = self._bn(y2, blk.bn2, training, ws)
        if blk.bypass is not None:
            yb = self._conv(x, blk.bypass)
            stb = self._bn(yb, blk.bnpass, training, ws)
            res, sr, br = yb, stb.scale.data_ptr(), stb.shift.data_ptr()
        else:
            yb = stb = None
            res, sr, br = x, 0, 0
        if out is None:
            out = y2.like()
        _ck(self.lib.ubr_f32_block_tail(y2.view(), st2.scale.data_ptr(), st2.shift.data_ptr(), res.view(), sr, br,
                                        out.view(), _stream()), "f32_block_tail")
        return out, (x, y1, st1, a1, y2, st2, yb, stb, out)

    def _block_bwd(self, blk, rec, dout, grads, ws, extra=None):
        x, y1, st1, a1, y2, st2, yb, stb, out = rec
        du = out.like()
        _ck(self.lib.ubr_f32_relu_mask(dout.view(), out.view(), du.view(), _stream()), "f32_relu_mask")
        dy2 = self._bn_act_bwd(du, y2, st2, True, blk.bn2, grads, ws)
        da1 = self._conv_bwd(a1, blk.conv2, dy2, grads, ws)
        dy1 = self._bn_act_bwd(da1, y1, st1, True, blk.bn1, grads, ws)
        dx = self._conv_bwd(x, blk.conv1, dy1, grads, ws)
        if blk.bypass is not None:
            dyb = self._bn_act_bwd(du, yb, stb, False, blk.bnpass, grads, ws)
            self._conv_bwd(x, blk.bypass, dyb, grads, ws, dx=dx, accumulate=True)
            return self._add(dx, extra, out=dx) if extra is not None else dx
        return self._add(dx, du, extra, out=dx)

    def _double_fwd(self, dbl, x, out, training, ws):
        mid, r1 = self._block_fwd(dbl.res1, x, None, training, ws)
        o, r2 = self._block_fwd(dbl.res2, mid, out, training, ws)
        return o, (r1, r2)

    def _double_bwd(self, dbl, recs, dout, grads, ws, extra=None):
        dmid = self._block_bwd(dbl.res2, recs[1], dout, grads, ws)
        return self._block_bwd(dbl.res1, recs[0], dmid, grads, ws, extra=extra)

    # ------------------------------------------------------------------ whole network (ub_uresnet.py:88-147)
    def forward(self, x):
        m = self.model
        require_cuda(x, "UResNet input")
        if x.dim() != 4 or x.shape[1] != m.conv1.in_channels:
            raise RuntimeError("UResNet: expected input (B,%d,H,W), got %s" % (m.conv1.in_channels, tuple(x.shape)))
        if x.shape[2] % 32 or x.shape[3] % 32:
            raise RuntimeError("UResNet: H and W must be multiples of 32 (5 stride-2 stages), got %dx%d" % tuple(x.shape[2:]))
        if m.conv1.weight.device != x.device:
            raise RuntimeError("UResNet: parameters on %s but input on %s" % (m.conv1.weight.device, x.device))
        self.lib = _lib.load()
        from .engine import model_params
        params = model_params(m)
        trainable = torch.is_grad_enabled() and any(p.requires_grad for p in params)
        if x.requires_grad and torch.is_grad_enabled():
            raise RuntimeError("ubresnet_b200: gradients with respect to the input image are not implemented")
        if trainable and not m.training:
            raise RuntimeError("ubresnet_b200: autograd through an eval()-mode forward is not implemented; call "
                               "model.train() for training or wrap inference in torch.no_grad()")
        with torch.cuda.device(x.device):
            if trainable:
                return _F32Fn.apply(self, x, *params)
            with torch.no_grad():
                return self._forward_impl(x.detach(), m.training, False)[0]

    def _forward_impl(self, x, training, need_grad):
        m = self.model
        p = m.inplanes
        n, cin, h, w = x.shape
        dev = x.device
        ws = self._workspace(dev)
        xin = ActF(x.float().permute(0, 2, 3, 1).contiguous())           # NCHW -> NHWC (free for 1 channel)
        cat = {}
        for i in (1, 2, 3, 4, 5):
            cd = p * 2 ** (i - 1)
            hh, ww = (h, w) if i == 1 else (h // 2 ** (i - 1), w // 2 ** (i - 1))
            cat[i] = torch.empty((n, hh, ww, 2 * cd), dtype=torch.float32, device=dev)
        y0 = self._conv(xin, m.conv1)
        st0 = self._bn(y0, m.bn1, training, ws)
        x0 = self._bn_act(y0, st0, True, out=ActF(cat[1], p, p))
        h0 = ActF.empty(n, h // 2, w // 2, p, dev)
        pool_idx = torch.empty((n, h // 2, w // 2, p), dtype=torch.uint8, device=dev)
        _ck(self.lib.ubr_f32_maxpool3s2_fwd(x0.view(), h0.view(), pool_idx.data_ptr(), _stream()), "f32_maxpool3s2_fwd")
        recs = {}
        cur = h0
        for i in range(1, 6):
            co = p * 2 ** i
            out = ActF(cat[i + 1], co, co) if i < 5 else None
            cur, recs["enc%d" % i] = self._double_fwd(getattr(m, "enc_layer%d" % i), cur, out, training, ws)
        dec_in = {}
        for i in (5, 4, 3, 2, 1):
            layer = getattr(m, "dec_layer%d" % i)
            cd = p * 2 ** (i - 1)
            dc = layer.deconv
            _ck(self.lib.ubr_f32_deconv_fwd(cur.view(), dc.weight.data_ptr(), cd, dc.kernel_size[0], dc.stride[0],
                                            dc.padding[0], ActF(cat[i], 0, cd).view(), _stream()), "f32_deconv_fwd")
            dec_in[i] = cur
            cur, recs["dec%d" % i] = self._double_fwd(layer.res, ActF(cat[i]), None, training, ws)
        y10 = self._conv(cur, m.conv10)
        st10 = self._bn(y10, m.bn10, training, ws)
        a10 = self._bn_act(y10, st10, True)
        ncls = m.conv11.out_channels
        logits = self._conv(a10, m.conv11)
        logp = logsoftmax_fwd(logits.buf.view(n * h * w, ncls), n, ncls, h, w)
        if self.keep_logits:
            self.last_logits = logits.buf.permute(0, 3, 1, 2).contiguous()
        if training:
            from .engine import Engine
            Engine._bump_batches_tracked(self)
        saved = None
        if need_grad:
            saved = dict(xin=xin, y0=y0, st0=st0, x0=x0, pool_idx=pool_idx, recs=recs, dec_in=dec_in, d1=cur, y10=y10,
                         st10=st10, a10=a10, logp=logp, shape=(n, h, w))
        return logp, saved

    def backward(self, saved, dlogp):
        m = self.model
        p = m.inplanes
        n, h, w = saved["shape"]
        dev = dlogp.device
        ws = self._workspace(dev)
        from .engine import model_params
        grads = {prm: torch.empty_like(prm, dtype=torch.float32) for prm in model_params(m)}
        ncls = m.conv11.out_channels
        dl = ActF.empty(n, h, w, ncls, dev)
        _ck(self.lib.ubr_f32_logsoftmax_bwd(saved["logp"].data_ptr(), dlogp.contiguous().float().data_ptr(), n, ncls, h * w,
                                            dl.view(), _stream()), "f32_logsoftmax_bwd")
        da10 = self._conv_bwd(saved["a10"], m.conv11, dl, grads, ws)
        dy10 = self._bn_act_bwd(da10, saved["y10"], saved["st10"], True, m.bn10, grads, ws)
        dcur = self._conv_bwd(saved["d1"], m.conv10, dy10, grads, ws)
        recs = saved["recs"]
        dskip = {}
        for i in (1, 2, 3, 4, 5):
            layer = getattr(m, "dec_layer%d" % i)
            cd = p * 2 ** (i - 1)
            dcat = self._double_bwd(layer.res, recs["dec%d" % i], dcur, grads, ws)
            ddec, dskip[i - 1] = dcat.slice(0, cd), dcat.slice(cd, cd)
            xin = saved["dec_in"][i]
            dc = layer.deconv
            k, s, pd = dc.kernel_size[0], dc.stride[0], dc.padding[0]
            _ck(self.lib.ubr_f32_deconv_wgrad(xin.view(), ddec.view(), k, s, pd, grads[dc.weight].data_ptr(), 0, ws.data_ptr(),
                                              ws.numel() * 8, _stream()), "f32_deconv_wgrad")
            dcur = xin.like()
            _ck(self.lib.ubr_f32_deconv_dgrad(ddec.view(), dc.weight.data_ptr(), dc.in_channels, k, s, pd, dcur.view(),
                                              _stream()), "f32_deconv_dgrad")
        for i in (5, 4, 3, 2, 1):
            extra = dskip[i - 1] if i > 1 else None
            dcur = self._double_bwd(getattr(m, "enc_layer%d" % i), recs["enc%d" % i], dcur, grads, ws, extra=extra)
        # stem: x0 gets the pool gradient plus the dec1 skip gradient
        dx0 = dskip[0]
        _ck(self.lib.ubr_f32_maxpool3s2_bwd(saved["pool_idx"].data_ptr(), dcur.view(), dx0.view(), 1, _stream()),
            "f32_maxpool3s2_bwd")
        dy0 = self._bn_act_bwd(dx0, saved["y0"], saved["st0"], True, m.bn1, grads, ws)
        self._conv_bwd(saved["xin"], m.conv1, dy0, grads, ws, need_dx=False)
        return grads


class _F32Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, engine, x, *params):
        logp, saved = engine._forward_impl(x.detach(), True, True)
        ctx.engine, ctx.saved, ctx.param_list = engine, saved, params
        return logp

    @staticmethod
    def backward(ctx, dlogp):
        saved, ctx.saved = ctx.saved, None
        with torch.cuda.device(dlogp.device):
            grads = ctx.engine.backward(saved, dlogp)
        return (None, None) + tuple(grads[p] if p.requires_grad else None for p in ctx.param_list)
