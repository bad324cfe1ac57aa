"""Thin torch-tensor wrappers over the C ABI.  Device memory, streams and allocation are PyTorch's;
every byte of arithmetic happens in libubresnet_b200.so.  No fallback paths.
"""
import os

import torch

from . import _lib
from ._lib import View

BN_EPS = 1e-5
BN_MOMENTUM = 0.1
# Reductions (BatchNorm statistics, BatchNorm-backward sums) are finished inside the producing launch by the last CTAs
# to arrive (csrc/tree_reduce.cuh) instead of by a separate finalize launch; UBR_TREE_FINALIZE=0 keeps the two-launch form.
TREE_FINALIZE = os.environ.get("UBR_TREE_FINALIZE", "1") != "0"
# ... and the same inside the CONV launches that emit statistics from their epilogue (ubr_conv2d_fwd_bn,
# ubr_conv2d_dgrad_bnreduce_final).  Off by default: the step time is the same either way (the finalize launch is hidden
# by programmatic dependent launch, the in-launch tail is a ~5-10 us single-CTA epilogue), and with the tail inside, a
# conv kernel's measured duration -- the denominator of its roofline fraction -- would carry the finalize as well.
TREE_FINALIZE_CONV = TREE_FINALIZE and os.environ.get("UBR_TREE_FINALIZE_CONV", "0") != "0"


def _tree(ws):
    """device pointer of the arrival-ticket tree that rides on a workspace made by bn_workspace(), or 0"""
    t = getattr(ws, "tree", None) if TREE_FINALIZE else None
    return 0 if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return 0 if t is None else t.data_ptr()


def _bump(*tensors):
    """the kernels write these tensors through raw pointers: advance their version counters like an in-place torch op
    would, so that caches keyed on `_version` (the eval-mode BatchNorm fold, engine._FoldedConv) see the update"""
    ts = [t for t in tensors if t is not None]
    if ts:
        torch.autograd.graph.increment_version(ts)


def require_cuda(t, name):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError("ubresnet_b200: %s must be a CUDA tensor (there is no CPU path)" % name)


class Act:
    """Channel slice [c0, c0+c) of an NHWC bf16 buffer (N,H,W,LD)."""
    __slots__ = ("buf", "c0", "c")

    def __init__(self, buf, c0=0, c=None):
        assert buf.dtype == torch.bfloat16 and buf.dim() == 4 and buf.is_contiguous()
        self.buf = buf
        self.c0 = c0
        self.c = buf.shape[3] - c0 if c is None else c
        assert 0 <= c0 and c0 + self.c <= buf.shape[3]

    @staticmethod
    def empty(n, h, w, c, device, ld=None):
        return Act(torch.empty((n, h, w, ld or c), dtype=torch.bfloat16, device=device), 0, c)

    @staticmethod
    def zeros(n, h, w, c, device, ld=None):
        return Act(torch.zeros((n, h, w, ld or c), dtype=torch.bfloat16, device=device), 0, c)

    @property
    def n(self): return self.buf.shape[0]

    @property
    def h(self): return self.buf.shape[1]

    @property
    def w(self): return self.buf.shape[2]

    @property
    def npix(self): return self.buf.shape[0] * self.buf.shape[1] * self.buf.shape[2]

    def slice(self, c0, c):
        return Act(self.buf, self.c0 + c0, c)

    def view(self):
        b = self.buf
        return View(b.data_ptr() + 2 * self.c0, b.shape[0], b.shape[1], b.shape[2], self.c, b.shape[3])

    def torch(self):
        return self.buf[..., self.c0:self.c0 + self.c]

    def nchw_float(self):
        return self.torch().permute(0, 3, 1, 2).float().contiguous()

    @staticmethod
    def from_nchw(x, ld=None, c0=0):
        n, c, h, w = x.shape
        a = Act.zeros(n, h, w, c, x.device, ld=ld) if ld is None else Act(
            torch.zeros((n, h, w, ld), dtype=torch.bfloat16, device=x.device), c0, c)
        a.torch().copy_(x.permute(0, 2, 3, 1))
        return a


_NULLVIEW = View(0, 0, 0, 0, 0, 0)


# ------------------------------------------------------------------------------------------------
def pack_weights(w, transposed, n_is_out):
    """fp32 OIHW (conv) / IOHW (deconv) parameter -> bf16 [taps+1][n_pad][k_pad] GEMM operand."""
    require_cuda(w, "weight")
    lib = _lib.load()
    w = w.detach().contiguous().float()
    if transposed:
        in_ch, out_ch, kh, kw = w.shape
    else:
        out_ch, in_ch, kh, kw = w.shape
    nn, kk = (out_ch, in_ch) if n_is_out else (in_ch, out_ch)
    n_pad, k_pad = (nn + 15) // 16 * 16, (kk + 15) // 16 * 16
    dst = torch.empty((kh * kw + 1, n_pad, k_pad), dtype=torch.bfloat16, device=w.device)
    _lib.check(lib.ubr_pack_weights(w.data_ptr(), 1 if transposed else 0, 1 if n_is_out else 0, out_ch, in_ch, kh, kw,
                                    n_pad, k_pad, dst.data_ptr(), _stream()), "pack_weights")
    return dst


def pack_weights_batch(table, njobs):
    """table: device int64 [njobs, 10] job records (see ubr_pack_weights_batch)."""
    lib = _lib.load()
    _lib.check(lib.ubr_pack_weights_batch(table.data_ptr(), int(njobs), _stream()), "pack_weights_batch")


def conv2d_fwd(x, wpk, bias, out_ch, ksize, stride, y=None, y_f32=None):
    lib = _lib.load()
    if y_f32 is not None:
        _lib.check(lib.ubr_conv2d_fwd(x.view(), wpk.data_ptr(), _ptr(bias), out_ch, ksize, stride, _NULLVIEW,
                                      y_f32.data_ptr(), y_f32.shape[-1], _stream()), "conv2d_fwd")
        return y_f32
    if y is None:
        y = Act.empty(x.n, x.h // stride, x.w // stride, out_ch, x.buf.device)
    _lib.check(lib.ubr_conv2d_fwd(x.view(), wpk.data_ptr(), _ptr(bias), out_ch, ksize, stride, y.view(), 0, 0,
                                  _stream()), "conv2d_fwd")
    return y


def conv2d_fwd_fused(x, wpk, bias, out_ch, ksize, stride, relu_pre, res=None, relu_post=False, y=None):
    """inference conv with BatchNorm folded into (wpk, bias): y = relu_post?(relu_pre?(conv(x) + bias) + res)"""
    lib = _lib.load()
    if y is None:
        y = Act.empty(x.n, x.h // stride, x.w // stride, out_ch, x.buf.device)
    _lib.check(lib.ubr_conv2d_fwd_fused(x.view(), wpk.data_ptr(), _ptr(bias), out_ch, ksize, stride, int(bool(relu_pre)),
                                        res.view() if res is not None else _NULLVIEW, int(bool(relu_post)), y.view(),
                                        _stream()), "conv2d_fwd_fused")
    return y


def block_chain_fwd(x, w1pk, b1, w2pk, b2, planes, res, y=None):
    """eval-mode residual block conv3x3 -> ReLU -> conv3x3 -> ReLU -> (+res) -> ReLU in one launch, the intermediate
    activation kept on chip (csrc/chain_halo.cu).  Returns None when the shape is not eligible (nothing launched)."""
    lib = _lib.load()
    if y is None:
        y = Act.empty(x.n, x.h, x.w, planes, x.buf.device)
    rc = lib.ubr_block_chain_fwd(x.view(), w1pk.data_ptr(), b1.data_ptr(), w2pk.data_ptr(), b2.data_ptr(), planes,
                                 res.view(), y.view(), _stream())
    if rc == 1:
        return None
    _lib.check(rc, "block_chain_fwd")
    return y


def conv2d_fwd_bn(x, wpk, bias, out_ch, ksize, stride, gamma, beta, running_mean, running_var, ws, y=None,
                  momentum=BN_MOMENTUM, eps=BN_EPS, state=None):
    """conv forward + training-mode BatchNorm statistics.  Where the conv kernel has a fused-statistics epilogue the
    stand-alone bn_stats pass over y is skipped.  Returns (y, BNState)."""
    import ctypes
    lib = _lib.load()
    if y is None:
        y = Act.empty(x.n, x.h // stride, x.w // stride, out_ch, x.buf.device)
    tree = _tree(ws) if TREE_FINALIZE_CONV else 0
    if tree:
        st = state or BNState(y.c, y.buf.device)
        st.count = float(y.npix)
        rc = lib.ubr_conv2d_fwd_bn(x.view(), wpk.data_ptr(), _ptr(bias), out_ch, ksize, stride, y.view(), ws.data_ptr(), tree,
                                   _ptr(gamma), _ptr(beta), _ptr(running_mean), _ptr(running_var), momentum, eps,
                                   st.scale.data_ptr(), st.shift.data_ptr(), st.mean.data_ptr(), st.invstd.data_ptr(),
                                   _stream())
        if rc == 1:
            conv2d_fwd(x, wpk, bias, out_ch, ksize, stride, y=y)
            return y, bn_train_stats(y, gamma, beta, running_mean, running_var, ws, state=st, momentum=momentum, eps=eps)
        _lib.check(rc, "conv2d_fwd_bn")
        _bump(running_mean, running_var)
        return y, st
    rows = ctypes.c_int(0)
    rc = lib.ubr_conv2d_fwd_stats(x.view(), wpk.data_ptr(), _ptr(bias), out_ch, ksize, stride, y.view(), ws.data_ptr(),
                                  ctypes.addressof(rows), _stream())
    if rc == 1:
        conv2d_fwd(x, wpk, bias, out_ch, ksize, stride, y=y)
        return y, bn_train_stats(y, gamma, beta, running_mean, running_var, ws, state=state, momentum=momentum, eps=eps)
    _lib.check(rc, "conv2d_fwd_stats")
    st = state or BNState(y.c, y.buf.device)
    st.count = float(y.npix)
    _lib.check(lib.ubr_bn_finalize_rows(ws.data_ptr(), rows.value, y.c, st.count, _ptr(gamma), _ptr(beta),
                                        _ptr(running_mean), _ptr(running_var), momentum, eps, st.scale.data_ptr(),
                                        st.shift.data_ptr(), st.mean.data_ptr(), st.invstd.data_ptr(), _stream()),
               "bn_finalize_rows")
    _bump(running_mean, running_var)
    return y, st


def conv2d_dgrad(dy, wpk, in_ch, ksize, stride, dx=None):
    lib = _lib.load()
    if dx is None:
        dx = Act.empty(dy.n, dy.h * stride, dy.w * stride, in_ch, dy.buf.device)
    _lib.check(lib.ubr_conv2d_dgrad(dy.view(), wpk.data_ptr(), in_ch, ksize, stride, dx.view(), _stream()),
               "conv2d_dgrad")
    return dx


def conv2d_dgrad_bn_bwd(dy, wpk, in_ch, ksize, stride, y_pre, st, relu, ws, dgamma=None, dbeta=None):
    """da = dgrad(dy) followed by the backward of a = relu?(bn(y_pre)): returns the gradient wrt y_pre.  Where the conv
    kernel has the fused reduction epilogue the stand-alone bn_act_bwd_reduce pass is skipped."""
    import ctypes
    lib = _lib.load()
    da = Act.empty(dy.n, dy.h * stride, dy.w * stride, in_ch, dy.buf.device)
    tree = _tree(ws) if TREE_FINALIZE_CONV else 0
    if tree:
        rc = lib.ubr_conv2d_dgrad_bnreduce_final(dy.view(), wpk.data_ptr(), in_ch, ksize, stride, da.view(), y_pre.view(),
                                                 st.scale.data_ptr(), st.shift.data_ptr(), st.mean.data_ptr(),
                                                 st.invstd.data_ptr(), int(relu), ws.data_ptr(), tree, _ptr(dgamma),
                                                 _ptr(dbeta), 0, _stream())
        if rc == 1:
            conv2d_dgrad(dy, wpk, in_ch, ksize, stride, dx=da)
            return bn_act_bwd(da, y_pre, st, relu, ws, dgamma, dbeta)
        _lib.check(rc, "conv2d_dgrad_bnreduce_final")
        dyp = Act.empty(y_pre.n, y_pre.h, y_pre.w, y_pre.c, y_pre.buf.device)
        _lib.check(lib.ubr_bn_act_bwd_apply_final(da.view(), y_pre.view(), st.scale.data_ptr(), st.shift.data_ptr(),
                                                  st.mean.data_ptr(), st.invstd.data_ptr(), int(relu), ws.data_ptr(),
                                                  st.count, dyp.view(), _stream()), "bn_act_bwd_apply_final")
        return dyp
    rows = ctypes.c_int(0)
    rc = lib.ubr_conv2d_dgrad_bnreduce(dy.view(), wpk.data_ptr(), in_ch, ksize, stride, da.view(), y_pre.view(),
                                       st.scale.data_ptr(), st.shift.data_ptr(), st.mean.data_ptr(), st.invstd.data_ptr(),
                                       int(relu), ws.data_ptr(), ctypes.addressof(rows), _stream())
    if rc == 1:
        conv2d_dgrad(dy, wpk, in_ch, ksize, stride, dx=da)
        return bn_act_bwd(da, y_pre, st, relu, ws, dgamma, dbeta)
    _lib.check(rc, "conv2d_dgrad_bnreduce")
    dyp = Act.empty(y_pre.n, y_pre.h, y_pre.w, y_pre.c, y_pre.buf.device)
    _lib.check(lib.ubr_bn_act_bwd_apply_rows(da.view(), y_pre.view(), st.scale.data_ptr(), st.shift.data_ptr(),
                                             st.mean.data_ptr(), st.invstd.data_ptr(), 0, int(relu), ws.data_ptr(),
                                             rows.value, st.count, dyp.view(), _ptr(dgamma), _ptr(dbeta), 0, _stream()),
               "bn_act_bwd_apply_rows")
    return dyp


def conv2d_dgrad_acc(dy, wpk, in_ch, ksize, stride, dx):
    """dx += dgrad(dy) fused in the conv epilogue.  Returns False (nothing done) if the shape has no fused path."""
    lib = _lib.load()
    rc = lib.ubr_conv2d_dgrad_acc(dy.view(), wpk.data_ptr(), in_ch, ksize, stride, dx.view(), _stream())
    if rc == 1:
        return False
    _lib.check(rc, "conv2d_dgrad_acc")
    return True


def wgrad_workspace_bytes(grid_n, grid_h, grid_w, m_ch, n_ch, ntaps):
    return int(_lib.load().ubr_wgrad_workspace_bytes(grid_n, grid_h, grid_w, m_ch, n_ch, ntaps))


def conv2d_wgrad(x, dy, out_ch, ksize, stride, dw, accumulate, ws):
    lib = _lib.load()
    assert dw.dtype == torch.float32 and dw.is_contiguous()
    _lib.check(lib.ubr_conv2d_wgrad(x.view(), dy.view(), out_ch, ksize, stride, dw.data_ptr(), int(accumulate),
                                    ws.data_ptr(), ws.numel() * ws.element_size(), _stream()), "conv2d_wgrad")
    return dw


def conv2d_dilated_fwd(x, wpk, bias, out_ch, dilation, y=None, relu=False):
    """3x3 stride-1 conv with padding == dilation (ASPP branches, reference models/ASPP_ResNet.py:203-214).
    relu=True: the inference form relu(conv + bias) for weights with the BatchNorm folded in."""
    lib = _lib.load()
    if y is None:
        y = Act.empty(x.n, x.h, x.w, out_ch, x.buf.device)
    fn = lib.ubr_conv2d_dilated_fwd_relu if relu else lib.ubr_conv2d_dilated_fwd
    _lib.check(fn(x.view(), wpk.data_ptr(), _ptr(bias), out_ch, dilation, y.view(), _stream()), "conv2d_dilated_fwd")
    return y


def conv2d_dilated_dgrad(dy, wpk, in_ch, dilation, dx=None):
    lib = _lib.load()
    if dx is None:
        dx = Act.empty(dy.n, dy.h, dy.w, in_ch, dy.buf.device)
    _lib.check(lib.ubr_conv2d_dilated_dgrad(dy.view(), wpk.data_ptr(), in_ch, dilation, dx.view(), _stream()),
               "conv2d_dilated_dgrad")
    return dx


def conv2d_dilated_wgrad(x, dy, out_ch, dilation, dw, accumulate, ws):
    lib = _lib.load()
    assert dw.dtype == torch.float32 and dw.is_contiguous()
    _lib.check(lib.ubr_conv2d_dilated_wgrad(x.view(), dy.view(), out_ch, dilation, dw.data_ptr(), int(accumulate),
                                            ws.data_ptr(), ws.numel() * ws.element_size(), _stream()),
               "conv2d_dilated_wgrad")
    return dw


def deconv_fwd(x, wpk, out_ch, y=None):
    lib = _lib.load()
    if y is None:
        y = Act.empty(x.n, x.h * 2, x.w * 2, out_ch, x.buf.device)
    _lib.check(lib.ubr_deconv4s2_fwd(x.view(), wpk.data_ptr(), out_ch, y.view(), _stream()), "deconv4s2_fwd")
    return y


def pack_weights_subpixel(w, n_is_out):
    """ConvTranspose2d(k4,s2,p1) IOHW parameter -> bf16 [9+1][n][k] operand of the equivalent 3x3 sub-pixel conv."""
    require_cuda(w, "weight")
    lib = _lib.load()
    w = w.detach().contiguous().float()
    in_ch, out_ch, kh, kw = w.shape
    assert kh == 4 and kw == 4
    pad = lambda v: (v + 15) // 16 * 16
    n_pad, k_pad = (pad(4 * out_ch), pad(in_ch)) if n_is_out else (pad(in_ch), pad(4 * out_ch))
    dst = torch.empty((10, n_pad, k_pad), dtype=torch.bfloat16, device=w.device)
    _lib.check(lib.ubr_pack_weights(w.data_ptr(), 2, 1 if n_is_out else 0, out_ch, in_ch, 3, 3, n_pad, k_pad,
                                    dst.data_ptr(), _stream()), "pack_weights(subpixel)")
    return dst


def deconv_fwd_subpixel(x, wpk3, out_ch, y):
    """Returns False (nothing done) when the shape has no sub-pixel path."""
    lib = _lib.load()
    rc = lib.ubr_deconv4s2_fwd_subpixel(x.view(), wpk3.data_ptr(), out_ch, y.view(), _stream())
    if rc == 1:
        return False
    _lib.check(rc, "deconv4s2_fwd_subpixel")
    return True


def pixel_unshuffle2(y):
    """(n, 2h, 2w, c) -> (n, h, w, 4c), channel (ph*2+pw)*c + co: ConvTranspose output gradient on the input grid."""
    lib = _lib.load()
    y4 = Act.empty(y.n, y.h // 2, y.w // 2, 4 * y.c, y.buf.device)
    _lib.check(lib.ubr_pixel_unshuffle2(y.view(), y4.view(), _stream()), "pixel_unshuffle2")
    return y4


def deconv_wgrad_gather(dw3, dw, accumulate=False):
    """dw3 fp32 [4*out_ch][in_ch][3][3] (sub-pixel weight gradient) -> ConvTranspose dw [in_ch][out_ch][4][4]."""
    lib = _lib.load()
    in_ch, out_ch = dw.shape[0], dw.shape[1]
    assert dw3.shape == (4 * out_ch, in_ch, 3, 3) and dw3.is_contiguous() and dw.is_contiguous()
    _lib.check(lib.ubr_deconv4s2_wgrad_gather(dw3.data_ptr(), in_ch, out_ch, dw.data_ptr(), int(accumulate), _stream()),
               "deconv4s2_wgrad_gather")
    return dw


def deconv_dgrad(dy, wpk, in_ch, dx=None):
    lib = _lib.load()
    if dx is None:
        dx = Act.empty(dy.n, dy.h // 2, dy.w // 2, in_ch, dy.buf.device)
    _lib.check(lib.ubr_deconv4s2_dgrad(dy.view(), wpk.data_ptr(), in_ch, dx.view(), _stream()), "deconv4s2_dgrad")
    return dx


def deconv_wgrad(x, dy, dw, accumulate, ws):
    lib = _lib.load()
    assert dw.dtype == torch.float32 and dw.is_contiguous()
    _lib.check(lib.ubr_deconv4s2_wgrad(x.view(), dy.view(), dw.data_ptr(), int(accumulate), ws.data_ptr(),
                                       ws.numel() * ws.element_size(), _stream()), "deconv4s2_wgrad")
    return dw


def stem_conv_fwd(x, w, b, y=None):
    """x (N,1,H,W) fp32 CUDA -> NHWC bf16 (N,H,W,OC)."""
    lib = _lib.load()
    require_cuda(x, "input")
    n, cin, h, wd = x.shape
    assert cin == 1 and x.dtype == torch.float32 and x.is_contiguous()
    oc = w.shape[0]
    if y is None:
        y = Act.empty(n, h, wd, oc, x.device)
    _lib.check(lib.ubr_stem_conv_fwd(x.data_ptr(), n, h, wd, w.data_ptr(), _ptr(b), oc, y.view(), _stream()),
               "stem_conv_fwd")
    return y


def stem_conv_fwd_relu(x, w, b, y=None):
    """relu(conv(x) + b) for folded inference weights; returns None when the shape has no fused path."""
    lib = _lib.load()
    require_cuda(x, "input")
    n, cin, h, wd = x.shape
    assert cin == 1 and x.dtype == torch.float32 and x.is_contiguous()
    oc = w.shape[0]
    if y is None:
        y = Act.empty(n, h, wd, oc, x.device)
    rc = lib.ubr_stem_conv_fwd_relu(x.data_ptr(), n, h, wd, w.data_ptr(), _ptr(b), oc, y.view(), _stream())
    if rc == 1:
        return None
    _lib.check(rc, "stem_conv_fwd_relu")
    return y


def stem_wgrad_workspace_bytes(n, h, w, oc):
    return int(_lib.load().ubr_stem_wgrad_workspace_bytes(n, h, w, oc))


def stem_conv_wgrad(x, dy, dw, db, accumulate, ws):
    lib = _lib.load()
    n, _, h, wd = x.shape
    _lib.check(lib.ubr_stem_conv_wgrad(x.data_ptr(), n, h, wd, dy.view(), _ptr(dw), _ptr(db), int(accumulate),
                                       ws.data_ptr(), ws.numel() * ws.element_size(), _stream()), "stem_conv_wgrad")


def bn_workspace(c, device):
    """reduction workspace for layers of up to `c` channels: partial rows + final sums, and riding on it as `.tree` the
    zero-initialised arrival tickets / second-level rows of the in-launch finalize (csrc/tree_reduce.cuh)"""
    lib = _lib.load()
    ws = torch.empty(int(lib.ubr_bn_workspace_bytes(c)) // 4, dtype=torch.float32, device=device)
    ws.tree = torch.zeros(int(lib.ubr_bn_tree_bytes(c)) // 4, dtype=torch.float32, device=device)
    return ws


class BNState:
    """Per-layer affine (scale, shift) plus the batch statistics the backward needs."""
    __slots__ = ("scale", "shift", "mean", "invstd", "count")

    def __init__(self, c, device):
        buf = torch.empty((4, c), dtype=torch.float32, device=device)
        self.scale, self.shift, self.mean, self.invstd = buf[0], buf[1], buf[2], buf[3]
        self.count = 0.0


def bn_train_stats(y, gamma, beta, running_mean, running_var, ws, state=None, momentum=BN_MOMENTUM, eps=BN_EPS):
    lib = _lib.load()
    st = state or BNState(y.c, y.buf.device)
    st.count = float(y.npix)
    tree = _tree(ws)
    if tree:
        rc = lib.ubr_bn_stats_final(y.view(), ws.data_ptr(), tree, _ptr(gamma), _ptr(beta), _ptr(running_mean),
                                    _ptr(running_var), momentum, eps, st.scale.data_ptr(), st.shift.data_ptr(),
                                    st.mean.data_ptr(), st.invstd.data_ptr(), _stream())
        if rc != 1:                                        # 1: too many channels for the in-launch tail
            _lib.check(rc, "bn_stats_final")
            _bump(running_mean, running_var)
            return st
    _lib.check(lib.ubr_bn_stats(y.view(), ws.data_ptr(), _stream()), "bn_stats")
    _lib.check(lib.ubr_bn_finalize(ws.data_ptr(), y.c, st.count, _ptr(gamma), _ptr(beta), _ptr(running_mean),
                                   _ptr(running_var), momentum, eps, st.scale.data_ptr(), st.shift.data_ptr(),
                                   st.mean.data_ptr(), st.invstd.data_ptr(), _stream()), "bn_finalize")
    _bump(running_mean, running_var)
    return st


def bn_eval_affine(gamma, beta, running_mean, running_var, state=None, eps=BN_EPS):
    lib = _lib.load()
    c = running_mean.numel()
    st = state or BNState(c, running_mean.device)
    _lib.check(lib.ubr_bn_eval_affine(c, _ptr(gamma), _ptr(beta), running_mean.data_ptr(), running_var.data_ptr(), eps,
                                      st.scale.data_ptr(), st.shift.data_ptr(), _stream()), "bn_eval_affine")
    return st


def bn_act(y, st, relu, out=None):
    lib = _lib.load()
    if out is None:
        out = Act.empty(y.n, y.h, y.w, y.c, y.buf.device)
    _lib.check(lib.ubr_bn_act(y.view(), st.scale.data_ptr(), st.shift.data_ptr(), int(relu), out.view(), _stream()),
               "bn_act")
    return out


def block_tail(y2, st2, res, st_res, out=None):
    lib = _lib.load()
    if out is None:
        out = Act.empty(y2.n, y2.h, y2.w, y2.c, y2.buf.device)
    sr = st_res.scale.data_ptr() if st_res is not None else 0
    br = st_res.shift.data_ptr() if st_res is not None else 0
    _lib.check(lib.ubr_block_tail(y2.view(), st2.scale.data_ptr(), st2.shift.data_ptr(), res.view(), sr, br, out.view(),
                                  _stream()), "block_tail")
    return out


def bn_act_bwd(dout, y, st, relu, ws, dgamma=None, dbeta=None, accumulate=False, dy=None):
    """Backward of out = relu?(bn(y)) in training mode.  Returns dy (grad wrt the conv output y)."""
    lib = _lib.load()
    if dy is None:
        dy = Act.empty(y.n, y.h, y.w, y.c, y.buf.device)
    s = _stream()
    tree = _tree(ws)
    if tree:
        rc = lib.ubr_bn_act_bwd_reduce_final(dout.view(), y.view(), st.scale.data_ptr(), st.shift.data_ptr(),
                                             st.mean.data_ptr(), st.invstd.data_ptr(), int(relu), ws.data_ptr(), tree,
                                             _ptr(dgamma), _ptr(dbeta), int(accumulate), s)
        if rc != 1:
            _lib.check(rc, "bn_act_bwd_reduce_final")
            _lib.check(lib.ubr_bn_act_bwd_apply_final(dout.view(), y.view(), st.scale.data_ptr(), st.shift.data_ptr(),
                                                      st.mean.data_ptr(), st.invstd.data_ptr(), int(relu), ws.data_ptr(),
                                                      st.count, dy.view(), s), "bn_act_bwd_apply_final")
            return dy
    _lib.check(lib.ubr_bn_act_bwd_reduce(dout.view(), y.view(), st.scale.data_ptr(), st.shift.data_ptr(),
                                         st.mean.data_ptr(), st.invstd.data_ptr(), int(relu), ws.data_ptr(), s),
               "bn_act_bwd_reduce")
    _lib.check(lib.ubr_bn_act_bwd_apply(dout.view(), y.view(), st.scale.data_ptr(), st.shift.data_ptr(),
                                        st.mean.data_ptr(), st.invstd.data_ptr(), 0, int(relu), ws.data_ptr(), st.count,
                                        dy.view(), _ptr(dgamma), _ptr(dbeta), int(accumulate), s), "bn_act_bwd_apply")
    return dy


def block_tail_bwd(dout, out, y2, st2, res, st_res, gamma_res, ws, dgamma2=None, dbeta2=None, dgamma_res=None,
                   dbeta_res=None, accumulate=False, dy2=None, dres=None):
    """Backward of the ReLU-add-ReLU block tail.  Returns (dy2, dres).  With a BatchNorm'd residual (`st_res`) the
    kernels recompute the final ReLU's mask from y2 and res, so `out` is not read and may be None."""
    lib = _lib.load()
    dev = y2.buf.device
    if dy2 is None:
        dy2 = Act.empty(y2.n, y2.h, y2.w, y2.c, dev)
    if dres is None:
        dres = Act.empty(y2.n, y2.h, y2.w, y2.c, dev)
    s = _stream()
    res_bn = st_res is not None
    rv = res.view() if res_bn else _NULLVIEW
    mr = st_res.mean.data_ptr() if res_bn else 0
    ir = st_res.invstd.data_ptr() if res_bn else 0
    sr = st_res.scale.data_ptr() if res_bn else 0
    br = st_res.shift.data_ptr() if res_bn else 0
    ov = out.view() if (out is not None and not res_bn) else _NULLVIEW
    tree = _tree(ws)
    if tree:
        rc = lib.ubr_block_tail_bwd_reduce_final(dout.view(), ov, y2.view(), st2.scale.data_ptr(),
                                                 st2.shift.data_ptr(), st2.mean.data_ptr(), st2.invstd.data_ptr(),
                                                 rv, mr, ir, sr, br, int(res_bn), ws.data_ptr(), tree, _ptr(dgamma2),
                                                 _ptr(dbeta2), _ptr(dgamma_res), _ptr(dbeta_res), int(accumulate), s)
        if rc != 1:
            _lib.check(rc, "block_tail_bwd_reduce_final")
            _lib.check(lib.ubr_block_tail_bwd_apply_final(dout.view(), ov, y2.view(), st2.scale.data_ptr(),
                                                          st2.shift.data_ptr(), st2.mean.data_ptr(), st2.invstd.data_ptr(), rv,
                                                          mr, ir, _ptr(gamma_res), sr, br, int(res_bn), ws.data_ptr(),
                                                          st2.count, dy2.view(), dres.view(), s), "block_tail_bwd_apply_final")
            return dy2, dres
    _lib.check(lib.ubr_block_tail_bwd_reduce(dout.view(), ov, y2.view(), st2.scale.data_ptr(),
                                             st2.shift.data_ptr(), st2.mean.data_ptr(), st2.invstd.data_ptr(), rv, mr,
                                             ir, sr, br, int(res_bn), ws.data_ptr(), s), "block_tail_bwd_reduce")
    _lib.check(lib.ubr_block_tail_bwd_apply(dout.view(), ov, y2.view(), st2.scale.data_ptr(),
                                            st2.shift.data_ptr(), st2.mean.data_ptr(), st2.invstd.data_ptr(), 0, rv, mr,
                                            ir, _ptr(gamma_res), sr, br, int(res_bn), ws.data_ptr(), st2.count, dy2.view(),
                                            dres.view(), _ptr(dgamma2), _ptr(dbeta2), _ptr(dgamma_res), _ptr(dbeta_res),
                                            int(accumulate), s), "block_tail_bwd_apply")
    return dy2, dres


def maxpool_fwd(x, y=None):
    lib = _lib.load()
    if y is None:
        y = Act.empty(x.n, x.h // 2, x.w // 2, x.c, x.buf.device)
    _lib.check(lib.ubr_maxpool3s2_fwd(x.view(), y.view(), _stream()), "maxpool3s2_fwd")
    return y


def maxpool_bwd(x, dy, dx, accumulate):
    lib = _lib.load()
    _lib.check(lib.ubr_maxpool3s2_bwd(x.view(), dy.view(), dx.view(), int(accumulate), _stream()), "maxpool3s2_bwd")
    return dx


def maxpool_fwd_idx(x, y=None):
    """MaxPool2d(3,2,1) that also returns the argmax taps (uint8, NHWC) for the gather backward."""
    lib = _lib.load()
    if y is None:
        y = Act.empty(x.n, x.h // 2, x.w // 2, x.c, x.buf.device)
    idx = torch.empty((x.n, x.h // 2, x.w // 2, x.c), dtype=torch.uint8, device=x.buf.device)
    _lib.check(lib.ubr_maxpool3s2_fwd_idx(x.view(), y.view(), idx.data_ptr(), _stream()), "maxpool3s2_fwd_idx")
    return y, idx


def maxpool_bwd_idx(idx, dy, dx, accumulate):
    lib = _lib.load()
    _lib.check(lib.ubr_maxpool3s2_bwd_idx(idx.data_ptr(), dy.view(), dx.view(), int(accumulate), _stream()),
               "maxpool3s2_bwd_idx")
    return dx


def maxpool3s1_fwd(x, y=None, want_idx=False):
    """MaxPool2d(3, stride=1, padding=1) (ASPP block 5); returns (y, idx|None)."""
    lib = _lib.load()
    if y is None:
        y = Act.empty(x.n, x.h, x.w, x.c, x.buf.device)
    idx = torch.empty((x.n, x.h, x.w, x.c), dtype=torch.uint8, device=x.buf.device) if want_idx else None
    _lib.check(lib.ubr_maxpool3s1_fwd(x.view(), y.view(), _ptr(idx), _stream()), "maxpool3s1_fwd")
    return y, idx


def maxpool3s1_bwd_idx(idx, dy, dx, accumulate):
    lib = _lib.load()
    _lib.check(lib.ubr_maxpool3s1_bwd_idx(idx.data_ptr(), dy.view(), dx.view(), int(accumulate), _stream()),
               "maxpool3s1_bwd_idx")
    return dx


def add(a, b=None, c=None, out=None):
    lib = _lib.load()
    if out is None:
        out = Act.empty(a.n, a.h, a.w, a.c, a.buf.device)
    _lib.check(lib.ubr_add(a.view(), b.view() if b is not None else _NULLVIEW,
                           c.view() if c is not None else _NULLVIEW, out.view(), _stream()), "add")
    return out


def logsoftmax_fwd(logits, n, ncls, h, w):
    """logits fp32 (N*H*W, LD) -> log-probs fp32 NCHW."""
    lib = _lib.load()
    logp = torch.empty((n, ncls, h, w), dtype=torch.float32, device=logits.device)
    _lib.check(lib.ubr_logsoftmax_fwd(logits.data_ptr(), logits.shape[-1], n, ncls, h * w, logp.data_ptr(), _stream()),
               "logsoftmax_fwd")
    return logp


def logsoftmax_bwd(logp, dlogp, dl=None):
    lib = _lib.load()
    n, ncls, h, w = logp.shape
    if dl is None:
        dl = Act.empty(n, h, w, 16, logp.device)
    dlogp = dlogp.contiguous()
    _lib.check(lib.ubr_logsoftmax_bwd(logp.data_ptr(), dlogp.data_ptr(), n, ncls, h * w, dl.view(), _stream()),
               "logsoftmax_bwd")
    return dl


def pixelwise_nll_fwd(logp, target, pixw, classw, ignore_index):
    lib = _lib.load()
    n, ncls, h, w = logp.shape
    ws = torch.empty(int(lib.ubr_nll_workspace_bytes(n, h * w)) // 4, dtype=torch.float32, device=logp.device)
    loss = torch.empty((), dtype=torch.float32, device=logp.device)
    _lib.check(lib.ubr_pixelwise_nll_fwd(logp.data_ptr(), target.data_ptr(), pixw.data_ptr(), _ptr(classw), n, ncls,
                                         h * w, ignore_index, ws.data_ptr(), loss.data_ptr(), _stream()),
               "pixelwise_nll_fwd")
    return loss


def pixelwise_nll_bwd(target, pixw, classw, gout, n, ncls, h, w, ignore_index):
    lib = _lib.load()
    dlogp = torch.empty((n, ncls, h, w), dtype=torch.float32, device=target.device)
    _lib.check(lib.ubr_pixelwise_nll_bwd(target.data_ptr(), pixw.data_ptr(), _ptr(classw), _ptr(gout), n, ncls, h * w,
                                         ignore_index, dlogp.data_ptr(), _stream()), "pixelwise_nll_bwd")
    return dlogp


def accuracy_counts(logp, target):
    """(2*C,) int64 device tensor: [pixels per class | correctly predicted pixels per class]."""
    lib = _lib.load()
    require_cuda(logp, "output")
    require_cuda(target, "target")
    n, ncls, h, w = logp.shape
    assert logp.dtype == torch.float32 and target.dtype == torch.int64 and target.numel() == n * h * w
    counts = torch.empty(2 * ncls, dtype=torch.int64, device=logp.device)
    _lib.check(lib.ubr_accuracy_counts(logp.contiguous().data_ptr(), target.contiguous().data_ptr(), n, ncls, h * w,
                                       counts.data_ptr(), _stream()), "accuracy_counts")
    return counts


def tile_gather(plane, origins, crop):
    """crop: int (square) or (crop_h, crop_w)"""
    lib = _lib.load()
    require_cuda(plane, "plane")
    rows, cols = plane.shape
    nt = origins.shape[0]
    ch, cw = (crop, crop) if isinstance(crop, int) else crop
    crops = torch.empty((nt, 1, ch, cw), dtype=torch.float32, device=plane.device)
    _lib.check(lib.ubr_tile_gather_rect(plane.data_ptr(), rows, cols, origins.data_ptr(), nt, ch, cw, crops.data_ptr(),
                                        _stream()), "tile_gather")
    return crops


def stitch_scatter(tiles, origins, rows, cols):
    lib = _lib.load()
    require_cuda(tiles, "tiles")
    nt, nch, ch, cw = tiles.shape
    out = torch.empty((nch, rows, cols), dtype=torch.float32, device=tiles.device)
    _lib.check(lib.ubr_stitch_scatter_rect(tiles.data_ptr(), nch, origins.data_ptr(), nt, ch, cw, rows, cols, out.data_ptr(),
                                           _stream()), "stitch_scatter")
    return out


def argmax_classes(logp, out=None):
    """(N,C,H,W) fp32 log-probs -> (N,H,W) uint8 classes (first maximum)"""
    lib = _lib.load()
    require_cuda(logp, "log-probs")
    n, ncls, h, w = logp.shape
    assert logp.dtype == torch.float32 and logp.is_contiguous()
    if out is None:
        out = torch.empty((n, h, w), dtype=torch.uint8, device=logp.device)
    _lib.check(lib.ubr_argmax_classes(logp.data_ptr(), n, ncls, h * w, out.data_ptr(), _stream()), "argmax_classes")
    return out


def stitch_scatter_u8(tiles, origins, rows, cols, fill=0, out=None):
    """(ntiles, crop, crop) uint8 class tiles -> (rows, cols) uint8"""
    lib = _lib.load()
    require_cuda(tiles, "tiles")
    nt, ch, cw = tiles.shape
    assert tiles.dtype == torch.uint8 and tiles.is_contiguous()
    if out is None:
        out = torch.empty((rows, cols), dtype=torch.uint8, device=tiles.device)
    _lib.check(lib.ubr_stitch_scatter_u8(tiles.data_ptr(), origins.data_ptr(), nt, ch, cw, rows, cols, int(fill),
                                         out.data_ptr(), _stream()), "stitch_scatter_u8")
    return out


def scale_f32_(buf, alpha):
    lib = _lib.load()
    _lib.check(lib.ubr_scale_f32(buf.data_ptr(), buf.numel(), float(alpha), _stream()), "scale_f32")
    return buf
