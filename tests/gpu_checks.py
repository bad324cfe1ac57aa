"""Per-op parity checks of the C-ABI kernels against plain torch fp32 CPU references.

Used by tests/test_gpu_ops.py (pytest -m gpu) and by tools/gpu_probe.py, which runs every check in its
own subprocess (a trapped kernel poisons the CUDA context) and dumps failing tensors for offline analysis.
Each check returns a list of (label, err, tol) triples; err <= tol means pass.
"""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from ubresnet_b200 import ops  # noqa: E402
from ubresnet_b200.ops import Act  # noqa: E402

DEV = "cuda:0"
DUMP_DIR = os.environ.get("UBR_DUMP_DIR", "")


def bf(x):
    """round through bf16, return fp32 CPU"""
    return x.to(torch.bfloat16).float()


def relerr(out, ref):
    out = out.detach().float().cpu()
    ref = ref.detach().float().cpu()
    den = ref.abs().max().item()
    return (out - ref).abs().max().item() / (den if den > 0 else 1.0)


def _dump(name, **tensors):
    if DUMP_DIR:
        os.makedirs(DUMP_DIR, exist_ok=True)
        torch.save({k: v.detach().cpu() for k, v in tensors.items()}, os.path.join(DUMP_DIR, name + ".pt"))


def gen(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(*shape, generator=g) * scale


def act_from(x_nchw, ld=None, c0=0):
    """CPU NCHW fp32 -> device Act (optionally a channel slice of a wider buffer)."""
    n, c, h, w = x_nchw.shape
    buf = torch.zeros((n, h, w, ld or c), dtype=torch.bfloat16, device=DEV)
    # poison the other channels so slice handling errors show up
    if ld and ld != c:
        buf.fill_(7.0)
    a = Act(buf, c0, c)
    a.torch().copy_(x_nchw.permute(0, 2, 3, 1).to(DEV))
    return a


def act_out(n, h, w, c, ld=None, c0=0):
    buf = torch.full((n, h, w, ld or c), 3.0, dtype=torch.bfloat16, device=DEV)
    return Act(buf, c0, c)


# ---------------------------------------------------------------------------------------------
def check_conv_fwd(cin, cout, k, stride, n, h, w, in_ld=None, out_ld=None, seed=0, bias=False, f32=False):
    name = "conv_fwd_ci%d_co%d_k%d_s%d_n%d_%dx%d" % (cin, cout, k, stride, n, h, w)
    x = bf(gen(n, cin, h, w, seed=seed))
    wt = bf(gen(cout, cin, k, k, seed=seed + 1, scale=(2.0 / (k * k * cin)) ** 0.5))
    b = gen(cout, seed=seed + 2) if bias else None
    ref = F.conv2d(x, wt, b, stride, (k - 1) // 2)
    xa = act_from(x, in_ld, 0 if not in_ld else (in_ld - cin))
    wpk = ops.pack_weights(wt.to(DEV), False, True)
    bd = b.to(DEV) if bias else None
    if f32:
        logits = torch.full((n * (h // stride) * (w // stride), 4), 5.0, dtype=torch.float32, device=DEV)
        ops.conv2d_fwd(xa, wpk, bd, cout, k, stride, y_f32=logits)
        torch.cuda.synchronize()
        out = logits.view(n, h // stride, w // stride, 4)[..., :cout].permute(0, 3, 1, 2)
        res = [(name + "_f32", relerr(out, ref), 2e-3)]
    else:
        ya = act_out(n, h // stride, w // stride, cout, out_ld, 0)
        ops.conv2d_fwd(xa, wpk, bd, cout, k, stride, y=ya)
        torch.cuda.synchronize()
        out = ya.nchw_float()
        res = [(name, relerr(out, ref), 1e-2)]
        if out_ld and out_ld != cout:
            untouched = ya.buf[..., cout:].float()
            res.append((name + "_slice_untouched", (untouched - 3.0).abs().max().item(), 0.0))
    if res[0][1] > res[0][2]:
        _dump(name, out=out, ref=ref, x=x, w=wt)
    return res


def check_conv_fwd_fused(cin, cout, k, stride, n, h, w, relu_pre=True, with_res=True, relu_post=True, res_ld=None,
                         out_ld=None, seed=0):
    """inference epilogue (ubr_conv2d_fwd_fused): relu_post(relu_pre(conv(x) + b) + res) against torch"""
    name = "conv_fused_ci%d_co%d_k%d_s%d_n%d_%dx%d_%d%d%d" % (cin, cout, k, stride, n, h, w, relu_pre, with_res, relu_post)
    ho, wo = h // stride, w // stride
    x = bf(gen(n, cin, h, w, seed=seed))
    wt = bf(gen(cout, cin, k, k, seed=seed + 1, scale=(2.0 / (k * k * cin)) ** 0.5))
    b = gen(cout, seed=seed + 2, scale=0.3)
    r = bf(gen(n, cout, ho, wo, seed=seed + 3)) if with_res else None
    ref = F.conv2d(x, wt, b, stride, (k - 1) // 2)
    if relu_pre:
        ref = F.relu(ref)
    if with_res:
        ref = ref + r
    if relu_post:
        ref = F.relu(ref)
    wpk = ops.pack_weights(wt.to(DEV), False, True)
    ra = act_from(r, res_ld, 0 if not res_ld else (res_ld - cout)) if with_res else None
    ya = act_out(n, ho, wo, cout, out_ld, 0)
    ops.conv2d_fwd_fused(act_from(x), wpk, b.to(DEV), cout, k, stride, relu_pre, ra, relu_post, y=ya)
    torch.cuda.synchronize()
    out = ya.nchw_float()
    res = [(name, relerr(out, ref), 1e-2)]
    if relu_post:
        res.append((name + "_nonneg", float((out < 0).any().item()), 0.0))
    if out_ld and out_ld != cout:
        res.append((name + "_slice_untouched", (ya.buf[..., cout:].float() - 3.0).abs().max().item(), 0.0))
    return res


def check_block_chain(cin, planes, n, h, w, in_ld=None, out_ld=None, seed=0):
    """eval residual-block chain (ubr_block_chain_fwd) against torch AND against the two-launch path it replaces
    (same accumulation order: bit-identical)."""
    name = "block_chain_ci%d_p%d_n%d_%dx%d" % (cin, planes, n, h, w)
    x = bf(gen(n, cin, h, w, seed=seed))
    w1 = bf(gen(planes, cin, 3, 3, seed=seed + 1, scale=(2.0 / (9 * cin)) ** 0.5))
    w2 = bf(gen(planes, planes, 3, 3, seed=seed + 2, scale=(2.0 / (9 * planes)) ** 0.5))
    b1 = gen(planes, seed=seed + 3, scale=0.3)
    b2 = gen(planes, seed=seed + 4, scale=0.3)
    identity = cin == planes
    r = x if identity else bf(gen(n, planes, h, w, seed=seed + 5))
    a1 = bf(F.relu(F.conv2d(x, w1, b1, 1, 1)))
    ref = F.relu(F.relu(F.conv2d(a1, w2, b2, 1, 1)) + r)
    w1pk = ops.pack_weights(w1.to(DEV), False, True)
    w2pk = ops.pack_weights(w2.to(DEV), False, True)
    xa = act_from(x, in_ld, 0)
    ra = xa if identity else act_from(r)
    ya = act_out(n, h, w, planes, out_ld, 0)
    got = ops.block_chain_fwd(xa, w1pk, b1.to(DEV), w2pk, b2.to(DEV), planes, ra, y=ya)
    torch.cuda.synchronize()
    res = [(name + "_eligible", 0.0 if got is not None else 1.0, 0.0)]
    if got is None:
        return res
    out = ya.nchw_float()
    res.append((name, relerr(out, ref), 1e-2))
    a1d = ops.conv2d_fwd_fused(xa, w1pk, b1.to(DEV), planes, 3, 1, True)
    y2 = ops.conv2d_fwd_fused(a1d, w2pk, b2.to(DEV), planes, 3, 1, True, ra, True)
    torch.cuda.synchronize()
    res.append((name + "_vs_two_launches", (out - y2.nchw_float()).abs().max().item(), 0.0))
    if out_ld and out_ld != planes:
        res.append((name + "_slice_untouched", (ya.buf[..., planes:].float() - 3.0).abs().max().item(), 0.0))
    return res


def check_conv_fwd_bn(cin, cout, k, n, h, w, seed=0, bias=False, expect_rows=True):
    """conv + fused BatchNorm statistics (ubr_conv2d_fwd_stats) against F.conv2d + batch statistics."""
    name = "conv_fwd_bn_ci%d_co%d_k%d_n%d_%dx%d" % (cin, cout, k, n, h, w)
    x = bf(gen(n, cin, h, w, seed=seed))
    wt = bf(gen(cout, cin, k, k, seed=seed + 1, scale=(2.0 / (k * k * cin)) ** 0.5))
    b = gen(cout, seed=seed + 2) if bias else None
    ref = bf(F.conv2d(x, wt, b, 1, (k - 1) // 2))             # statistics are taken over the stored bf16 values
    gamma, beta = gen(cout, seed=seed + 3).abs() + 0.5, gen(cout, seed=seed + 4)
    rm, rv = torch.zeros(cout, device=DEV), torch.ones(cout, device=DEV)
    ws = ops.bn_workspace(cout, DEV)
    ws.fill_(float("nan"))
    wpk = ops.pack_weights(wt.to(DEV), False, True)
    y, st = ops.conv2d_fwd_bn(act_from(x), wpk, b.to(DEV) if bias else None, cout, k, 1, gamma.to(DEV), beta.to(DEV), rm, rv, ws)
    torch.cuda.synchronize()
    mean = ref.mean((0, 2, 3))
    var = ref.var((0, 2, 3), unbiased=False)
    cnt = n * h * w
    return [(name + "_y", relerr(y.nchw_float(), ref), 1e-2),
            (name + "_mean", (st.mean.cpu() - mean).abs().max().item() / max(ref.abs().max().item(), 1e-6), 2e-3),
            (name + "_invstd", relerr(st.invstd, 1.0 / torch.sqrt(var + 1e-5)), 5e-3),
            (name + "_running_var", relerr(rv, 0.9 + 0.1 * var * cnt / (cnt - 1)), 5e-3)]


def check_conv_dgrad(cin, cout, k, stride, n, h, w, seed=0):
    """h, w are the INPUT sizes of the conv (dx size)."""
    name = "conv_dgrad_ci%d_co%d_k%d_s%d_n%d_%dx%d" % (cin, cout, k, stride, n, h, w)
    ho, wo = h // stride, w // stride
    dy = bf(gen(n, cout, ho, wo, seed=seed))
    wt = bf(gen(cout, cin, k, k, seed=seed + 1, scale=(2.0 / (k * k * cout)) ** 0.5))
    ref = torch.nn.grad.conv2d_input((n, cin, h, w), wt, dy, stride, (k - 1) // 2)
    dya = act_from(dy)
    wpk = ops.pack_weights(wt.to(DEV), False, False)
    dxa = act_out(n, h, w, cin)
    ops.conv2d_dgrad(dya, wpk, cin, k, stride, dx=dxa)
    torch.cuda.synchronize()
    out = dxa.nchw_float()
    e = relerr(out, ref)
    if e > 1e-2:
        _dump(name, out=out, ref=ref, dy=dy, w=wt)
    return [(name, e, 1e-2)]


def check_conv_dgrad_acc(cin, cout, k, n, h, w, seed=0, expect_fused=True):
    """dx += dgrad(dy) fused in the epilogue (ubr_conv2d_dgrad_acc); falls back (returns False) on unfused shapes."""
    name = "conv_dgrad_acc_ci%d_co%d_k%d_n%d_%dx%d" % (cin, cout, k, n, h, w)
    dy = bf(gen(n, cout, h, w, seed=seed))
    wt = bf(gen(cout, cin, k, k, seed=seed + 1, scale=(2.0 / (k * k * cout)) ** 0.5))
    base = bf(gen(n, cin, h, w, seed=seed + 2))
    ref = torch.nn.grad.conv2d_input((n, cin, h, w), wt, dy, 1, (k - 1) // 2) + base
    dxa = act_from(base, 2 * cin, cin)
    wpk = ops.pack_weights(wt.to(DEV), False, False)
    fused = ops.conv2d_dgrad_acc(act_from(dy), wpk, cin, k, 1, dxa)
    torch.cuda.synchronize()
    res = [(name + "_fused_path", 0.0 if fused == expect_fused else 1.0, 0.0)]
    if fused:
        res.append((name, relerr(dxa.nchw_float(), ref), 1e-2))
        res.append((name + "_slice_untouched", (dxa.buf[..., :cin].float() - 7.0).abs().max().item(), 0.0))
    else:
        res.append((name + "_untouched", relerr(dxa.nchw_float(), base), 0.0))
    return res


def check_conv_wgrad(cin, cout, k, stride, n, h, w, seed=0, cout_pad=None):
    name = "conv_wgrad_ci%d_co%d_k%d_s%d_n%d_%dx%d" % (cin, cout, k, stride, n, h, w)
    ho, wo = h // stride, w // stride
    x = bf(gen(n, cin, h, w, seed=seed))
    dy = bf(gen(n, cout, ho, wo, seed=seed + 1))
    ref = torch.nn.grad.conv2d_weight(x, (cout, cin, k, k), dy, stride, (k - 1) // 2)
    xa = act_from(x)
    cp = cout_pad or cout
    dyp = torch.zeros(n, cp, ho, wo)
    dyp[:, :cout] = dy
    dya = act_from(dyp)
    nbytes = ops.wgrad_workspace_bytes(n, ho, wo, cp, cin, k * k)
    ws = torch.empty(nbytes // 4 + 4, dtype=torch.float32, device=DEV)
    dw = torch.full((cout, cin, k, k), 9.0, dtype=torch.float32, device=DEV)
    ops.conv2d_wgrad(xa, dya, cout, k, stride, dw, False, ws)
    torch.cuda.synchronize()
    e = relerr(dw, ref)
    res = [(name, e, 2e-3)]
    ops.conv2d_wgrad(xa, dya, cout, k, stride, dw, True, ws)
    torch.cuda.synchronize()
    res.append((name + "_accum", relerr(dw, 2 * ref), 2e-3))
    if e > 2e-3:
        _dump(name, out=dw, ref=ref, x=x, dy=dy)
    return res


def check_deconv(cin, cout, n, h, w, out_ld=None, seed=0):
    """ConvTranspose2d(cin, cout, 4, 2, 1): fwd, dgrad, wgrad.  h, w = input size."""
    name = "deconv_ci%d_co%d_n%d_%dx%d" % (cin, cout, n, h, w)
    x = bf(gen(n, cin, h, w, seed=seed))
    wt = bf(gen(cin, cout, 4, 4, seed=seed + 1, scale=(2.0 / (16 * cout)) ** 0.5))
    ref = F.conv_transpose2d(x, wt, None, 2, 1)
    xa = act_from(x)
    wpk = ops.pack_weights(wt.to(DEV), True, True)
    ya = act_out(n, 2 * h, 2 * w, cout, out_ld, 0)
    ops.deconv_fwd(xa, wpk, cout, y=ya)
    torch.cuda.synchronize()
    out = ya.nchw_float()
    res = [(name + "_fwd", relerr(out, ref), 1e-2)]
    if res[0][1] > 1e-2:
        _dump(name + "_fwd", out=out, ref=ref, x=x, w=wt)
    # sub-pixel path (3x3 conv over the input grid + pixel shuffle), where the shape supports it
    yb = act_out(n, 2 * h, 2 * w, cout, out_ld, 0)
    if ops.deconv_fwd_subpixel(xa, ops.pack_weights_subpixel(wt.to(DEV), True), cout, yb):
        torch.cuda.synchronize()
        res.append((name + "_fwd_subpixel", relerr(yb.nchw_float(), ref), 1e-2))
        if out_ld and out_ld != cout:
            res.append((name + "_fwd_subpixel_slice_untouched", (yb.buf[..., cout:].float() - 3.0).abs().max().item(), 0.0))
    elif w % 128 == 0 and cin <= 64:
        res.append((name + "_fwd_subpixel_missing", 1.0, 0.0))
    # dgrad: dx = conv2d(dy, W viewed as OIHW with O=cin), stride 2, pad 1
    dy = bf(gen(n, cout, 2 * h, 2 * w, seed=seed + 2))
    ref_dx = F.conv2d(dy, wt, None, 2, 1)
    dya = act_from(dy, out_ld, 0)
    wpk_d = ops.pack_weights(wt.to(DEV), True, False)
    dxa = act_out(n, h, w, cin)
    ops.deconv_dgrad(dya, wpk_d, cin, dx=dxa)
    torch.cuda.synchronize()
    o2 = dxa.nchw_float()
    res.append((name + "_dgrad", relerr(o2, ref_dx), 1e-2))
    if res[-1][1] > 1e-2:
        _dump(name + "_dgrad", out=o2, ref=ref_dx, dy=dy, w=wt)
    # wgrad
    xg = x.clone().requires_grad_(False)
    wg = wt.clone().requires_grad_(True)
    F.conv_transpose2d(xg, wg, None, 2, 1).backward(dy)
    ref_dw = wg.grad
    nbytes = ops.wgrad_workspace_bytes(n, h, w, cin, cout, 16)
    ws = torch.empty(nbytes // 4 + 4, dtype=torch.float32, device=DEV)
    dw = torch.full((cin, cout, 4, 4), 9.0, dtype=torch.float32, device=DEV)
    ops.deconv_wgrad(xa, dya, dw, False, ws)
    torch.cuda.synchronize()
    res.append((name + "_wgrad", relerr(dw, ref_dw), 2e-3))
    if res[-1][1] > 2e-3:
        _dump(name + "_wgrad", out=dw, ref=ref_dw, x=x, dy=dy)
    # sub-pixel backward (pixel-unshuffle + 3x3 conv / 3x3 wgrad on the input grid), as Engine._deconv_bwd runs it
    if 4 * cout <= 64 and cin <= 64:
        dy4 = ops.pixel_unshuffle2(dya)
        ref4 = torch.cat([dy[:, :, ph::2, pw::2] for ph in (0, 1) for pw in (0, 1)], 1)
        res.append((name + "_unshuffle", relerr(dy4.nchw_float(), ref4), 0.0))
        dx3 = ops.conv2d_fwd(dy4, ops.pack_weights_subpixel(wt.to(DEV), False), None, cin, 3, 1)
        ws3 = torch.empty(ops.wgrad_workspace_bytes(n, h, w, 4 * cout, cin, 9) // 4 + 4, dtype=torch.float32, device=DEV)
        dw3 = torch.empty((4 * cout, cin, 3, 3), dtype=torch.float32, device=DEV)
        ops.conv2d_wgrad(xa, dy4, 4 * cout, 3, 1, dw3, False, ws3)
        dwb = torch.full((cin, cout, 4, 4), 9.0, dtype=torch.float32, device=DEV)
        ops.deconv_wgrad_gather(dw3, dwb)
        torch.cuda.synchronize()
        res.append((name + "_dgrad_subpixel", relerr(dx3.nchw_float(), ref_dx), 1e-2))
        res.append((name + "_wgrad_subpixel", relerr(dwb, ref_dw), 2e-3))
        ops.deconv_wgrad_gather(dw3, dwb, accumulate=True)
        torch.cuda.synchronize()
        res.append((name + "_wgrad_subpixel_accum", relerr(dwb, 2 * ref_dw), 2e-3))
    return res


def check_stem(n, h, w, oc=16, seed=0):
    name = "stem_n%d_%dx%d_oc%d" % (n, h, w, oc)
    x = gen(n, 1, h, w, seed=seed, scale=30.0)
    wt = gen(oc, 1, 7, 7, seed=seed + 1, scale=(2.0 / (49 * oc)) ** 0.5)
    b = gen(oc, seed=seed + 2, scale=0.1)
    ref = F.conv2d(x, wt, b, 1, 3)
    ya = act_out(n, h, w, oc, 2 * oc, oc)
    ops.stem_conv_fwd(x.to(DEV), wt.to(DEV), b.to(DEV), y=ya)
    torch.cuda.synchronize()
    res = [(name + "_fwd", relerr(ya.nchw_float(), ref), 6e-3),
           # the stem accumulates fp32-accurate products (stem_tc.cu splits x and W into two bf16 terms each), so
           # nearly every output equals the bf16 rounding of the fp32 reference; plain bf16 operands would not
           (name + "_fwd_not_bf16_of_fp32_ref", 1.0 - (ya.nchw_float().cpu() == bf(ref)).float().mean().item(), 0.05)]
    # inference form relu(conv + b) into a channel slice (the dec1 skip half of the concat buffer); None = no fused path
    yr = act_out(n, h, w, oc, 2 * oc, oc)
    got = ops.stem_conv_fwd_relu(x.to(DEV), wt.to(DEV), b.to(DEV), y=yr)
    torch.cuda.synchronize()
    if oc == 16 and w % 128 == 0 and h % 8 == 0:
        res.append((name + "_fwd_relu", relerr(yr.nchw_float(), F.relu(ref)) if got is not None else 1.0, 6e-3))
        res.append((name + "_fwd_relu_slice_untouched", (yr.buf[..., :oc].float() - 3.0).abs().max().item(), 0.0))
    else:
        res.append((name + "_fwd_relu_declined", 0.0 if got is None else 1.0, 0.0))
    dy = bf(gen(n, oc, h, w, seed=seed + 3))
    ref_dw = torch.nn.grad.conv2d_weight(x, (oc, 1, 7, 7), dy, 1, 3)
    ref_db = dy.sum((0, 2, 3))
    dya = act_from(dy)
    ws = torch.empty(ops.stem_wgrad_workspace_bytes(n, h, w, oc) // 4, dtype=torch.float32, device=DEV)
    dw = torch.full((oc, 1, 7, 7), 9.0, dtype=torch.float32, device=DEV)
    db = torch.full((oc,), 9.0, dtype=torch.float32, device=DEV)
    ops.stem_conv_wgrad(x.to(DEV), dya, dw, db, False, ws)
    torch.cuda.synchronize()
    res.append((name + "_wgrad", relerr(dw, ref_dw), 1e-4))
    res.append((name + "_bgrad", relerr(db, ref_db), 1e-4))
    ops.stem_conv_wgrad(x.to(DEV), dya, dw, db, True, ws)
    torch.cuda.synchronize()
    res.append((name + "_wgrad_accumulate", relerr(dw, 2 * ref_dw), 1e-4))
    res.append((name + "_bgrad_accumulate", relerr(db, 2 * ref_db), 1e-4))
    return res


def check_optim(seed=0):
    """multi-tensor SGD / Adam (training/optim.py) against torch.optim on the same parameters and gradients, 4 steps;
    odd sizes, an unaligned view and a tensor larger than one chunk are in the set"""
    from ubresnet_b200.training import optim as fused
    res = []
    g = torch.Generator().manual_seed(seed)
    shapes = [(16, 1, 7, 7), (16,), (3,), (512, 513), (33, 7, 3, 3), (1,), (4097,)]
    base = [torch.randn(s, generator=g) for s in shapes]
    flat = torch.randn(1001, generator=g)                       # parameter living at a 4-byte-aligned offset
    cases = [("sgd_m0.9_wd", fused.SGD, torch.optim.SGD, dict(lr=1e-2, momentum=0.9, weight_decay=1e-3)),
             ("sgd_plain", fused.SGD, torch.optim.SGD, dict(lr=1e-2)),
             ("sgd_nesterov", fused.SGD, torch.optim.SGD, dict(lr=1e-2, momentum=0.8, nesterov=True, weight_decay=1e-2)),
             ("sgd_dampening", fused.SGD, torch.optim.SGD, dict(lr=1e-2, momentum=0.9, dampening=0.3)),
             ("adam_wd", fused.Adam, torch.optim.Adam, dict(lr=1e-3, weight_decay=1e-4)),
             ("adam_betas", fused.Adam, torch.optim.Adam, dict(lr=3e-3, betas=(0.8, 0.95), eps=1e-6))]
    for name, mine, theirs, kw in cases:
        def make():
            ps = [torch.nn.Parameter(b.clone().to(DEV)) for b in base]
            store = flat.clone().to(DEV)
            ps.append(torch.nn.Parameter(store[1:]))                # data_ptr % 16 == 4
            return ps
        pa, pb = make(), make()
        oa, ob = mine(pa, **kw), theirs(pb, **kw)
        gg = torch.Generator().manual_seed(seed + 1)
        for step in range(4):
            for a, b in zip(pa, pb):
                gr = torch.randn(a.shape, generator=gg).to(DEV)
                a.grad, b.grad = gr.clone(), gr.clone()
            oa.step()
            ob.step()
        torch.cuda.synchronize()
        # every step must bump the parameters' version counters (the engine's packed-weight cache is keyed on them)
        res.append((name + "_versions_bumped", float(sum(1 for a in pa if a._version < 4)), 0.0))
        res.append((name + "_params", max(relerr(a, b) for a, b in zip(pa, pb)), 2e-6))
        sa, sb = oa.state_dict()["state"], ob.state_dict()["state"]
        worst = 0.0
        for k in sb:
            for nm, v in sb[k].items():
                if torch.is_tensor(v) and v.numel() > 1:
                    worst = max(worst, relerr(sa[k][nm], v))
                elif nm == "step":
                    worst = max(worst, abs(float(sa[k][nm]) - float(v)))
        res.append((name + "_state", worst, 2e-6))
        # the torch class accepts our state_dict (checkpoint round trip, train_ubresnet2018_wlarcv2.py:260-266)
        oc = theirs(make(), **kw)
        oc.load_state_dict(oa.state_dict())
        res.append((name + "_state_dict_loads_into_torch", 0.0, 0.0))
    return res


def _bn_ref(y, gamma, beta, eps=1e-5):
    mean = y.mean((0, 2, 3), keepdim=True)
    var = y.var((0, 2, 3), unbiased=False, keepdim=True)
    return (y - mean) / torch.sqrt(var + eps) * gamma.view(1, -1, 1, 1) + beta.view(1, -1, 1, 1)


def check_bn(c, n, h, w, seed=0):
    name = "bn_c%d_n%d_%dx%d" % (c, n, h, w)
    y = bf(gen(n, c, h, w, seed=seed) * 2 + 0.5)
    gamma = gen(c, seed=seed + 1).abs() + 0.5
    beta = gen(c, seed=seed + 2)
    rm = torch.zeros(c)
    rv = torch.ones(c)
    ref = F.batch_norm(y, rm, rv, gamma, beta, True, 0.1, 1e-5)
    ya = act_from(y, 2 * c, c)
    ws = ops.bn_workspace(c, DEV)
    rmd, rvd = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
    st = ops.bn_train_stats(ya, gamma.to(DEV), beta.to(DEV), rmd, rvd, ws)
    out = ops.bn_act(ya, st, True)
    torch.cuda.synchronize()
    res = [(name + "_mean", relerr(st.mean, y.mean((0, 2, 3))), 1e-5 + 1e-4),
           (name + "_running_mean", relerr(rmd, rm), 1e-4),
           (name + "_running_var", relerr(rvd, rv), 1e-4),
           (name + "_act", relerr(out.nchw_float(), F.relu(ref)), 6e-3)]
    # backward of relu(bn(y)) wrt y, gamma, beta
    yg = y.clone().double().requires_grad_(True)
    gg = gamma.clone().double().requires_grad_(True)
    bg = beta.clone().double().requires_grad_(True)
    dout = bf(gen(n, c, h, w, seed=seed + 3))
    F.relu(_bn_ref(yg, gg, bg)).backward(dout.double())
    da = act_from(dout)
    dgam = torch.zeros(c, device=DEV)
    dbet = torch.zeros(c, device=DEV)
    dy = ops.bn_act_bwd(da, ya, st, True, ws, dgam, dbet)
    torch.cuda.synchronize()
    res += [(name + "_bwd_dy", relerr(dy.nchw_float(), yg.grad), 1e-2),
            (name + "_bwd_dgamma", relerr(dgam, gg.grad), 2e-3),
            (name + "_bwd_dbeta", relerr(dbet, bg.grad), 2e-3)]
    # eval affine
    st2 = ops.bn_eval_affine(gamma.to(DEV), beta.to(DEV), rmd, rvd)
    out2 = ops.bn_act(ya, st2, False)
    ref2 = F.batch_norm(y, rmd.cpu(), rvd.cpu(), gamma, beta, False, 0.1, 1e-5)
    torch.cuda.synchronize()
    res.append((name + "_eval", relerr(out2.nchw_float(), ref2), 6e-3))
    return res


def check_block_tail(c, n, h, w, bypass, seed=0):
    name = "tail_c%d_n%d_%dx%d_%s" % (c, n, h, w, "bn" if bypass else "id")
    y2 = bf(gen(n, c, h, w, seed=seed))
    r = bf(gen(n, c, h, w, seed=seed + 1))
    g2, b2 = gen(c, seed=seed + 2).abs() + 0.5, gen(c, seed=seed + 3) * 0.3
    gr, br = gen(c, seed=seed + 4).abs() + 0.5, gen(c, seed=seed + 5) * 0.3
    dout = bf(gen(n, c, h, w, seed=seed + 6))

    def fwd(y2_, r_, g2_, b2_, gr_, br_):
        a = F.relu(_bn_ref(y2_, g2_, b2_))
        s = _bn_ref(r_, gr_, br_) if bypass else r_
        return F.relu(a + s)

    t = [v.clone().double().requires_grad_(True) for v in (y2, r, g2, b2, gr, br)]
    ref = fwd(*t)
    ref.backward(dout.double())
    ws = ops.bn_workspace(c, DEV)
    y2a, ra = act_from(y2), act_from(r, 2 * c, c)
    st2 = ops.bn_train_stats(y2a, g2.to(DEV), b2.to(DEV), None, None, ws)
    str_ = ops.bn_train_stats(ra, gr.to(DEV), br.to(DEV), None, None, ws) if bypass else None
    out = ops.block_tail(y2a, st2, ra, str_)
    torch.cuda.synchronize()
    res = [(name + "_fwd", relerr(out.nchw_float(), ref), 8e-3)]
    # identity residual: the STORED (bf16) out gives the outer mask; BatchNorm'd residual: the kernels recompute it from
    # y2 and res with the forward arithmetic and never read `out` (None here)
    dg2, db2 = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    dgr, dbr = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    dy2, dres = ops.block_tail_bwd(act_from(dout), None if bypass else out, y2a, st2, ra, str_, gr.to(DEV) if bypass else None, ws, dg2, db2,
                                   dgr if bypass else None, dbr if bypass else None)
    torch.cuda.synchronize()
    res += [(name + "_dy2", relerr(dy2.nchw_float(), t[0].grad), 1.5e-2),
            (name + "_dres", relerr(dres.nchw_float(), t[1].grad), 1.5e-2),
            (name + "_dgamma2", relerr(dg2, t[2].grad), 5e-3),
            (name + "_dbeta2", relerr(db2, t[3].grad), 5e-3)]
    if bypass:
        res += [(name + "_dgammar", relerr(dgr, t[4].grad), 5e-3), (name + "_dbetar", relerr(dbr, t[5].grad), 5e-3)]
    return res


def check_maxpool(c, n, h, w, seed=0):
    name = "maxpool_c%d_n%d_%dx%d" % (c, n, h, w)
    x = F.relu(bf(gen(n, c, h, w, seed=seed)))          # post-ReLU input: many exact ties at 0
    xg = x.clone().requires_grad_(True)
    ref = F.max_pool2d(xg, 3, 2, 1)
    dy = bf(gen(n, c, h // 2, w // 2, seed=seed + 1))
    ref.backward(dy)
    xa = act_from(x, 2 * c, c)
    ya = ops.maxpool_fwd(xa)
    dx = act_out(n, h, w, c)
    ops.maxpool_bwd(xa, act_from(dy), dx, False)
    torch.cuda.synchronize()
    res = [(name + "_fwd", relerr(ya.nchw_float(), ref), 0.0), (name + "_bwd", relerr(dx.nchw_float(), xg.grad), 8e-3)]
    # recorded-argmax variant (what the engine uses): identical results
    yb, idx = ops.maxpool_fwd_idx(xa)
    dx2 = act_out(n, h, w, c)
    ops.maxpool_bwd_idx(idx, act_from(dy), dx2, False)
    torch.cuda.synchronize()
    res += [(name + "_fwd_idx", relerr(yb.nchw_float(), ref), 0.0),
            (name + "_bwd_idx", (dx2.nchw_float() - dx.nchw_float()).abs().max().item(), 0.0)]
    return res


def check_dilated(cin, cout, dil, n, h, w, in_ld=None, seed=0):
    """3x3 stride-1 conv with padding == dilation (ASPP branches): fwd (+bias), dgrad, wgrad against torch."""
    name = "dilated_ci%d_co%d_d%d_n%d_%dx%d" % (cin, cout, dil, n, h, w)
    x = bf(gen(n, cin, h, w, seed=seed))
    wt = bf(gen(cout, cin, 3, 3, seed=seed + 1, scale=(2.0 / (9 * cin)) ** 0.5))
    b = gen(cout, seed=seed + 2)
    dy = bf(gen(n, cout, h, w, seed=seed + 3))
    xa = act_from(x, in_ld, 0 if not in_ld else (in_ld - cin))
    ya = act_out(n, h, w, cout, 2 * cout, cout)
    ops.conv2d_dilated_fwd(xa, ops.pack_weights(wt.to(DEV), False, True), b.to(DEV), cout, dil, y=ya)
    dxa = act_out(n, h, w, cin)
    ops.conv2d_dilated_dgrad(act_from(dy), ops.pack_weights(wt.to(DEV), False, False), cin, dil, dx=dxa)
    ws = torch.empty(ops.wgrad_workspace_bytes(n, h, w, cout, cin, 9) // 4 + 4, dtype=torch.float32, device=DEV)
    dw = torch.full((cout, cin, 3, 3), 9.0, dtype=torch.float32, device=DEV)
    ops.conv2d_dilated_wgrad(xa, act_from(dy), cout, dil, dw, False, ws)
    yr = act_out(n, h, w, cout, 2 * cout, cout)
    ops.conv2d_dilated_fwd(xa, ops.pack_weights(wt.to(DEV), False, True), b.to(DEV), cout, dil, y=yr, relu=True)
    torch.cuda.synchronize()
    return [(name + "_fwd", relerr(ya.nchw_float(), F.conv2d(x, wt, b, 1, dil, dil)), 1e-2),
            (name + "_fwd_relu", relerr(yr.nchw_float(), F.relu(F.conv2d(x, wt, b, 1, dil, dil))), 1e-2),
            (name + "_fwd_slice_untouched", (ya.buf[..., :cout].float() - 3.0).abs().max().item(), 0.0),
            (name + "_dgrad", relerr(dxa.nchw_float(), torch.nn.grad.conv2d_input((n, cin, h, w), wt, dy, 1, dil, dil)), 1e-2),
            (name + "_wgrad", relerr(dw, torch.nn.grad.conv2d_weight(x, (cout, cin, 3, 3), dy, 1, dil, dil)), 2e-3)]


def check_maxpool_s1(c, n, h, w, seed=0):
    """MaxPool2d(3, 1, 1) (ASPP block 5): bit-exact forward into a channel slice, gather backward with accumulate."""
    name = "maxpool_s1_c%d_n%d_%dx%d" % (c, n, h, w)
    x = F.relu(bf(gen(n, c, h, w, seed=seed)))          # post-ReLU input: many exact ties at 0
    xg = x.clone().requires_grad_(True)
    ref = F.max_pool2d(xg, 3, 1, 1)
    dy = bf(gen(n, c, h, w, seed=seed + 1))
    ref.backward(dy)
    base = bf(gen(n, c, h, w, seed=seed + 2))
    xa = act_from(x, 2 * c, c)
    ya, idx = ops.maxpool3s1_fwd(xa, y=act_out(n, h, w, c, c + 64, 64), want_idx=True)
    yb, none = ops.maxpool3s1_fwd(xa)
    dx = act_from(base, 2 * c, 0)
    ops.maxpool3s1_bwd_idx(idx, act_from(dy, c + 16, 16), dx, True)
    torch.cuda.synchronize()
    return [(name + "_fwd", relerr(ya.nchw_float(), ref), 0.0), (name + "_fwd_noidx", relerr(yb.nchw_float(), ref), 0.0),
            (name + "_noidx_is_none", 0.0 if none is None else 1.0, 0.0),
            (name + "_bwd_acc", relerr(dx.nchw_float(), bf(xg.grad + base)), 8e-3)]


def check_add(c, n, h, w, seed=0):
    a, b, d = (bf(gen(n, c, h, w, seed=seed + i)) for i in range(3))
    out = ops.add(act_from(a), act_from(b, 2 * c, c), act_from(d))
    torch.cuda.synchronize()
    return [("add_c%d" % c, relerr(out.nchw_float(), a + b + d), 8e-3)]


def check_loss(n, ncls, h, w, seed=0, classw=False):
    name = "loss_n%d_c%d_%dx%d%s" % (n, ncls, h, w, "_cw" if classw else "")
    logits = gen(n, ncls, h, w, seed=seed, scale=3.0)
    g = torch.Generator().manual_seed(seed + 1)
    target = torch.randint(0, ncls, (n, h, w), generator=g)
    target[0, 0, :5] = -100
    pw = torch.rand(n, h, w, generator=g) * 100 + 1
    cw = (torch.rand(ncls, generator=g) + 0.5) if classw else None
    lg = logits.clone().requires_grad_(True)
    logp_ref = F.log_softmax(lg, 1)
    loss_ref = torch.mean(F.nll_loss(logp_ref, target, cw, ignore_index=-100, reduction="none") * pw)
    loss_ref.backward()
    flat = torch.zeros(n * h * w, 4 if ncls <= 4 else 16)
    flat[:, :ncls] = logits.permute(0, 2, 3, 1).reshape(-1, ncls)
    logp = ops.logsoftmax_fwd(flat.to(DEV), n, ncls, h, w)
    cwd = cw.to(DEV) if classw else None
    loss = ops.pixelwise_nll_fwd(logp, target.to(DEV), pw.to(DEV), cwd, -100)
    gout = torch.ones((), device=DEV)
    dlogp = ops.pixelwise_nll_bwd(target.to(DEV), pw.to(DEV), cwd, gout, n, ncls, h, w, -100)
    dl = ops.logsoftmax_bwd(logp, dlogp)
    torch.cuda.synchronize()
    dl_nchw = dl.nchw_float()
    return [(name + "_logp", relerr(logp, logp_ref), 1e-5),
            (name + "_loss", abs(loss.item() - loss_ref.item()) / abs(loss_ref.item()), 1e-5),
            (name + "_dlogits", relerr(dl_nchw[:, :ncls], lg.grad), 6e-3),
            (name + "_dlogits_pad", dl_nchw[:, ncls:].abs().max().item(), 0.0)]


def check_accuracy(n, ncls, h, w, seed=0):
    """fused argmax + per-class counters (ubr_accuracy_counts) against the oracle restatement: integer exact."""
    from oracle import uresnet_port as port
    from ubresnet_b200.training.accuracy import accuracy
    name = "accuracy_n%d_c%d_%dx%d" % (n, ncls, h, w)
    g = torch.Generator().manual_seed(seed)
    logp = torch.log_softmax(bf(torch.randn(n, ncls, h, w, generator=g)), 1)      # bf16-rounded: exact ties occur
    tgt = torch.randint(0, ncls, (n, h, w), generator=g)
    tgt[0, :2, :5] = -100
    cnt = ops.accuracy_counts(logp.to(DEV), tgt.to(DEV)).cpu().tolist()
    ref = port.accuracy_counts(logp, tgt)
    acc = accuracy(logp.to(DEV), tgt.to(DEV), nonzero=True)
    acc_ref = port.accuracy(logp, tgt, nonzero=True)
    return [(name + "_counts", float(sum(abs(a - b) for a, b in zip(cnt, ref))), 0.0),
            (name + "_percent", max(abs(a - b) for a, b in zip(acc, acc_ref)), 1e-9)]


def check_tiling(rows=208, cols=400, crop=64, seed=0):
    import numpy as np
    from oracle import uresnet_port as port
    plan = [(r, c) for r in port.tile_starts(rows, crop) for c in port.tile_starts(cols, crop)]
    plane = gen(rows, cols, seed=seed)
    ref_tiles = port.tile_gather_np(plane.numpy(), plan, crop)
    org = torch.tensor(plan, dtype=torch.int32, device=DEV)
    crops = ops.tile_gather(plane.to(DEV), org, crop)
    torch.cuda.synchronize()
    e1 = float(np.abs(crops[:, 0].cpu().numpy() - ref_tiles).max())
    tiles = gen(len(plan), 3, crop, crop, seed=seed + 1)
    ref_st = port.stitch_scatter_np(tiles.numpy(), plan, rows, cols)
    st = ops.stitch_scatter(tiles.to(DEV), org, rows, cols)
    torch.cuda.synchronize()
    e2 = float(np.abs(st.cpu().numpy() - ref_st).max())
    return [("tile_gather_bitexact", e1, 0.0), ("stitch_scatter_bitexact", e2, 0.0)]


def check_pack(seed=0):
    w = gen(24, 32, 3, 3, seed=seed)
    p = ops.pack_weights(w.to(DEV), False, True)
    torch.cuda.synchronize()
    ref = torch.zeros(10, 32, 32)
    ref[:9, :24, :] = w.permute(2, 3, 0, 1).reshape(9, 24, 32)
    e1 = (p.float().cpu() - bf(ref)).abs().max().item()
    p2 = ops.pack_weights(w.to(DEV), False, False)
    ref2 = torch.zeros(10, 32, 32)
    ref2[:9, :, :24] = w.permute(2, 3, 1, 0).reshape(9, 32, 24)
    e2 = (p2.float().cpu() - bf(ref2)).abs().max().item()
    wt = gen(32, 16, 4, 4, seed=seed + 1)
    p3 = ops.pack_weights(wt.to(DEV), True, True)
    ref3 = torch.zeros(17, 16, 32)
    ref3[:16] = wt.permute(2, 3, 1, 0).reshape(16, 16, 32)
    e3 = (p3.float().cpu() - bf(ref3)).abs().max().item()
    return [("pack_oihw_fwd", e1, 0.0), ("pack_oihw_dgrad", e2, 0.0), ("pack_iohw_fwd", e3, 0.0)]


CHECKS = {
    "pack": lambda: check_pack(),
    "pointwise": lambda: (check_bn(16, 2, 32, 32) + check_bn(64, 3, 16, 16, seed=5) + check_bn(512, 4, 4, 4, seed=9)
                          + check_block_tail(32, 2, 16, 16, True) + check_block_tail(32, 2, 16, 16, False, seed=3)
                          + check_block_tail(256, 2, 8, 8, True, seed=7)
                          + check_maxpool(16, 2, 32, 32) + check_add(32, 2, 8, 8)),
    "stem": lambda: check_stem(2, 64, 64) + check_stem(1, 40, 72, oc=32, seed=4),
    "stem_tc": lambda: (check_stem(2, 64, 128, seed=5) + check_stem(1, 72, 256, seed=6) + check_stem(3, 128, 384, seed=7)
                        + check_stem(32, 40, 512, seed=8)),
    "loss": lambda: check_loss(2, 3, 32, 32) + check_loss(2, 4, 16, 48, seed=3, classw=True),
    "tiling": lambda: check_tiling(),
    "optim": lambda: check_optim(),
    # inference epilogue on each output path: halo direct stores (Cout <= 32), halo TMA-store (Cout 64), implicit GEMM
    # (C >= 128, stride 2, partial tiles), residual / output living in channel slices of wider buffers
    "conv_fused_halo_direct": lambda: (check_conv_fwd_fused(16, 16, 3, 1, 2, 32, 128)
                                       + check_conv_fwd_fused(32, 32, 3, 1, 1, 16, 256, res_ld=64, out_ld=64, seed=3)
                                       + check_conv_fwd_fused(16, 16, 7, 1, 1, 16, 128, with_res=False, relu_post=False, seed=5)),
    "conv_fused_halo_tma": lambda: (check_conv_fwd_fused(64, 64, 3, 1, 2, 16, 128, seed=7)
                                    + check_conv_fwd_fused(32, 64, 1, 1, 1, 8, 128, relu_pre=False, with_res=False,
                                                           relu_post=False, seed=9)
                                    + check_conv_fwd_fused(64, 64, 3, 1, 1, 8, 128, res_ld=128, out_ld=128, seed=11)),
    "conv_fused_gemm": lambda: (check_conv_fwd_fused(128, 128, 3, 1, 2, 16, 16, seed=13)
                                + check_conv_fwd_fused(64, 128, 3, 2, 2, 32, 32, with_res=False, relu_post=False, seed=15)
                                + check_conv_fwd_fused(128, 256, 1, 2, 1, 24, 40, relu_pre=False, with_res=False,
                                                       relu_post=False, seed=17)
                                + check_conv_fwd_fused(256, 256, 3, 1, 3, 12, 20, res_ld=512, seed=19)
                                + check_conv_fwd_fused(32, 32, 3, 1, 2, 24, 40, seed=21)),
    "accuracy": lambda: check_accuracy(2, 3, 64, 96) + check_accuracy(3, 4, 33, 17, seed=5),
    # tensor-core forward: each swizzle mode / K-chunk count / N-tile count / stride / kernel size
    "conv_fwd_c64": lambda: check_conv_fwd(64, 64, 3, 1, 2, 16, 16),
    "conv_fwd_c16": lambda: check_conv_fwd(16, 32, 3, 1, 2, 32, 32),
    "conv_fwd_c32": lambda: check_conv_fwd(32, 32, 3, 1, 2, 32, 32, seed=2),
    "conv_fwd_big": lambda: check_conv_fwd(128, 256, 3, 1, 2, 16, 16, seed=3) + check_conv_fwd(512, 512, 3, 1, 4, 8, 8, seed=4),
    "conv_fwd_wide": lambda: check_conv_fwd(16, 16, 3, 1, 1, 8, 256, seed=5) + check_conv_fwd(64, 32, 3, 1, 3, 24, 40, seed=6),
    "conv_fwd_s2": lambda: check_conv_fwd(32, 64, 3, 2, 2, 32, 32, seed=7) + check_conv_fwd(64, 128, 1, 2, 2, 16, 16, seed=8),
    "conv_fwd_1x1": lambda: check_conv_fwd(32, 16, 1, 1, 2, 32, 32, seed=9) + check_conv_fwd(512, 256, 1, 1, 2, 8, 8, seed=10),
    "conv_fwd_7x7": lambda: check_conv_fwd(16, 16, 7, 1, 2, 32, 32, seed=11, bias=True)
                            + check_conv_fwd(16, 3, 7, 1, 2, 32, 32, seed=12, bias=True, f32=True),
    "conv_fwd_slices": lambda: check_conv_fwd(32, 16, 3, 1, 2, 16, 16, in_ld=64, out_ld=32, seed=13),
    "conv_dgrad": lambda: (check_conv_dgrad(64, 64, 3, 1, 2, 16, 16) + check_conv_dgrad(16, 32, 3, 1, 2, 32, 32, seed=2)
                           + check_conv_dgrad(256, 128, 3, 1, 2, 8, 8, seed=3) + check_conv_dgrad(16, 16, 7, 1, 1, 32, 32, seed=4)),
    "conv_dgrad_s2": lambda: (check_conv_dgrad(32, 64, 3, 2, 2, 32, 32, seed=5)
                              + check_conv_dgrad(64, 128, 1, 2, 2, 16, 16, seed=6)),
    "deconv": lambda: check_deconv(32, 16, 2, 16, 16) + check_deconv(512, 256, 2, 4, 4, out_ld=512, seed=3),
    "deconv_subpixel": lambda: (check_deconv(32, 16, 2, 32, 128, out_ld=32, seed=71) + check_deconv(64, 32, 1, 32, 128, seed=72)
                                + check_deconv(16, 16, 1, 16, 256, seed=73)),
    "conv_wgrad_c64": lambda: check_conv_wgrad(64, 64, 3, 1, 2, 16, 16),
    "conv_wgrad_c16": lambda: check_conv_wgrad(16, 32, 3, 1, 2, 32, 32, seed=2) + check_conv_wgrad(32, 16, 3, 1, 2, 32, 32, seed=3),
    "conv_wgrad_big": lambda: check_conv_wgrad(128, 256, 3, 1, 4, 8, 8, seed=4) + check_conv_wgrad(256, 128, 1, 1, 4, 8, 8, seed=5),
    "conv_wgrad_s2": lambda: check_conv_wgrad(32, 64, 3, 2, 2, 32, 32, seed=6) + check_conv_wgrad(64, 128, 1, 2, 2, 16, 16, seed=7),
    "conv_wgrad_7x7": lambda: check_conv_wgrad(16, 16, 7, 1, 2, 32, 32, seed=8) + check_conv_wgrad(16, 3, 7, 1, 2, 32, 32, seed=9, cout_pad=16),
    # halo-resident kernels (W % 128 == 0, C <= 64, stride 1): every swizzle span, tap set, output mode
    "halo_fwd_c16": lambda: check_conv_fwd(16, 16, 3, 1, 2, 64, 128, seed=21) + check_conv_fwd(16, 32, 3, 1, 1, 32, 256, seed=22),
    "halo_fwd_c32": lambda: check_conv_fwd(32, 32, 3, 1, 2, 32, 128, seed=23) + check_conv_fwd(32, 16, 1, 1, 2, 40, 128, seed=24),
    "halo_fwd_c64": lambda: check_conv_fwd(64, 64, 3, 1, 1, 32, 128, seed=25) + check_conv_fwd(64, 32, 1, 1, 1, 16, 256, seed=26),
    "halo_fwd_7x7": lambda: (check_conv_fwd(16, 16, 7, 1, 1, 64, 128, seed=27, bias=True)
                             + check_conv_fwd(16, 3, 7, 1, 1, 32, 256, seed=28, bias=True, f32=True)),
    "halo_fwd_slices": lambda: check_conv_fwd(32, 16, 3, 1, 2, 32, 128, in_ld=64, out_ld=32, seed=29),
    "halo_dgrad": lambda: (check_conv_dgrad(16, 32, 3, 1, 2, 64, 128, seed=31) + check_conv_dgrad(128, 64, 3, 1, 1, 32, 128, seed=32)
                           + check_conv_dgrad(16, 16, 7, 1, 1, 32, 256, seed=33) + check_conv_dgrad(64, 32, 1, 1, 1, 32, 128, seed=34)),
    "halo_fwd_bn": lambda: (check_conv_fwd_bn(16, 16, 3, 2, 64, 128, seed=61) + check_conv_fwd_bn(32, 32, 3, 1, 32, 256, seed=62)
                            + check_conv_fwd_bn(16, 16, 7, 1, 32, 128, seed=63, bias=True) + check_conv_fwd_bn(64, 32, 1, 1, 32, 128, seed=64)
                            + check_conv_fwd_bn(64, 64, 3, 1, 32, 128, seed=65) + check_conv_fwd_bn(16, 32, 3, 2, 32, 32, seed=66)),
    # statistics fused in the implicit-GEMM epilogue (C >= 128 / narrow layers): one partial row per 128-pixel tile
    "gemm_fwd_bn": lambda: (check_conv_fwd_bn(128, 256, 3, 2, 16, 16, seed=67) + check_conv_fwd_bn(512, 512, 1, 4, 8, 8, seed=68, bias=True)
                            + check_conv_fwd_bn(64, 128, 3, 8, 32, 32, seed=69) + check_conv_fwd_bn(256, 16, 3, 1, 24, 40, seed=70)),
    "halo_dgrad_acc": lambda: (check_conv_dgrad_acc(16, 16, 3, 2, 64, 128, seed=51) + check_conv_dgrad_acc(32, 16, 3, 1, 32, 256, seed=52)
                               + check_conv_dgrad_acc(64, 64, 3, 1, 32, 128, seed=53)          # halo, TMA-store path
                               + check_conv_dgrad_acc(16, 16, 3, 1, 32, 32, seed=54)           # implicit GEMM
                               + check_conv_dgrad_acc(128, 256, 3, 3, 12, 20, seed=55)         # GEMM, partial tiles
                               + check_conv_dgrad_acc(256, 128, 1, 2, 16, 16, seed=56)),
    "halo_wgrad_3x3": lambda: (check_conv_wgrad(16, 16, 3, 1, 2, 64, 128, seed=41) + check_conv_wgrad(32, 32, 3, 1, 1, 32, 256, seed=42)
                               + check_conv_wgrad(64, 64, 3, 1, 1, 32, 128, seed=43) + check_conv_wgrad(16, 32, 3, 1, 1, 32, 128, seed=44)
                               + check_conv_wgrad(64, 32, 3, 1, 1, 32, 128, seed=45)),
    "halo_wgrad_1x1": lambda: check_conv_wgrad(32, 16, 1, 1, 2, 32, 128, seed=46) + check_conv_wgrad(64, 32, 1, 1, 1, 32, 128, seed=47),
    "halo_wgrad_7x7": lambda: (check_conv_wgrad(16, 16, 7, 1, 1, 64, 128, seed=48)
                               + check_conv_wgrad(16, 3, 7, 1, 1, 32, 256, seed=49, cout_pad=16)),
}


def _with_conv_tails(fn):
    """run a check with the in-launch finalize of the conv epilogues switched on (ops.TREE_FINALIZE_CONV, off by default)"""
    def run():
        old = ops.TREE_FINALIZE_CONV
        ops.TREE_FINALIZE_CONV = ops.TREE_FINALIZE
        try:
            return [(lbl + " [in-launch finalize]", err, tol) for lbl, err, tol in fn()]
        finally:
            ops.TREE_FINALIZE_CONV = old
    return run


# the same statistics checks with the reduction finished inside the conv launch (ubr_conv2d_fwd_bn, tree_reduce.cuh)
CHECKS.update({
    # channel counts beyond the in-launch tail's shared memory (c = 2048: the *_final entry points answer 1 and the
    # two-launch form runs) and at its edge (c = 1024), many partial rows (two ticket levels)
    "pointwise_wide": lambda: (check_bn(2048, 2, 4, 4, seed=21) + check_bn(1024, 2, 8, 8, seed=22) + check_bn(16, 4, 128, 128, seed=23)
                               + check_block_tail(1024, 2, 4, 4, True, seed=24) + check_block_tail(2048, 1, 4, 4, False, seed=25)),
    "halo_fwd_bn_tail": _with_conv_tails(CHECKS["halo_fwd_bn"]),
    "gemm_fwd_bn_tail": _with_conv_tails(CHECKS["gemm_fwd_bn"]),
})


# ASPP_ResNet operators (reference models/ASPP_ResNet.py:188-286): dilated branches, stride-1 max-pool, and the
# 192/320/576/384/768-channel concat operands (64-wide N tiles in conv_gemm / wgrad_gemm)
CHECKS.update({
    # eval conv->ReLU->conv chains, one launch (N1): the three resolutions / channel widths of the U-ResNet and odd bands
    "block_chain_512": lambda: (check_block_chain(32, 16, 2, 64, 512, seed=201) + check_block_chain(16, 16, 1, 96, 512, seed=202)),
    "block_chain_256": lambda: (check_block_chain(16, 32, 2, 64, 256, seed=203) + check_block_chain(32, 32, 3, 40, 256, seed=204)
                                + check_block_chain(64, 32, 2, 32, 256, in_ld=128, out_ld=64, seed=205)),
    "block_chain_128": lambda: (check_block_chain(64, 64, 2, 64, 128, seed=206) + check_block_chain(16, 16, 1, 8, 128, seed=207)
                                + check_block_chain(32, 64, 5, 16, 128, seed=208)),
})

CHECKS.update({
    "aspp_dilated_d3": lambda: check_dilated(128, 16, 3, 2, 16, 16, seed=81) + check_dilated(256, 16, 3, 1, 8, 8, in_ld=512, seed=82),
    "aspp_dilated_d5": lambda: check_dilated(128, 16, 5, 2, 16, 16, seed=83) + check_dilated(512, 16, 5, 3, 4, 4, seed=84),
    "aspp_dilated_d1": lambda: check_dilated(64, 32, 1, 1, 16, 24, seed=85),
    "aspp_maxpool_s1": lambda: check_maxpool_s1(128, 2, 16, 16, seed=86) + check_maxpool_s1(16, 1, 5, 9, seed=87),
    "aspp_combine_1x1": lambda: (check_conv_fwd(192, 128, 1, 1, 2, 16, 16, seed=88, bias=True)
                                 + check_conv_dgrad(192, 128, 1, 1, 2, 16, 16, seed=89)
                                 + check_conv_wgrad(192, 128, 1, 1, 2, 16, 16, seed=90)
                                 + check_conv_dgrad(320, 256, 1, 1, 2, 8, 8, seed=91)
                                 + check_conv_wgrad(576, 512, 1, 1, 4, 4, 4, seed=92)),
    "aspp_decoder_convs": lambda: (check_conv_fwd(384, 256, 3, 1, 1, 16, 16, seed=93) + check_conv_dgrad(384, 256, 3, 1, 1, 16, 16, seed=94)
                                   + check_conv_wgrad(384, 256, 3, 1, 1, 16, 16, seed=95)
                                   + check_conv_wgrad(768, 512, 1, 1, 2, 8, 8, seed=96)
                                   + check_deconv(1024, 256, 1, 4, 4, out_ld=768, seed=97)),
})


def run(name):
    res = CHECKS[name]()
    ok = True
    for label, err, tol in res:
        good = err <= tol
        ok &= good
        print("%-4s %-60s err=%.3e tol=%.1e" % ("ok" if good else "FAIL", label, err, tol), flush=True)
    return ok


if __name__ == "__main__":
    names = sys.argv[1:] or list(CHECKS)
    allok = True
    for nm in names:
        allok &= run(nm)
    sys.exit(0 if allok else 1)
