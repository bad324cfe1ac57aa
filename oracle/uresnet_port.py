"""TEST INFRASTRUCTURE ONLY -- CPU fp32 restatement of the ubresnet hot path.

A functional (state_dict driven) plain-PyTorch statement of what the reference
computes, so it can travel to the GPU box where `/root/reference` is absent.
It is the parity oracle and the `cpu_baseline`/`--impl reference` arm of
`bench.py` (kind "port").  It must never be imported from `ubresnet_b200/`.

Pinned (in this container) against the unmodified reference via
`oracle/reference_shim.py` -- see `tests/test_oracle_pin.py` -- and against the
fixtures under `tests/golden/` that `oracle/make_golden.py` produced FROM the
reference.  Reference tests/golden vectors for this path: none exist
(SURVEY.md section 4), so the pin is "reference outputs generated here".

What each function follows (paths relative to /root/reference):
  stem / head / forward order ........ models/ub_uresnet.py:88-147
  residual block (ReLU-add-ReLU) ..... models/common_layers.py:39-58
  double block ....................... models/common_layers.py:109-120
  deconv + [deconv, skip] concat ..... models/common_layers.py:122-132
  init ............................... models/ub_uresnet.py:73-79
  loss ............................... training/pixelwise_nllloss.py:41-61
"""
import math
from collections import OrderedDict

import torch
import torch.nn.functional as F

BN_EPS = 1e-5        # nn.BatchNorm2d default used at ub_uresnet.py:42, common_layers.py:24
BN_MOMENTUM = 0.1


# ----------------------------------------------------------------------------------------------
# parameter inventory (names == reference state_dict keys)
# ----------------------------------------------------------------------------------------------
def _block_entries(prefix, cin, cout, stride):
    ents = [(prefix + "conv1.weight", ("conv", cout, cin, 3)),
            (prefix + "bn1", ("bn", cout)),
            (prefix + "conv2.weight", ("conv", cout, cout, 3)),
            (prefix + "bn2", ("bn", cout))]
    if cin != cout or stride > 1:
        ents += [(prefix + "bypass.weight", ("conv", cout, cin, 1)),
                 (prefix + "bnpass", ("bn", cout))]
    return ents


def inventory(num_classes=3, input_channels=1, inplanes=16, final_conv_kernels=16):
    """Ordered list of (name, spec) in the reference's module registration order."""
    p = inplanes
    ents = [("conv1.weight", ("conv", p, input_channels, 7)), ("conv1.bias", ("bias", p, input_channels * 49)),
            ("bn1", ("bn", p))]
    enc = [(p, 2 * p, 1), (2 * p, 4 * p, 2), (4 * p, 8 * p, 2), (8 * p, 16 * p, 2), (16 * p, 32 * p, 2)]
    for i, (ci, co, s) in enumerate(enc, 1):
        ents += _block_entries("enc_layer%d.res1." % i, ci, co, s)
        ents += _block_entries("enc_layer%d.res2." % i, co, co, 1)
    dec = [(5, 32 * p, 16 * p), (4, 16 * p, 8 * p), (3, 8 * p, 4 * p), (2, 4 * p, 2 * p), (1, 2 * p, p)]
    for i, ci, cd in dec:
        ents += [("dec_layer%d.deconv.weight" % i, ("deconv", ci, cd, 4))]
        ents += _block_entries("dec_layer%d.res.res1." % i, 2 * cd, cd, 1)
        ents += _block_entries("dec_layer%d.res.res2." % i, cd, cd, 1)
    k = final_conv_kernels
    ents += [("conv10.weight", ("conv", k, p, 7)), ("conv10.bias", ("bias", k, p * 49)), ("bn10", ("bn", k)),
             ("conv11.weight", ("conv", num_classes, k, 7)), ("conv11.bias", ("bias", num_classes, k * 49))]
    return ents


class _RoundBF16(torch.autograd.Function):
    """bf16 storage emulation: round the value forward and the gradient backward (used only by the
    `quant=True` diagnostic mode that mimics where the CUDA pipeline stores bf16)."""

    @staticmethod
    def forward(ctx, t):
        return t.to(torch.bfloat16).to(torch.float32)

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).to(torch.float32)


_QUANT = [False]


def _q(t):
    return _RoundBF16.apply(t) if _QUANT[0] else t


def _bn(x, sd, name, training):
    return F.batch_norm(x, sd[name + ".running_mean"], sd[name + ".running_var"],
                        sd[name + ".weight"], sd[name + ".bias"], training, BN_MOMENTUM, BN_EPS)


def _qw(w):
    """weights are used as bf16 GEMM operands by the CUDA path (no gradient rounding: wgrad is fp32)"""
    return (w + (w.detach().to(torch.bfloat16).to(torch.float32) - w.detach())) if _QUANT[0] else w


def _bump(sd, name, training):
    if training and (name + ".num_batches_tracked") in sd:
        sd[name + ".num_batches_tracked"] += 1


def _res_block(x, sd, pre, stride, training, taps):
    # common_layers.py:39-58 -- note relu2 BEFORE the add, and a second ReLU after it
    y = _q(F.conv2d(x, _qw(sd[pre + "conv1.weight"]), None, stride, 1))
    y = _q(F.relu(_bn(y, sd, pre + "bn1", training))); _bump(sd, pre + "bn1", training)
    y = _q(F.conv2d(y, _qw(sd[pre + "conv2.weight"]), None, 1, 1))
    y = F.relu(_bn(y, sd, pre + "bn2", training)); _bump(sd, pre + "bn2", training)
    if (pre + "bypass.weight") in sd:
        s = _q(F.conv2d(x, _qw(sd[pre + "bypass.weight"]), None, stride, 0))
        s = _bn(s, sd, pre + "bnpass", training); _bump(sd, pre + "bnpass", training)
    else:
        s = x
    out = _q(F.relu(y + s))
    if taps is not None:
        taps[pre[:-1]] = out
    return out


def _double(x, sd, pre, stride, training, taps):
    x = _res_block(x, sd, pre + "res1.", stride, training, taps)
    return _res_block(x, sd, pre + "res2.", 1, training, taps)


def forward(sd, x, training=True, taps=None, return_logits=False, quant=False):
    _QUANT[0] = bool(quant)
    try:
        return _forward(sd, x, training, taps, return_logits)
    finally:
        _QUANT[0] = False


def _forward(sd, x, training=True, taps=None, return_logits=False):
    """log-probabilities (B,C,H,W) exactly as UResNet.forward (ub_uresnet.py:88-147).

    `sd` maps reference state_dict names to tensors; BN running buffers are updated in place
    when `training`.  `taps` (optional dict) receives intermediate activations by name.
    """
    y = _q(F.conv2d(x, sd["conv1.weight"], sd["conv1.bias"], 1, 3))
    x0 = _q(F.relu(_bn(y, sd, "bn1", training))); _bump(sd, "bn1", training)
    h = F.max_pool2d(x0, 3, 2, 1)
    if taps is not None:
        taps["x0"] = x0
    skips = [x0]
    for i, s in zip(range(1, 6), (1, 2, 2, 2, 2)):
        h = _double(h, sd, "enc_layer%d." % i, s, training, taps)
        skips.append(h)
    for i in (5, 4, 3, 2, 1):
        skip = skips[i - 1]
        # ConvTranspose2d(k4,s2,p1) with output_size == skip size -> output_padding 0 (common_layers.py:128)
        up = _q(F.conv_transpose2d(h, _qw(sd["dec_layer%d.deconv.weight" % i]), None, 2, 1))
        assert up.shape[-2:] == skip.shape[-2:]
        h = torch.cat([up, skip], 1)                           # [deconv, skip]  (common_layers.py:130)
        h = _double(h, sd, "dec_layer%d.res." % i, 1, training, taps)
    h = _q(F.conv2d(h, _qw(sd["conv10.weight"]), sd["conv10.bias"], 1, 3))
    h = _q(F.relu(_bn(h, sd, "bn10", training))); _bump(sd, "bn10", training)
    logits = F.conv2d(h, _qw(sd["conv11.weight"]), sd["conv11.bias"], 1, 3)
    if taps is not None:
        taps["logits"] = logits
    logp = F.log_softmax(logits, dim=1)
    return (logp, logits) if return_logits else logp


def pixelwise_nll(logp, target, pixelweights, class_weight=None, ignore_index=-100):
    """mean_{b,h,w}( pw * cw[t] * -logp[t] ); ignored pixels stay in the denominator
    (pixelwise_nllloss.py:51-59: nll_loss(reduce=False) * pixelweights, then torch.mean)."""
    per_pixel = F.nll_loss(logp, target, class_weight, ignore_index=ignore_index, reduction="none")
    return torch.mean(per_pixel * pixelweights)


def init_state_dict(num_classes=3, input_channels=1, inplanes=16, final_conv_kernels=16, seed=None,
                    generator=None):
    """Reference-style init (ub_uresnet.py:73-79): conv/deconv W ~ N(0, sqrt(2/(k*k*out_channels))),
    BN gamma=1 beta=0; conv biases U(-1/sqrt(fan_in), 1/sqrt(fan_in)) (torch default).
    NOTE: draws are NOT bit-identical to constructing the reference module (different RNG order);
    tests that need identical weights copy a state_dict instead."""
    g = generator
    if g is None:
        g = torch.Generator().manual_seed(0 if seed is None else seed)
    sd = OrderedDict()
    for name, spec in inventory(num_classes, input_channels, inplanes, final_conv_kernels):
        kind = spec[0]
        if kind == "conv":
            _, co, ci, k = spec
            sd[name] = torch.randn(co, ci, k, k, generator=g) * math.sqrt(2.0 / (k * k * co))
        elif kind == "deconv":
            _, ci, co, k = spec
            # nn.ConvTranspose2d.out_channels == co
            sd[name] = torch.randn(ci, co, k, k, generator=g) * math.sqrt(2.0 / (k * k * co))
        elif kind == "bias":
            _, co, fan_in = spec
            b = 1.0 / math.sqrt(fan_in)
            sd[name] = (torch.rand(co, generator=g) * 2 - 1) * b
        else:
            c = spec[1]
            sd[name + ".weight"] = torch.ones(c)
            sd[name + ".bias"] = torch.zeros(c)
            sd[name + ".running_mean"] = torch.zeros(c)
            sd[name + ".running_var"] = torch.ones(c)
            sd[name + ".num_batches_tracked"] = torch.zeros((), dtype=torch.long)
    return sd


def clone_state_dict(sd, requires_grad=False):
    out = OrderedDict()
    for k, v in sd.items():
        t = v.detach().clone()
        if requires_grad and t.is_floating_point() and "running_" not in k:
            t.requires_grad_(True)
        out[k] = t
    return out


def train_step_grads(sd, x, target, pixelweights, class_weight=None, quant=False):
    """fwd + loss + bwd; returns (loss, logp, grads dict, updated-buffer state_dict).
    quant=True: diagnostic bf16-storage emulation (NOT the parity oracle)."""
    work = clone_state_dict(sd, requires_grad=True)
    logp = forward(work, x, training=True, quant=quant)
    loss = pixelwise_nll(logp, target, pixelweights, class_weight)
    loss.backward()
    grads = OrderedDict((k, v.grad.detach()) for k, v in work.items() if v.requires_grad)
    return loss.detach(), logp.detach(), grads, work


def accuracy_counts(output, target):
    """[pixels per class | correct pixels per class] (2*C ints) behind `accuracy()`
    (training/train_ubresnet2018_wlarcv2.py:519-547: pred = output.max(1), correct = pred.eq(target), per-class sums)."""
    pred = output.max(1)[1]
    tgt = target.reshape(pred.shape)
    ncls = output.size(1)
    num = [int((tgt == c).sum()) for c in range(ncls)]
    corr = [int(((pred == tgt) & (tgt == c)).sum()) for c in range(ncls)]
    return num + corr


def accuracy(output, target, nonzero=False):
    """percent accuracies: per class (0.0 if absent), total (:552-561); nonzero adds the track+shower value of
    training/train_ubresnet2018_wlarcv1.py:583."""
    ncls = output.size(1)
    cnt = accuracy_counts(output, target)
    num, corr = cnt[:ncls], cnt[ncls:]
    res = [100.0 * corr[c] / num[c] if num[c] > 0 else 0.0 for c in range(ncls)]
    res.append(100.0 * float(sum(corr)) / sum(num))
    if nonzero:
        res.append(100.0 * float(corr[1] + corr[2]) / (num[1] + num[2]))
    return res


# ----------------------------------------------------------------------------------------------
# whole-view tiling (numpy restatement; SURVEY.md 8c: builder-defined plan, slice + overwrite)
# ----------------------------------------------------------------------------------------------
def tile_starts(extent, crop, ntiles=None):
    """Evenly spread crop origins covering [0, extent): first at 0, last at extent-crop."""
    if extent <= crop:
        return [0]
    if ntiles is None:
        ntiles = -(-extent // crop)          # ceil
        # require some overlap for context at seams (matches SURVEY plan: 1008 -> 2 rows, 3456 -> 7 cols)
        while ntiles * crop - extent < (ntiles - 1) * min(16, crop // 4):
            ntiles += 1
    span = extent - crop
    return [int(math.floor(i * span / (ntiles - 1) + 0.5)) for i in range(ntiles)]


def tile_plan(rows=1008, cols=3456, crop=512):
    """[(row0, col0)] in row-major tile order."""
    return [(r, c) for r in tile_starts(rows, crop) for c in tile_starts(cols, crop)]


def tile_gather_np(plane, plan, crop=512):
    """plane (rows, cols) -> (ntiles, crop, crop): numpy slicing as run_ubresnet_wholeview.py:238-239."""
    import numpy as np
    out = np.zeros((len(plan), crop, crop), dtype=plane.dtype)
    for i, (r0, c0) in enumerate(plan):
        out[i] = plane[r0:r0 + crop, c0:c0 + crop]
    return out


def stitch_scatter_np(tiles, plan, rows, cols):
    """(ntiles, C, crop, crop) -> (C, rows, cols); sequential overwrite in ascending tile index
    == Image2D::overlay(kOverWrite) applied tile by tile (Image2D.cxx:362-367)."""
    import numpy as np
    c = tiles.shape[1]
    crop = tiles.shape[-1]
    out = np.zeros((c, rows, cols), dtype=tiles.dtype)
    for i, (r0, c0) in enumerate(plan):
        out[:, r0:r0 + crop, c0:c0 + crop] = tiles[i]
    return out
