"""TEST INFRASTRUCTURE ONLY -- CPU fp32 restatement of the reference ASPP_ResNet forward (config 5).

Functional, state_dict driven, plain PyTorch; travels to the GPU box where `/root/reference` is absent.  Only
`tests/` may import it.  Pinned in this container against the unmodified reference class (loaded through
`oracle/reference_shim.load_reference_aspp`) and against `tests/golden/aspp_b2_128.pt`, which `oracle/make_golden.py`
generated FROM the reference.  The reference has no tests or golden vectors of its own for this model.

What each function follows (paths relative to /root/reference):
  ASPP branches + pooled input, concat order .... models/ASPP_ResNet.py:188-262
  ASPP_post (1x1 -> BN -> ReLU) ................. models/ASPP_ResNet.py:265-286
  forward order, skip = cat(aspp, enc) .......... models/ASPP_ResNet.py:416-523
  blocks / deconv+concat ........................ models/ASPP_ResNet.py:62-186 (same as common_layers.py)
"""
import torch
import torch.nn.functional as F

from . import uresnet_port as _up
from .uresnet_port import _bn, _bump, _double, _q, _qw, pixelwise_nll, clone_state_dict  # noqa: F401

_BRANCHES = (("B1", 0, 1), ("B2", 1, 1), ("B3", 3, 3), ("B4", 5, 5))      # (prefix, padding, dilation)


def _aspp(x, sd, pre, training):
    outs = []
    for b, pad, dil in _BRANCHES:
        y = _q(F.conv2d(x, _qw(sd[pre + b + "_conv.weight"]), sd[pre + b + "_conv.bias"], 1, pad, dil))
        y = _q(F.relu(_bn(y, sd, pre + b + "_bn", training))); _bump(sd, pre + b + "_bn", training)
        outs.append(y)
    outs.append(F.max_pool2d(x, 3, 1, 1))                              # block 5 (:217-219, 242)
    return torch.cat(outs, 1)


def _aspp_post(x, sd, pre, training):
    y = _q(F.conv2d(x, _qw(sd[pre + "ASPP_conv.weight"]), sd[pre + "ASPP_conv.bias"]))
    y = _q(F.relu(_bn(y, sd, pre + "ASPP_bn", training))); _bump(sd, pre + "ASPP_bn", training)
    return y


def forward(sd, x, training=True, taps=None, quant=False):
    """log-probabilities (B,C,H,W) as ASPP_ResNet.forward (models/ASPP_ResNet.py:416-523).
    quant=True: diagnostic bf16-storage emulation (same points as uresnet_port), NOT the parity oracle."""
    _up._QUANT[0] = bool(quant)
    try:
        return _forward(sd, x, training, taps)
    finally:
        _up._QUANT[0] = False


def _forward(sd, x, training, taps):
    y = _q(F.conv2d(x, sd["conv1.weight"], sd["conv1.bias"], 1, 3))
    x0 = _q(F.relu(_bn(y, sd, "bn1", training))); _bump(sd, "bn1", training)
    h = F.max_pool2d(x0, 3, 2, 1)
    e = {}
    for i, s in zip(range(1, 6), (1, 2, 2, 2, 2)):
        h = _double(h, sd, "enc_layer%d." % i, s, training, taps)
        e[i] = h
    skip = {1: x0, 2: e[1], 3: e[2]}
    for lvl in (3, 4, 5):
        a = _aspp(e[lvl], sd, "ASPP_layer_enc%d." % lvl, training)
        a = _aspp_post(a, sd, "ASPP_combine_enc%d." % lvl, training)
        if taps is not None:
            taps["aspp%d" % lvl] = a
        skip[lvl + 1] = torch.cat((a, e[lvl]), 1)                      # (:459, 471, 483)
    h = skip[6]
    for i in (5, 4, 3, 2, 1):
        up = _q(F.conv_transpose2d(h, _qw(sd["dec_layer%d.deconv.weight" % i]), None, 2, 1))
        h = torch.cat([up, skip[i]], 1)
        h = _double(h, sd, "dec_layer%d.res." % i, 1, training, taps)
    h = _q(F.conv2d(h, _qw(sd["conv10.weight"]), sd["conv10.bias"], 1, 3))
    h = _q(F.relu(_bn(h, sd, "bn10", training))); _bump(sd, "bn10", training)
    logits = F.conv2d(h, _qw(sd["conv11.weight"]), sd["conv11.bias"], 1, 3)
    if taps is not None:
        taps["logits"] = logits
    return F.log_softmax(logits, dim=1)


def train_step_grads(sd, x, target, pixelweights, class_weight=None, quant=False):
    """fwd + PixelWiseNLLLoss + bwd; returns (loss, logp, grads dict, updated-buffer state_dict)."""
    from collections import OrderedDict
    work = clone_state_dict(sd, requires_grad=True)
    logp = forward(work, x, training=True, quant=quant)
    loss = pixelwise_nll(logp, target, pixelweights, class_weight)
    loss.backward()
    grads = OrderedDict((k, v.grad.detach()) for k, v in work.items() if v.requires_grad)
    return loss.detach(), logp.detach(), grads, work
