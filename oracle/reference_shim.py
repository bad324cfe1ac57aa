"""TEST INFRASTRUCTURE ONLY -- loader for the *unmodified* reference modules.

Imports `/root/reference/models/{common_layers,ub_uresnet}.py` and
`/root/reference/training/pixelwise_nllloss.py` under Python 3 / torch 2.x by
patching the source TEXT at load time (no reference file is edited or copied):

  1. py2 `print a, b` statements            -> `print(a, b)`      (ub_uresnet.py:91 ff.)
  2. `out += outbp` / `out += x`             -> `out = out + ...`  (common_layers.py:52,54;
     value-identical, but legal for modern autograd which saves the ReLU output)
  3. `criterion.size_average = True`          (pixelwise_nllloss.py:51 reads an attribute
     that `_Loss` no longer stores)

This only works inside the build container where `/root/reference` is mounted.
It is used by `oracle/make_golden.py` (fixture generation) and by the CPU tests
that pin `oracle/uresnet_port.py` against the real reference.  Nothing in the
product path (`ubresnet_b200/`) may import this module.
"""
import os
import re
import sys
import types
import warnings

REFERENCE_ROOT = os.environ.get("UBRESNET_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "models", "ub_uresnet.py"))


_PRINT_RE = re.compile(r"^(\s*)print (.*)$", re.M)
_IADD_RE = re.compile(r"^(\s*)out \+= (\w+)\s*$", re.M)


def _load(name: str, relpath: str, inject=None) -> types.ModuleType:
    with open(os.path.join(REFERENCE_ROOT, relpath), "r") as fh:
        text = fh.read()
    text = _PRINT_RE.sub(r"\1print(\2)", text)
    text = _IADD_RE.sub(r"\1out = out + \2", text)
    mod = types.ModuleType(name)
    mod.__file__ = os.path.join(REFERENCE_ROOT, relpath)
    if inject:
        mod.__dict__.update(inject)
    code = compile(text, mod.__file__, "exec")
    exec(code, mod.__dict__)
    return mod


_cache = {}


def load_reference():
    """Returns (ub_uresnet module, common_layers module, pixelwise_nllloss module)."""
    if "mods" in _cache:
        return _cache["mods"]
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    saved = {k: sys.modules.get(k) for k in ("common_layers", "ub_uresnet")}
    try:
        common = _load("common_layers", "models/common_layers.py")
        sys.modules["common_layers"] = common  # `from common_layers import *` in ub_uresnet.py:6
        net = _load("ub_uresnet", "models/ub_uresnet.py")
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            loss = _load("pixelwise_nllloss", "training/pixelwise_nllloss.py")
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    _cache["mods"] = (net, common, loss)
    return _cache["mods"]


def make_reference_model(num_classes=3, input_channels=1, inplanes=16):
    net, _, _ = load_reference()
    return net.UResNet(num_classes=num_classes, input_channels=input_channels,
                       inplanes=inplanes, showsizes=False)


def make_reference_criterion(weight=None, ignore_index=-100):
    _, _, loss = load_reference()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        crit = loss.PixelWiseNLLLoss(weight=weight, ignore_index=ignore_index)
    crit.size_average = True  # shim 3
    return crit


def reference_loss(crit, logp, target, pixelweights):
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return crit(logp, target, pixelweights)


def load_reference_aspp():
    """The unmodified reference `models/ASPP_ResNet.py` module.  It imports ROOT, larcv and the py2-only `commands`
    module at file scope (:22,31-32) without using them in the model: empty stub modules satisfy those imports."""
    if "aspp" in _cache:
        return _cache["aspp"]
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    stubs = {}
    for name in ("ROOT", "larcv", "commands"):
        if name not in sys.modules:
            stubs[name] = types.ModuleType(name)
    stubs.get("larcv", sys.modules.get("larcv")).larcv = types.SimpleNamespace()
    sys.modules.update(stubs)
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            mod = _load("ASPP_ResNet", "models/ASPP_ResNet.py")
    finally:
        for name in stubs:
            sys.modules.pop(name, None)
    _cache["aspp"] = mod
    return mod


def make_reference_aspp_model(num_classes=3, in_channels=1, inplanes=16):
    return load_reference_aspp().ASPP_ResNet(num_classes=num_classes, in_channels=in_channels, inplanes=inplanes,
                                             showsizes=False)


def load_reference_accuracy(script="training/train_ubresnet2018_wlarcv2.py"):
    """The reference `accuracy()` function, cut out of the (un-importable: ROOT, larcv, py2) training script TEXT
    (train_ubresnet2018_wlarcv2.py:509-566 / train_ubresnet2018_wlarcv1.py:528-586) and exec'd with the print shim."""
    key = "acc:" + script
    if key in _cache:
        return _cache[key]
    import time
    import torch
    with open(os.path.join(REFERENCE_ROOT, script), "r") as fh:
        text = fh.read()
    m = re.search(r"^def accuracy\(.*?(?=^def )", text, re.S | re.M)
    body = _PRINT_RE.sub(r"\1print(\2)", m.group(0))
    ns = {"torch": torch, "time": time}
    exec(compile(body, os.path.join(REFERENCE_ROOT, script), "exec"), ns)
    _cache[key] = ns["accuracy"]
    return _cache[key]
