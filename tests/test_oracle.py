"""CPU tests of the oracle: (1) pinned against the UNMODIFIED reference when /root/reference is mounted (build
container), (2) pinned against the committed golden fixtures that were generated from the reference (anywhere)."""
import os
from collections import OrderedDict

import pytest
import torch

from oracle import reference_shim as shim
from oracle import uresnet_port as port
from ubresnet_b200 import synthetic

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
needs_ref = pytest.mark.skipif(not shim.available(), reason="/root/reference not mounted")


def _rel(a, b):
    return (a - b).abs().max().item() / max(b.abs().max().item(), 1e-30)


@needs_ref
def test_port_matches_reference_forward_backward():
    torch.manual_seed(11)
    ref = shim.make_reference_model(3, 1, 16)
    sd = OrderedDict((k, v.detach().clone()) for k, v in ref.state_dict().items())
    img, lab, wgt = synthetic.make_batch(2, 64, 64)
    ref.train()
    crit = shim.make_reference_criterion()
    logp_ref = ref(img)
    loss_ref = shim.reference_loss(crit, logp_ref, lab, wgt)
    loss_ref.backward()
    loss, logp, grads, after = port.train_step_grads(sd, img, lab, wgt)
    assert _rel(logp, logp_ref.detach()) < 1e-5
    assert abs(loss.item() - loss_ref.item()) / abs(loss_ref.item()) < 1e-6
    for name, p in ref.named_parameters():
        if name in ("conv1.bias", "conv10.bias"):
            continue            # exact zero in theory; both sides are round-off noise
        assert _rel(grads[name], p.grad) < 2e-4, name
    rsd = ref.state_dict()
    for k, v in after.items():
        if "running_" in k:
            assert _rel(v.detach(), rsd[k]) < 1e-5, k
        if k.endswith("num_batches_tracked"):
            assert int(v) == int(rsd[k])


@needs_ref
def test_port_matches_reference_eval_4class():
    torch.manual_seed(12)
    ref = shim.make_reference_model(4, 1, 16).eval()
    sd = OrderedDict((k, v.detach().clone()) for k, v in ref.state_dict().items())
    img, _, _ = synthetic.make_batch(1, 64, 96)
    with torch.no_grad():
        a = ref(img)
        b = port.forward(sd, img, training=False)
    assert a.shape == (1, 4, 64, 96)
    assert _rel(b, a) < 1e-5


@needs_ref
def test_module_init_is_bit_identical_to_reference():
    from ubresnet_b200.models import UResNet
    torch.manual_seed(5)
    mine = UResNet(3, 1, 16).state_dict()
    torch.manual_seed(5)
    ref = shim.make_reference_model(3, 1, 16).state_dict()
    assert list(mine.keys()) == list(ref.keys())
    assert len(mine) == 321
    for k in mine:
        assert torch.equal(mine[k], ref[k]), k


@needs_ref
def test_loss_port_matches_reference_with_class_weights():
    g = torch.Generator().manual_seed(3)
    logp = torch.log_softmax(torch.randn(2, 3, 16, 16, generator=g), 1)
    tgt = torch.randint(0, 3, (2, 16, 16), generator=g)
    tgt[0, 0, :3] = -100
    pw = torch.rand(2, 16, 16, generator=g) * 10
    cw = torch.tensor([0.5, 1.0, 3.0])
    crit = shim.make_reference_criterion(weight=cw)
    a = shim.reference_loss(crit, logp, tgt, pw)
    b = port.pixelwise_nll(logp, tgt, pw, cw)
    assert abs(a.item() - b.item()) < 1e-6 * abs(a.item())


def test_port_matches_golden_uresnet():
    gold = torch.load(os.path.join(GOLD, "uresnet_b2_128.pt"), weights_only=False)
    from ubresnet_b200.models import UResNet
    torch.manual_seed(gold["seed"])
    sd = OrderedDict((k, v.detach().clone()) for k, v in UResNet(3, 1, 16).state_dict().items())
    chk = sum(float(v.double().abs().sum()) for v in sd.values() if v.is_floating_point())
    assert abs(chk - gold["weight_abs_sum"]) <= 1e-9 * gold["weight_abs_sum"]
    loss, logp, grads, after = port.train_step_grads(sd, gold["image"], gold["label"], gold["weight"])
    assert _rel(logp, gold["logp"]) < 1e-4
    assert abs(loss.item() - gold["loss"]) / abs(gold["loss"]) < 1e-5
    for name, g in gold["grads"].items():
        if name in ("conv1.bias", "conv10.bias"):
            continue
        assert _rel(grads[name], g) < 1e-3, name
    for k, v in gold["running"].items():
        assert _rel(after[k].detach(), v) < 1e-4, k
    with torch.no_grad():
        le = port.forward(after, gold["image"], training=False)
    assert _rel(le, gold["logp_eval"]) < 1e-3


def test_loss_port_matches_golden():
    gold = torch.load(os.path.join(GOLD, "pixelwise_nll.pt"), weights_only=False)
    for tag, case in gold["cases"].items():
        lp = gold["logp"].clone().requires_grad_(True)
        loss = port.pixelwise_nll(lp, gold["target"], gold["pixelweights"], case["classw"])
        loss.backward()
        assert abs(loss.item() - case["loss"]) < 1e-6 * abs(case["loss"]), tag
        assert _rel(lp.grad, case["dlogp"]) < 1e-6, tag


def test_tile_plan_matches_survey():
    assert port.tile_starts(1008, 512) == [0, 496]
    assert port.tile_starts(3456, 512) == [0, 491, 981, 1472, 1963, 2453, 2944]
    assert len(port.tile_plan()) == 14
    from ubresnet_b200.deploy import wholeview
    assert wholeview.tile_plan() == port.tile_plan()


def test_stitch_is_highest_index_wins():
    import numpy as np
    plan = [(0, 0), (0, 2), (2, 0), (2, 2)]
    tiles = np.stack([np.full((1, 4, 4), i + 1.0, dtype=np.float32) for i in range(4)])
    out = port.stitch_scatter_np(tiles, plan, 6, 6)[0]
    assert out[0, 0] == 1 and out[0, 3] == 2 and out[3, 0] == 3 and out[3, 3] == 4 and out[2, 2] == 4


# ------------------------------------------------------------------------------------------------
# ASPP_ResNet (BASELINE.json configs[4]; reference models/ASPP_ResNet.py)
# ------------------------------------------------------------------------------------------------
@needs_ref
def test_aspp_port_matches_reference_forward_backward():
    from oracle import aspp_port
    torch.manual_seed(13)
    ref = shim.make_reference_aspp_model(3, 1, 16)
    sd = OrderedDict((k, v.detach().clone()) for k, v in ref.state_dict().items())
    img, lab, wgt = synthetic.make_batch(2, 64, 64)
    ref.train()
    crit = shim.make_reference_criterion()
    logp_ref = ref(img)
    loss_ref = shim.reference_loss(crit, logp_ref, lab, wgt)
    loss_ref.backward()
    loss, logp, grads, after = aspp_port.train_step_grads(sd, img, lab, wgt)
    assert _rel(logp, logp_ref.detach()) < 1e-5
    assert abs(loss.item() - loss_ref.item()) / abs(loss_ref.item()) < 1e-6
    for name, p in ref.named_parameters():
        if name.endswith("_conv.bias") or name in ("conv1.bias", "conv10.bias"):
            continue            # bias in front of a train-mode BN: exact zero in theory, round-off on both sides
        assert _rel(grads[name], p.grad) < 2e-4, name
    rsd = ref.state_dict()
    for k, v in after.items():
        if "running_" in k:
            assert _rel(v.detach(), rsd[k]) < 1e-5, k
        if k.endswith("num_batches_tracked"):
            assert int(v) == int(rsd[k])


@needs_ref
def test_aspp_module_init_is_bit_identical_to_reference():
    from ubresnet_b200.models import ASPP_ResNet
    torch.manual_seed(6)
    mine = ASPP_ResNet(3, 1, 16, showsizes=False).state_dict()
    torch.manual_seed(6)
    ref = shim.make_reference_aspp_model(3, 1, 16).state_dict()
    assert list(mine.keys()) == list(ref.keys())
    assert len(mine) == 426
    for k in mine:
        assert torch.equal(mine[k], ref[k]), k


def test_aspp_port_matches_golden():
    from oracle import aspp_port
    from ubresnet_b200.models import ASPP_ResNet
    gold = torch.load(os.path.join(GOLD, "aspp_b2_128.pt"), weights_only=False)
    torch.manual_seed(gold["seed"])
    model = ASPP_ResNet(3, 1, 16, showsizes=False)
    sd = OrderedDict((k, v.detach().clone()) for k, v in model.state_dict().items())
    assert len(sd) == gold["n_state"] == 426
    assert sum(p.numel() for p in model.parameters()) == 31575683
    chk = sum(float(v.double().abs().sum()) for v in sd.values() if v.is_floating_point())
    assert abs(chk - gold["weight_abs_sum"]) <= 1e-9 * gold["weight_abs_sum"]
    loss, logp, grads, after = aspp_port.train_step_grads(sd, gold["image"], gold["label"], gold["weight"])
    assert _rel(logp, gold["logp"]) < 1e-4
    assert abs(loss.item() - gold["loss"]) / abs(gold["loss"]) < 1e-5
    for name, g in gold["grads"].items():
        if name.endswith("_conv.bias"):
            continue
        assert _rel(grads[name], g) < 1e-3, name
    for k, v in gold["running"].items():
        assert _rel(after[k].detach(), v) < 1e-4, k
    with torch.no_grad():
        le = aspp_port.forward(after, gold["image"], training=False)
    assert _rel(le, gold["logp_eval"]) < 1e-3


@needs_ref
@pytest.mark.parametrize("script,nonzero", [("training/train_ubresnet2018_wlarcv2.py", False),
                                            ("training/train_ubresnet2018_wlarcv1.py", True)])
def test_accuracy_port_matches_reference(script, nonzero):
    ref_acc = shim.load_reference_accuracy(script)
    g = torch.Generator().manual_seed(21)
    out = torch.log_softmax(torch.randn(2, 3, 24, 40, generator=g), 1)
    tgt = torch.randint(0, 3, (2, 24, 40), generator=g)
    a = ref_acc(out, tgt.clone(), None) if nonzero else ref_acc(out, tgt.clone())
    b = port.accuracy(out, tgt, nonzero=nonzero)
    assert len(a) == len(b) == (5 if nonzero else 4)
    for x, y in zip(a, b):
        assert abs(float(x) - y) < 1e-4, (a, b)
    # absent class -> 0.0 (reference :553-556)
    tgt2 = tgt.clamp(max=1)
    a2 = ref_acc(out, tgt2.clone(), None) if nonzero else ref_acc(out, tgt2.clone())
    assert float(a2[2]) == port.accuracy(out, tgt2, nonzero=nonzero)[2] == 0.0


def test_port_matches_trained_fixture():
    """the oracle port on the reference-TRAINED weights reproduces the outputs the reference itself produced on them
    (tests/golden/uresnet_trained.pt, oracle/make_golden.py::trained_fixture): eval + train log-probs, raw conv11
    logits, losses and gradients"""
    fx = torch.load(os.path.join(GOLD, "uresnet_trained.pt"), weights_only=False)
    sd = OrderedDict((k, (v.float() if v.is_floating_point() else v.clone()))
                     for k, v in fx["checkpoint"]["state_dict"].items())
    assert set(fx["checkpoint"]) == {"iter", "epoch", "state_dict", "best_prec1", "optimizer"}
    img, lab, wgt = fx["image"], fx["label"], fx["weight"]
    with torch.no_grad():
        logp, logits = port.forward(port.clone_state_dict(sd), img, training=False, return_logits=True)
    assert _rel(logp, fx["logp_eval"]) < 1e-5 and _rel(logits, fx["logits_eval"]) < 1e-5
    assert abs(port.pixelwise_nll(logp, lab, wgt).item() - fx["loss_eval"]) / fx["loss_eval"] < 1e-5
    nz = img[:, 0] != 0
    acc = (logp.argmax(1)[nz] == lab[nz]).float().mean().item()
    assert acc > 0.95, "fixture is not trained-like: accuracy on non-zero pixels %.3f" % acc
    loss, logp_t, grads, _ = port.train_step_grads(sd, img, lab, wgt)
    assert _rel(logp_t, fx["logp_train"]) < 1e-5
    assert abs(loss.item() - fx["loss_train"]) / fx["loss_train"] < 1e-5
    for k, g in fx["grads"].items():
        if k in ("conv1.bias", "conv10.bias"):
            continue
        assert _rel(grads[k], g) < 5e-4, k
