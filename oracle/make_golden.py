"""TEST INFRASTRUCTURE -- generates tests/golden/*.pt FROM THE UNMODIFIED REFERENCE (via reference_shim).

Run inside the build container only (needs /root/reference):  python -m oracle.make_golden
The fixtures travel to the GPU box, the reference does not.
"""
import os
import sys
from collections import OrderedDict

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import reference_shim as shim  # noqa: E402
from ubresnet_b200 import synthetic  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
GRAD_KEYS = ["conv1.weight", "conv1.bias", "bn1.weight", "bn1.bias", "enc_layer1.res1.conv1.weight",
             "enc_layer1.res1.bypass.weight", "enc_layer2.res1.bnpass.weight", "enc_layer3.res2.bn2.bias",
             "enc_layer5.res2.bn1.weight", "dec_layer1.deconv.weight", "dec_layer2.res.res1.bn1.weight",
             "dec_layer1.res.res2.conv2.weight", "conv10.weight", "conv10.bias", "bn10.weight", "conv11.weight",
             "conv11.bias"]


def uresnet_fixture(seed=20180721, batch=2, size=128):
    torch.manual_seed(seed)
    model = shim.make_reference_model(3, 1, 16)
    chk = sum(float(v.double().abs().sum()) for v in model.state_dict().values() if v.is_floating_point())
    img, lab, wgt = synthetic.make_batch(batch, size, size)
    model.train()
    crit = shim.make_reference_criterion()
    logp = model(img)
    loss = shim.reference_loss(crit, logp, lab, wgt)
    loss.backward()
    params = dict(model.named_parameters())
    grads = OrderedDict((k, params[k].grad.clone()) for k in GRAD_KEYS)
    gnorm = OrderedDict((k, float(p.grad.abs().max())) for k, p in params.items())
    sd = model.state_dict()
    running = OrderedDict((k, sd[k].clone()) for k in ("bn1.running_mean", "bn1.running_var",
                                                        "enc_layer5.res2.bn2.running_var", "bn10.running_mean"))
    model.eval()
    with torch.no_grad():
        logp_eval = model(img)
    return dict(seed=seed, image=img, label=lab, weight=wgt, logp=logp.detach(), loss=float(loss), grads=grads,
                grad_absmax=gnorm, weight_abs_sum=chk, running=running, logp_eval=logp_eval,
                note="generated from /root/reference (commit 3491bf2) through oracle/reference_shim.py")


ASPP_GRAD_KEYS = ["conv1.weight", "ASPP_layer_enc3.B1_conv.weight", "ASPP_layer_enc3.B3_conv.weight",
                  "ASPP_layer_enc3.B3_conv.bias", "ASPP_layer_enc4.B2_conv.weight", "ASPP_layer_enc4.B4_bn.weight",
                  "ASPP_layer_enc5.B4_conv.weight", "ASPP_layer_enc5.B1_bn.bias", "ASPP_combine_enc3.ASPP_conv.weight",
                  "ASPP_combine_enc4.ASPP_bn.weight", "ASPP_combine_enc5.ASPP_conv.bias", "enc_layer3.res2.bn2.weight",
                  "enc_layer5.res2.bn2.bias", "dec_layer4.res.res1.bn1.weight", "dec_layer5.res.res1.bnpass.weight",
                  "dec_layer2.deconv.weight", "conv10.weight", "conv11.weight", "conv11.bias"]


def aspp_fixture(seed=20180722, batch=2, size=128):
    """ASPP_ResNet(3,1,16) (reference models/ASPP_ResNet.py) train step + eval forward on seeded synthetic crops."""
    torch.manual_seed(seed)
    model = shim.make_reference_aspp_model(3, 1, 16)
    chk = sum(float(v.double().abs().sum()) for v in model.state_dict().values() if v.is_floating_point())
    img, lab, wgt = synthetic.make_batch(batch, size, size)
    model.train()
    crit = shim.make_reference_criterion()
    logp = model(img)
    loss = shim.reference_loss(crit, logp, lab, wgt)
    loss.backward()
    params = dict(model.named_parameters())
    grads = OrderedDict((k, params[k].grad.clone()) for k in ASPP_GRAD_KEYS)
    gnorm = OrderedDict((k, float(p.grad.abs().max())) for k, p in params.items())
    sd = model.state_dict()
    running = OrderedDict((k, sd[k].clone()) for k in ("bn1.running_mean", "ASPP_layer_enc3.B3_bn.running_var",
                                                        "ASPP_combine_enc5.ASPP_bn.running_mean", "bn10.running_var"))
    model.eval()
    with torch.no_grad():
        logp_eval = model(img)
    return dict(seed=seed, image=img, label=lab, weight=wgt, logp=logp.detach(), loss=float(loss), grads=grads,
                grad_absmax=gnorm, weight_abs_sum=chk, running=running, logp_eval=logp_eval, n_state=len(sd),
                note="generated from /root/reference models/ASPP_ResNet.py (commit 3491bf2) through "
                     "oracle/reference_shim.load_reference_aspp")


TRAINED_SCHEDULE = [(300, 1e-2), (150, 2e-3), (100, 4e-4)]   # (steps, lr): SGD m=0.9 wd=1e-3 (train_ubresnet2018_wlarcv1.py:127-129)


def trained_fixture(seed=20180723, batch=4, size=128, schedule=TRAINED_SCHEDULE, log=True):
    """"Trained-like" weights (SURVEY 8c "Weights"): the SHIMMED REFERENCE model trained with torch.optim.SGD on the
    synthetic generator until the class logits separate (non-zero-pixel accuracy ~98 %).  The weights are rounded to
    fp16 for storage (36 MB instead of 72) BEFORE the reference outputs below are computed, so the fp32 upcast of the
    stored tensors is exactly the state the goldens belong to.  Layout = the reference checkpoint dict
    (train_ubresnet2018_wlarcv2.py:260-266)."""
    import time
    torch.manual_seed(seed)
    model = shim.make_reference_model(3, 1, 16)
    crit = shim.make_reference_criterion()
    opt = torch.optim.SGD(model.parameters(), lr=schedule[0][1], momentum=0.9, weight_decay=1e-3)
    model.train()
    it, t0 = 0, time.time()
    for nsteps, lr in schedule:
        for g in opt.param_groups:
            g["lr"] = lr
        for _ in range(nsteps):
            img, lab, wgt = synthetic.make_batch(batch, size, size, step=it)
            opt.zero_grad()
            logp = model(img)
            loss = shim.reference_loss(crit, logp, lab, wgt)
            loss.backward()
            opt.step()
            if log and it % 50 == 0:
                nz = img[:, 0] != 0
                acc = (logp.argmax(1)[nz] == lab[nz]).float().mean().item()
                print("  trained fixture: iter %d lr %g loss %.4f acc(nonzero) %.3f  %.0fs" %
                      (it, lr, float(loss), acc, time.time() - t0), flush=True)
            it += 1
    # fp16 storage round trip, then reload into the reference model
    sd16 = OrderedDict((k, (v.half() if v.is_floating_point() else v.clone())) for k, v in model.state_dict().items())
    model.load_state_dict(OrderedDict((k, (v.float() if v.is_floating_point() else v)) for k, v in sd16.items()))
    # reference outputs on the stored state: eval forward (log-probs + raw conv11 logits via a forward hook,
    # ub_uresnet.py:141) and one train-mode step (loss, log-probs, a few gradients)
    img, lab, wgt = synthetic.make_batch(2, size, size, step=100000)
    grabbed = {}
    hook = model.conv11.register_forward_hook(lambda m, i, o: grabbed.__setitem__("logits", o.detach().clone()))
    model.eval()
    with torch.no_grad():
        logp_eval = model(img)
    logits_eval = grabbed["logits"]
    model.train()
    model.zero_grad()
    logp = model(img)
    logits_train = grabbed["logits"]
    hook.remove()
    loss = shim.reference_loss(crit, logp, lab, wgt)
    loss.backward()
    params = dict(model.named_parameters())
    grads = OrderedDict((k, params[k].grad.clone()) for k in GRAD_KEYS)
    ckpt = {"iter": it, "epoch": 0.0, "state_dict": sd16, "best_prec1": 0.0, "optimizer": {}}
    return dict(checkpoint=ckpt, seed=seed, image=img, label=lab, weight=wgt, logp_eval=logp_eval,
                logits_eval=logits_eval, logp_train=logp.detach(), logits_train=logits_train, loss_train=float(loss),
                loss_eval=float(shim.reference_loss(crit, logp_eval, lab, wgt)), grads=grads, train_iters=it,
                note="weights: /root/reference UResNet(3,1,16) trained %d SGD steps on ubresnet_b200.synthetic "
                     "crops through oracle/reference_shim.py, stored fp16; outputs: the reference on the stored state"
                     % it)


def loss_fixture(seed=7):
    g = torch.Generator().manual_seed(seed)
    logits = torch.randn(2, 3, 24, 40, generator=g) * 2
    logp = torch.log_softmax(logits, 1).requires_grad_(True)
    tgt = torch.randint(0, 3, (2, 24, 40), generator=g)
    tgt[1, 3, :7] = -100
    pw = torch.rand(2, 24, 40, generator=g) * 50 + 1
    out = {}
    for tag, cw in (("plain", None), ("classw", torch.tensor([0.3, 1.0, 2.5]))):
        crit = shim.make_reference_criterion(weight=cw)
        if logp.grad is not None:
            logp.grad = None
        loss = shim.reference_loss(crit, logp, tgt, pw)
        loss.backward()
        out[tag] = dict(loss=float(loss), dlogp=logp.grad.clone(), classw=cw)
    return dict(logp=logp.detach(), target=tgt, pixelweights=pw, cases=out)


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.save(uresnet_fixture(), os.path.join(GOLD, "uresnet_b2_128.pt"))
    torch.save(loss_fixture(), os.path.join(GOLD, "pixelwise_nll.pt"))
    torch.save(aspp_fixture(), os.path.join(GOLD, "aspp_b2_128.pt"))
    if "--trained" in sys.argv or not os.path.isfile(os.path.join(GOLD, "uresnet_trained.pt")):
        torch.save(trained_fixture(), os.path.join(GOLD, "uresnet_trained.pt"))
    for f in sorted(os.listdir(GOLD)):
        print(f, os.path.getsize(os.path.join(GOLD, f)))


if __name__ == "__main__":
    main()
