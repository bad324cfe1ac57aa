"""-m gpu: whole-network parity (forward, loss, backward, BN buffers) through the public nn.Module API."""
import pytest
import torch

from tests import model_checks

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", list(model_checks.CHECKS))
def test_model(name):
    results, _ = model_checks.run_check(name)
    bad = [(lbl, err, tol) for lbl, err, tol in results if not err <= tol]
    assert not bad, "parity failures: %s" % bad


def test_cpu_input_raises():
    from ubresnet_b200.models import UResNet
    m = UResNet(3, 1, 16).cuda()
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 1, 64, 64))


def test_checkpoint_roundtrip_and_deploy_loader(tmp_path):
    """reference checkpoint dict layout (train_ubresnet2018_wlarcv2.py:260-266) incl. DataParallel `module.` prefix"""
    from collections import OrderedDict
    from ubresnet_b200.deploy.ubresnet_funcs import load_cosmic_retrain_model
    from ubresnet_b200.models import UResNet
    torch.manual_seed(3)
    m = UResNet(num_classes=4, input_channels=1, inplanes=16)
    sd = OrderedDict(("module." + k, v) for k, v in m.state_dict().items())
    path = str(tmp_path / "checkpoint.0th.tar")
    torch.save({"iter": 0, "epoch": 0, "state_dict": sd, "best_prec1": 0.0, "optimizer": {}}, path)
    m2 = load_cosmic_retrain_model(path, device="cuda:0").to("cuda:0").eval()
    x = torch.randn(1, 1, 64, 64, device="cuda:0")
    with torch.no_grad():
        a = m.to("cuda:0").eval()(x)
        b = m2(x)
    assert a.shape == (1, 4, 64, 64)
    assert torch.equal(a, b)
    assert torch.allclose(a.exp().sum(1), torch.ones_like(a[:, 0]), atol=1e-4)


def test_wholeview_deploy_bitexact_tiling():
    from oracle import uresnet_port as port
    from ubresnet_b200 import synthetic
    from ubresnet_b200.deploy import wholeview
    ev = synthetic.make_wholeview(rows=256, cols=448, planes=1)
    plan = wholeview.tile_plan(256, 448, 128)
    assert plan == [(r, c) for r in port.tile_starts(256, 128) for c in port.tile_starts(448, 128)]
    crops = wholeview.gather(ev[0].cuda(), plan, 128)
    ref = port.tile_gather_np(ev[0].numpy(), plan, 128)
    assert (crops[:, 0].cpu().numpy() == ref).all()
    tiles = torch.randn(len(plan), 3, 128, 128)
    st = wholeview.stitch(tiles.cuda(), plan, 256, 448)
    assert (st.cpu().numpy() == port.stitch_scatter_np(tiles.numpy(), plan, 256, 448)).all()


def test_wholeview_1008x3456_crop512_bitexact():
    """BASELINE.json configs[3] geometry: 3 planes 1008x3456 -> 42 crops of 512^2 (SURVEY 8c tile plan), gather and
    stitch bit-exact against the numpy restatement"""
    from oracle import uresnet_port as port
    from ubresnet_b200 import synthetic
    from ubresnet_b200.deploy import wholeview
    ev = synthetic.make_wholeview(rows=1008, cols=3456, planes=3)
    plan = wholeview.tile_plan(1008, 3456, 512)
    assert plan == port.tile_plan(1008, 3456, 512) and len(plan) == 14
    assert [r for r, _ in plan[::7]] == [0, 496] and [c for _, c in plan[:7]] == [0, 491, 981, 1472, 1963, 2453, 2944]
    g = torch.Generator().manual_seed(5)
    for p in range(3):
        crops = wholeview.gather(ev[p].cuda(), plan, 512)
        assert (crops[:, 0].cpu().numpy() == port.tile_gather_np(ev[p].numpy(), plan, 512)).all()
        tiles = torch.randn(len(plan), 3, 512, 512, generator=g)
        st = wholeview.stitch(tiles.cuda(), plan, 1008, 3456)
        assert (st.cpu().numpy() == port.stitch_scatter_np(tiles.numpy(), plan, 1008, 3456)).all()


def test_wholeview_run_event_matches_per_crop_pipeline():
    """deploy/wholeview.run_event (all planes' crops in one forward) == gather -> per-chunk forward -> stitch, bit for bit,
    for any crops-per-forward setting"""
    from ubresnet_b200 import synthetic
    from ubresnet_b200.deploy import wholeview
    from ubresnet_b200.models import UResNet
    torch.manual_seed(7)
    model = UResNet(num_classes=3, input_channels=1, inplanes=16).cuda().eval()
    ev = synthetic.make_wholeview(rows=256, cols=448, planes=2)
    plan = wholeview.tile_plan(256, 448, 128)
    with torch.no_grad():
        ref = []
        for p in range(2):
            crops = wholeview.gather(ev[p].cuda(), plan, 128)
            outs = torch.cat([model(crops[i:i + 3]) for i in range(0, crops.shape[0], 3)], 0)
            ref.append(wholeview.stitch(outs, plan, 256, 448))
        ref = torch.stack(ref, 0)
    full = wholeview.run_event(model, ev, crop=128)
    chunked = wholeview.run_event(model, ev, crop=128, batch=5)
    assert full.shape == (2, 3, 256, 448)
    assert torch.equal(full, ref) and torch.equal(chunked, ref)


def test_wholeview_run_events_pipeline_matches_run_event():
    """the overlapped deploy loop yields, in order, exactly argmax(run_event) of each event"""
    from ubresnet_b200 import synthetic
    from ubresnet_b200.deploy import wholeview
    from ubresnet_b200.models import UResNet
    torch.manual_seed(11)
    model = UResNet(num_classes=3, input_channels=1, inplanes=16).cuda().eval()
    events = [synthetic.make_wholeview(seed=100 + i, rows=256, cols=448, planes=2).pin_memory() for i in range(5)]
    refs = [wholeview.run_event(model, e, crop=128).argmax(1).to(torch.uint8).cpu() for e in events]
    got = [m.clone() for m in wholeview.run_events(model, events, crop=128)]
    assert len(got) == len(refs)
    for g, r in zip(got, refs):
        assert g.shape == (2, 256, 448) and torch.equal(g, r)
    assert list(wholeview.run_events(model, [], crop=128)) == []
    one = [m.clone() for m in wholeview.run_events(model, events[:1], crop=128)]
    assert len(one) == 1 and torch.equal(one[0], refs[0])


def test_nn_dataparallel_wrapping():
    """`nn.DataParallel(model)` (train_ubresnet2018_wlarcv2.py:99,103): needs >= 2 GPUs in one process"""
    import os
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "dataparallel_check.py")], capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stdout[-500:] + r.stderr[-500:]


@pytest.mark.parametrize("arch", ["uresnet", "aspp"])
def test_training_step_is_bit_deterministic(arch):
    """tools/determinism_check.py: repeated steps under allocator perturbation give bit-identical losses and gradients
    (fixed-order reductions; catches lifetime bugs of the weight-gradient side stream)"""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "determinism_check.py"), arch, "4"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-500:] + r.stderr[-500:]
