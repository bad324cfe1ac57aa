"""CPU tests of the boundary: the C-ABI library loads and exports every declared symbol; host-side module logic."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "ubresnet_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ubr_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from ubresnet_b200 import _lib
    lib = ctypes.CDLL(_lib.LIBPATH)
    names = _declared_symbols()
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), "missing export %s" % n
    # and the Python binding table covers the header one to one
    assert sorted(_lib.SIGNATURES) == names


def test_library_refuses_without_gpu():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from ubresnet_b200 import _lib
    lib = _lib.load()
    assert lib.ubr_abi_version() == 1
    assert lib.ubr_device_check() == -2          # UBR_E_UNSUPPORTED, no CPU fallback
    assert b"no CPU fallback" in lib.ubr_last_error()


def test_model_cpu_input_raises_no_fallback():
    from ubresnet_b200.models import UResNet
    from ubresnet_b200.training.pixelwise_nllloss import PixelWiseNLLLoss
    m = UResNet(3, 1, 16)
    with pytest.raises(RuntimeError, match="no CPU path"):
        m(torch.zeros(1, 1, 64, 64))
    with pytest.raises(RuntimeError, match="no CPU path"):
        PixelWiseNLLLoss()(torch.zeros(1, 3, 8, 8), torch.zeros(1, 8, 8, dtype=torch.long), torch.ones(1, 8, 8))


def test_state_dict_layout():
    from oracle import uresnet_port as port
    from ubresnet_b200.models import UResNet
    m = UResNet(num_classes=3, input_channels=1, inplanes=16)
    sd = m.state_dict()
    assert len(sd) == 321
    assert sum(p.numel() for p in m.parameters()) == 18100931
    ref = port.init_state_dict(3, 1, 16)
    assert set(sd.keys()) == set(ref.keys())
    for k in sd:
        assert tuple(sd[k].shape) == tuple(ref[k].shape), k
    assert sd["dec_layer5.deconv.weight"].shape == (512, 256, 4, 4)
    for attr in ("conv1", "bn1", "relu1", "pool1", "enc_layer1", "enc_layer5", "dec_layer5", "dec_layer1", "conv10",
                 "bn10", "relu10", "conv11", "softmax", "inplanes", "nkernels", "_showsizes"):
        assert hasattr(m, attr), attr
    m4 = UResNet(inplanes=32, input_channels=1, num_classes=4)
    assert m4.conv11.weight.shape == (4, 16, 7, 7) and m4.enc_layer5.res2.conv2.weight.shape == (1024, 1024, 3, 3)


def test_dropin_import_paths():
    import importlib.util
    for rel, sym in (("models/ub_uresnet.py", "UResNet"), ("models/common_layers.py", "DoubleResNet"),
                     ("training/pixelwise_nllloss.py", "PixelWiseNLLLoss"),
                     ("deploy/ubresnet_funcs.py", "load_cosmic_retrain_model")):
        spec = importlib.util.spec_from_file_location("dropin_" + sym, os.path.join(ROOT, rel))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        assert hasattr(mod, sym)


def test_synthetic_generator_is_deterministic_and_sane():
    from ubresnet_b200 import synthetic
    a = synthetic.make_batch(2, 128, 128, rank=1, step=3)
    b = synthetic.make_batch(2, 128, 128, rank=1, step=3)
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    img, lab, wgt = a
    assert img.shape == (2, 1, 128, 128) and lab.dtype == torch.long and wgt.dtype == torch.float32
    assert set(lab.unique().tolist()) <= {0, 1, 2}
    assert (img[:, 0][lab > 0] >= synthetic.ADC_THRESH).all()
    assert wgt.min() >= 1.0 and wgt.max() <= synthetic.WEIGHT_MAX


def test_stage_map_covers_every_parameter():
    from ubresnet_b200 import parallel
    from ubresnet_b200.models import UResNet
    m = UResNet(3, 1, 16)
    stages = {parallel.stage_of_param(n) for n, _ in m.named_parameters()}
    assert stages == set(parallel.STAGE_ORDER)


def test_fused_optimizers_refuse_cpu_parameters():
    """training/optim.py has no CPU path: a CPU parameter with a gradient raises instead of falling back"""
    import torch
    from ubresnet_b200.training import optim
    for cls in (optim.SGD, optim.Adam):
        p = torch.nn.Parameter(torch.zeros(4))
        p.grad = torch.ones(4)
        opt = cls([p], lr=0.1)
        with pytest.raises(RuntimeError):
            opt.step()
        assert isinstance(opt, torch.optim.Optimizer) and opt.state_dict()["param_groups"][0]["lr"] == 0.1


def test_default_ctor_three_input_channels_selects_fp32():
    """the reference's default ctor (input_channels=3, models/ub_uresnet.py:31) constructs and gets the fp32 schedule;
    an explicit bf16 request for it is refused at construction, not at the first forward"""
    import warnings
    import pytest
    from ubresnet_b200.models import UResNet
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        m = UResNet()
    assert m.conv1.in_channels == 3 and m.precision == "fp32" and any("fp32" in str(x.message) for x in w)
    assert len(m.state_dict()) == 321
    with pytest.raises(ValueError):
        UResNet(3, 3, 16, precision="bf16")
    assert UResNet(3, 1, 16).precision == "bf16"


def test_bench_per_launch_roofline_split():
    """bench.conv_bound_split: a launch is tensor bound when FLOPs / peak >= bytes / peak (arithmetic intensity above the
    ridge); fractions are achieved / peak per class, FLOP-weighted for the tensor class"""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    peaks = dict(hbm_gbs=6500.0, tf_sustained=1300.0, tf_burst=1600.0, source="test")
    x16 = (32, 512, 512, 16)
    # 3x3 16->16 at 512^2: 72 FLOP/B (HBM bound); 7x7 16->16: 392 FLOP/B (tensor bound); ridge = 200
    calls = [("ubr_conv2d_fwd_bn", (x16, 0, 0, 16, 3, 1, x16), 0.10), ("ubr_conv2d_fwd", (x16, 0, 0, 16, 7, 1, x16), 0.20)]
    out = bench.conv_bound_split({"conv_halo_kernel": calls, "bn_act_kernel": [("ubr_bn_act", (x16,), 1.0)]}, 1, peaks)
    assert abs(out["ridge_flop_per_byte"] - 200.0) < 1e-9
    assert out["hbm_bound"]["launches_per_step"] == 1 and out["tensor_bound"]["launches_per_step"] == 1
    npx = 32 * 512 * 512
    gbs = 2.0 * npx * 32 / 0.10e-3 / 1e9
    tfl = 2.0 * npx * 16 * 16 * 49 / 0.20e-3 / 1e12
    assert abs(out["hbm_bound"]["achieved_gbs"] - gbs) < 1e-6 * gbs
    assert abs(out["tensor_bound"]["achieved_tflops"] - tfl) < 1e-6 * tfl
    assert abs(out["tensor_bound"]["frac_flop_weighted"] - tfl / 1300.0) < 1e-9       # one launch: weighted == plain
    roof_ms = 2.0 * npx * 32 / 6500e9 * 1e3 + 2.0 * npx * 16 * 16 * 49 / 1300e12 * 1e3
    assert abs(out["mixed_frac"] - roof_ms / 0.30) < 1e-9
