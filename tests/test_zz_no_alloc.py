"""Allocation-free entry point `ubr_channel_sum_static` (include/ubresnet_b200.h): no counterpart in the reference.
Called through the C ABI and compared with `ubr_channel_sum` (caller workspace)
on the same input, twice in a row with different inputs (the built-in scratch keeps the previous call's contents).
Both entry points run the same two kernels with the same grids, so the results are bit-identical; a float64 torch sum
bounds both.  Collected last (file name) so that a failure here hides no other test under `pytest -x`."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _both(lib, _lib, ops, a, nvalid, accumulate, seed_out):
    c = a.buf.shape[3]
    ws = torch.empty(lib.ubr_bn_workspace_bytes(c) // 4, dtype=torch.float32, device=DEV)
    out_ws = seed_out.clone()
    out_st = seed_out.clone()
    _lib.check(lib.ubr_channel_sum(a.view(), nvalid, ws.data_ptr(), out_ws.data_ptr(), accumulate, ops._stream()),
               "channel_sum")
    before = torch.cuda.memory_allocated(DEV)
    _lib.check(lib.ubr_channel_sum_static(a.view(), nvalid, out_st.data_ptr(), accumulate, ops._stream()),
               "channel_sum_static")
    assert torch.cuda.memory_allocated(DEV) == before
    torch.cuda.synchronize()
    return out_ws, out_st


@pytest.mark.parametrize("c,n,h,w", [(16, 2, 64, 96), (512, 1, 16, 16), (64, 3, 33, 47)])
def test_static_scratch_matches_workspace_entry_point(c, n, h, w):
    from ubresnet_b200 import _lib, ops
    from ubresnet_b200.ops import Act
    lib = _lib.load()
    g = torch.Generator(device=DEV).manual_seed(1234 + c)
    for call in range(2):                      # two calls in a row: the second sees the first call's scratch contents
        x = (torch.randn(n, h, w, c, device=DEV, generator=g) * (1.0 + call) + 0.25 * call).to(torch.bfloat16)
        a = Act(x)
        nvalid = c if call == 0 else max(1, c - 5)
        accumulate = call                      # second call adds onto a non-zero output
        seed = torch.full((c,), 3.0, dtype=torch.float32, device=DEV)
        out_ws, out_st = _both(lib, _lib, ops, a, nvalid, accumulate, seed)
        assert torch.equal(out_ws, out_st), (call, (out_ws - out_st).abs().max().item())
        ref = x.double().sum(dim=(0, 1, 2)).float()[:nvalid] + (3.0 if accumulate else 0.0)
        tol = 1e-4 * x.float().abs().sum(dim=(0, 1, 2))[:nvalid] + 1e-4
        assert bool(((out_st[:nvalid] - ref).abs() <= tol).all())
        assert bool((out_st[nvalid:] == 3.0).all())          # entries past nvalid are not written


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    from ubresnet_b200 import _lib, ops
    from ubresnet_b200.ops import Act
    lib = _lib.load()
    a = Act(torch.zeros(1, 4, 4, 1024, dtype=torch.bfloat16, device=DEV))
    out = torch.zeros(1024, dtype=torch.float32, device=DEV)
    rc = lib.ubr_channel_sum_static(a.view(), 1024, out.data_ptr(), 0, ops._stream())
    assert rc == -1
    assert "max 512" in lib.ubr_last_error().decode()
