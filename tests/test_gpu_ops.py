"""-m gpu: per-op parity of every C-ABI kernel against torch fp32 CPU references (see gpu_checks.py)."""
import pytest

from tests import gpu_checks

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", list(gpu_checks.CHECKS))
def test_op(name):
    results = gpu_checks.CHECKS[name]()
    bad = [(lbl, err, tol) for lbl, err, tol in results if not err <= tol]
    assert not bad, "parity failures: %s" % bad
