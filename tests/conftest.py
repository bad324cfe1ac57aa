import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session", autouse=True)
def _built_library():
    from ubresnet_b200 import build
    build.build(force=False)
    yield


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_cuda = torch.cuda.is_available()
    except Exception:
        has_cuda = False
    if has_cuda:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def pytest_terminal_summary(terminalreporter):
    """print every measured parity error of the run (GPUTEST records carry the numbers, not only a pass count)"""
    try:
        from tests import model_checks
    except Exception:
        return
    if not model_checks.MEASURED:
        return
    tr = terminalreporter
    tr.write_line("")
    tr.write_line("measured parity errors (check:label err / bound):")
    for name, label, err, tol in model_checks.MEASURED:
        tr.write_line("  %-4s %s:%s  %.3e / %.1e" % ("ok" if err <= tol else "FAIL", name, label, err, tol))
