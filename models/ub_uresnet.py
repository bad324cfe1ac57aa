"""Drop-in path for reference scripts: `sys.path.append($UBRESNET_MODELDIR); from ub_uresnet import UResNet`
(reference deploy/ubresnet_funcs.py:29-36).  Re-exports the B200 implementation."""
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ubresnet_b200.models.ub_uresnet import UResNet  # noqa: E402,F401
