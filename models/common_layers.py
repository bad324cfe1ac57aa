"""Drop-in path for `from common_layers import *` (reference models/ub_uresnet.py:6)."""
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ubresnet_b200.models.common_layers import BasicBlock, ConvTransposeLayer, DoubleResNet, conv3x3  # noqa: E402,F401
