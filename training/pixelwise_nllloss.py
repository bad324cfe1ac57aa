"""Drop-in path for `from pixelwise_nllloss import PixelWiseNLLLoss` (reference training scripts)."""
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ubresnet_b200.training.pixelwise_nllloss import PixelWiseNLLLoss  # noqa: E402,F401
