"""Drop-in path for `from ubresnet_funcs import load_cosmic_retrain_model` (reference deploy scripts)."""
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ubresnet_b200.deploy.ubresnet_funcs import load_cosmic_retrain_model  # noqa: E402,F401
