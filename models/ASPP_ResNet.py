"""Drop-in path for reference scripts: `from ASPP_ResNet import ASPP_ResNet` with the models
directory on sys.path (the reference's training/Sem_Seg_ASPP_ResNet1.py:43,188 does this for its ASPP_ResNet1 twin).  Re-exports the B200 implementation."""
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ubresnet_b200.models.ASPP_ResNet import ASPP, ASPP_post, ASPP_ResNet  # noqa: E402,F401
from ubresnet_b200.models.common_layers import BasicBlock, ConvTransposeLayer, DoubleResNet, conv3x3  # noqa: E402,F401
