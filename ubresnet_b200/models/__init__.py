from .ub_uresnet import UResNet  # noqa: F401
from .common_layers import BasicBlock, ConvTransposeLayer, DoubleResNet, conv3x3  # noqa: F401
from .ASPP_ResNet import ASPP, ASPP_post, ASPP_ResNet  # noqa: F401
