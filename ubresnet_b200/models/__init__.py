from .ub_uresnet import UResNet  # noqa: F401
from .common_layers import BasicBlock, ConvTransposeLayer, DoubleResNet, conv3x3  # noqa: F401
