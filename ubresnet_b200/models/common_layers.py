"""Building blocks of the U-ResNet -- drop-in for the reference `models/common_layers.py`.

Same public names, constructor signatures and sub-module / parameter names as the reference
(conv3x3 :13-15, BasicBlock :18-58, DoubleResNet :109-120, ConvTransposeLayer :122-132), so
`state_dict()` keys and shapes are identical and reference checkpoints load unchanged.  The modules are
parameter containers: the arithmetic runs in the sm_100a kernels scheduled by `ubresnet_b200.engine`
(whole-network forward/backward); calling a block on its own goes through the same kernels.
"""
import torch
import torch.nn as nn


def conv3x3(in_planes, out_planes, stride=1):
    """3x3 convolution, padding 1, no bias."""
    return nn.Conv2d(in_planes, out_planes, kernel_size=3, stride=stride, padding=1, bias=False)


def _standalone(module, *inputs):
    from .. import engine
    return engine.run_submodule(module, *inputs)


class BasicBlock(nn.Module):
    """conv-bn-relu-conv-bn-relu, then (+ bnpass(bypass(x)) | + x), then relu  (ReLU-add-ReLU)."""
    expansion = 1

    def __init__(self, inplanes, planes, stride=1):
        super().__init__()
        self.stride = stride
        self.conv1 = conv3x3(inplanes, planes, stride)
        self.bn1 = nn.BatchNorm2d(planes)
        self.relu1 = nn.ReLU(inplace=False)
        self.conv2 = conv3x3(planes, planes)
        self.bn2 = nn.BatchNorm2d(planes)
        self.relu2 = nn.ReLU(inplace=False)
        needs_projection = (inplanes != planes) or stride > 1
        self.bypass = nn.Conv2d(inplanes, planes, kernel_size=1, stride=stride, padding=0, bias=False) \
            if needs_projection else None
        self.bnpass = nn.BatchNorm2d(planes) if needs_projection else None
        self.relu = nn.ReLU(inplace=False)

    def forward(self, x):
        return _standalone(self, x)


class DoubleResNet(nn.Module):
    def __init__(self, inplanes, planes, stride=1):
        super().__init__()
        self.res1 = BasicBlock(inplanes, planes, stride)
        self.res2 = BasicBlock(planes, planes, 1)

    def forward(self, x):
        return _standalone(self, x)


class ConvTransposeLayer(nn.Module):
    """deconv(k4,s2,p1) -> cat([deconv_out, skip], 1) -> DoubleResNet."""

    def __init__(self, deconv_inplanes, deconv_outplanes, res_outplanes):
        super().__init__()
        self.deconv = nn.ConvTranspose2d(deconv_inplanes, deconv_outplanes, kernel_size=4, stride=2, padding=1,
                                         bias=False)
        self.res = DoubleResNet(res_outplanes + deconv_outplanes, res_outplanes, stride=1)

    def forward(self, x, skip_x):
        return _standalone(self, x, skip_x)
