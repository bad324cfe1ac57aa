"""U-ResNet with three Atrous-Spatial-Pyramid-Pooling skip layers -- drop-in for the reference
`models/ASPP_ResNet.py` (ASPP :188-262, ASPP_post :265-286, ASPP_ResNet :289-523).

Constructor signature `ASPP_ResNet(num_classes=3, in_channels=3, inplanes=16, showsizes=True)`, the attribute tree
(`conv1, bn1, relu1, pool1, enc_layer1..5, ASPP_layer_enc{3,4,5}, ASPP_combine_enc{3,4,5}, dec_layer5..1, conv10,
bn10, relu10, conv11, softmax`), parameter names / shapes (426 state_dict entries at inplanes=16) and the
initialisation (conv/deconv W ~ N(0, sqrt(2/(kh*kw*out_channels))), BN gamma=1 beta=0, :394-402) are the
reference's, so its checkpoints load unchanged.  `forward` returns log-probabilities (B, num_classes, H, W) fp32;
the arithmetic runs in the sm_100a kernel schedule of `ubresnet_b200.engine_aspp` (bf16 NHWC activations, fp32
accumulation and statistics).  There is no CPU path: a CPU input raises.

The reference file imports ROOT / larcv / `commands` at module scope (:22,31-32) without using them in the model;
they are not needed here.
"""
import math

import torch
import torch.nn as nn

from .common_layers import BasicBlock, ConvTransposeLayer, DoubleResNet, conv3x3  # noqa: F401  (re-exported like the reference)


class ASPP(nn.Module):
    """Four conv branches to `outplanes` channels (1x1; 3x3 d1; 3x3 d3; 3x3 d5 -- all biased, BN, ReLU) plus a
    MaxPool2d(3,1,1) of the input; concatenated along channels (reference :188-262)."""

    def __init__(self, inplanes, outplanes=16, nkernels=16, showsizes=False):
        super().__init__()
        stride = 1
        self.inplanes = inplanes
        self.outplanes = outplanes
        self.nkernels = nkernels
        self.showsizes = showsizes
        self.B1_conv = nn.Conv2d(inplanes, outplanes, kernel_size=1, stride=stride, padding=0, dilation=1, bias=True)
        self.B1_bn = nn.BatchNorm2d(nkernels)
        self.B1_relu = nn.ReLU(inplace=False)
        self.B2_conv = nn.Conv2d(inplanes, outplanes, kernel_size=3, stride=stride, padding=1, dilation=1, bias=True)
        self.B2_bn = nn.BatchNorm2d(nkernels)
        self.B2_relu = nn.ReLU(inplace=False)
        self.B3_conv = nn.Conv2d(inplanes, outplanes, kernel_size=3, stride=stride, padding=3, dilation=3, bias=True)
        self.B3_bn = nn.BatchNorm2d(nkernels)
        self.B3_relu = nn.ReLU(inplace=False)
        self.B4_conv = nn.Conv2d(inplanes, outplanes, kernel_size=3, stride=stride, padding=5, dilation=5, bias=True)
        self.B4_bn = nn.BatchNorm2d(nkernels)
        self.B4_relu = nn.ReLU(inplace=False)
        self.B5_gp = nn.MaxPool2d(kernel_size=3, stride=stride, padding=1, dilation=1, return_indices=False,
                                  ceil_mode=False)

    def forward(self, x):
        from .. import engine_aspp
        return engine_aspp.run_aspp_submodule(self, x)


class ASPP_post(nn.Module):
    """biased 1x1 conv -> BN -> ReLU over the ASPP concat (reference :265-286)."""

    def __init__(self, inplanes, outplanes):
        super().__init__()
        self.inplanes = inplanes
        self.outplanes = outplanes
        self.nkernels = outplanes
        self.ASPP_conv = nn.Conv2d(inplanes, outplanes, kernel_size=1, stride=1, padding=0, bias=True)
        self.ASPP_bn = nn.BatchNorm2d(self.nkernels)
        self.ASPP_relu = nn.ReLU(inplace=False)

    def forward(self, x):
        from .. import engine_aspp
        return engine_aspp.run_aspp_submodule(self, x)


class ASPP_ResNet(nn.Module):

    def __init__(self, num_classes=3, in_channels=3, inplanes=16, showsizes=True):
        self.inplanes = inplanes
        super().__init__()
        self.nkernels = 16
        self.in_channels = in_channels
        self.num_classes = num_classes
        self.showsizes = showsizes
        p = inplanes

        self.conv1 = nn.Conv2d(in_channels, p, kernel_size=7, stride=1, padding=3, bias=True)
        self.bn1 = nn.BatchNorm2d(p)
        self.relu1 = nn.ReLU(inplace=False)
        self.pool1 = nn.MaxPool2d(3, stride=2, padding=1)

        for i, stride in zip(range(1, 6), (1, 2, 2, 2, 2)):
            setattr(self, "enc_layer%d" % i, self._make_encoding_layer(p * 2 ** (i - 1), p * 2 ** i, stride=stride))

        # ASPP on enc3 / enc4 / enc5 (registration order as the reference :342-352)
        self.ASPP_layer_enc3 = self.ASPP_layer(p * 8)
        self.ASPP_combine_enc3 = self.ASPP_combine(p * 12, p * 8)
        self.ASPP_layer_enc4 = self.ASPP_layer(p * 16)
        self.ASPP_combine_enc4 = self.ASPP_combine(p * 20, p * 16)
        self.ASPP_layer_enc5 = self.ASPP_layer(p * 32)
        self.ASPP_combine_enc5 = self.ASPP_combine(p * 36, p * 32)

        self.dec_layer5 = self._make_decoding_layer(p * 64, p * 16, p * 32)
        self.dec_layer4 = self._make_decoding_layer(p * 32, p * 8, p * 16)
        self.dec_layer3 = self._make_decoding_layer(p * 16, p * 4, p * 4)
        self.dec_layer2 = self._make_decoding_layer(p * 4, p * 2, p * 2)
        self.dec_layer1 = self._make_decoding_layer(p * 2, p, p)

        self.nkernels = 16
        self.conv10 = nn.Conv2d(p, self.nkernels, kernel_size=7, stride=1, padding=3, bias=True)
        self.bn10 = nn.BatchNorm2d(self.nkernels)
        self.relu10 = nn.ReLU(inplace=False)
        # the reference declares conv11 with in_channels == inplanes (:385), equal to nkernels only at inplanes=16
        self.conv11 = nn.Conv2d(p, num_classes, kernel_size=7, stride=1, padding=3, bias=True)
        self.softmax = nn.LogSoftmax(dim=1)

        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
                fan = m.kernel_size[0] * m.kernel_size[1] * m.out_channels
                m.weight.data.normal_(0, math.sqrt(2. / fan))
            elif isinstance(m, nn.BatchNorm2d):
                m.weight.data.fill_(1)
                m.bias.data.zero_()

        self._engine = None

    def _make_encoding_layer(self, inplanes, planes, stride=2):
        return DoubleResNet(inplanes, planes, stride=stride)

    def _make_decoding_layer(self, inplanes, deconvplanes, resnetplanes):
        return ConvTransposeLayer(inplanes, deconvplanes, resnetplanes)

    def ASPP_layer(self, inplanes):
        return ASPP(inplanes)

    def ASPP_combine(self, inplanes, outplanes):
        return ASPP_post(inplanes, outplanes)

    def engine(self):
        if self._engine is None:
            from ..engine_aspp import ASPPEngine
            self._engine = ASPPEngine(self)
        return self._engine

    def forward(self, x):
        if self.showsizes:
            print("input: ", x.size(), " is_cuda=", x.is_cuda)
        out = self.engine().forward(x)
        if self.showsizes:
            print("  softmax: ", out.size())
        return out
