"""U-ResNet (MicroBooNE track/shower semantic segmentation) -- drop-in for the reference
`models/ub_uresnet.py` (UResNet.__init__ :31-79, forward :88-147).

Constructor signature, attribute tree (`conv1, bn1, relu1, pool1, enc_layer1..5, dec_layer5..1, conv10, bn10,
relu10, conv11, softmax`, `inplanes`, `nkernels`, `_showsizes`), parameter names/shapes/dtypes and the
initialisation (conv/deconv W ~ N(0, sqrt(2/(kh*kw*out_channels))), BN gamma=1, beta=0) are the reference's,
so ssnet2018-style checkpoints and the training/deploy scripts work unchanged.  `forward` returns
log-probabilities (B, num_classes, H, W) fp32 like the reference; the arithmetic is executed by the sm_100a
kernel schedule in `ubresnet_b200.engine` (bf16 NHWC activations, fp32 accumulation and statistics).
There is no CPU path: a CPU input raises.
"""
import math
import os

import torch
import torch.nn as nn

from .common_layers import ConvTransposeLayer, DoubleResNet


class UResNet(nn.Module):

    def __init__(self, num_classes=3, input_channels=3, inplanes=16, final_conv_kernels=16, showsizes=False,
                 precision=None):
        """`precision` (extra, keyword; default from $UBR_PRECISION, else "bf16"): "bf16" = tensor-core schedule with
        bf16 NHWC storage and fp32 accumulation; "fp32" = the reference's arithmetic type end to end (parity mode,
        CUDA cores; also the only mode for input_channels != 1).  May be changed later through `model.precision`."""
        self.inplanes = inplanes
        super().__init__()
        self._showsizes = showsizes
        self.precision = precision or os.environ.get("UBR_PRECISION", "bf16")
        if self.precision not in ("bf16", "fp32"):
            raise ValueError("precision must be 'bf16' or 'fp32', got %r" % (self.precision,))
        if input_channels != 1 and self.precision == "bf16":
            # the tensor-core stem kernels are single-channel (wire-plane crops); the reference's default ctor
            # (input_channels=3, models/ub_uresnet.py:31) still has to construct and run
            if precision == "bf16":
                raise ValueError("UResNet: precision='bf16' takes input_channels=1; other channel counts run in "
                                 "precision='fp32'")
            import warnings
            warnings.warn("ubresnet_b200: input_channels=%d -> precision='fp32' (the bf16 tensor-core schedule is "
                          "single-channel)" % input_channels)
            self.precision = "fp32"
        p = self.inplanes

        # stem
        self.conv1 = nn.Conv2d(input_channels, p, kernel_size=7, stride=1, padding=3, bias=True)
        self.bn1 = nn.BatchNorm2d(p)
        self.relu1 = nn.ReLU(inplace=False)
        self.pool1 = nn.MaxPool2d(3, stride=2, padding=1)

        # encoder: widths p*2^i, first stage keeps resolution
        for i, stride in zip(range(1, 6), (1, 2, 2, 2, 2)):
            setattr(self, "enc_layer%d" % i, self._make_encoding_layer(p * 2 ** (i - 1), p * 2 ** i, stride=stride))
        # decoder, deepest first
        for i in (5, 4, 3, 2, 1):
            setattr(self, "dec_layer%d" % i,
                    self._make_decoding_layer(p * 2 ** i, p * 2 ** (i - 1), p * 2 ** (i - 1)))

        # head: 7x7 -> BN -> ReLU -> 7x7 (no BN / ReLU on the last layer)
        self.nkernels = final_conv_kernels
        self.conv10 = nn.Conv2d(p, self.nkernels, kernel_size=7, stride=1, padding=3, bias=True)
        self.bn10 = nn.BatchNorm2d(self.nkernels)
        self.relu10 = nn.ReLU(inplace=False)
        self.conv11 = nn.Conv2d(self.nkernels, num_classes, kernel_size=7, stride=1, padding=3, bias=True)
        self.softmax = nn.LogSoftmax(dim=1)

        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
                fan = m.kernel_size[0] * m.kernel_size[1] * m.out_channels
                m.weight.data.normal_(0, math.sqrt(2. / fan))
            elif isinstance(m, nn.BatchNorm2d):
                m.weight.data.fill_(1)
                m.bias.data.zero_()

        self._engine = None
        self._engine_f32 = None

    def _make_encoding_layer(self, inplanes, planes, stride=2):
        return DoubleResNet(inplanes, planes, stride=stride)

    def _make_decoding_layer(self, inplanes, deconvplanes, resnetplanes):
        return ConvTransposeLayer(inplanes, deconvplanes, resnetplanes)

    def engine(self):
        # `nn.DataParallel` (training/train_ubresnet2018_wlarcv2.py:99,103) shallow-copies the module per device on every
        # forward: a replica must not run through the original's engine (its caches belong to the original's parameters
        # on another device), so an engine whose `.model` is not this object is replaced
        if self.precision == "fp32":
            if self._engine_f32 is None or self._engine_f32.model is not self:
                from ..engine_f32 import EngineF32
                self._engine_f32 = EngineF32(self)
            return self._engine_f32
        if self._engine is None or self._engine.model is not self:
            from ..engine import Engine
            self._engine = Engine(self)
        return self._engine

    def forward(self, x):
        if self._showsizes:
            print("input: ", x.size(), " is_cuda=", x.is_cuda)
        out = self.engine().forward(x)
        if self._showsizes:
            print("  softmax: ", out.size())
        return out
