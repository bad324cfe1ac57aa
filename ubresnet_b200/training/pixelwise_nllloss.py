"""PixelWiseNLLLoss -- drop-in for the reference `training/pixelwise_nllloss.py:34-61`.

loss = mean_{b,h,w}( pixelweights[b,h,w] * weight[target] * -predict[b,target,h,w] ), where `predict` holds
log-probabilities; pixels whose target equals `ignore_index` contribute 0 but stay in the mean's denominator
(F.nll_loss(reduce=False) * pixelweights, then torch.mean -- reference :51-59).  One fused CUDA pass each way.
"""
import torch
import torch.nn as nn

from .. import ops


def _assert_no_grad(variable):
    assert not variable.requires_grad, \
        "nn criterions don't compute the gradient w.r.t. targets - please " \
        "mark these variables as not requiring gradients"


class _PixelWiseNLLFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, predict, target, pixelweights, classw, ignore_index):
        with torch.cuda.device(predict.device):
            loss = ops.pixelwise_nll_fwd(predict, target, pixelweights, classw, ignore_index)
        ctx.save_for_backward(target, pixelweights, classw if classw is not None else torch.empty(0))
        ctx.has_cw = classw is not None
        ctx.shape = tuple(predict.shape)
        ctx.ignore_index = ignore_index
        return loss

    @staticmethod
    def backward(ctx, gout):
        target, pixw, cw = ctx.saved_tensors
        n, c, h, w = ctx.shape
        with torch.cuda.device(target.device):
            g = gout.detach().to(torch.float32).contiguous()
            dlogp = ops.pixelwise_nll_bwd(target, pixw, cw if ctx.has_cw else None, g, n, c, h, w, ctx.ignore_index)
        return dlogp, None, None, None, None


class PixelWiseNLLLoss(nn.modules.loss._WeightedLoss):
    def __init__(self, weight=None, size_average=True, ignore_index=-100):
        super().__init__(weight, reduction="mean" if size_average else "sum")
        self.size_average = size_average
        self.ignore_index = ignore_index
        self.reduce = False

    def forward(self, predict, target, pixelweights):
        """predict (b,c,h,w) log-softmax output; target (b,h,w) int64 class; pixelweights (b,h,w) float."""
        _assert_no_grad(target)
        _assert_no_grad(pixelweights)
        ops.require_cuda(predict, "PixelWiseNLLLoss predict")
        ops.require_cuda(target, "PixelWiseNLLLoss target")
        ops.require_cuda(pixelweights, "PixelWiseNLLLoss pixelweights")
        if predict.dim() != 4 or target.shape != (predict.shape[0],) + tuple(predict.shape[2:]) \
                or pixelweights.shape != target.shape:
            raise RuntimeError("PixelWiseNLLLoss: shape mismatch predict %s target %s pixelweights %s" %
                               (tuple(predict.shape), tuple(target.shape), tuple(pixelweights.shape)))
        cw = None
        if self.weight is not None:
            cw = self.weight.to(device=predict.device, dtype=torch.float32).contiguous()
        return _PixelWiseNLLFn.apply(predict.float().contiguous(), target.long().contiguous(),
                                     pixelweights.float().contiguous(), cw, int(self.ignore_index))
