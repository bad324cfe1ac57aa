"""ubresnet_b200 -- B200-native (sm_100a) implementation of the NuTufts/ubresnet training/inference hot path.

Host side is PyTorch (memory, streams, autograd plumbing, torch.distributed); all arithmetic runs in
hand-written CUDA kernels behind the C ABI declared in include/ubresnet_b200.h.
"""
__version__ = "0.1.0"
