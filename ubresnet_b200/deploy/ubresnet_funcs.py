"""Checkpoint loader -- drop-in for the reference `deploy/ubresnet_funcs.py:41-68`.

Builds UResNet(inplanes=16, input_channels=1, num_classes=4), remaps the checkpoint's GPU id, strips the
`module.` prefix that nn.DataParallel checkpoints carry (:52-64) and loads the state_dict strictly.
"""
from collections import OrderedDict

import torch

from ..models.ub_uresnet import UResNet


def load_cosmic_retrain_model(checkpointfile, device="cuda:0", checkpointgpu=0, num_classes=4, inplanes=16):
    model = UResNet(inplanes=inplanes, input_channels=1, num_classes=num_classes, showsizes=False)
    map_location = None
    if device != "cuda:%d" % (checkpointgpu):
        map_location = {"cuda:%d" % (checkpointgpu): device}
    checkpoint = torch.load(checkpointfile, map_location=map_location, weights_only=False)
    state = checkpoint["state_dict"]
    if any("module." in k for k in state):
        state = OrderedDict((k[7:] if k.startswith("module.") else k, v) for k, v in state.items())
        checkpoint["state_dict"] = state
    model.load_state_dict(state)
    return model
