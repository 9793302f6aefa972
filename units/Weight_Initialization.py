"""units.Weight_Initialization — drop-in for the reference's units/Weight_Initialization.py:7-13 (DCGAN init applied
by `.apply(weights_init)`; host-side, construction time only)."""
import torch.nn as nn


def weights_init(m):
    classname = m.__class__.__name__
    if classname.find('Conv') != -1:
        nn.init.normal_(m.weight.data, 0.0, 0.02)
    elif classname.find('BatchNorm') != -1:
        nn.init.normal_(m.weight.data, 1.0, 0.02)
        nn.init.constant_(m.bias.data, 0)
