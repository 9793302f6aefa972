"""discriminator_1 — drop-in for the reference's discriminators/discriminator_1.py:5-24 (BASELINE config 1): an MLP
3*H*W -> 256 -> 64 -> 1 with LeakyReLU(0.2) and a Sigmoid head (so train/gan.py's BCELoss is valid), on libganb200."""
import torch.nn as nn

from gan_b200 import layers as L


class Discriminator(nn.Module):
    def __init__(self, image_size):
        """
        :param image_size: tuple (3, h, w)
        """
        super().__init__()
        self.in_image_size = image_size[0] * image_size[1] * image_size[2]
        self.discriminator = nn.Sequential()
        self.discriminator.add_module(name="0", module=L.Linear(in_features=self.in_image_size, out_features=256))
        self.discriminator.add_module(name="1", module=L.LeakyReLU(0.2))
        self.discriminator.add_module(name="2", module=L.Linear(in_features=256, out_features=64))
        self.discriminator.add_module(name="3", module=L.LeakyReLU(0.2))
        self.discriminator.add_module(name="4", module=L.Linear(in_features=64, out_features=1))
        self.discriminator.add_module(name="5", module=L.Sigmoid())

    def forward(self, x):
        x = x.reshape(x.size(0), -1)
        return self.discriminator(x)
