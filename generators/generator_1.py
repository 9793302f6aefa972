"""generator_1 — drop-in for the reference's generators/generator_1.py:5-28 (BASELINE config 1): an MLP
z -> 256 -> 512 -> 3*H*W with LeakyReLU(0.2) and Tanh, on libganb200 kernels (linear = the GEMM engine, fused bias;
LeakyReLU / Tanh = pointwise kernels).  Same module tree (`generator.0 ... generator.5`), so state_dicts interchange."""
import torch.nn as nn

from gan_b200 import layers as L


class Generator(nn.Module):
    def __init__(self, z_dim, target_image_size):
        """
        :param z_dim: latent z dim
        :param target_image_size: tuple (3, h, w)
        """
        super().__init__()
        self.view_image_size = target_image_size[0] * target_image_size[1] * target_image_size[2]
        self.out_image_size = target_image_size
        self.z_dim = z_dim
        self.generator = nn.Sequential()
        self.generator.add_module(name="0", module=L.Linear(in_features=self.z_dim, out_features=256))
        self.generator.add_module(name="1", module=L.LeakyReLU(0.2))
        self.generator.add_module(name="2", module=L.Linear(in_features=256, out_features=512))
        self.generator.add_module(name="3", module=L.LeakyReLU(0.2))
        self.generator.add_module(name="4", module=L.Linear(in_features=512, out_features=self.view_image_size))
        self.generator.add_module(name="5", module=L.Tanh())

    def forward(self, x):
        x = x.reshape(x.size(0), -1)
        out = self.generator(x)
        return out.reshape(x.size(0), *self.out_image_size)
