"""train.ganR2 — drop-in for the reference's train/ganR2.py: class Train(trainunits.Units) with the same constructor
and the same generator_trainstep / discriminator_trainstep / compute_grad2 / gradient_penalty / train methods.
The step itself is gan_b200.steps (recipe "ganR2"), running on libganb200 kernels."""
import torch

from gan_b200 import layers, optim
from gan_b200.steps import RECIPES, LoopMixin, StepMixin
from train import trainunits


class Train(StepMixin, LoopMixin, trainunits.Units):
    recipe = RECIPES["ganR2"]

    def __init__(self, dataloader, device, num_epochs, nz, generator, generator_name, discriminator,
                 discriminator_name):
        rc = self.recipe
        super(Train, self).__init__(generator, generator_name, discriminator, discriminator_name,
                                    torch.randn(rc.fixed_noise, nz, 1, 1, device=device), len(dataloader))
        self.dataloader = dataloader
        self.device = device
        self.num_epochs = num_epochs
        self.nz = nz
        if rc.bce:
            self.criterion = layers.BCELoss().to(device)
        if rc.adamw:
            self.optimizer_G = optim.AdamW(generator.parameters(), lr=0.0001, betas=(0.5, 0.999))
            self.optimizer_D = optim.AdamW(discriminator.parameters(), lr=0.0004, betas=(0.5, 0.999))
