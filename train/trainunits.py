"""train.trainunits — base class of every training loop (drop-in for the reference's train/trainunits.py:11-153).

Same constructor, attributes (`optimizer_G`, `optimizer_D`, `record`, `epoch`, `i`, ...) and helper methods.
The two optimisers are libganb200's fused multi-tensor Adam with the reference's hyper-parameters
(G lr 1e-4 betas (0.5,0.99); D lr 4e-4 betas (0.0,0.99) — reference :18-19).  The preview window
(matplotlib, reference :20,44-51) and torchvision image grids are optional: on a headless box without them the
calls are no-ops / plain tensor dumps, the training step itself never depends on them.
"""
import math
import os
from pathlib import Path

import torch

from gan_b200 import checkpoint as ckpt_io
from gan_b200 import optim

try:  # no graphics stack on the GPU boxes; the preview is a convenience, never part of the step
    import matplotlib.pyplot as plt
except Exception:  # pragma: no cover
    plt = None
try:
    import torchvision.utils as vutils
except Exception:  # pragma: no cover
    vutils = None


def _ckpt_stem(units, train_name, epoch, i):
    e = epoch + units.epoch + (i + units.i) // units.epoch_len
    return (units.generator_name + ' ' + units.discriminator_name + ' ' + train_name + ' epoch_' + str(e) + ' i_'
            + str((i + units.i) % units.epoch_len)), e, (i + units.i) % units.epoch_len


class Units():
    def __init__(self, generator, generator_name, discriminator, discriminator_name, fixed_noise, epoch_len):
        self.generator = generator
        self.generator_name = generator_name
        self.fixed_noise = fixed_noise
        self.discriminator = discriminator
        self.discriminator_name = discriminator_name
        self.optimizer_G = optim.Adam(generator.parameters(), lr=0.0001, betas=(0.5, 0.99))
        self.optimizer_D = optim.Adam(discriminator.parameters(), lr=0.0004, betas=(0.0, 0.99))
        if plt is not None:
            plt.ion()
        self.show_model()
        self.epoch = 0
        self.epoch_len = epoch_len
        self.i = 0
        self.record = {}
        self.record.update({'epoch': self.epoch, 'i': self.i})

    # ---- bookkeeping -----------------------------------------------------------------------------------
    def make_record(self, record_set):
        for name, value in record_set:
            self.record.setdefault(name, []).append(value)

    def write_record_to_txt(self, file_name, record_set):
        with open(file_name + '.txt', 'a', encoding='utf-8') as file:
            file.write(''.join(name + ' ' + str(value) + ' ' for name, value in record_set) + '\n')

    def show_model(self):
        print('# generator parameters:', sum(param.numel() for param in self.generator.parameters()))
        print('# discriminator parameters:', sum(param.numel() for param in self.discriminator.parameters()))

    # ---- previews --------------------------------------------------------------------------------------
    def show_images(self, gen_imgs, b_size):
        if plt is None or vutils is None:
            return
        with torch.no_grad():
            plt.clf()
            grid = vutils.make_grid(gen_imgs.cpu(), nrow=math.ceil(math.sqrt(b_size)), padding=2, normalize=True)
            plt.imshow(grid.squeeze().permute(1, 2, 0))

    def save_images(self, path, epoch, i):
        """reference :52-56.  The generator runs in train mode (batch statistics, running stats updated: SURVEY B4).
        Min/max normalisation, grid assembly and uint8 quantisation happen on the device, the PNG is written by a worker
        thread once an asynchronous copy has landed (gan_b200/preview.py): same file bytes-for-pixels as the reference's
        make_grid(padding=2, normalize=True) + save_image, but no device synchronisation in the training loop.
        GANB200_PREVIEW=sync restores the reference's blocking host path (needs torchvision)."""
        with torch.no_grad():
            gen_imgs = self.generator(self.fixed_noise)
        target = path + str(epoch) + "-" + str(i)
        if os.environ.get("GANB200_PREVIEW") == "sync" and vutils is not None:
            vutils.save_image(vutils.make_grid(gen_imgs.cpu(), padding=2, normalize=True).squeeze(), target + ".png")
            return
        from gan_b200 import preview
        preview.save_grid_async(gen_imgs, target + ".png", nrow=8, padding=2)

    # ---- checkpoints (whole-module pickles, reference :58-131; interop rules in gan_b200/checkpoint.py) -----------
    # Opt-in departures from the reference (environment, read at call time):
    #   GANB200_CKPT_OPTIMIZER=1  save_ckpt adds both optimiser states; the loaders restore them
    #   GANB200_CKPT_INPLACE=1    the loaders copy the file's weights INTO the live modules, so the optimisers built in
    #                             __init__ keep training them (the reference replaces self.generator / self.discriminator
    #                             with the unpickled object and leaves the optimisers bound to the old parameters:
    #                             after a resume its Adam steps update modules nobody runs — kept by default)
    def save_ckpt(self, train_name, epoch, i):
        print('Saving.....')
        stem, e, ii = _ckpt_stem(self, train_name, epoch, i)
        state = {'generator': self.generator, 'generator_name': self.generator_name,
                 'discriminator': self.discriminator, 'discriminator_name': self.discriminator_name,
                 'method': train_name, 'epoch': e, 'i': ii,
                 # extra keys (the reference's loader ignores them): same storages as the modules, no extra bytes
                 'generator_state': self.generator.state_dict(),
                 'discriminator_state': self.discriminator.state_dict()}
        if os.environ.get('GANB200_CKPT_OPTIMIZER') == '1':
            state['optimizer_G'] = self.optimizer_G.state_dict()
            state['optimizer_D'] = self.optimizer_D.state_dict()
        from gan_b200 import rng
        state['ganb200_noise_rng'] = rng.state_dict()   # {device: [seed, forward counter]} of the in-kernel noise stream
        os.makedirs('checkpoint', exist_ok=True)
        torch.save(state, './checkpoint/' + stem + '_ckpt.pth')

    def save_record(self, train_name, epoch, i):
        stem, e, ii = _ckpt_stem(self, train_name, epoch, i)
        state = {'epoch': e, 'i': ii, 'generator_name': self.generator_name,
                 'discriminator_name': self.discriminator_name, 'method': train_name, 'record': self.record}
        os.makedirs('checkpoint records', exist_ok=True)
        torch.save(state, './checkpoint records/' + stem + '_record.pth')
        self.record = {'epoch': self.epoch, 'i': self.i}

    def _load(self, name, what):
        print('==> Resuming ' + what + ' from checkpoint..')
        os.makedirs('checkpoint', exist_ok=True)
        ckpt = Path('./checkpoint/' + name + '.pth')
        if not ckpt.is_file():
            print('not find ckpt')
            return None
        print('find ckpt')
        checkpoint = torch.load(ckpt, weights_only=False)  # whole-module pickle (torch>=2.6 defaults to True)
        print(what + ' ' + checkpoint[what + '_name'])
        print(what + ' epoch: ', checkpoint['epoch'])
        print(what + ' i: ', checkpoint['i'])
        if what == 'generator' and checkpoint.get('ganb200_noise_rng'):
            from gan_b200 import rng
            try:
                rng.load_state_dict(checkpoint['ganb200_noise_rng'])   # a resume continues the same noise stream
            except Exception:
                pass   # written on another device layout: start a fresh stream
        return checkpoint

    def _adopt(self, checkpoint, what, live, optimizer, opt_key):
        """The module to run after a load: never the stock-torch hybrid a reference pickle unpickles into."""
        inplace = os.environ.get('GANB200_CKPT_INPLACE') == '1'
        module = ckpt_io.load_into(checkpoint[what], live) if inplace else ckpt_io.adopt(checkpoint[what], live)
        if inplace and opt_key in checkpoint and os.environ.get('GANB200_CKPT_OPTIMIZER') == '1':
            optimizer.load_state_dict(checkpoint[opt_key])
        return module

    def load_generator_ckpt(self, name):
        checkpoint = self._load(name, 'generator')
        if checkpoint is not None:
            self.generator = self._adopt(checkpoint, 'generator', self.generator, self.optimizer_G, 'optimizer_G')
            self.epoch = checkpoint['epoch']
            self.i = checkpoint['i']
            self.record.update({'epoch': self.epoch, 'i': self.i})

    def load_discriminator_ckpt(self, name):
        checkpoint = self._load(name, 'discriminator')
        if checkpoint is not None:
            self.discriminator = self._adopt(checkpoint, 'discriminator', self.discriminator, self.optimizer_D,
                                             'optimizer_D')

    def draw_plt_record(self, name_png, per_iteration):
        if plt is None:
            return
        fig1 = plt.figure(num=name_png, figsize=(16, 9), clear=True)
        ax = fig1.add_subplot(111)
        styles = [(0, (3, 3, 1, 2)), (0, (1, 1)), (0, (5, 5)), (0, (5, 3, 1, 2)), (0, (3, 1, 1, 1, 2, 1)),
                  (0, (3, 4, 1, 2, 1, 2))]
        curves = [k for k in self.record if k not in ('epoch', 'i')]
        for style, key in zip(styles, curves):
            ax.plot(self.record[key], label=key, linewidth=1.0, linestyle=style)
        ax.legend(fontsize='x-large', loc=1)
        ax.set_xlabel('every ' + str(per_iteration) + ' iterations')
        ax.set_ylabel('loss amount')
        plt.savefig(name_png + '.png', dpi=960)
