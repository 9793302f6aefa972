"""Sampling / preview path (reference train/trainunits.py:52-56): the on-device grid equals
torchvision's make_grid(padding=2, normalize=True) + save_image quantisation bit for bit on uint8, and save_images writes
the same PNG pixels the reference's call writes — asynchronously."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

vutils = pytest.importorskip("torchvision.utils")

CASES = [(64, 3, 64, 64, 8), (8, 3, 64, 64, 8), (16, 3, 64, 64, 8), (5, 3, 9, 7, 8), (1, 3, 64, 64, 8), (10, 1, 28, 28, 4)]


def torchvision_bytes(x, nrow, padding=2):
    grid = vutils.make_grid(x.cpu(), nrow=nrow, padding=padding, normalize=True)
    return grid.mul(255).add_(0.5).clamp_(0, 255).permute(1, 2, 0).to("cpu", torch.uint8)


@pytest.mark.parametrize("case", CASES)
def test_oracle_grid_equals_torchvision(case):
    from oracle.kernels_ref import RefBackend
    n, c, h, w, nrow = case
    x = torch.randn(n, c, h, w, generator=torch.Generator().manual_seed(n + h)) * 1.7 + 0.3
    assert torch.equal(RefBackend().make_grid_u8(x, nrow, 2), torchvision_bytes(x, nrow))


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_device_grid_equals_torchvision_bit_exact(case):
    from gan_b200 import kernels as KN
    n, c, h, w, nrow = case
    x = torch.randn(n, c, h, w, generator=torch.Generator().manual_seed(n + h)) * 1.7 + 0.3
    got = KN.CudaBackend().make_grid_u8(x.cuda(), nrow, 2).cpu()
    want = torchvision_bytes(x, nrow)
    assert got.shape == want.shape
    assert torch.equal(got, want), int((got != want).sum())


def _save_images_roundtrip(tmp_path, device):
    from PIL import Image
    from generators import generator_1
    from discriminators import discriminator_1
    from train import gan
    from gan_b200 import preview
    torch.manual_seed(0)
    G = generator_1.Generator(32, (3, 64, 64)).to(device)
    D = discriminator_1.Discriminator((3, 64, 64)).to(device)
    T = gan.Train([(torch.zeros(4, 3, 64, 64), torch.zeros(4))], device, 1, 32, G, "g", D, "d")
    out = str(tmp_path) + "/"
    T.save_images(out, 3, 7)
    preview.flush()
    img = np.asarray(Image.open(out + "3-7.png"))
    with torch.no_grad():
        want = torchvision_bytes(G(T.fixed_noise), 8).numpy()
    assert img.shape == want.shape and (img == want).all()


def test_save_images_writes_the_reference_pixels_cpu(tmp_path, ref_backend):
    _save_images_roundtrip(tmp_path, torch.device("cpu"))


@pytest.mark.gpu
def test_save_images_writes_the_reference_pixels_gpu(tmp_path):
    _save_images_roundtrip(tmp_path, torch.device("cuda"))
