"""units.mixup — drop-in for the reference's units/mixup.py:9-27 (two blend helpers that no reference loop calls;
kept for API parity).  The Beta draw stays a host-side numpy call on numpy's global RNG, so the mixing
coefficients are bit-identical to the reference's for the same np.random.seed; the blend runs on libganb200.

Quirks reproduced: `mixup_data_wgan` multiplies a float64 (B,1) label with (B,3,64,64) images — it broadcasts only when the
last image dimension equals B (or B == 1), blends per image ROW rather than per sample, and returns float64."""
import numpy as np
import torch

from gan_b200 import kernels as KN


def mixup_data(true, fake, alpha, device):
    lam = np.random.beta(alpha, alpha) if alpha > 0 else 1
    batch_size = true.size(0)
    label = torch.full((batch_size, 1), lam, dtype=torch.float32, device=device)
    mixed = KN.backend().mixup(true.contiguous(), fake.contiguous(), float(lam))
    return mixed, label, lam


def mixup_data_wgan(true, fake, alpha, device):
    batch_size = true.size(0)
    if alpha > 0:
        label = np.random.beta(alpha, alpha, size=(batch_size, 1))
    else:
        label = np.ones(shape=(batch_size, 1))
    label = torch.from_numpy(label).to(device)
    if true.dim() != 4 or (batch_size != 1 and true.shape[-2] != batch_size):
        # same failure as torch's broadcast of (B,1) against (B,C,H,W) in the reference
        raise RuntimeError(f"The size of tensor a ({batch_size}) must match the size of tensor b ({true.shape[-2]}) at "
                           f"non-singleton dimension {true.dim() - 2}")
    mixed = KN.backend().mixup_wgan_f64(label.contiguous(), true.contiguous(), fake.contiguous())
    return mixed, label
