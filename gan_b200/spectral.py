"""spectral_norm — the drop-in for `torch.nn.utils.parametrizations.spectral_norm` as the reference's *sn critics use it
(discriminators/discriminator_7_wgangpsn.py:14,53-76,112,121; also discriminator_3/5_wgangpsn.py).

Same registration mechanics as torch (torch.nn.utils.parametrize), hence the same state_dict layout
(`<conv>.parametrizations.weight.original`, `<conv>.parametrizations.weight.0._u/_v`), the same RNG consumption at
construction (u, v ~ N(0,1), 15 warm-up iterations, +1 from register_parametrization's consistency check) and the same
train/eval behaviour: every evaluation of `<conv>.weight` in train mode advances u and v by one power iteration in
place (so one wgangpsnR2 iteration advances them four times: three critic forwards in the D step, one in the G step).

What differs is where the arithmetic runs: the power iteration, sigma = u.(W v), W/sigma and its backward
(G/sigma - <G,W>/sigma^2 u v^T) are libganb200 kernels (csrc/misc.cu)."""
import torch
from torch.nn.utils import parametrize
from torch.nn.utils.parametrizations import _SpectralNorm as _TorchSpectralNorm

from . import kernels as KN
from . import ops


class _SpectralNorm(_TorchSpectralNorm):
    _registering = False  # True only inside spectral_norm(): torch's consistency check evaluates forward() once on the
    # freshly constructed (CPU) weight; that single call is part of initialisation, like torch.nn.init, and uses torch.

    def forward(self, weight):
        if self._registering:
            return super().forward(weight)
        if weight.ndim == 1 or self.dim != 0:
            raise NotImplementedError("gan_b200 spectral_norm: conv / linear weights normalised over dim 0 only")
        w = weight if weight.is_contiguous() else weight.contiguous()
        w2d = w.detach().reshape(w.shape[0], -1)
        n_iter = self.n_power_iterations if self.training else 0
        sigma = KN.backend().spectral_power_iter(w2d, self._u, self._v, n_iter, self.eps)
        # u, v are advanced in place by the next forward: the backward of THIS forward needs today's values
        return ops.spectral_scale(w, self._u.clone(), self._v.clone(), sigma)


def spectral_norm(module, name="weight", n_power_iterations=1, eps=1e-12, dim=None):
    """Signature and defaults of torch.nn.utils.parametrizations.spectral_norm."""
    weight = getattr(module, name, None)
    if not isinstance(weight, torch.Tensor):
        raise ValueError(f"Module '{module}' has no parameter or buffer with name '{name}'")
    if dim is None:
        transposed = (torch.nn.ConvTranspose1d, torch.nn.ConvTranspose2d, torch.nn.ConvTranspose3d)
        dim = 1 if isinstance(module, transposed) else 0
    sn = _SpectralNorm(weight, n_power_iterations, dim, eps)
    sn._registering = True
    try:
        parametrize.register_parametrization(module, name, sn)
    finally:
        sn._registering = False
    return module
