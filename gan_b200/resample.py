"""1-D linear operators for the separable resampling kernel (ganb200_resample2d).

Smooth, bicubic x2 / x0.5, AdaptiveAvgPool2d and ReplicationPad2d are all separable linear maps on the
H and W axes.  Each is built here as a small dense matrix (host, float64), compositions are matrix
products, adjoints are transposes, and the kernel consumes them as padded per-output tap tables.

Reference semantics restated:
  Smooth            discriminators/discriminator_9_4.py:56-72, generators/generator_13_5.py:134-150
  bicubic up x2     generators/generator_13_5.py:160  (nn.Upsample, align_corners=False, A=-0.75)
  bicubic down x0.5 discriminators/discriminator_9_4.py:81 (F.interpolate size=h//2, no antialias)
  AdaptiveAvgPool2d discriminators/discriminator_9_4.py:86, generators/generator_13_5.py:44
"""
import functools
import math

import numpy as np
import torch


class LinOp1d:
    """Dense [n_out, n_in] operator + cached device tap tables."""

    def __init__(self, mat, name=""):
        self.mat = np.ascontiguousarray(mat, dtype=np.float64)
        self.n_out, self.n_in = self.mat.shape
        self.name = name
        self._T = None
        self._taps = {}
        self._dense = {}

    @property
    def T(self):
        if self._T is None:
            self._T = LinOp1d(self.mat.T, self.name + "^T")
            self._T._T = self
        return self._T

    def __matmul__(self, other):
        """(self @ other)(x) = self(other(x))"""
        return LinOp1d(self.mat @ other.mat, f"({self.name}.{other.name})")

    def taps(self, device):
        """(idx int32 [n_out,T], w float32 [n_out,T], T) on `device`."""
        key = str(device)
        if key not in self._taps:
            nz = [np.nonzero(row)[0] for row in self.mat]
            T = max(1, max(len(z) for z in nz))
            idx = np.zeros((self.n_out, T), dtype=np.int32)
            w = np.zeros((self.n_out, T), dtype=np.float32)
            for o, z in enumerate(nz):
                idx[o, :len(z)] = z
                w[o, :len(z)] = self.mat[o, z]
            self._taps[key] = (torch.from_numpy(idx).to(device), torch.from_numpy(w).to(device), T)
        return self._taps[key]

    def dense(self, device, dtype):
        key = (str(device), dtype)
        if key not in self._dense:
            self._dense[key] = torch.from_numpy(self.mat).to(device=device, dtype=dtype)
        return self._dense[key]


@functools.lru_cache(maxsize=None)
def identity(n):
    return LinOp1d(np.eye(n), f"I{n}")


@functools.lru_cache(maxsize=None)
def replicate_pad(n, p):
    m = np.zeros((n + 2 * p, n))
    for o in range(n + 2 * p):
        m[o, min(max(o - p, 0), n - 1)] = 1.0
    return LinOp1d(m, f"pad{p}_{n}")


@functools.lru_cache(maxsize=None)
def reflect_pad(n, p):
    """padding_mode='reflect' of nn.Conv2d / nn.ReflectionPad2d: the border is mirrored WITHOUT repeating the edge sample
    (generators/generator_7.py:13-14, discriminators/discriminator_6_wgangp.py:12-13)."""
    assert 0 <= p < n, "reflect padding needs pad < size"
    m = np.zeros((n + 2 * p, n))
    for o in range(n + 2 * p):
        i = o - p
        i = -i if i < 0 else (2 * (n - 1) - i if i > n - 1 else i)
        m[o, i] = 1.0
    return LinOp1d(m, f"reflect{p}_{n}")


@functools.lru_cache(maxsize=None)
def smooth3(n):
    """[1,2,1]/4 with replicate padding (the 2-D kernel [[1,2,1],[2,4,2],[1,2,1]]/16 is its outer product)."""
    m = np.zeros((n, n))
    for o in range(n):
        for t, wt in zip((-1, 0, 1), (0.25, 0.5, 0.25)):
            m[o, min(max(o + t, 0), n - 1)] += wt
    return LinOp1d(m, f"smooth{n}")


def _cubic_coeffs(t, A=-0.75):
    def c1(x):  # |x| <= 1
        return ((A + 2) * x - (A + 3)) * x * x + 1

    def c2(x):  # 1 < |x| < 2
        return ((A * x - 5 * A) * x + 8 * A) * x - 4 * A

    return [c2(t + 1.0), c1(t), c1(1.0 - t), c2(2.0 - t)]


@functools.lru_cache(maxsize=None)
def bicubic(n_in, n_out, scale=None):
    """torch upsample_bicubic2d along one axis, align_corners=False.  `scale` = the src/dst ratio torch uses
    (n_in/n_out when a size is given; 1/scale_factor when scale_factor is given)."""
    if scale is None:
        scale = n_in / n_out
    m = np.zeros((n_out, n_in))
    for o in range(n_out):
        src = scale * (o + 0.5) - 0.5
        ix = math.floor(src)
        t = src - ix
        for k, wt in enumerate(_cubic_coeffs(t)):
            m[o, min(max(ix - 1 + k, 0), n_in - 1)] += wt
    return LinOp1d(m, f"bicubic{n_in}->{n_out}")


@functools.lru_cache(maxsize=None)
def adaptive_avg(n_in, n_out):
    m = np.zeros((n_out, n_in))
    for o in range(n_out):
        s = (o * n_in) // n_out
        e = -((-(o + 1) * n_in) // n_out)
        m[o, s:e] = 1.0 / (e - s)
    return LinOp1d(m, f"aavg{n_in}->{n_out}")


@functools.lru_cache(maxsize=None)
def smooth_then_bicubic_down2(n):
    """DownSample of discriminator_9_4.py:74-81 fused into one operator."""
    return bicubic(n, n // 2) @ smooth3(n)


@functools.lru_cache(maxsize=None)
def bicubic_up2_then_smooth(n):
    """smooth(up_sample(x)) of generator_13_5.py:167 fused into one operator."""
    return smooth3(2 * n) @ bicubic(n, 2 * n, 0.5)
