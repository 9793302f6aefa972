"""TEST INFRASTRUCTURE — CPU restatement of every libganb200 kernel, in plain torch ops.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may import this file.  It is the checker
for the CUDA kernels (same method names and argument meaning as gan_b200.kernels.CudaBackend) and it lets
the autograd formulas in gan_b200/ops.py be grad-checked in float64 without a GPU.  Each method states the
math of the kernel it mirrors (include/ganb200.h) using stock ATen on the CPU; nothing here is shipped.
"""
import torch
import torch.nn.functional as F

U_SIGMOID, U_TANH, U_SQUARE, U_AFFINE, U_RSQRT_EPS, U_SQRT, U_LEAKY, U_RECIP, U_LEAKY_MASK, U_COPY, U_POW = range(11)
U_GELU, U_GELU_GRAD, U_GELU_GRAD2 = 11, 12, 13
B_ADD, B_SUB, B_MUL, B_DIV, B_SIGMOID_BWD, B_TANH_BWD, B_LEAKY_BWD = range(7)
V_NC, V_C, V_N, V_SCALAR = range(4)
R_HW, R_NHW, R_CHW, R_ALL = range(4)


def _v(x, v, vmode):
    n, c = (x.shape[0], x.shape[1]) if x.dim() > 1 else (1, x.shape[0])
    tail = (1,) * (x.dim() - 2)
    if vmode == V_NC:
        return v.reshape(n, c, *tail)
    if vmode == V_C:
        return v.reshape(1, c, *tail) if x.dim() > 1 else v
    if vmode == V_N:
        return v.reshape(n, 1, *tail)
    return v.reshape(())


def _pad(x, pad, pad_mode):
    if pad == 0:
        return x
    return F.pad(x, (pad,) * 4, mode="replicate" if pad_mode == 1 else "constant")


class RefBackend:
    name = "ref"

    # ---- convolution family (include/ganb200.h: fprop / dgrad / wgrad) ---------------------------------
    def conv_fprop(self, x, w, bias, stride, pad, pad_mode, alpha, iscale=None, oscale=None, engine=None, keep_x=False):
        with torch.no_grad():
            if iscale is not None:
                x = x * iscale[:, :, None, None]
            y = F.conv2d(_pad(x, pad, pad_mode), w, None, stride) * alpha
            if oscale is not None:
                y = y * oscale[:, :, None, None]
            if bias is not None:
                y = y + bias[None, :, None, None]
            return y

    def conv_dgrad(self, dy, w, in_hw, stride, pad, pad_mode, alpha, iscale=None, oscale=None, engine=None,
                   keep_dy=False):
        N, K = dy.shape[:2]
        C = w.shape[1]
        with torch.enable_grad():
            x = torch.zeros((N, C, *in_hw), dtype=dy.dtype, requires_grad=True)
            y = F.conv2d(_pad(x, pad, pad_mode), w.detach(), None, stride)
            g = dy.detach() if oscale is None else dy.detach() * oscale.detach()[:, :, None, None]
            (dx,) = torch.autograd.grad(y, x, g)
        dx = dx * alpha
        if iscale is not None:
            dx = dx * iscale[:, :, None, None]
        return dx.detach()

    def conv_wgrad(self, x, dy, rs, stride, pad, pad_mode, alpha, iscale=None, oscale=None, engine=None,
                   keep_x=False, keep_dy=False):
        K, C = dy.shape[1], x.shape[1]
        with torch.enable_grad():
            w = torch.zeros((K, C, *rs), dtype=x.dtype, requires_grad=True)
            xs = x.detach() if iscale is None else x.detach() * iscale.detach()[:, :, None, None]
            y = F.conv2d(_pad(xs, pad, pad_mode), w, None, stride)
            g = dy.detach() if oscale is None else dy.detach() * oscale.detach()[:, :, None, None]
            (dw,) = torch.autograd.grad(y, w, g)
        return (dw * alpha).detach()

    # ---- separable resampling: y = Mh x Mw^T on the last two axes -------------------------------------
    def resample2d(self, x, op_h, op_w):
        mh = op_h.dense(x.device, x.dtype)
        mw = op_w.dense(x.device, x.dtype)
        return torch.einsum("oh,...hw,pw->...op", mh, x, mw).contiguous()

    # ---- pointwise -----------------------------------------------------------------------------------
    def ew_unary(self, op, x, p0=0.0, p1=0.0):
        if op == U_SIGMOID: return torch.sigmoid(x)
        if op == U_TANH: return torch.tanh(x)
        if op == U_SQUARE: return x * x
        if op == U_AFFINE: return p0 * x + p1
        if op == U_RSQRT_EPS: return torch.rsqrt(x + p0)
        if op == U_SQRT: return torch.sqrt(x)
        if op == U_LEAKY: return torch.where(x > 0, x, p0 * x)
        if op == U_RECIP: return 1.0 / x
        if op == U_LEAKY_MASK: return torch.where(x > 0, torch.ones_like(x), torch.full_like(x, p0))
        if op == U_COPY: return x.clone()
        if op == U_POW: return torch.pow(x, p0)
        if op == U_GELU: return F.gelu(x)                      # exact (erf) form, nn.GELU() — generators/generator_6.py:12
        if op in (U_GELU_GRAD, U_GELU_GRAD2):
            phi = torch.exp(-0.5 * x * x) * 0.3989422804014327
            if op == U_GELU_GRAD2: return (2.0 - x * x) * phi
            return 0.5 * (1.0 + torch.erf(x * 0.7071067811865476)) + x * phi
        raise ValueError(op)

    def ew_binary(self, op, a, b, p0=0.0):
        if op == B_ADD: return a + b
        if op == B_SUB: return a - b
        if op == B_MUL: return a * b
        if op == B_DIV: return a / b
        if op == B_SIGMOID_BWD: return a * b * (1 - b)
        if op == B_TANH_BWD: return a * (1 - b * b)
        if op == B_LEAKY_BWD: return torch.where(b > 0, a, a * p0)
        raise ValueError(op)

    def ew_bcast(self, op, x, v, vmode):
        vv = _v(x, v, vmode)
        return (x + vv if op == B_ADD else x * vv).contiguous()

    def ew_fma_bcast(self, x, v, z, vmode):
        return (x + _v(x, v, vmode) * z).contiguous()

    def bmm(self, a, b, ta=False, tb=False, alpha=1.0):
        a = a.transpose(1, 2) if ta else a
        b = b.transpose(1, 2) if tb else b
        return (alpha * torch.bmm(a, b)).contiguous()

    def softmax_rows_fwd(self, x):
        return torch.softmax(x, dim=-1).contiguous()

    def softmax_rows_bwd(self, gy, y):
        return (y * (gy - (gy * y).sum(-1, keepdim=True))).contiguous()

    def pixnorm_fwd(self, x, eps):
        """generators/generator_11.py:9-16 — x / sqrt(eps + mean over channels of x^2); also returns 1 / norm per pixel"""
        n, c = x.shape[:2]
        r = torch.rsqrt(eps + (x * x).sum(1, keepdim=True) / c)
        return (x * r).contiguous(), r.reshape(n, -1).contiguous()

    def pixnorm_bwd(self, gy, y, r):
        c = y.shape[1]
        d = (gy * y).sum(1, keepdim=True) / c
        return (r.reshape(y.shape[0], 1, *y.shape[2:]) * (gy - y * d)).contiguous()

    def pool_sum2_ok(self, t0, t1, out):
        return (t0.dim() == 4 and t0.shape == t1.shape and out == 5 and t0.shape[-1] == t0.shape[-2]
                and t0.shape[-1] in (32, 64))

    def pool_sum2(self, t0, t1, out):
        return torch.nn.functional.adaptive_avg_pool2d(t0 + t1, out).contiguous()

    def sk_merge2_fwd(self, t0, t1, att):
        a = att.reshape(att.shape[0], 2, -1)
        sh = (t0.shape[0], t0.shape[1]) + (1,) * (t0.dim() - 2)
        return (a[:, 0].reshape(sh) * t0 + a[:, 1].reshape(sh) * t1).contiguous()

    def sk_bwd_finish(self, g, att, gp):
        x = torch.zeros_like(g).requires_grad_()
        with torch.enable_grad():
            y = torch.nn.functional.adaptive_avg_pool2d(x, gp.shape[-1])
        (up,) = torch.autograd.grad(y, x, gp)
        a = att.reshape(att.shape[0], 2, -1)
        sh = (g.shape[0], g.shape[1], 1, 1)
        return (a[:, 0].reshape(sh) * g + up).contiguous(), (a[:, 1].reshape(sh) * g + up).contiguous()

    def sk_merge2_bwd(self, g, t0, t1, att, att_only=False):
        a = att.reshape(att.shape[0], 2, -1)
        sh = (t0.shape[0], t0.shape[1]) + (1,) * (t0.dim() - 2)
        dims = tuple(range(2, t0.dim()))
        s0 = (g * t0).sum(dims) if dims else g * t0
        s1 = (g * t1).sum(dims) if dims else g * t1
        if att_only:
            return None, None, torch.stack([s0, s1], 1).contiguous()
        return ((a[:, 0].reshape(sh) * g).contiguous(), (a[:, 1].reshape(sh) * g).contiguous(),
                torch.stack([s0, s1], 1).contiguous())

    def reduce(self, x, y, rmode, scale=1.0):
        t = x if y is None else x * y
        if t.dim() == 1:
            t = t[None, :]
        t = t.reshape(t.shape[0], t.shape[1], -1)
        if rmode == R_HW: out = t.sum(2)
        elif rmode == R_NHW: out = t.sum((0, 2))
        elif rmode == R_CHW: out = t.sum((1, 2))
        else: out = t.sum().reshape(1)
        return out * scale

    # ---- PReLU ---------------------------------------------------------------------------------------
    @staticmethod
    def _slope(x, a):
        if a.numel() == 1:
            return a.reshape(())
        return a.reshape(1, -1, *((1,) * (x.dim() - 2)))

    def prelu_fwd(self, x, a):
        return torch.where(x > 0, x, self._slope(x, a) * x)

    def maskscale(self, g, x, a):
        return torch.where(x > 0, g, self._slope(x, a) * g)

    def _asum(self, t, a):
        if a.numel() == 1:
            return t.sum().reshape(1)
        dims = [d for d in range(t.dim()) if d != 1]
        return t.sum(dims)

    def prelu_bwd(self, gy, x, a, need_gx=True, need_ga=True):
        gx = self.maskscale(gy, x, a) if need_gx else None
        ga = self._asum(gy * torch.clamp(x, max=0), a) if need_ga else None
        return gx, ga

    def masked_dot(self, u, v, x, a_n):
        t = u * v * (x <= 0).to(u.dtype)
        if a_n == 1:
            return t.sum().reshape(1)
        return t.sum([d for d in range(t.dim()) if d != 1])

    # ---- BatchNorm (training) --------------------------------------------------------------------------
    @staticmethod
    def _cview(x, v):
        return v.reshape(1, -1, *((1,) * (x.dim() - 2)))

    def bn_stats(self, x):
        dims = [d for d in range(x.dim()) if d != 1]
        mean = x.mean(dims)
        var = ((x - self._cview(x, mean)) ** 2).mean(dims)
        return mean, var

    def bn_apply(self, x, mean, var, gamma, beta, running_mean, running_var, eps, momentum, count_total, nbt=None):
        if nbt is not None:
            nbt += 1
        invstd = torch.rsqrt(var + eps)
        y = (x - self._cview(x, mean)) * self._cview(x, invstd)
        if gamma is not None:
            y = y * self._cview(x, gamma)
        if beta is not None:
            y = y + self._cview(x, beta)
        if running_mean is not None:
            running_mean.mul_(1 - momentum).add_(momentum * mean)
        if running_var is not None:
            n = float(count_total)
            running_var.mul_(1 - momentum).add_(momentum * (var * n / (n - 1) if n > 1 else var))
        return y, invstd

    def bn_small_ok(self, x):
        n = x.numel() // x.shape[1] if x.dim() > 1 else 1
        return n <= 8192

    def bn_small_fwd(self, x, gamma, beta, running_mean, running_var, eps, momentum, nbt=None, slope=None):
        mean, var = self.bn_stats(x)
        count = x.numel() // x.shape[1]
        y, invstd = self.bn_apply(x, mean, var, gamma, beta, running_mean, running_var, eps, momentum, count, nbt)
        if slope is not None:
            y = self.prelu_fwd(y, slope)
        return y, mean, invstd

    def bn_small_bwd(self, gy, x, mean, invstd, gamma, beta=None, slope=None):
        count = x.numel() // x.shape[1]
        gs = None
        if slope is not None:  # recompute the batch-norm output, then the PReLU adjoint
            u = (x - self._cview(x, mean)) * self._cview(x, invstd)
            if gamma is not None:
                u = u * self._cview(x, gamma)
            if beta is not None:
                u = u + self._cview(x, beta)
            gy, gs = self.prelu_bwd(gy, u, slope)
        s1, s2 = self.bn_bwd_reduce(gy, x, mean, invstd)
        out = self.bn_bwd_apply(gy, x, mean, invstd, gamma, s1, s2, count)
        return out if slope is None else (*out, gs)

    def bn_bwd_reduce(self, gy, x, mean, invstd):
        dims = [d for d in range(x.dim()) if d != 1]
        xhat = (x - self._cview(x, mean)) * self._cview(x, invstd)
        return gy.sum(dims), (gy * xhat).sum(dims)

    def bn_bwd_apply(self, gy, x, mean, invstd, gamma, sum_gy, sum_gy_xhat, count_total):
        xhat = (x - self._cview(x, mean)) * self._cview(x, invstd)
        n = float(count_total)
        g = gamma if gamma is not None else torch.ones_like(mean)
        gx = self._cview(x, g * invstd) * (gy - self._cview(x, sum_gy) / n - xhat * self._cview(x, sum_gy_xhat) / n)
        return gx, sum_gy_xhat.clone(), sum_gy.clone()

    # ---- softmax over branches, MiniBatchStdDev, BCE --------------------------------------------------------
    def softmax_m_fwd(self, x):
        return torch.softmax(x, dim=1)

    def softmax_m_bwd(self, gy, y):
        return y * (gy - (gy * y).sum(1, keepdim=True))

    def mbstd_fwd(self, x, G, eps):
        g = x.reshape(G, -1)
        return torch.sqrt(g.var(dim=0) + eps).mean().reshape(1)

    def _mbstd_parts(self, x, G, eps):
        g = x.reshape(G, -1)
        d = g - g.mean(0, keepdim=True)
        sigma = torch.sqrt((d * d).sum(0, keepdim=True) / (G - 1) + eps)
        return d, sigma, g.shape[1]

    def mbstd_bwd(self, g, x, G, eps):
        d, sigma, L = self._mbstd_parts(x, G, eps)
        return (g.reshape(()) * d / ((G - 1) * L * sigma)).reshape(x.shape)

    def mbstd_bwd_bwd(self, g, x, ggx, G, eps):
        d, sigma, L = self._mbstd_parts(x, G, eps)
        gg = ggx.reshape(G, -1)
        dot = (gg * d).sum(0, keepdim=True)
        out_gg = (dot / ((G - 1) * L * sigma)).sum().reshape(1)
        out_gx = g.reshape(()) / ((G - 1) * L) * ((gg - gg.mean(0, keepdim=True)) / sigma - d * dot / ((G - 1) * sigma ** 3))
        return out_gg, out_gx.reshape(x.shape)

    def bce_fwd(self, p, t):
        bad = ((p < 0) | (p > 1) | torch.isnan(p)).any().to(torch.int32).reshape(1)
        lp = torch.clamp(torch.log(p), min=-100)
        lq = torch.clamp(torch.log(1 - p), min=-100)
        return (-(t * lp + (1 - t) * lq)).mean().reshape(1), bad

    def bce_bwd(self, g, p, t):
        return g.reshape(()) / p.numel() * (p - t) / torch.clamp((1 - p) * p, min=1e-12)

    # ---- fused StyleConv layer (gan_b200/csrc/style.cu) --------------------------------------------------------
    def noise_fill(self, shape, rng_state, layer, device):
        """deterministic in (seed, forward counter, layer) like the CUDA kernel; the values themselves come from a
        torch CPU generator (the GPU tests compare the fused kernels with `noise_fill` of the same backend)."""
        gen = torch.Generator().manual_seed((int(rng_state[0]) * 1000003 + int(rng_state[1]) * 7919 + int(layer)) % (2**63))
        return torch.randn(shape, generator=gen, dtype=torch.float32).to(torch.get_default_dtype())

    def demod(self, s, wsq, alpha, eps):
        return torch.rsqrt(alpha * alpha * (s * s) @ wsq.t() + eps)

    def demod_bwd_s(self, gs1, s, Ad, sinv, wsq, alpha, out=None):
        dd = -0.5 * Ad * sinv * sinv
        gs = 2 * s * alpha * alpha * (dd @ wsq)
        gs = gs if gs1 is None else gs + gs1
        if out is not None:
            out.copy_(gs)
            return out
        return gs

    # ---- grouped launches (gan_b200/csrc/group.cu; job lists from gan_b200/group.py) ---------------------------
    @staticmethod
    def _mat(flat, off, n_i, n_r, ld, red_is_row):
        """(n_i, n_r) view with element (i, r) at flat[off + r*ld + i] (red_is_row) or flat[off + i*ld + r]"""
        return torch.as_strided(flat, (n_i, n_r), (1, ld) if red_is_row else (ld, 1), off)

    def linear_group(self, plan, a_base, b_base, o_base):
        of = o_base.reshape(-1)
        for j in plan.jobs:
            af, ao = (j.a.detach().reshape(-1), 0) if j.a is not None else (a_base.reshape(-1), j.a_off)
            bf, bo = (j.b.detach().reshape(-1), 0) if j.b is not None else (b_base.reshape(-1), j.b_off)
            A = self._mat(af, ao, j.ni, j.nr, j.lda, j.a_red_is_row)
            Bm = self._mat(bf, bo, j.nj, j.nr, j.ldb, j.b_red_is_row)
            y = (A @ Bm.t()) * j.alpha
            if j.bias is not None:
                y = y + j.bias.detach()
            torch.as_strided(of, (j.ni, j.nj), (j.ldo, 1), j.out_off).copy_(y)
        return o_base

    @staticmethod
    def _rows(flat, j, N):
        return torch.as_strided(flat.reshape(-1), (N, j.C), (j.C, 1), j.x_off)

    def bn1d_group_stats(self, plan, x_base, N, total_c):
        mean, var = x_base.new_empty(total_c), x_base.new_empty(total_c)
        for j in plan.jobs:
            x = self._rows(x_base, j, N)
            mean[j.stat_off:j.stat_off + j.C] = x.mean(0)
            var[j.stat_off:j.stat_off + j.C] = x.var(0, unbiased=False)
        return mean, var

    def bn1d_group_apply(self, plan, x_base, N, mean, var, eps, momentum, count_total):
        y_base, invstd = torch.empty_like(x_base), torch.empty_like(mean)
        with torch.no_grad():
            for j in plan.jobs:
                sl = slice(j.stat_off, j.stat_off + j.C)
                x = self._rows(x_base, j, N)
                is_ = torch.rsqrt(var[sl] + eps)
                u = (x - mean[sl]) * is_ * j.gamma.detach() + j.beta.detach()
                if j.slope is not None:
                    u = torch.where(u > 0, u, j.slope.detach() * u)
                self._rows(y_base, j, N).copy_(u)
                invstd[sl] = is_
                unb = var[sl] * (count_total / (count_total - 1.0)) if count_total > 1 else var[sl]
                j.running_mean.mul_(1 - momentum).add_(momentum * mean[sl])
                j.running_var.mul_(1 - momentum).add_(momentum * unb)
                j.nbt.add_(1)
        return y_base, invstd

    def _bn_gu(self, j, gy_base, x_base, N, mean, invstd):
        sl = slice(j.stat_off, j.stat_off + j.C)
        xh = (self._rows(x_base, j, N) - mean[sl]) * invstd[sl]
        g = self._rows(gy_base, j, N)
        ps = None
        if j.slope is not None:
            u = xh * j.gamma.detach() + j.beta.detach()
            neg = ~(u > 0)
            ps = (g * u * neg).sum(0)
            g = torch.where(neg, g * j.slope.detach(), g)
        return sl, xh, g, ps

    def bn1d_group_bwd_reduce(self, plan, gy_base, x_base, N, mean, invstd, total_c, has_slope):
        s1, s2 = torch.empty_like(mean), torch.empty_like(mean)
        gsl = torch.empty_like(mean) if has_slope else None
        for j in plan.jobs:
            sl, xh, g, ps = self._bn_gu(j, gy_base, x_base, N, mean, invstd)
            s1[sl], s2[sl] = g.sum(0), (g * xh).sum(0)
            if has_slope:
                gsl[sl] = ps
        return s1, s2, gsl

    def bn1d_group_bwd_apply(self, plan, gy_base, x_base, N, mean, invstd, s1, s2, count_total, total_c):
        gx_base, gbias = torch.empty_like(x_base), torch.empty_like(mean)
        for j in plan.jobs:
            sl, xh, g, _ = self._bn_gu(j, gy_base, x_base, N, mean, invstd)
            gx = j.gamma.detach() * invstd[sl] * (g - s1[sl] / count_total - xh * s2[sl] / count_total)
            self._rows(gx_base, j, N).copy_(gx)
            gbias[sl] = gx.sum(0)
        return gx_base, gbias

    def demod_bwd_w_(self, gw, w, Ad, sinv, s, alpha):
        dd = -0.5 * Ad * sinv * sinv
        dwsq = alpha * alpha * dd.t() @ (s * s)
        gw += 2 * w * dwsq.reshape(dwsq.shape + (1,) * (w.dim() - 2))
        return gw

    def style_tail_fwd(self, y, ns, slope, rng_state, layer):
        u = y
        if ns is not None:
            z = self.noise_fill(y.shape, rng_state, layer, y.device).to(y.dtype)
            u = y.add_(self._cview(y, ns) * z)
        a = None
        if slope is not None:
            a = torch.where(u > 0, u, self._cview(u, slope) * u)
        return u, a

    def style_tail_bwd(self, ga, u, ns, slope, rng_state, layer, need_Ad=True):
        N, K = u.shape[:2]
        gu = ga if slope is None else ga * torch.where(u > 0, torch.ones_like(u), self._cview(u, slope) * torch.ones_like(u))
        yconv, Z = u, None
        if ns is not None:
            z = self.noise_fill(u.shape, rng_state, layer, u.device).to(u.dtype)
            yconv = u - self._cview(u, ns) * z
            Z = (gu * z).reshape(N, K, -1).sum(-1)
        Ad = (gu * yconv).reshape(N, K, -1).sum(-1) if need_Ad else None
        Pm = (ga * u.clamp(max=0)).reshape(N, K, -1).sum(-1) if slope is not None else None
        return gu, Ad, Z, Pm

    def gather_flip(self, cache, idx, flips):
        out = cache.index_select(0, idx).clone()
        sel = flips.to(torch.bool).nonzero().flatten()
        out[sel] = out[sel].flip(-1)
        return out

    def make_grid_u8(self, x, nrow=8, padding=2):
        """torchvision.utils.make_grid(normalize=True) + save_image's quantisation, restated op by op (the GPU test and
        tests/test_units.py compare it with torchvision itself)."""
        t = x.detach().clone().float()
        if t.shape[1] == 1:
            t = torch.cat((t, t, t), 1)
        lo, hi = float(t.min()), float(t.max())
        t.clamp_(min=lo, max=hi)
        t.sub_(lo).div_(max(hi - lo, 1e-5))
        n, c, h, w = t.shape
        if n == 1:
            grid = t[0]
        else:
            xm = min(nrow, n)
            ym = (n + xm - 1) // xm
            grid = t.new_zeros((c, (h + padding) * ym + padding, (w + padding) * xm + padding))
            for k in range(n):
                y0, x0 = (k // xm) * (h + padding) + padding, (k % xm) * (w + padding) + padding
                grid[:, y0:y0 + h, x0:x0 + w] = t[k]
        return grid.mul(255).add_(0.5).clamp_(0, 255).permute(1, 2, 0).to(torch.uint8).contiguous()

    def scale_reduce(self, dxs, s, x, need_gx=True, need_gs=True):
        N, C = dxs.shape[:2]
        gx = dxs * s.reshape(N, C, *([1] * (dxs.dim() - 2))) if need_gx else None
        gs1 = (dxs * x).reshape(N, C, -1).sum(-1) if need_gs else None
        return gx, gs1

    # ---- misc ------------------------------------------------------------------------------------------------
    def chan_copy(self, src, src_c0, dst, dst_c0, ncopy, accumulate=False):
        if accumulate:
            dst[:, dst_c0:dst_c0 + ncopy] += src[:, src_c0:src_c0 + ncopy]
        else:
            dst[:, dst_c0:dst_c0 + ncopy] = src[:, src_c0:src_c0 + ncopy]
        return dst

    def counter_add(self, ctr, inc):
        ctr += inc

    def adam_reference(self, *a, **k):
        return adam_reference(*a, **k)

    def mixup(self, t, f, lam):
        return lam * t + (1 - lam) * f

    def mixup_wgan_f64(self, label, t, f):
        return label * t + (1 - label) * f

    def clamp_(self, x, lo, hi):
        return x.clamp_(lo, hi)

    def spectral_power_iter(self, w2d, u, v, n_iter, eps):
        """torch/nn/utils/parametrizations.py _SpectralNorm._power_method: u first, then v."""
        for _ in range(n_iter):
            u.copy_(F.normalize(torch.mv(w2d, v), dim=0, eps=eps))
            v.copy_(F.normalize(torch.mv(w2d.t(), u), dim=0, eps=eps))
        return torch.dot(u, torch.mv(w2d, v)).reshape(1)

    def spectral_scale(self, w, sigma):
        return w / sigma

    def spectral_scale_bwd(self, g, w, u, v, sigma):
        dot = (g * w).sum()
        uv = torch.outer(u, v).reshape(w.shape)
        return g / sigma - dot / (sigma * sigma) * uv

    # -- depthwise 3x3 conv, channel shuffle (discriminator_7_wgangpsn.py) ----------------------------------
    def dwconv_fprop(self, x, w, bias, stride, pad, alpha):
        with torch.no_grad():
            y = F.conv2d(x, w, None, stride, pad, groups=x.shape[1]) * alpha
            return y if bias is None else y + bias[None, :, None, None]

    def dwconv_dgrad(self, dy, w, in_hw, stride, pad, alpha):
        N, C = dy.shape[:2]
        with torch.enable_grad():
            x = torch.zeros((N, C, *in_hw), dtype=dy.dtype, requires_grad=True)
            y = F.conv2d(x, w.detach(), None, stride, pad, groups=C)
            (dx,) = torch.autograd.grad(y, x, dy.detach())
        return (dx * alpha).detach()

    def dwconv_wgrad(self, x, dy, rs, stride, pad, alpha):
        C = x.shape[1]
        with torch.enable_grad():
            w = torch.zeros((C, 1, *rs), dtype=x.dtype, requires_grad=True)
            y = F.conv2d(x.detach(), w, None, stride, pad, groups=C)
            (dw,) = torch.autograd.grad(y, w, dy.detach())
        return (dw * alpha).detach()

    def chan_shuffle(self, x, groups):
        n, c = x.shape[:2]
        rest = x.shape[2:]
        return x.reshape(n, groups, c // groups, *rest).transpose(1, 2).reshape(n, c, *rest).contiguous()


def adam_reference(p, g, m, v, lr, b1, b2, eps, wd, decoupled, step, grad_scale=1.0):
    """single-tensor torch.optim.Adam / AdamW math (torch/optim/adam.py _single_tensor_adam), in place."""
    g = g * grad_scale
    if wd != 0:
        if decoupled:
            p.mul_(1 - lr * wd)
        else:
            g = g + wd * p
    m.lerp_(g, 1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    bc1 = 1 - b1 ** step
    bc2 = 1 - b2 ** step
    denom = (v.sqrt() / (bc2 ** 0.5)).add_(eps)
    p.addcdiv_(m, denom, value=-(lr / bc1))
