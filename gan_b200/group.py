"""Grouped evaluation of the 519 per-layer style MLPs of generator_13_5 (gan_b200/csrc/group.cu).

Every modulated convolution owns  to_style = Sequential(MappingNetwork(d_latent, 1), EqualizedLinear(d_latent, Cin),
BatchNorm1d(Cin))  with  MappingNetwork(.., 1) = Sequential(EqualizedLinear, BatchNorm1d, PReLU)
(/root/reference/generators/generator_13_5.py:205-216, 223-227), and all of them read the same `w`.  `GroupedStyles`
evaluates the whole family right after the mapping network in 6 launches forward (grouped linear, batch-norm statistics,
batch-norm apply + PReLU, twice) and 11 backward, instead of 4 resp. 8 launches PER LAYER — same arithmetic per layer
(fp32, sums in a fixed order), same parameters, same running-statistics updates.  Parameter gradients come back as views
into one flat buffer per kind; the gradient of each style is written by its consumer straight into its segment of one
ragged buffer (`tensor._gb_grad_out`, see ops.StyleLayerFn.backward), so the backward needs no gather.

Data parallel: the two batch-norm families exchange their statistics through gan_b200.ops._sync_bn_hooks — ONE all-reduce
per family and direction instead of one per layer.
"""
import ctypes

import torch
from torch.autograd import Function

from . import kernels as KN
from . import ops


class GlinJob:
    """out[i][j] = alpha * sum_r A(i,r) B(j,r) (+ bias[j]); an operand is a tensor (static: a parameter) or an element
    offset into the base pointer given at launch; element (i,r) sits at [r*ld + i] when *_red_is_row else [i*ld + r]."""
    __slots__ = ("a", "a_off", "b", "b_off", "out_off", "bias", "ni", "nj", "nr", "lda", "ldb", "ldo", "a_red_is_row",
                 "b_red_is_row", "alpha")

    def __init__(self, **kw):
        self.a = self.b = self.bias = None
        self.a_off = self.b_off = self.out_off = 0
        self.a_red_is_row = self.b_red_is_row = 0
        for k, v in kw.items():
            setattr(self, k, v)


class GbnJob:
    __slots__ = ("gamma", "beta", "slope", "running_mean", "running_var", "nbt", "x_off", "stat_off", "C")

    def __init__(self, **kw):
        self.slope = None
        for k, v in kw.items():
            setattr(self, k, v)


class Plan:
    """A list of jobs + the device table the CUDA backend builds from it once (cached per device)."""

    def __init__(self, jobs):
        self.jobs = jobs
        self._dev = {}


class StylePlans:
    def __init__(self, mods, B):
        """mods: the to_style Sequentials; B: batch rows"""
        D = mods[0][0].net[0].weight.raw.shape[1]
        self.B, self.D, self.L = B, D, len(mods)
        self.C = [m[1].weight.raw.shape[0] for m in mods]
        self.off = [0]
        for c in self.C:
            self.off.append(self.off[-1] + c)
        self.sumC = self.off[-1]
        l1f, l2f, l2d, l2w, l1d, l1w, bn1, bn2 = [], [], [], [], [], [], [], []
        for j, m in enumerate(mods):
            lin1, n1, act1 = m[0].net[0], m[0].net[1], m[0].net[2]
            lin2, n2 = m[1], m[2]
            W1, W2, C, o = lin1.weight.raw, lin2.weight.raw, self.C[j], self.off[j]
            c1, c2 = lin1.weight.gain, lin2.weight.gain
            hb = j * B * D
            l1f.append(GlinJob(a_off=0, b=W1, out_off=hb, bias=lin1.bias, ni=B, nj=D, nr=D, lda=D, ldb=D, ldo=D, alpha=c1))
            l2f.append(GlinJob(a_off=hb, b=W2, out_off=B * o, bias=lin2.bias, ni=B, nj=C, nr=D, lda=D, ldb=D, ldo=C,
                               alpha=c2))
            l2d.append(GlinJob(a_off=B * o, b=W2, b_red_is_row=1, out_off=hb, ni=B, nj=D, nr=C, lda=C, ldb=D, ldo=D,
                               alpha=c2))
            l2w.append(GlinJob(a_off=B * o, a_red_is_row=1, b_off=hb, b_red_is_row=1, out_off=o * D, ni=C, nj=D, nr=B,
                               lda=C, ldb=D, ldo=D, alpha=c2))
            l1d.append(GlinJob(a_off=hb, b=W1, b_red_is_row=1, out_off=hb, ni=B, nj=D, nr=D, lda=D, ldb=D, ldo=D, alpha=c1))
            l1w.append(GlinJob(a_off=hb, a_red_is_row=1, b_off=0, b_red_is_row=1, out_off=j * D * D, ni=D, nj=D, nr=B,
                               lda=D, ldb=D, ldo=D, alpha=c1))
            bn1.append(GbnJob(gamma=n1.weight, beta=n1.bias, slope=act1.weight, running_mean=n1.running_mean,
                              running_var=n1.running_var, nbt=n1.num_batches_tracked, x_off=hb, stat_off=j * D, C=D))
            bn2.append(GbnJob(gamma=n2.weight, beta=n2.bias, running_mean=n2.running_mean, running_var=n2.running_var,
                              nbt=n2.num_batches_tracked, x_off=B * o, stat_off=o, C=C))
        self.l1_fwd, self.l2_fwd, self.l2_dgrad, self.l2_wgrad = Plan(l1f), Plan(l2f), Plan(l2d), Plan(l2w)
        self.l1_dgrad, self.l1_wgrad, self.bn1, self.bn2 = Plan(l1d), Plan(l1w), Plan(bn1), Plan(bn2)
        self.last_gS = None
        n1, n2 = mods[0][0].net[1], mods[0][2]
        self.eps1, self.mom1, self.eps2, self.mom2 = n1.eps, n1.momentum, n2.eps, n2.momentum


def _flat_params(mods):
    out = []
    for m in mods:
        lin1, n1, act1 = m[0].net[0], m[0].net[1], m[0].net[2]
        out += [lin1.weight.raw, lin1.bias, n1.weight, n1.bias, act1.weight, m[1].weight.raw, m[1].bias, m[2].weight,
                m[2].bias]
    return out


class GroupedStyles:
    """All `to_style(w)` of a generator in grouped launches.  `eligible()` checks the structure the kernels assume."""

    def __init__(self, to_styles):
        self.mods = list(to_styles)
        self._plans = {}
        self._sig = None

    def eligible(self):
        from . import layers as L
        if not self.mods:
            return False
        eps, mom = set(), set()
        for m in self.mods:
            try:
                net = m[0].net
                ok = (len(m) == 3 and len(net) == 3 and isinstance(net[1], L.BatchNorm1d) and isinstance(net[2], L.PReLU)
                      and isinstance(m[2], L.BatchNorm1d) and hasattr(net[0], "weight") and hasattr(m[1], "weight"))
            except Exception:
                return False
            if not ok:
                return False
            for bn in (net[1], m[2]):
                if not (bn.training and bn.track_running_stats and bn.affine and bn.momentum is not None):
                    return False
            if net[2].weight.numel() != net[1].num_features:
                return False
            for sub in (m, m[0], net, net[0], net[1], net[2], m[1], m[2]):
                if sub._forward_hooks or sub._forward_pre_hooks or sub._backward_hooks:
                    return False   # somebody observes per-layer activations: run layer by layer
            eps.add((net[1].eps, m[2].eps))
            mom.add((net[1].momentum, m[2].momentum))
        return len(eps) == 1 and len(mom) == 1

    def plans(self, B):
        params = _flat_params(self.mods)
        sig = (B, params[0].device, tuple(p.data_ptr() for p in params))
        if self._sig != sig:   # first call, another batch size, or the parameters moved (.to(device), new storage)
            self._plans = StylePlans(self.mods, B)
            self._sig = sig
        return self._plans, params

    def __call__(self, w):
        plans, params = self.plans(w.shape[0])
        outs = StyleBankFn.apply(plans, w, *params)
        gS = plans.last_gS
        plans.last_gS = None
        if gS is not None:
            B = plans.B
            for j, s in enumerate(outs):
                # the consumer's backward writes d(loss)/d(style) here (ops.StyleLayerFn.backward): no gather later
                s._gb_grad_out = gS[B * plans.off[j]:B * plans.off[j + 1]].view(B, plans.C[j])
        return outs


class StyleBankFn(Function):
    @staticmethod
    def forward(ctx, plans, w, *params):
        k = ops.K()
        w = ops._c(w)
        B, D, L = plans.B, plans.D, plans.L
        dev = w.device
        hooks = ops._sync_bn_hooks
        H1 = torch.empty(L * B * D, dtype=w.dtype, device=dev)
        k.linear_group(plans.l1_fwd, w, None, H1)
        mean1, var1 = k.bn1d_group_stats(plans.bn1, H1, B, L * D)
        count = B
        if hooks["stats"] is not None:
            mean1, var1, count = hooks["stats"](mean1, var1, B)
        H2, invstd1 = k.bn1d_group_apply(plans.bn1, H1, B, mean1, var1, plans.eps1, plans.mom1, count)
        H3 = torch.empty(B * plans.sumC, dtype=w.dtype, device=dev)
        k.linear_group(plans.l2_fwd, H2, None, H3)
        mean2, var2 = k.bn1d_group_stats(plans.bn2, H3, B, plans.sumC)
        if hooks["stats"] is not None:
            mean2, var2, _ = hooks["stats"](mean2, var2, B)
        S, invstd2 = k.bn1d_group_apply(plans.bn2, H3, B, mean2, var2, plans.eps2, plans.mom2, count)
        ctx.save_for_backward(w, H1, H2, H3, mean1, invstd1, mean2, invstd2)
        ctx.plans, ctx.count = plans, count
        ctx.gS = None
        if any(ctx.needs_input_grad):
            ctx.gS = torch.empty_like(S)
        plans.last_gS = ctx.gS
        return tuple(S[B * plans.off[j]:B * plans.off[j + 1]].view(B, plans.C[j]) for j in range(L))

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, *gouts):
        w, H1, H2, H3, mean1, invstd1, mean2, invstd2 = ctx.saved_tensors
        plans, count, gS = ctx.plans, ctx.count, ctx.gS
        k = ops.K()
        B, D, L = plans.B, plans.D, plans.L
        dev = w.device
        hooks = ops._sync_bn_hooks
        for j, g in enumerate(gouts):
            seg = gS[B * plans.off[j]:B * plans.off[j + 1]].view(B, plans.C[j])
            if g is None:
                seg.zero_()
            elif g.data_ptr() != seg.data_ptr():      # a consumer that did not use the segment (unfused path)
                k.chan_copy(ops._c(g), 0, seg, 0, plans.C[j])
        s1_2, s2_2, _ = k.bn1d_group_bwd_reduce(plans.bn2, gS, H3, B, mean2, invstd2, plans.sumC, False)
        t1, t2 = hooks["grads"](s1_2, s2_2) if hooks["grads"] is not None else (s1_2, s2_2)
        gH3, gb2 = k.bn1d_group_bwd_apply(plans.bn2, gS, H3, B, mean2, invstd2, t1, t2, count, plans.sumC)
        gH2 = torch.empty(L * B * D, dtype=w.dtype, device=dev)
        k.linear_group(plans.l2_dgrad, gH3, None, gH2)
        gW2 = torch.empty(plans.sumC * D, dtype=w.dtype, device=dev)
        k.linear_group(plans.l2_wgrad, gH3, H2, gW2)
        s1_1, s2_1, gsl1 = k.bn1d_group_bwd_reduce(plans.bn1, gH2, H1, B, mean1, invstd1, L * D, True)
        t1, t2 = hooks["grads"](s1_1, s2_1) if hooks["grads"] is not None else (s1_1, s2_1)
        gH1, gb1 = k.bn1d_group_bwd_apply(plans.bn1, gH2, H1, B, mean1, invstd1, t1, t2, count, L * D)
        parts = torch.empty(L * B * D, dtype=w.dtype, device=dev)
        k.linear_group(plans.l1_dgrad, gH1, None, parts)
        gw = k.reduce(parts.view(L, B * D, 1), None, KN.R_NHW, 1.0).view(B, D)
        gW1 = torch.empty(L * D * D, dtype=w.dtype, device=dev)
        k.linear_group(plans.l1_wgrad, gH1, w, gW1)
        grads = [None, gw]
        for j in range(L):
            o, C = plans.off[j], plans.C[j]
            a, b = j * D, (j + 1) * D
            grads += [gW1[j * D * D:(j + 1) * D * D].view(D, D), gb1[a:b], s2_1[a:b], s1_1[a:b], gsl1[a:b],
                      gW2[o * D:(o + C) * D].view(C, D), gb2[o:o + C], s2_2[o:o + C], s1_2[o:o + C]]
        return tuple(grads)


# ---------------------------------------------------------------------------------------------------------------
# device tables (CUDA backend)
# ---------------------------------------------------------------------------------------------------------------
def _ptr(t):
    return t.data_ptr() if t is not None else None


def glin_table(plan, device):
    """-> (device uint8 tensor of ganb200_glin_job[], n_jobs, n_blocks, keep-alive list)"""
    from ._lib import GlinJobC
    hit = plan._dev.get(device)
    if hit is not None:
        return hit
    arr = (GlinJobC * len(plan.jobs))()
    blk = 0
    for i, j in enumerate(plan.jobs):
        c = arr[i]
        c.a, c.b, c.out, c.bias = _ptr(j.a), _ptr(j.b), None, _ptr(j.bias)
        c.a_off, c.b_off, c.out_off = j.a_off, j.b_off, j.out_off
        c.ni, c.nj, c.nr, c.lda, c.ldb, c.ldo = j.ni, j.nj, j.nr, j.lda, j.ldb, j.ldo
        c.a_red_is_row, c.b_red_is_row, c.alpha, c.first_block = j.a_red_is_row, j.b_red_is_row, j.alpha, blk
        blk += ((j.ni + 31) // 32) * ((j.nj + 31) // 32)
    host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
    hit = (host.to(device), len(plan.jobs), blk)
    plan._dev[device] = hit
    return hit


def gbn_table(plan, device):
    from ._lib import GbnJobC
    hit = plan._dev.get(device)
    if hit is not None:
        return hit
    arr = (GbnJobC * len(plan.jobs))()
    blk = 0
    for i, j in enumerate(plan.jobs):
        c = arr[i]
        c.gamma, c.beta, c.slope = _ptr(j.gamma), _ptr(j.beta), _ptr(j.slope)
        c.running_mean, c.running_var, c.nbt = _ptr(j.running_mean), _ptr(j.running_var), _ptr(j.nbt)
        c.x_off, c.stat_off, c.C, c.first_block = j.x_off, j.stat_off, j.C, blk
        blk += (j.C + 31) // 32
    host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
    hit = (host.to(device), len(plan.jobs), blk)
    plan._dev[device] = hit
    return hit
