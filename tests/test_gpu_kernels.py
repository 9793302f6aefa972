"""GPU parity: every libganb200 kernel, called through the C-ABI (ctypes), against the oracle
(oracle/kernels_ref.py — the CPU restatement in stock torch ops) on the same seeded inputs.

Tolerances (BASELINE.json north_star): rel. L2 <= 1e-5 for fp32 paths, <= 1e-2 for the bf16 tcgen05 engine.
"""
import ctypes
import os

import pytest
import torch

from gan_b200 import kernels as KN
from gan_b200 import resample as RS

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-5
BF16_TOL = 1e-2


@pytest.fixture(scope="module")
def cu():
    assert torch.cuda.is_available()
    return KN.CudaBackend()


@pytest.fixture(scope="module")
def ref():
    from oracle.kernels_ref import RefBackend
    return RefBackend()


def rel(a, b):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    den = b.norm().item()
    return (a - b).norm().item() / (den if den > 0 else 1.0)


def g(t):
    return None if t is None else t.cuda()


def rnd(*shape):
    return torch.randn(*shape)


CONV_CASES = [
    # N, C, H, W, K, R, stride, pad, pad_mode
    (4, 3, 16, 16, 64, 3, 1, 1, 1),      # stem
    (4, 64, 16, 16, 64, 3, 1, 1, 1),     # same-width 3x3
    (2, 64, 16, 16, 128, 3, 1, 1, 1),    # C -> 2C
    (2, 128, 18, 18, 128, 3, 2, 1, 1),   # stride-2 down conv
    (8, 1025, 4, 4, 1025, 3, 1, 1, 1),   # odd channel count
    (4, 64, 5, 5, 64, 3, 1, 0, 0),       # SE conv, pad 0
    (4, 48, 16, 16, 54, 5, 1, 2, 1),     # k5 modconv shape
    (4, 390, 8, 8, 96, 1, 1, 0, 1),      # 1x1 on wide concat
    (3, 7, 9, 11, 5, 3, 1, 1, 0),        # ragged, zero pad
    (4, 256, 1, 1, 4100, 1, 1, 0, 0),    # linear
    (4, 24, 16, 16, 32, 4, 2, 1, 0),     # the conv whose dgrad is ConvTranspose2d(k4, s2, p1)
    (4, 108, 16, 16, 3, 5, 1, 2, 1),     # ToRGB: K = 3
    (4, 64, 10, 10, 64, 3, 2, 1, 1),     # stride 2, replicate pad, even padded size
    (4, 40, 7, 9, 24, 3, 2, 1, 0),       # stride 2, zero pad, odd sizes
    (4, 64, 64, 64, 64, 3, 2, 1, 0),     # cfg4 critic: stride-2 3x3, zero pad, 64 -> 32
    (4, 32, 64, 64, 64, 3, 1, 1, 0),     # cfg4 critic: 3x3 zero pad at full resolution
    (4, 513, 4, 4, 512, 3, 1, 1, 0),     # cfg4 critic: conv after the batch-stddev channel
    (4, 512, 4, 4, 512, 4, 1, 0, 0),     # cfg4 critic: 4x4 valid conv -> 1x1
    (4, 48, 64, 64, 48, 3, 1, 1, 1),     # dominant G13_5 modconv shape (k3 @ 64x64)
    (2, 96, 64, 64, 96, 5, 1, 2, 1),     # dominant G13_5 modconv shape (k5 @ 64x64)
    (4, 96, 32, 32, 96, 5, 1, 2, 1),     # k5 @ 32x32
    (4, 192, 16, 16, 96, 3, 1, 1, 1),    # k3 @ 16x16
    (3, 72, 20, 36, 200, 3, 1, 1, 0),    # halo kernel: ragged plane, two output-channel tiles, zero pad
    (2, 130, 33, 17, 70, 5, 1, 2, 1),    # halo kernel: odd plane, three channel chunks, k5
    (2, 64, 64, 64, 64, 3, 1, 1, 1),     # flat kernel: the critic's full-resolution 3x3
    (2, 130, 33, 40, 70, 5, 1, 2, 1),    # flat kernel: ragged plane, three channel chunks, K = 70 in a 128-wide tile
    (1, 48, 40, 36, 200, 3, 1, 1, 1),    # flat kernel: two output-channel tiles
    (2, 390, 64, 64, 96, 1, 1, 0, 1),    # flat kernel: 1x1 on the wide concat at full resolution
    (2, 24, 40, 24, 16, 7, 1, 3, 1),     # halo kernel: k7
    (2, 64, 32, 32, 32, 4, 2, 1, 0),     # ConvTranspose2d(k4, s2, p1) as dgrad: 2x2 tap phases at 32x32 (halo eligible)
    (3, 40, 21, 64, 72, 5, 1, 2, 1),     # row-ring wgrad: 63 rows over 3 images (ranges cross image boundaries), K > 64
    (5, 130, 9, 32, 200, 3, 1, 1, 0),    # row-ring wgrad: 32-pixel rows, zero pad, three channel tiles, two filter tiles
    (2, 16, 70, 32, 16, 7, 1, 3, 1),     # row-ring wgrad: k7 (seven accumulators)
    (2, 20, 24, 24, 3, 9, 1, 0, 0),      # tree generators' LastConv: 9x9 on the reflect-padded map (CUDA-core engine only)
    (2, 16, 36, 36, 16, 5, 2, 0, 0),     # tree critic's down-sampling conv: 5x5 stride 2 on the reflect-padded map
]


@pytest.mark.parametrize("case", CONV_CASES)
@pytest.mark.parametrize("scales", [False, True])
def test_conv_family_fp32(cu, ref, case, scales):
    N, C, H, W, K, R, stride, pad, pm = case
    torch.manual_seed(1)
    x, w, b = rnd(N, C, H, W), rnd(K, C, R, R), rnd(K)
    isc = torch.rand(N, C) + 0.5 if scales else None
    osc = torch.rand(N, K) + 0.5 if scales else None
    y_ref = ref.conv_fprop(x, w, b, stride, pad, pm, 0.1, isc, osc)
    y = cu.conv_fprop(g(x), g(w), g(b), stride, pad, pm, 0.1, g(isc), g(osc), engine=1)
    assert rel(y, y_ref) < FP32_TOL
    dy = rnd(*y_ref.shape)
    dx_ref = ref.conv_dgrad(dy, w, (H, W), stride, pad, pm, 0.1, isc, osc)
    dx = cu.conv_dgrad(g(dy), g(w), (H, W), stride, pad, pm, 0.1, g(isc), g(osc), engine=1)
    assert rel(dx, dx_ref) < FP32_TOL
    dw_ref = ref.conv_wgrad(x, dy, (R, R), stride, pad, pm, 0.1, isc, osc)
    dw = cu.conv_wgrad(g(x), g(dy), (R, R), stride, pad, pm, 0.1, g(isc), g(osc), engine=1)
    assert rel(dw, dw_ref) < FP32_TOL


@pytest.mark.parametrize("case", CONV_CASES)
@pytest.mark.parametrize("scales", [False, True])
def test_conv_family_tcgen05(cu, ref, case, scales):
    """bf16 operands, fp32 accumulate in TMEM; only the (shape, op) pairs the engine declares eligible."""
    N, C, H, W, K, R, stride, pad, pm = case
    torch.manual_seed(2)
    x, w, b = rnd(N, C, H, W), rnd(K, C, R, R), rnd(K)
    isc = torch.rand(N, C) + 0.5 if scales else None
    osc = torch.rand(N, K) + 0.5 if scales else None
    P, Q = KN.out_hw(H, W, R, R, stride, pad)
    d = KN.ConvDesc(N, C, H, W, K, R, R, stride, pad, pm, P, Q, 0.1, 0)
    dy = rnd(N, K, P, Q)
    ran = 0
    if cu.lib.ganb200_conv2d_tc_eligible(ctypes.byref(d), KN.FPROP):
        y = cu.conv_fprop(g(x), g(w), g(b), stride, pad, pm, 0.1, g(isc), g(osc), engine=2)
        assert rel(y, ref.conv_fprop(x, w, b, stride, pad, pm, 0.1, isc, osc)) < BF16_TOL
        ran += 1
    if cu.lib.ganb200_conv2d_tc_eligible(ctypes.byref(d), KN.DGRAD):
        dx = cu.conv_dgrad(g(dy), g(w), (H, W), stride, pad, pm, 0.1, g(isc), g(osc), engine=2)
        assert rel(dx, ref.conv_dgrad(dy, w, (H, W), stride, pad, pm, 0.1, isc, osc)) < BF16_TOL
        ran += 1
    if cu.lib.ganb200_conv2d_tc_eligible(ctypes.byref(d), KN.WGRAD):
        dw = cu.conv_wgrad(g(x), g(dy), (R, R), stride, pad, pm, 0.1, g(isc), g(osc), engine=2)
        assert rel(dw, ref.conv_wgrad(x, dy, (R, R), stride, pad, pm, 0.1, isc, osc)) < BF16_TOL
        ran += 1
    if ran == 0:
        pytest.skip("shape not eligible for the tcgen05 engine")


@pytest.mark.parametrize("N,C,K", [(256, 256, 256), (64, 256, 48), (5, 7, 3), (33, 1590, 97), (4, 2048, 2048)])
@pytest.mark.parametrize("scales", [False, True])
def test_linear_one_launch_kernels(cu, ref, N, C, K, scales):
    """engine 0 routes nn.Linear shapes (1x1 filter, 1x1 map) to the fp32 one-launch kernels: fp32 tolerance"""
    torch.manual_seed(9)
    x, w, b, dy = rnd(N, C, 1, 1), rnd(K, C, 1, 1), rnd(K), rnd(N, K, 1, 1)
    isc = torch.rand(N, C) + 0.5 if scales else None
    osc = torch.rand(N, K) + 0.5 if scales else None
    d = KN.ConvDesc(N, C, 1, 1, K, 1, 1, 1, 0, 0, 1, 1, 0.3, 0)
    assert cu.lib.ganb200_conv2d_route(ctypes.byref(d), KN.FPROP) == 3
    assert rel(cu.conv_fprop(g(x), g(w), g(b), 1, 0, 0, 0.3, g(isc), g(osc), engine=0),
               ref.conv_fprop(x, w, b, 1, 0, 0, 0.3, isc, osc)) < FP32_TOL
    assert rel(cu.conv_dgrad(g(dy), g(w), (1, 1), 1, 0, 0, 0.3, g(isc), g(osc), engine=0),
               ref.conv_dgrad(dy, w, (1, 1), 1, 0, 0, 0.3, isc, osc)) < FP32_TOL
    assert rel(cu.conv_wgrad(g(x), g(dy), (1, 1), 1, 0, 0, 0.3, g(isc), g(osc), engine=0),
               ref.conv_wgrad(x, dy, (1, 1), 1, 0, 0, 0.3, isc, osc)) < FP32_TOL


def bf16r(t):
    return t.to(torch.bfloat16).to(torch.float32)


@pytest.fixture(params=[(1, 2), (1, 0), (0, 0)], ids=["flat", "halo", "per-tap"])
def halo_mode(cu, request):
    """which tcgen05 forward kernel serves the problems that qualify for more than one: (halo mode, flat mode)"""
    prev = cu.lib.ganb200_conv_halo_mode(request.param[0])
    prev_flat = cu.lib.ganb200_conv_flat_mode(request.param[1])
    yield request.param
    cu.lib.ganb200_conv_halo_mode(prev)
    cu.lib.ganb200_conv_flat_mode(prev_flat)


def test_tcgen05_operand_start_may_be_shifted_by_single_rows(cu):
    """the flat convolution kernel reads every filter tap from one shared-memory box through an A descriptor whose start is
    shifted by whole 128-byte rows — not multiples of the 8-row SWIZZLE_128B atom.  Hardware check (csrc/probe.cu)."""
    torch.manual_seed(0)
    rows, n = 256, 64
    a = torch.randn(rows, 64).to(torch.bfloat16)
    b = torch.randn(n, 64).to(torch.bfloat16)
    ag, bg = a.cuda(), b.cuda()
    out = torch.empty(128, n, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for shift in (0, 8, 1, 3, 7, 9, 66, 69, 127, 128):
        out.zero_()
        rc = cu.lib.ganb200_tc_shift_probe(ctypes.c_void_p(ag.data_ptr()), ctypes.c_void_p(bg.data_ptr()),
                                           ctypes.c_void_p(out.data_ptr()), rows, n, shift, 0, st)
        assert rc == 0
        want = a[shift:shift + 128].float() @ b.float().t()
        assert rel(out, want) < 1e-5, shift


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_family_tcgen05_exact_on_rounded_operands(cu, ref, case, halo_mode):
    """The engine rounds (x*iscale), (dy*oscale) and w to bf16 and accumulates in fp32: against the oracle fed the SAME
    rounded operands only the summation order differs, so any indexing / border / tap error shows (the 1e-2 test above
    would hide a wrong border pixel)."""
    N, C, H, W, K, R, stride, pad, pm = case
    torch.manual_seed(7)
    x, w, b = rnd(N, C, H, W), rnd(K, C, R, R), rnd(K)
    isc, osc = torch.rand(N, C) + 0.5, torch.rand(N, K) + 0.5
    P, Q = KN.out_hw(H, W, R, R, stride, pad)
    d = KN.ConvDesc(N, C, H, W, K, R, R, stride, pad, pm, P, Q, 0.1, 0)
    dy = rnd(N, K, P, Q)
    xr, wr = bf16r(x * isc[:, :, None, None]), bf16r(w)
    dyr = bf16r(dy * osc[:, :, None, None])
    tol, ran = 2e-4, 0
    if cu.lib.ganb200_conv2d_tc_eligible(ctypes.byref(d), KN.FPROP):
        y = cu.conv_fprop(g(x), g(w), g(b), stride, pad, pm, 0.1, g(isc), g(osc), engine=2)
        assert rel(y, ref.conv_fprop(xr, wr, b, stride, pad, pm, 0.1, None, osc)) < tol
        ran += 1
    if cu.lib.ganb200_conv2d_tc_eligible(ctypes.byref(d), KN.DGRAD):
        dx = cu.conv_dgrad(g(dy), g(w), (H, W), stride, pad, pm, 0.1, g(isc), g(osc), engine=2)
        assert rel(dx, ref.conv_dgrad(dyr, wr, (H, W), stride, pad, pm, 0.1, isc, None)) < tol
        ran += 1
    if cu.lib.ganb200_conv2d_tc_eligible(ctypes.byref(d), KN.WGRAD):
        dw = cu.conv_wgrad(g(x), g(dy), (R, R), stride, pad, pm, 0.1, g(isc), g(osc), engine=2)
        assert rel(dw, ref.conv_wgrad(xr, dyr, (R, R), stride, pad, pm, 0.1, None, None)) < tol
        ran += 1
    if ran == 0:
        pytest.skip("shape not eligible for the tcgen05 engine")


def test_tcgen05_engine_covers_the_dominant_shapes(cu):
    """The same-width 3x3 stride-1 convs of D9_4 (66 % of its FLOPs) must be tensor-core eligible."""
    for (N, C, HW) in [(8, 64, 64), (8, 128, 32), (8, 256, 16), (8, 512, 8), (8, 1025, 4)]:
        d = KN.ConvDesc(N, C, HW, HW, C, 3, 3, 1, 1, 1, HW, HW, 1.0, 0)
        assert cu.lib.ganb200_conv2d_tc_eligible(ctypes.byref(d), KN.FPROP) == 1, (N, C, HW)


@pytest.mark.parametrize("n", [64, 32, 5, 4])
def test_resample_ops(cu, ref, n):
    torch.manual_seed(3)
    x = rnd(3, 5, n, n)
    ops = [RS.smooth3(n), RS.adaptive_avg(n, 5) if n >= 5 else RS.identity(n), RS.replicate_pad(n, 2)]
    if n % 2 == 0:
        ops += [RS.smooth_then_bicubic_down2(n), RS.bicubic(n, 2 * n, 0.5), RS.bicubic_up2_then_smooth(n)]
    for op in ops:
        for o in (op, op.T):
            xi = rnd(3, 5, o.n_in, o.n_in)
            assert rel(cu.resample2d(g(xi), o, o), ref.resample2d(xi, o, o)) < FP32_TOL, o.name
    del x


def test_pointwise(cu, ref):
    torch.manual_seed(4)
    x = rnd(6, 10, 7, 9)
    xp = x.abs() + 0.1
    for op, inp, p0, p1 in [(KN.U_SIGMOID, x, 0, 0), (KN.U_TANH, x, 0, 0), (KN.U_SQUARE, x, 0, 0),
                            (KN.U_AFFINE, x, 1.7, -0.3), (KN.U_RSQRT_EPS, xp, 1e-8, 0), (KN.U_SQRT, xp, 0, 0),
                            (KN.U_LEAKY, x, 0.2, 0), (KN.U_RECIP, xp, 0, 0), (KN.U_LEAKY_MASK, x, 0.1, 0),
                            (KN.U_COPY, x, 0, 0), (KN.U_POW, xp, 1.5, 0),
                            (KN.U_GELU, x * 3, 0, 0), (KN.U_GELU_GRAD, x * 3, 0, 0), (KN.U_GELU_GRAD2, x * 3, 0, 0)]:
        assert rel(cu.ew_unary(op, g(inp), p0, p1), ref.ew_unary(op, inp, p0, p1)) < FP32_TOL, op
    a, b = rnd(6, 10, 7, 9), torch.rand(6, 10, 7, 9) * 0.8 + 0.1
    for op, p0 in [(KN.B_ADD, 0), (KN.B_SUB, 0), (KN.B_MUL, 0), (KN.B_DIV, 0), (KN.B_SIGMOID_BWD, 0),
                   (KN.B_TANH_BWD, 0), (KN.B_LEAKY_BWD, 0.2)]:
        bb = b if op != KN.B_LEAKY_BWD else x
        assert rel(cu.ew_binary(op, g(a), g(bb), p0), ref.ew_binary(op, a, bb, p0)) < FP32_TOL, op
    N, C = 6, 10
    for vmode, v in [(KN.V_NC, rnd(N, C)), (KN.V_C, rnd(C)), (KN.V_N, rnd(N)), (KN.V_SCALAR, rnd(1))]:
        for op in (KN.B_ADD, KN.B_MUL):
            assert rel(cu.ew_bcast(op, g(x), g(v), vmode), ref.ew_bcast(op, x, v, vmode)) < FP32_TOL
        assert rel(cu.ew_fma_bcast(g(x), g(v), g(a), vmode), ref.ew_fma_bcast(x, v, a, vmode)) < FP32_TOL
    for rmode in (KN.R_HW, KN.R_NHW, KN.R_CHW, KN.R_ALL):
        for y in (None, a):
            assert rel(cu.reduce(g(x), g(y), rmode, 0.5), ref.reduce(x, y, rmode, 0.5)) < FP32_TOL
    big = rnd(16, 3, 64, 64)
    assert rel(cu.reduce(g(big), g(big), KN.R_CHW, 1.0), ref.reduce(big, big, KN.R_CHW, 1.0)) < FP32_TOL


@pytest.mark.parametrize("an", [1, 12])
def test_prelu_family(cu, ref, an):
    torch.manual_seed(5)
    x, gy, u = rnd(5, 12, 6, 6), rnd(5, 12, 6, 6), rnd(5, 12, 6, 6)
    a = torch.rand(an) * 0.5
    assert rel(cu.prelu_fwd(g(x), g(a)), ref.prelu_fwd(x, a)) < FP32_TOL
    assert rel(cu.maskscale(g(gy), g(x), g(a)), ref.maskscale(gy, x, a)) < FP32_TOL
    gx, ga = cu.prelu_bwd(g(gy), g(x), g(a))
    gx_r, ga_r = ref.prelu_bwd(gy, x, a)
    assert rel(gx, gx_r) < FP32_TOL and rel(ga, ga_r) < FP32_TOL
    assert rel(cu.masked_dot(g(u), g(gy), g(x), an), ref.masked_dot(u, gy, x, an)) < FP32_TOL


@pytest.mark.parametrize("shape", [(16, 20, 8, 8), (32, 256), (8, 3, 64, 64)])
def test_batchnorm(cu, ref, shape):
    torch.manual_seed(6)
    x, gy = rnd(*shape) * 2 + 0.5, rnd(*shape)
    C = shape[1]
    gamma, beta = rnd(C), rnd(C)
    m, v = cu.bn_stats(g(x))
    m_r, v_r = ref.bn_stats(x)
    assert rel(m, m_r) < FP32_TOL and rel(v, v_r) < FP32_TOL
    rm, rv, nbt = torch.zeros(C), torch.ones(C), torch.zeros((), dtype=torch.int64)
    rm_g, rv_g, nbt_g = g(rm.clone()), g(rv.clone()), g(nbt.clone())
    n, _, hw = KN._nchw(x)
    y, inv = cu.bn_apply(g(x), m, v, g(gamma), g(beta), rm_g, rv_g, 1e-5, 0.1, n * hw, nbt_g)
    y_r, inv_r = ref.bn_apply(x, m_r, v_r, gamma, beta, rm, rv, 1e-5, 0.1, n * hw, nbt)
    assert rel(y, y_r) < FP32_TOL and rel(inv, inv_r) < FP32_TOL
    assert rel(rm_g, rm) < FP32_TOL and rel(rv_g, rv) < FP32_TOL and int(nbt_g) == int(nbt) == 1
    s1, s2 = cu.bn_bwd_reduce(g(gy), g(x), m, inv)
    s1_r, s2_r = ref.bn_bwd_reduce(gy, x, m_r, inv_r)
    assert rel(s1, s1_r) < 1e-4 and rel(s2, s2_r) < 1e-4
    out = cu.bn_bwd_apply(g(gy), g(x), m, inv, g(gamma), s1, s2, n * hw)
    out_r = ref.bn_bwd_apply(gy, x, m_r, inv_r, gamma, s1_r, s2_r, n * hw)
    for a, b in zip(out, out_r):
        assert rel(a, b) < 1e-4


@pytest.mark.parametrize("shape", [(256, 256), (4, 48), (256, 12, 4, 4), (7, 33, 5, 5), (1024, 3, 2, 2), (6, 5, 1, 1),
                                   (256, 12, 5, 5), (8192, 2)])
@pytest.mark.parametrize("affine", [True, False])
def test_batchnorm_fused_small(cu, ref, shape, affine):
    """one-launch training-mode batch norm (N*HW <= 8192; 256-thread variant above 4096) against the oracle's stats + apply / reduce + apply"""
    torch.manual_seed(16)
    x, gy = rnd(*shape) * 2 + 0.5, rnd(*shape)
    C = shape[1]
    gamma, beta = (rnd(C), rnd(C)) if affine else (None, None)
    assert cu.bn_small_ok(g(x)) and ref.bn_small_ok(x)
    rm, rv, nbt = rnd(C), torch.rand(C) + 0.5, torch.full((), 3, dtype=torch.int64)
    rm_g, rv_g, nbt_g = g(rm.clone()), g(rv.clone()), g(nbt.clone())
    y, m, inv = cu.bn_small_fwd(g(x), g(gamma), g(beta), rm_g, rv_g, 1e-5, 0.1, nbt_g)
    y_r, m_r, inv_r = ref.bn_small_fwd(x, gamma, beta, rm, rv, 1e-5, 0.1, nbt)
    assert rel(y, y_r) < FP32_TOL and rel(m, m_r) < FP32_TOL and rel(inv, inv_r) < FP32_TOL
    assert rel(rm_g, rm) < FP32_TOL and rel(rv_g, rv) < FP32_TOL and int(nbt_g) == int(nbt) == 4
    out = cu.bn_small_bwd(g(gy), g(x), m, inv, g(gamma))
    out_r = ref.bn_small_bwd(gy, x, m_r, inv_r, gamma)
    for a, b in zip(out, out_r):
        assert rel(a, b) < 1e-4
    big = rnd(64, 3, 16, 16)
    assert not cu.bn_small_ok(g(big)) and not ref.bn_small_ok(big)
    if affine:  # with the following nn.PReLU(C) fused in
        slope = torch.rand(C) * 0.5
        rm_g, rv_g, nbt_g = g(rm.clone()), g(rv.clone()), g(nbt.clone())
        y, m, inv = cu.bn_small_fwd(g(x), g(gamma), g(beta), rm_g, rv_g, 1e-5, 0.1, nbt_g, g(slope))
        y_r, m_r, inv_r = ref.bn_small_fwd(x, gamma, beta, rm.clone(), rv.clone(), 1e-5, 0.1, nbt.clone(), slope)
        assert rel(y, y_r) < FP32_TOL and rel(m, m_r) < FP32_TOL and rel(inv, inv_r) < FP32_TOL
        out = cu.bn_small_bwd(g(gy), g(x), m, inv, g(gamma), g(beta), g(slope))
        out_r = ref.bn_small_bwd(gy, x, m_r, inv_r, gamma, beta, slope)
        assert len(out) == len(out_r) == 4
        for a, b in zip(out, out_r):
            assert rel(a, b) < 1e-4


@pytest.mark.parametrize("shape", [(4, 48, 64, 64), (3, 7, 5, 5), (2, 5, 3, 3), (5, 6, 1, 1), (2, 9, 34, 30)])
def test_sk_merge_two_branches_fused(cu, ref, shape):
    """att0 * t0 + att1 * t1 and its adjoints in one launch each (ganb200_sk_merge2_fwd / _bwd) against the oracle, and
    the layer-level path against the mul_bc + fma_bc chain it replaces"""
    torch.manual_seed(21)
    t0, t1, gy = rnd(*shape), rnd(*shape), rnd(*shape)
    att = torch.softmax(rnd(shape[0], 2, shape[1]), dim=1).contiguous()
    out = cu.sk_merge2_fwd(g(t0), g(t1), g(att))
    assert rel(out, ref.sk_merge2_fwd(t0, t1, att)) < FP32_TOL
    got = cu.sk_merge2_bwd(g(gy), g(t0), g(t1), g(att))
    want = ref.sk_merge2_bwd(gy, t0, t1, att)
    for a, b in zip(got, want):
        assert a.shape == b.shape and rel(a, b) < 1e-5
    from gan_b200 import layers as L
    grads = {}
    for fused in ("1", "0"):
        os.environ["GANB200_SK_MERGE_FUSED"] = fused
        leaves = [t.clone().cuda().requires_grad_() for t in (t0, t1, att)]
        y = L.sk_merge([leaves[0], leaves[1]], leaves[2].view(shape[0], 2, shape[1], 1, 1))
        y.backward(g(gy))
        grads[fused] = [y.detach()] + [t.grad for t in leaves]
    os.environ.pop("GANB200_SK_MERGE_FUSED")
    for a, b in zip(grads["1"], grads["0"]):
        assert rel(a, b) < 1e-5


@pytest.mark.parametrize("shape", [(4, 48, 64, 64), (3, 7, 32, 32), (1, 1, 64, 64)])
def test_pooled_branch_sum(cu, ref, shape):
    """AdaptiveAvgPool2d(5)(t0 + t1) in one pass (ganb200_pool_sum2) against torch's pooling of the sum, and the op's
    gradient against the add + pooling chain"""
    torch.manual_seed(22)
    t0, t1 = rnd(*shape), rnd(*shape)
    assert cu.pool_sum2_ok(g(t0), g(t1), 5) and ref.pool_sum2_ok(t0, t1, 5)
    assert not cu.pool_sum2_ok(g(rnd(2, 3, 16, 16)), g(rnd(2, 3, 16, 16)), 5)
    assert rel(cu.pool_sum2(g(t0), g(t1), 5), ref.pool_sum2(t0, t1, 5)) < FP32_TOL
    from gan_b200 import ops
    a, b = g(t0).requires_grad_(), g(t1).requires_grad_()
    gy = g(rnd(shape[0], shape[1], 5, 5))
    ops.pooled_branch_sum([a, b], 5).backward(gy)
    a2, b2 = g(t0).requires_grad_(), g(t1).requires_grad_()
    ops.adaptive_avg_pool(ops.add(a2, b2), 5).backward(gy)
    assert rel(a.grad, a2.grad) < FP32_TOL and rel(b.grad, b2.grad) < FP32_TOL


@pytest.mark.parametrize("shape", [(4, 48, 64, 64), (3, 7, 32, 32)])
def test_sk_block_backward_in_one_pass(cu, ref, shape):
    """ganb200_sk_bwd_finish (att_i * g + adjoint of the 5x5 pooling, both branches) against the oracle, and a whole SK
    block — pooled sum -> attention weights -> merge — with the linked backward against the unlinked one"""
    torch.manual_seed(23)
    N, C = shape[:2]
    gy, gp = rnd(*shape), rnd(N, C, 5, 5)
    att = torch.softmax(rnd(N, 2, C), dim=1).contiguous()
    got = cu.sk_bwd_finish(g(gy), g(att), g(gp))
    want = ref.sk_bwd_finish(gy, att, gp)
    for a, b in zip(got, want):
        assert rel(a, b) < FP32_TOL
    _, _, gatt = cu.sk_merge2_bwd(g(gy), g(rnd(*shape)), g(rnd(*shape)), g(att), att_only=True)
    assert gatt.shape == (N, 2, C)
    from gan_b200 import layers as L
    from gan_b200 import ops
    t0, t1, wmix = rnd(*shape), rnd(*shape), rnd(2, 25)
    grads = {}
    for linked in ("1", "0"):
        os.environ["GANB200_SK_BWD_FUSED"] = linked
        a, b, wm = g(t0).requires_grad_(), g(t1).requires_grad_(), g(wmix).requires_grad_()
        branches = [a, b]
        pooled = ops.pooled_branch_sum(branches, 5)                       # [N, C, 5, 5]
        logits = torch.einsum("ncp,mp->nmc", pooled.reshape(N, C, 25), wm)  # a stand-in for the attention network
        y = L.sk_merge(branches, torch.softmax(logits, dim=1).reshape(N, 2, C, 1, 1))
        y.backward(g(gy))
        grads[linked] = [y.detach(), a.grad, b.grad, wm.grad]
    os.environ.pop("GANB200_SK_BWD_FUSED")
    for u, v in zip(grads["1"], grads["0"]):
        assert rel(u, v) < 1e-5


@pytest.mark.parametrize("shape", [(3, 70, 45, 33), (2, 4096, 32, 130), (1, 5, 7, 1), (4, 64, 64, 64)])
def test_bmm_all_transposes(cu, ref, shape):
    """ganb200_bmm (strided batched fp32 product) against torch.bmm for the four transpose combinations"""
    torch.manual_seed(24)
    B, M, K, N = shape
    for ta in (False, True):
        for tb in (False, True):
            a = rnd(B, K, M) if ta else rnd(B, M, K)
            b = rnd(B, N, K) if tb else rnd(B, K, N)
            got = cu.bmm(g(a), g(b), ta, tb, 0.5)
            assert got.shape == (B, M, N)
            assert rel(got, ref.bmm(a, b, ta, tb, 0.5)) < FP32_TOL, (ta, tb)


@pytest.mark.parametrize("shape", [(2, 37, 4096), (3, 5, 1), (1, 300, 257)])
def test_softmax_rows(cu, ref, shape):
    torch.manual_seed(25)
    x, gy = rnd(*shape) * 3, rnd(*shape)
    y = cu.softmax_rows_fwd(g(x))
    y_r = ref.softmax_rows_fwd(x)
    assert rel(y, y_r) < FP32_TOL
    assert rel(cu.softmax_rows_bwd(g(gy), y), ref.softmax_rows_bwd(gy, y_r)) < 1e-5


def test_self_attention_double_backward(cu):
    """layers.self_attention (bmm -> row softmax -> bmm) against the reference's formulation in torch ops, through a
    gradient penalty (create_graph): discriminators/discriminator_9_3.py:70-77"""
    from gan_b200 import layers as L
    torch.manual_seed(26)
    B, C, H, W = 2, 6, 8, 8
    f, k, h = (rnd(B, C, H, W) for _ in range(3))

    def ref_attention(f, k, h):
        n = H * W
        s = torch.bmm(f.view(B, -1, n).permute(0, 2, 1), k.view(B, -1, n))
        return torch.bmm(h.view(B, -1, n), torch.softmax(s, dim=1)).view(B, -1, H, W)

    res = {}
    for tag, fn, dev in (("ours", L.self_attention, "cuda"), ("ref", ref_attention, "cpu")):
        ff, kk, hh = (t.clone().to(dev).requires_grad_() for t in (f, k, h))
        out = fn(ff, kk, hh)
        (gf,) = torch.autograd.grad((out * out).sum(), ff, create_graph=True)
        pen = (gf * gf).sum()
        res[tag] = [out.detach()] + list(torch.autograd.grad(pen, [ff, kk, hh]))
    for a, b in zip(res["ours"], res["ref"]):
        assert rel(a, b) < 1e-4


def test_softmax_mbstd_bce(cu, ref):
    torch.manual_seed(7)
    x, gy = rnd(6, 2, 48), rnd(6, 2, 48)
    y = cu.softmax_m_fwd(g(x))
    y_r = ref.softmax_m_fwd(x)
    assert rel(y, y_r) < FP32_TOL
    assert rel(cu.softmax_m_bwd(g(gy), y), ref.softmax_m_bwd(gy, y_r)) < FP32_TOL
    xs, gg, ggx = rnd(8, 64, 4, 4), rnd(1), rnd(8, 64, 4, 4)
    assert rel(cu.mbstd_fwd(g(xs), 4, 1e-8), ref.mbstd_fwd(xs, 4, 1e-8)) < FP32_TOL
    assert rel(cu.mbstd_bwd(g(gg), g(xs), 4, 1e-8), ref.mbstd_bwd(gg, xs, 4, 1e-8)) < FP32_TOL
    o1, o2 = cu.mbstd_bwd_bwd(g(gg), g(xs), g(ggx), 4, 1e-8)
    r1, r2 = ref.mbstd_bwd_bwd(gg, xs, ggx, 4, 1e-8)
    assert rel(o1, r1) < 1e-4 and rel(o2, r2) < 1e-4
    p, t = torch.rand(64, 1) * 0.98 + 0.01, torch.rand(64, 1)
    out, flag = cu.bce_fwd(g(p), g(t))
    out_r, flag_r = ref.bce_fwd(p, t)
    assert rel(out, out_r) < FP32_TOL and int(flag) == int(flag_r) == 0
    _, flag = cu.bce_fwd(g(p + 1.0), g(t))
    assert int(flag) == 1
    assert rel(cu.bce_bwd(g(gg), g(p), g(t)), ref.bce_bwd(gg, p, t)) < FP32_TOL


def test_misc(cu, ref):
    torch.manual_seed(8)
    src, dst = rnd(4, 10, 5, 5), rnd(4, 16, 5, 5)
    for acc in (False, True):
        want = ref.chan_copy(src, 2, dst.clone(), 3, 7, acc)
        got = cu.chan_copy(g(src), 2, g(dst.clone()), 3, 7, acc)
        assert rel(got, want) < FP32_TOL
    t, f = rnd(64, 3, 64, 64), rnd(64, 3, 64, 64)
    assert rel(cu.mixup(g(t), g(f), 0.3), ref.mixup(t, f, 0.3)) < FP32_TOL
    label = torch.rand(64, 1, dtype=torch.float64)
    got = cu.mixup_wgan_f64(g(label), g(t), g(f))
    want = ref.mixup_wgan_f64(label, t, f)
    assert got.dtype == torch.float64 and rel(got, want) < 1e-12
    x = rnd(1000)
    assert torch.equal(cu.clamp_(g(x.clone()), -0.5, 0.5).cpu(), x.clamp(-0.5, 0.5))
    w2d, u, v = rnd(40, 27), torch.nn.functional.normalize(rnd(40), dim=0), torch.nn.functional.normalize(rnd(27), dim=0)
    ug, vg = g(u.clone()), g(v.clone())
    s = cu.spectral_power_iter(g(w2d), ug, vg, 3, 1e-12)
    s_r = ref.spectral_power_iter(w2d, u, v, 3, 1e-12)
    assert rel(s, s_r) < 1e-4 and rel(ug, u) < 1e-4 and rel(vg, v) < 1e-4


@pytest.mark.parametrize("decoupled,wd,b1", [(False, 0.0, 0.0), (False, 0.0, 0.5), (True, 0.01, 0.5)])
def test_fused_adam_matches_torch(decoupled, wd, b1):
    """FusedAdam (one launch over all tensors) against torch.optim.Adam/AdamW on CPU, 3 steps, grad=None skipped."""
    from gan_b200 import optim
    torch.manual_seed(9)
    shapes = [(64, 64, 3, 3), (1025,), (7,), (4100, 33), (1,)]
    ps_ref = [torch.nn.Parameter(rnd(*s)) for s in shapes] + [torch.nn.Parameter(rnd(5))]
    ps = [torch.nn.Parameter(p.detach().clone().cuda()) for p in ps_ref]
    cls = torch.optim.AdamW if decoupled else torch.optim.Adam
    o_ref = cls(ps_ref, lr=4e-4, betas=(b1, 0.99), weight_decay=wd)
    o = optim.FusedAdam(ps, lr=4e-4, betas=(b1, 0.99), weight_decay=wd, decoupled=decoupled)
    for _ in range(3):
        for pr, p in zip(ps_ref[:-1], ps[:-1]):
            gr = rnd(*pr.shape)
            pr.grad, p.grad = gr, gr.cuda()
        o_ref.step()
        o.step()
    for pr, p in zip(ps_ref, ps):
        assert rel(p, pr) < 1e-6
    assert len(o.state[ps[-1]]) == 0  # grad is None -> no state


def test_philox_noise(cu):
    st = torch.tensor([987654321, 5], dtype=torch.int64, device="cuda")
    z = cu.noise_fill((8, 16, 64, 64), st, 3, torch.device("cuda"))
    assert abs(float(z.mean())) < 5e-3 and abs(float(z.var()) - 1.0) < 1e-2
    assert abs(float((z ** 4).mean()) - 3.0) < 0.1                      # Gaussian kurtosis
    assert torch.equal(z, cu.noise_fill((8, 16, 64, 64), st, 3, torch.device("cuda")))   # pure function of the key
    z2 = cu.noise_fill((8, 16, 64, 64), st, 4, torch.device("cuda"))    # another layer: uncorrelated
    assert abs(float((z * z2).mean())) < 5e-3
    st2 = torch.tensor([987654321, 6], dtype=torch.int64, device="cuda")  # next forward: uncorrelated
    assert abs(float((z * cu.noise_fill((8, 16, 64, 64), st2, 3, torch.device("cuda"))).mean())) < 5e-3
    flat = z.flatten()
    assert abs(float((flat[:-1] * flat[1:]).mean())) < 5e-3             # neighbours uncorrelated


@pytest.mark.parametrize("hw", [(4, 4), (64, 64)])
@pytest.mark.parametrize("noise,act", [(True, True), (False, True), (True, False)])
def test_style_tail(cu, hw, noise, act):
    torch.manual_seed(10)
    N, K = 6, 12
    st = torch.tensor([42, 3], dtype=torch.int64, device="cuda")
    y = rnd(N, K, *hw).cuda()
    ns = (torch.rand(K) * 0.3).cuda() if noise else None
    slope = (torch.rand(K) * 0.5).cuda() if act else None
    z = cu.noise_fill(tuple(y.shape), st, 7, y.device)
    u_want = y + ns.view(1, K, 1, 1) * z if noise else y
    a_want = torch.where(u_want > 0, u_want, slope.view(1, K, 1, 1) * u_want) if act else None
    u, a = cu.style_tail_fwd(y.clone(), ns, slope, st, 7)
    assert rel(u, u_want) < FP32_TOL
    if act:
        assert rel(a, a_want) < FP32_TOL
    ga = rnd(N, K, *hw).cuda()
    gu, Ad, Z, Pm = cu.style_tail_bwd(ga, u, ns, slope, st, 7)
    gu_want = ga * torch.where(u > 0, torch.ones_like(u), slope.view(1, K, 1, 1).expand_as(u)) if act else ga
    assert rel(gu, gu_want) < FP32_TOL
    assert rel(Ad, (gu_want * y).sum((2, 3))) < 1e-4
    if noise:
        assert rel(Z, (gu_want * z).sum((2, 3))) < 1e-4
    if act:
        assert rel(Pm, (ga * u.clamp(max=0)).sum((2, 3))) < 1e-4


def test_demod_and_scale_reduce(cu, ref):
    torch.manual_seed(11)
    B, C, K, RS = 8, 48, 54, 9
    s, wsq, Ad = rnd(B, C) + 1.5, torch.rand(K, C) * 9, rnd(B, K)
    w, gw, gs1 = rnd(K, C, 3, 3), rnd(K, C, 3, 3), rnd(B, C)
    sinv_r = ref.demod(s, wsq, 0.2, 1e-8)
    sinv = cu.demod(g(s), g(wsq), 0.2, 1e-8)
    assert rel(sinv, sinv_r) < FP32_TOL
    assert rel(cu.demod_bwd_s(g(gs1), g(s), g(Ad), sinv, g(wsq), 0.2), ref.demod_bwd_s(gs1, s, Ad, sinv_r, wsq, 0.2)) < FP32_TOL
    assert rel(cu.demod_bwd_s(None, g(s), g(Ad), sinv, g(wsq), 0.2), ref.demod_bwd_s(None, s, Ad, sinv_r, wsq, 0.2)) < FP32_TOL
    assert rel(cu.demod_bwd_w_(g(gw.clone()), g(w), g(Ad), sinv, g(s), 0.2),
               ref.demod_bwd_w_(gw.clone(), w, Ad, sinv_r, s, 0.2)) < FP32_TOL
    for hw in [(4, 4), (64, 64)]:
        dxs, x = rnd(B, C, *hw), rnd(B, C, *hw)
        gx, gsr = cu.scale_reduce(g(dxs), g(s), g(x))
        gx_r, gsr_r = ref.scale_reduce(dxs, s, x)
        assert rel(gx, gx_r) < FP32_TOL and rel(gsr, gsr_r) < 1e-4


@pytest.mark.parametrize("B,C,K", [(12, 1590, 396), (3, 2049, 17), (256, 780, 396)])
def test_demod_wide_inputs(cu, ref, B, C, K):
    """root 1x1 modconvs read the whole concat (780 ... 1590 channels): the staged s^2 tile exceeds 48 KB"""
    torch.manual_seed(12)
    s, wsq = rnd(B, C) + 1.5, torch.rand(K, C) * 9
    assert rel(cu.demod(g(s), g(wsq), 0.2, 1e-8), ref.demod(s, wsq, 0.2, 1e-8)) < FP32_TOL


def test_fused_style_layer_on_gpu_matches_unfused_ops(cu):
    """StyleLayerFn through libganb200 (fp32 engine and tcgen05 engine) against the unfused op chain on the same noise."""
    from gan_b200 import ops
    torch.manual_seed(12)
    B, C, Ko, H, k = 8, 48, 54, 16, 5
    st = torch.tensor([77, 2], dtype=torch.int64, device="cuda")
    for engine, tol in ((1, 1e-4), (2, BF16_TOL)):
        leaves = [rnd(B, C, H, H).cuda().requires_grad_(), rnd(Ko, C, k, k).cuda().requires_grad_(),
                  (rnd(B, C) + 1.5).cuda().requires_grad_(), (torch.rand(Ko) * 0.3).cuda().requires_grad_(),
                  (torch.rand(Ko) * 0.5).cuda().requires_grad_()]
        x, w, s, ns, slope = leaves
        out = ops.style_layer(x, w, s, ns, slope, 2, 0.03, True, 1e-8, engine, st, 5)
        y = ops.modulated_conv2d(x, w, s, 2, 0.03, True, 1e-8, engine=1)
        y = ops.prelu(ops.noise_add(y, ns, cu.noise_fill(tuple(y.shape), st, 5, y.device)), slope)
        assert rel(out, y) < tol
        gout = torch.randn_like(out)
        for a, b in zip(torch.autograd.grad(out, leaves, gout), torch.autograd.grad(y, leaves, gout)):
            assert rel(a, b) < max(tol, 1e-4) * 3


# ---------------------------------------------------------------------------------------------------------------------
# kernels of the other configs: depthwise 3x3 family, channel shuffle, spectral norm, batch-wide stddev channel
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", [(4, 32, 64, 64, 2), (4, 244, 32, 32, 1), (3, 122, 16, 16, 1), (5, 976, 4, 4, 2),
                                  (2, 7, 9, 11, 2), (2, 7, 9, 11, 1), (64, 16, 1, 1, 1),
                                  # discriminator_6_wgangp.py:12 — 5x5 / 7x7 on a map the caller has reflect-padded (pad 0)
                                  (4, 6, 36, 36, 1, 5, 0), (4, 12, 22, 22, 1, 7, 0), (4, 16, 10, 10, 1, 3, 0),
                                  (2, 5, 13, 9, 2, 5, 2), (2, 5, 13, 9, 1, 7, 3), (3, 24, 8, 8, 2, 7, 4), (2, 3, 7, 7, 1, 7, 0)])
def test_depthwise_conv_family(cu, ref, case):
    N, C, H, W, stride = case[:5]
    k, pad = case[5:] if len(case) > 5 else (3, 1)
    torch.manual_seed(11)
    x, w, b = rnd(N, C, H, W), rnd(C, 1, k, k), rnd(C)
    y_r = ref.dwconv_fprop(x, w, b, stride, pad, 0.7)
    assert rel(cu.dwconv_fprop(g(x), g(w), g(b), stride, pad, 0.7), y_r) < FP32_TOL
    assert rel(cu.dwconv_fprop(g(x), g(w), None, stride, pad, 1.0), ref.dwconv_fprop(x, w, None, stride, pad, 1.0)) < FP32_TOL
    dy = rnd(*y_r.shape)
    assert rel(cu.dwconv_dgrad(g(dy), g(w), (H, W), stride, pad, 0.7), ref.dwconv_dgrad(dy, w, (H, W), stride, pad, 0.7)) < FP32_TOL
    assert rel(cu.dwconv_wgrad(g(x), g(dy), (k, k), stride, pad, 0.7), ref.dwconv_wgrad(x, dy, (k, k), stride, pad, 0.7)) < 2e-5


@pytest.mark.parametrize("shape", [(4, 7, 5, 3), (2, 1024, 4, 4), (3, 20, 64, 64), (5, 130, 1, 1)])
def test_pixnorm(cu, ref, shape):
    """Pixnorm of generators/generator_11.py:9-16 and its backward"""
    torch.manual_seed(14)
    x, gy = rnd(*shape) * 2, rnd(*shape)
    y_r, r_r = ref.pixnorm_fwd(x, 1e-8)
    y, r = cu.pixnorm_fwd(g(x), 1e-8)
    assert rel(y, y_r) < FP32_TOL and rel(r, r_r) < FP32_TOL
    assert rel(cu.pixnorm_bwd(g(gy), y, r), ref.pixnorm_bwd(gy, y_r, r_r)) < FP32_TOL
    zero = torch.zeros(2, 6, 3, 3)                      # eps keeps an all-zero pixel finite
    assert torch.isfinite(cu.pixnorm_fwd(g(zero), 1e-8)[0]).all()


def test_depthwise_conv_rejects_other_kernels(cu):
    from gan_b200._lib import GanB200Error
    for k, pad in ((4, 1), (9, 4), (1, 0)):           # even, larger than 7, 1x1 (that is a pointwise scale, not this family)
        with pytest.raises(GanB200Error):
            cu.dwconv_fprop(g(rnd(1, 4, 12, 12)), g(rnd(4, 1, k, k)), None, 1, pad, 1.0)
    with pytest.raises(GanB200Error):                  # 3x5: square kernels only
        cu.dwconv_fprop(g(rnd(1, 4, 12, 12)), g(rnd(4, 1, 3, 5)), None, 1, 1, 1.0)


@pytest.mark.parametrize("shape,groups", [((4, 244, 32, 32), 2), ((3, 12, 5, 7), 3), ((2, 1952, 4, 4), 2)])
def test_channel_shuffle(cu, ref, shape, groups):
    torch.manual_seed(12)
    x = rnd(*shape)
    y = cu.chan_shuffle(g(x), groups)
    assert torch.equal(y.cpu(), ref.chan_shuffle(x, groups))          # a permutation: bit exact
    back = cu.chan_shuffle(y, shape[1] // groups)                      # the adjoint is the inverse
    assert torch.equal(back.cpu(), x)


@pytest.mark.parametrize("shape", [(32, 3, 3, 3), (244, 244, 1, 1), (976, 976, 1, 1), (1, 1952, 4, 4), (122, 1, 3, 3)])
def test_spectral_norm_kernels(cu, ref, shape):
    torch.manual_seed(13)
    F = torch.nn.functional
    w = rnd(*shape)
    rows, cols = shape[0], w.numel() // shape[0]
    u, v = F.normalize(rnd(rows), dim=0), F.normalize(rnd(cols), dim=0)
    for n_iter in (1, 0, 3):
        ug, vg = g(u.clone()), g(v.clone())
        ur, vr = u.clone(), v.clone()
        s = cu.spectral_power_iter(g(w).reshape(rows, cols), ug, vg, n_iter, 1e-12)
        s_r = ref.spectral_power_iter(w.reshape(rows, cols), ur, vr, n_iter, 1e-12)
        assert rel(s, s_r) < 1e-5 and rel(ug, ur) < 1e-5 and rel(vg, vr) < 1e-5
    assert rel(cu.spectral_scale(g(w), s), ref.spectral_scale(w, s_r)) < FP32_TOL
    gw = rnd(*shape)
    assert rel(cu.spectral_scale_bwd(g(gw), g(w), ug, vg, s), ref.spectral_scale_bwd(gw, w, ur, vr, s_r)) < 2e-5


@pytest.mark.parametrize("B", [4, 16, 17, 64])
def test_batch_stddev_any_group_count(cu, ref, B):
    """discriminator_3_wgangp_progan.StandardDeviation uses G = batch: the register path (G <= 16) and the re-read
    path must agree with the oracle, including the double backward."""
    torch.manual_seed(14)
    xs, gg, ggx = rnd(B, 32, 4, 4), rnd(1), rnd(B, 32, 4, 4)
    assert rel(cu.mbstd_fwd(g(xs), B, 1e-7), ref.mbstd_fwd(xs, B, 1e-7)) < FP32_TOL
    assert rel(cu.mbstd_bwd(g(gg), g(xs), B, 1e-7), ref.mbstd_bwd(gg, xs, B, 1e-7)) < FP32_TOL
    o1, o2 = cu.mbstd_bwd_bwd(g(gg), g(xs), g(ggx), B, 1e-7)
    r1, r2 = ref.mbstd_bwd_bwd(gg, xs, ggx, B, 1e-7)
    assert rel(o1, r1) < 1e-4 and rel(o2, r2) < 1e-4


def test_spectral_norm_module_on_gpu_matches_torch(monkeypatch):
    """gan_b200.spectral.spectral_norm vs torch.nn.utils.parametrizations.spectral_norm on the same seeded conv:
    same buffers after construction, same normalised weight, same u/v after three train-mode forwards, same
    gradient of the original weight."""
    import copy
    from torch.nn.utils.parametrizations import spectral_norm as torch_sn
    from gan_b200 import layers as L
    from gan_b200.spectral import spectral_norm as my_sn
    monkeypatch.setenv("GANB200_PRECISION", "fp32")
    torch.manual_seed(15)
    base = torch.nn.Conv2d(24, 40, 3, padding=1)
    mine = L.Conv2d(24, 40, 3, padding=1)
    mine.load_state_dict(copy.deepcopy(base.state_dict()))
    torch.manual_seed(16)
    a = torch_sn(base)
    torch.manual_seed(16)
    b = my_sn(mine).cuda()
    assert list(a.state_dict()) == list(b.state_dict())
    x = rnd(4, 24, 8, 8)
    for _ in range(3):
        ya, yb = a(x), b(x.cuda())
    assert rel(yb, ya) < 1e-4
    for k, va in a.state_dict().items():
        assert rel(b.state_dict()[k], va) < 1e-5, k
    ya.square().sum().backward()
    yb.square().sum().backward()
    pa = a.parametrizations.weight.original.grad
    pb = b.parametrizations.weight.original.grad
    assert rel(pb, pa) < 1e-3
    b.eval()
    u0 = b.parametrizations.weight[0]._u.clone()
    b(x.cuda())
    assert torch.equal(u0, b.parametrizations.weight[0]._u)     # eval: no power iteration


# ---------------------------------------------------------------------------------------------------------------
# grouped launches (csrc/group.cu): the CUDA kernels against the oracle's job-by-job loop on the same job lists
# ---------------------------------------------------------------------------------------------------------------
def _bank_modules(device, widths=(3, 48, 33, 70, 256), d=64, seed=0):
    from generators.generator_13_5 import Conv2dWeightModulate
    torch.manual_seed(seed)
    mods = [Conv2dWeightModulate(d, c, 8, 3).to_style for c in widths]
    for m in mods:
        for p in m.parameters():
            p.data.add_(0.1 * torch.randn_like(p))
        m.to(device)
    return mods


@pytest.mark.parametrize("B", [4, 37, 256])
def test_grouped_style_bank_cuda_matches_oracle(cu, ref, B):
    """GroupedStyles forward + backward on the CUDA kernels == the same Python on the oracle backend (fp32 <= 1e-5):
    covers linear_group in its fprop / dgrad / wgrad operand layouts with ragged widths, and the four BatchNorm1d
    family kernels incl. PReLU, running statistics and the bias-gradient column."""
    import copy
    from gan_b200.group import GroupedStyles
    cpu_mods = _bank_modules("cpu")
    gpu_mods = [copy.deepcopy(m).cuda() for m in cpu_mods]
    torch.manual_seed(3)
    w = torch.randn(B, 64)
    gs = [torch.randn(B, m[1].weight.raw.shape[0]) for m in cpu_mods]
    out = {}
    for leg, mods, dev in (("ref", cpu_mods, "cpu"), ("cuda", gpu_mods, "cuda")):
        prev = KN.set_backend(ref if leg == "ref" else cu)
        try:
            w_ = w.detach().clone().to(dev).requires_grad_()
            bank = GroupedStyles(mods)
            assert bank.eligible()
            outs = bank(w_)
            torch.autograd.backward(list(outs), [t.to(dev) for t in gs])
            out[leg] = ([o.detach().cpu() for o in outs], w_.grad.cpu(), [p.grad.cpu() for m in mods for p in m.parameters()],
                        [b.detach().cpu() for m in mods for b in m.buffers()])
        finally:
            KN.set_backend(prev)
    a, b = out["cuda"], out["ref"]
    for u, v in zip(a[0], b[0]):
        assert rel(u, v) < FP32_TOL
    assert rel(a[1], b[1]) < FP32_TOL
    n_par = len(a[2]) // len(cpu_mods)
    for i, (u, v) in enumerate(zip(a[2], b[2])):
        # the two linear biases feed a BatchNorm: analytically zero gradient, what both sides compute is rounding noise
        # (|g| ~ 1e-6 next to 1e+1 for the weights) -> measured on the scale of the module's largest gradient
        m0 = (i // n_par) * n_par
        scale = max(t.norm().item() for t in b[2][m0:m0 + n_par])
        assert (u - v).norm().item() <= FP32_TOL * max(v.norm().item(), scale), (i, rel(u, v))
    for u, v in zip(a[3], b[3]):
        assert rel(u.double(), v.double()) < FP32_TOL


def test_grouped_style_bank_cuda_equals_layer_by_layer(cu):
    """the grouped launches against the per-layer kernels they replace (linear_small + bn_small), both on the GPU"""
    import copy
    from gan_b200.group import GroupedStyles
    B = 32
    mods_a = _bank_modules("cuda")
    mods_b = [copy.deepcopy(m) for m in mods_a]
    w = torch.randn(B, 64, device="cuda")
    gs = [torch.randn(B, m[1].weight.raw.shape[0], device="cuda") for m in mods_a]
    wa, wb = w.clone().requires_grad_(), w.clone().requires_grad_()
    oa = GroupedStyles(mods_a)(wa)
    ob = [m(wb) for m in mods_b]
    torch.autograd.backward(list(oa), gs)
    torch.autograd.backward(ob, gs)
    for u, v in zip(oa, ob):
        assert rel(u, v) < FP32_TOL
    assert rel(wa.grad, wb.grad) < FP32_TOL
    for ma, mb in zip(mods_a, mods_b):
        scale = max(q.grad.norm().item() for q in mb.parameters())
        for (n, p), (_, q) in zip(ma.named_parameters(), mb.named_parameters()):
            assert (p.grad - q.grad).norm().item() <= FP32_TOL * max(q.grad.norm().item(), scale), n
        for (n, p), (_, q) in zip(ma.named_buffers(), mb.named_buffers()):
            assert rel(p.double(), q.double()) < FP32_TOL, n


# ---------------------------------------------------------------------------------------------------------------
# scatter epilogue (round 2): dgrad of replicate padding straight into the unpadded plane, style-gradient reduce fused
# ---------------------------------------------------------------------------------------------------------------
SCATTER_CASES = [
    # N, C, H, W, K, R, stride, pad
    (4, 96, 64, 64, 96, 5, 1, 2),    # halo kernel, k5
    (4, 48, 64, 64, 54, 3, 1, 1),    # halo kernel, K not a multiple of 32
    (4, 64, 64, 64, 64, 3, 1, 1),    # per-tap kernel (the critic's)
    (8, 128, 16, 16, 128, 3, 1, 1),  # small plane: padded domain 18 x 18
    (8, 192, 8, 8, 192, 5, 1, 2),    # 8 x 8 plane (several images per tile)
    (16, 1025, 4, 4, 1025, 3, 1, 1), # 4 x 4 plane, odd channels
    (4, 64, 18, 18, 64, 3, 2, 1),    # stride 2: four output phases scatter into one plane
    (4, 3, 32, 32, 64, 3, 1, 1),     # C = 3 outputs (the stem's dgrad)
]


@pytest.mark.parametrize("case", SCATTER_CASES)
def test_dgrad_scatter_epilogue_equals_fold_pass(cu, ref, case):
    N, C, H, W, K, R, stride, pad = case
    torch.manual_seed(21)
    P, Q = KN.out_hw(H, W, R, R, stride, pad)
    dy, w = rnd(N, K, P, Q).cuda(), rnd(K, C, R, R).cuda()
    isc = (torch.rand(N, C) + 0.5).cuda()
    out = {}
    for mode in (0, 1):
        prev = cu.lib.ganb200_conv_scatter_mode(mode)
        try:
            out[mode] = (cu.conv_dgrad(dy, w, (H, W), stride, pad, 1, 0.05, engine=2),
                         cu.conv_dgrad(dy, w, (H, W), stride, pad, 1, 0.05, iscale=isc, engine=2))
        finally:
            cu.lib.ganb200_conv_scatter_mode(prev)
    for a, b in zip(out[1], out[0]):
        assert rel(a, b) < FP32_TOL   # the same bf16 products, fp32 sums re-ordered on the border pixels only
    want = ref.conv_dgrad(dy.cpu(), w.cpu(), (H, W), stride, pad, 1, 0.05)
    assert rel(out[1][0], want) < BF16_TOL


@pytest.mark.parametrize("case", [(4, 96, 64, 96, 5), (4, 48, 64, 54, 3), (4, 390, 32, 96, 1), (8, 96, 16, 96, 3),
                                  (8, 192, 8, 192, 5), (16, 384, 4, 384, 3)])
@pytest.mark.parametrize("needs", [(True, True, True), (False, True, True), (True, False, False)])
def test_modconv_bwd_fused_scatter_equals_unfused(cu, case, needs):
    """ganb200_modconv_bwd with the scatter epilogue (gx = s * fold(dxs) and gs1 = sum_hw dxs * x from the MMA epilogue)
    against the round-1 sequence (padded domain, then fold_scale_reduce)."""
    N, C, HW, K, R = case
    need_x, need_s, need_w = needs
    torch.manual_seed(22)
    x, w = rnd(N, C, HW, HW).cuda(), rnd(K, C, R, R).cuda()
    s, sinv = (torch.rand(N, C) + 0.5).cuda(), (torch.rand(N, K) + 0.5).cuda()
    gu = rnd(N, K, HW, HW).cuda()
    out = {}
    for mode in (0, 1):
        prev = cu.lib.ganb200_conv_scatter_mode(mode)
        try:
            out[mode] = cu.modconv_bwd(x, w, s, sinv, gu, R // 2, 0.02, need_x, need_s, need_w, engine=2)
        finally:
            cu.lib.ganb200_conv_scatter_mode(prev)
    assert out[0] is not None and out[1] is not None
    for a, b, name in zip(out[1], out[0], ("gx", "gs1", "gw")):
        assert (a is None) == (b is None), name
        if a is not None:
            assert rel(a, b) < 2e-5, (name, rel(a, b))
