"""Tensor-level wrappers over the C-ABI (torch is used for device memory and the current stream only).

`backend()` returns the object whose methods launch kernels.  The product backend is `CudaBackend`
(libganb200.so); it refuses CPU tensors.  tests/ may install a different backend with `set_backend`
(the CPU restatement in oracle/kernels_ref.py) to check the autograd formulas without a GPU — the
product itself never imports oracle/.
"""
import ctypes
import os

import torch

from . import _lib
from ._lib import ConvDesc

# enums from include/ganb200.h
U_SIGMOID, U_TANH, U_SQUARE, U_AFFINE, U_RSQRT_EPS, U_SQRT, U_LEAKY, U_RECIP, U_LEAKY_MASK, U_COPY, U_POW = range(11)
U_GELU, U_GELU_GRAD, U_GELU_GRAD2 = 11, 12, 13
B_ADD, B_SUB, B_MUL, B_DIV, B_SIGMOID_BWD, B_TANH_BWD, B_LEAKY_BWD = range(7)
V_NC, V_C, V_N, V_SCALAR = range(4)
R_HW, R_NHW, R_CHW, R_ALL = range(4)
FPROP, DGRAD, WGRAD = range(3)

# engine: 0 auto (tcgen05 when eligible), 1 fp32 CUDA cores, 2 force tcgen05
_ENGINE = {"auto": 0, "fp32": 1, "simt": 1, "bf16": 0, "tc": 2}


def default_engine():
    v = os.environ.get("GANB200_PRECISION", "auto")
    if v not in _ENGINE:
        raise ValueError(f"GANB200_PRECISION={v!r}: expected one of {sorted(_ENGINE)}")
    return _ENGINE[v]


LAUNCH_COUNT = [0]  # number of libganb200 kernel-launching calls (bench.py reports it as gpu_launches)


class ConvTimer:
    """bench.py: CUDA events around every convolution-engine call on the launching stream.  Off by default."""

    def __init__(self):
        self.records = []  # (engine "tc"|"simt", flops, start_event, end_event)

    def totals(self):
        """-> {engine: (flops, seconds, calls)}; call after a device synchronize."""
        out = {}
        for eng, fl, e0, e1 in self.records:
            f, t, n = out.get(eng, (0.0, 0.0, 0))
            out[eng] = (f + fl, t + e0.elapsed_time(e1) * 1e-3, n + 1)
        return out


# Packed-operand caches (tcgen05 engine).  PARAM_EPOCH counts in-place parameter updates that do not go through torch
# (the fused Adam, weight clipping, CUDA-graph replays): a packed weight is valid for the (tensor._version, PARAM_EPOCH) it
# was made at.  GANB200_PACK_CACHE=0 switches both caches off (every pass packs its own operands, as in round 1).
PARAM_EPOCH = [0]
PACK_CACHE = os.environ.get("GANB200_PACK_CACHE", "1") != "0"


def bump_param_epoch():
    PARAM_EPOCH[0] += 1


CONV_TIMER = None
OP_CENSUS = None    # tools/kernel_profile.py: dict (kernel family, operator / shape ...) -> number of calls
CONV_CENSUS = None  # bench.py: dict (op, desc fields..., has_iscale, has_oscale, has_bias, engine) -> number of calls


def _nchw(t):
    """logical [N, C, HW] of a >=1-D tensor."""
    if t.dim() == 1:
        return 1, t.shape[0], 1
    n, c = t.shape[0], t.shape[1]
    hw = 1
    for s in t.shape[2:]:
        hw *= s
    return n, c, hw


def out_hw(h, w, r, s, stride, pad):
    return (h + 2 * pad - r) // stride + 1, (w + 2 * pad - s) // stride + 1


class CudaBackend:
    name = "cuda"

    def __init__(self):
        self.lib = _lib.load()
        self._checked = set()
        self._ws = {}

    # -- plumbing -------------------------------------------------------------------------------
    def _prep(self, *tensors, allow_f64=False):
        dev = None
        for t in tensors:
            if t is None:
                continue
            if not t.is_cuda:
                raise _lib.GanB200Error("gan_b200 kernels need CUDA tensors (there is no CPU fallback)")
            if t.dtype != torch.float32 and t.dtype != torch.int32 and not (allow_f64 and t.dtype == torch.float64):
                raise _lib.GanB200Error(f"gan_b200 kernels are float32, got {t.dtype}")
            if not t.is_contiguous():
                raise _lib.GanB200Error("gan_b200 kernels need contiguous tensors")
            dev = t.device
        idx = dev.index if dev.index is not None else torch.cuda.current_device()
        if idx not in self._checked:
            _lib.check(self.lib.ganb200_device_check(idx))
            self._checked.add(idx)
        LAUNCH_COUNT[0] += 1
        return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    @staticmethod
    def _p(t):
        return ctypes.c_void_p(t.data_ptr()) if t is not None else None

    def _workspace(self, nbytes, device):
        if nbytes == 0:
            return None
        # a fresh tensor from the caching allocator: stream-ordered reuse is handled by torch
        return torch.empty(nbytes, dtype=torch.uint8, device=device)

    # -- packed operands of the tcgen05 engine -----------------------------------------------------
    def _aligned_buffer(self, nbytes, device):
        buf = torch.empty(nbytes + 1024, dtype=torch.uint8, device=device)
        off = (-buf.data_ptr()) % 1024
        return buf[off:off + nbytes]

    def packed_weight(self, w, d, op):
        """bf16 packed form of weight `w` for (desc, op), cached on the tensor until it changes: w._version moves with
        every torch in-place op, PARAM_EPOCH with the updates torch does not see.  None when caching is off."""
        if not PACK_CACHE:
            return None
        key = (op, d.N if op == DGRAD else 0, d.C, d.H, d.W, d.K, d.R, d.S, d.stride, d.pad, d.pad_mode)
        stamp = (w._version, PARAM_EPOCH[0], w.data_ptr())
        cache = getattr(w, "_gb_wpack", None)
        if cache is None:
            cache = {}
            try:
                w._gb_wpack = cache
            except Exception:
                return None
        hit = cache.get(key)
        if hit is not None and hit[0] == stamp:
            return hit[1]
        nb = self.lib.ganb200_conv2d_packed_weight_bytes(ctypes.byref(d), op)
        if nb == 0:
            return None
        buf = hit[1] if (hit is not None and hit[1].numel() == nb) else self._aligned_buffer(nb, w.device)
        LAUNCH_COUNT[0] += 1
        st = ctypes.c_void_p(torch.cuda.current_stream(w.device).cuda_stream)
        _lib.check(self.lib.ganb200_conv2d_pack_weight(ctypes.byref(d), op, self._p(w), self._p(buf), st))
        cache[key] = (stamp, buf)
        return buf

    def packed_act(self, t, scale, pad, keep, transient=False):
        """bf16 NHWC packed form of activation `t` (scale folded in, replicate padding `pad` materialised).  With `keep`
        the buffer stays on the tensor (t._gb_pack) for the other passes that need the same operand — x: the wgrad of the
        layer after its fprop; dy: wgrad after dgrad — and dies with it.  None when there is nothing cached and keep is
        False (the engine then packs into its own workspace)."""
        if not PACK_CACHE:
            return None
        cache = getattr(t, "_gb_pack", None)
        key = (pad, 0 if scale is None else scale.data_ptr())
        if cache is not None and cache[0] == t._version:
            hit = cache[1].get(key)
            if hit is not None and hit[1] is scale:
                return hit[0]
        if not keep and not transient:
            return None
        N, C, H, W = t.shape
        nb = self.lib.ganb200_tc_packed_act_bytes(N, C, H, W, pad)
        buf = self._aligned_buffer(nb, t.device)
        LAUNCH_COUNT[0] += 1
        st = ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)
        _lib.check(self.lib.ganb200_tc_pack_act(self._p(t), self._p(scale), self._p(buf), N, C, H, W, pad, st))
        if not keep:
            return buf                 # transient: used by this call only
        if cache is None or cache[0] != t._version:
            cache = (t._version, {})
            try:
                t._gb_pack = cache
            except Exception:
                return buf
        cache[1][key] = (buf, scale)   # the reference to `scale` keeps its address from being reused while cached
        return buf

    def _is_tc(self, d, op):
        return PACK_CACHE and self.lib.ganb200_conv2d_route(ctypes.byref(d), op) == 2

    # -- convolution family ---------------------------------------------------------------------
    def _t0(self, d, op, flags=(False, False, False)):
        if CONV_CENSUS is not None:
            route = {1: "simt", 2: "tc", 3: "linear"}[self.lib.ganb200_conv2d_route(ctypes.byref(d), op)]
            key = (op, d.N, d.C, d.H, d.W, d.K, d.R, d.S, d.stride, d.pad, d.pad_mode) + tuple(flags) + (route,)
            CONV_CENSUS[key] = CONV_CENSUS.get(key, 0) + 1
        if CONV_TIMER is None:
            return None
        route = {1: "simt", 2: "tc", 3: "linear"}[self.lib.ganb200_conv2d_route(ctypes.byref(d), op)]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        return (route, 2.0 * d.N * d.K * d.P * d.Q * d.C * d.R * d.S, e0, e1)

    def _t1(self, tm):
        if tm is not None:
            tm[3].record()
            CONV_TIMER.records.append(tm)

    def _desc(self, N, C, H, W, K, R, S, stride, pad, pad_mode, alpha, engine):
        P, Q = out_hw(H, W, R, S, stride, pad)
        return ConvDesc(N, C, H, W, K, R, S, stride, pad, pad_mode, P, Q, float(alpha),
                        default_engine() if engine is None else engine)

    def conv_fprop(self, x, w, bias, stride, pad, pad_mode, alpha, iscale=None, oscale=None, engine=None,
                   keep_x=False):
        """keep_x: the packed x stays cached on the tensor for the wgrad of the same layer (callers that save x)."""
        N, C, H, W = x.shape
        K, C2, R, S = w.shape
        assert C == C2, (x.shape, w.shape)
        d = self._desc(N, C, H, W, K, R, S, stride, pad, pad_mode, alpha, engine)
        st = self._prep(x, w, bias, iscale, oscale)
        y = torch.empty((N, K, d.P, d.Q), dtype=torch.float32, device=x.device)
        tm = self._t0(d, FPROP, (iscale is not None, oscale is not None, bias is not None))
        if self._is_tc(d, FPROP):
            wp = self.packed_weight(w, d, FPROP)
            xp = None
            if wp is not None:   # (the _pk entry takes both operands packed: x is packed here, kept only on request)
                xp = self.packed_act(x, iscale, pad if (pad_mode == 1 and pad > 0) else 0, keep_x, transient=True)
            if xp is not None and wp is not None:
                _lib.check(self.lib.ganb200_conv2d_fprop_pk(ctypes.byref(d), self._p(xp), self._p(wp), self._p(bias),
                                                            self._p(oscale), self._p(y), st))
                self._t1(tm)
                return y
        nb = self.lib.ganb200_conv2d_workspace_bytes(ctypes.byref(d), FPROP)
        ws = self._workspace(nb, x.device)
        _lib.check(self.lib.ganb200_conv2d_fprop(ctypes.byref(d), self._p(x), self._p(w), self._p(bias),
                                                 self._p(iscale), self._p(oscale), self._p(y), self._p(ws), nb, st))
        self._t1(tm)
        return y

    def conv_dgrad(self, dy, w, in_hw, stride, pad, pad_mode, alpha, iscale=None, oscale=None, engine=None,
                   keep_dy=False):
        """keep_dy: the packed dy stays cached on the tensor for the wgrad that shares it."""
        N, K, P, Q = dy.shape
        K2, C, R, S = w.shape
        assert K == K2, (dy.shape, w.shape)
        H, W = in_hw
        d = self._desc(N, C, H, W, K, R, S, stride, pad, pad_mode, alpha, engine)
        assert (d.P, d.Q) == (P, Q), ((d.P, d.Q), (P, Q))
        st = self._prep(dy, w, iscale, oscale)
        dx = torch.empty((N, C, H, W), dtype=torch.float32, device=dy.device)
        nb = self.lib.ganb200_conv2d_workspace_bytes(ctypes.byref(d), DGRAD)
        ws = self._workspace(nb, dy.device)
        tm = self._t0(d, DGRAD, (iscale is not None, oscale is not None, False))
        if self._is_tc(d, DGRAD):
            wp = self.packed_weight(w, d, DGRAD)
            dyp = self.packed_act(dy, oscale, 0, keep_dy)
            if wp is not None or dyp is not None:
                _lib.check(self.lib.ganb200_conv2d_dgrad_pk(ctypes.byref(d), self._p(dy), self._p(dyp), self._p(w),
                                                            self._p(wp), self._p(iscale), self._p(oscale), self._p(dx),
                                                            self._p(ws), nb, st))
                self._t1(tm)
                return dx
        _lib.check(self.lib.ganb200_conv2d_dgrad(ctypes.byref(d), self._p(dy), self._p(w), self._p(iscale),
                                                 self._p(oscale), self._p(dx), self._p(ws), nb, st))
        self._t1(tm)
        return dx

    def conv_wgrad(self, x, dy, rs, stride, pad, pad_mode, alpha, iscale=None, oscale=None, engine=None,
                   keep_x=False, keep_dy=False):
        N, C, H, W = x.shape
        N2, K, P, Q = dy.shape
        R, S = rs
        d = self._desc(N, C, H, W, K, R, S, stride, pad, pad_mode, alpha, engine)
        assert N == N2 and (d.P, d.Q) == (P, Q), (x.shape, dy.shape)
        st = self._prep(x, dy, iscale, oscale)
        dw = torch.empty((K, C, R, S), dtype=torch.float32, device=x.device)
        nb = self.lib.ganb200_conv2d_workspace_bytes(ctypes.byref(d), WGRAD)
        ws = self._workspace(nb, x.device)
        tm = self._t0(d, WGRAD, (iscale is not None, oscale is not None, False))
        if self._is_tc(d, WGRAD):
            xp = self.packed_act(x, iscale, pad if (pad_mode == 1 and pad > 0) else 0, keep_x)
            dyp = self.packed_act(dy, oscale, 0, keep_dy)
            if xp is not None or dyp is not None:
                _lib.check(self.lib.ganb200_conv2d_wgrad_pk(ctypes.byref(d), self._p(x), self._p(xp), self._p(dy),
                                                            self._p(dyp), self._p(iscale), self._p(oscale), self._p(dw),
                                                            self._p(ws), nb, st))
                self._t1(tm)
                return dw
        _lib.check(self.lib.ganb200_conv2d_wgrad(ctypes.byref(d), self._p(x), self._p(dy), self._p(iscale),
                                                 self._p(oscale), self._p(dw), self._p(ws), nb, st))
        self._t1(tm)
        return dw

    # -- depthwise 3x3 / 5x5 / 7x7 conv (discriminator_7_wgangpsn.py:55,61,72; discriminator_6_wgangp.py:12) -----------------------------------
    def dwconv_fprop(self, x, w, bias, stride, pad, alpha):
        N, C, H, W = x.shape
        assert w.shape[0] == C and w.shape[1] == 1, (x.shape, w.shape)
        R, S = w.shape[2:]
        P, Q = out_hw(H, W, R, S, stride, pad)
        st = self._prep(x, w, bias)
        y = torch.empty((N, C, P, Q), dtype=torch.float32, device=x.device)
        _lib.check(self.lib.ganb200_dwconv2d_fprop(self._p(x), self._p(w), self._p(bias), self._p(y), N, C, H, W, R, S,
                                                   stride, pad, alpha, st))
        return y

    def dwconv_dgrad(self, dy, w, in_hw, stride, pad, alpha):
        N, C, P, Q = dy.shape
        R, S = w.shape[2:]
        H, W = in_hw
        assert out_hw(H, W, R, S, stride, pad) == (P, Q) and w.shape[0] == C, (dy.shape, w.shape, in_hw)
        st = self._prep(dy, w)
        dx = torch.empty((N, C, H, W), dtype=torch.float32, device=dy.device)
        _lib.check(self.lib.ganb200_dwconv2d_dgrad(self._p(dy), self._p(w), self._p(dx), N, C, H, W, R, S, stride, pad,
                                                   alpha, st))
        return dx

    def dwconv_wgrad(self, x, dy, rs, stride, pad, alpha):
        N, C, H, W = x.shape
        R, S = rs
        assert dy.shape[:2] == (N, C) and out_hw(H, W, R, S, stride, pad) == tuple(dy.shape[2:]), (x.shape, dy.shape)
        st = self._prep(x, dy)
        LAUNCH_COUNT[0] += 1
        dw = torch.empty((C, 1, R, S), dtype=torch.float32, device=x.device)
        nb = self.lib.ganb200_dwconv2d_wgrad_workspace_bytes_rs(N, C, R, S)
        ws = self._workspace(nb, x.device)
        _lib.check(self.lib.ganb200_dwconv2d_wgrad(self._p(x), self._p(dy), self._p(dw), self._p(ws), nb, N, C, H, W,
                                                   R, S, stride, pad, alpha, st))
        return dw

    def chan_shuffle(self, x, groups):
        N, C, HW = _nchw(x)
        st = self._prep(x)
        y = torch.empty_like(x)
        _lib.check(self.lib.ganb200_chan_shuffle(self._p(x), self._p(y), N, C, HW, groups, st))
        return y

    # -- resampling -----------------------------------------------------------------------------
    def resample2d(self, x, op_h, op_w):
        *lead, Hi, Wi = x.shape
        assert Hi == op_h.n_in and Wi == op_w.n_in, (x.shape, op_h.name, op_w.name)
        st = self._prep(x)
        nc = 1
        for s in lead:
            nc *= s
        y = torch.empty((*lead, op_h.n_out, op_w.n_out), dtype=torch.float32, device=x.device)
        ih, wh, th = op_h.taps(x.device)
        iw, ww, tw = op_w.taps(x.device)
        if OP_CENSUS is not None:
            key = ("resample2d", op_h.name, nc, Hi, Wi, op_h.n_out, op_w.n_out, th)
            OP_CENSUS[key] = OP_CENSUS.get(key, 0) + 1
        _lib.check(self.lib.ganb200_resample2d(self._p(x), self._p(y), nc, Hi, Wi, op_h.n_out, op_w.n_out,
                                               self._p(ih), self._p(wh), th, self._p(iw), self._p(ww), tw, st))
        return y

    # -- pointwise ------------------------------------------------------------------------------
    def ew_unary(self, op, x, p0=0.0, p1=0.0):
        st = self._prep(x)
        y = torch.empty_like(x)
        _lib.check(self.lib.ganb200_ew_unary(op, self._p(x), self._p(y), x.numel(), p0, p1, st))
        return y

    def ew_binary(self, op, a, b, p0=0.0):
        assert a.shape == b.shape, (a.shape, b.shape)
        st = self._prep(a, b)
        y = torch.empty_like(a)
        _lib.check(self.lib.ganb200_ew_binary(op, self._p(a), self._p(b), self._p(y), a.numel(), p0, st))
        return y

    def ew_bcast(self, op, x, v, vmode):
        N, C, HW = _nchw(x)
        st = self._prep(x, v)
        y = torch.empty_like(x)
        _lib.check(self.lib.ganb200_ew_bcast(op, self._p(x), self._p(v), self._p(y), N, C, HW, vmode, st))
        return y

    def ew_fma_bcast(self, x, v, z, vmode):
        N, C, HW = _nchw(x)
        st = self._prep(x, v, z)
        y = torch.empty_like(x)
        _lib.check(self.lib.ganb200_ew_fma_bcast(self._p(x), self._p(v), self._p(z), self._p(y), N, C, HW, vmode, st))
        return y

    def bmm(self, a, b, ta=False, tb=False, alpha=1.0):
        """batched op(a) @ op(b) for contiguous 3-D fp32 tensors; ta / tb transpose the last two dimensions of a / b
        (by strides: nothing is copied)"""
        if a.dim() != 3 or b.dim() != 3 or a.shape[0] != b.shape[0]:
            raise _lib.GanB200Error(f"bmm: need two 3-D tensors with one batch size, got {tuple(a.shape)} {tuple(b.shape)}")
        st = self._prep(a, b)
        Bn = a.shape[0]
        M, Ka = (a.shape[2], a.shape[1]) if ta else (a.shape[1], a.shape[2])
        Kb, N = (b.shape[2], b.shape[1]) if tb else (b.shape[1], b.shape[2])
        if Ka != Kb:
            raise _lib.GanB200Error(f"bmm: inner dimensions differ ({Ka} vs {Kb})")
        sam, sak = (1, a.shape[2]) if ta else (a.shape[2], 1)
        sbk, sbn = (1, b.shape[2]) if tb else (b.shape[2], 1)
        c = torch.empty((Bn, M, N), dtype=torch.float32, device=a.device)
        _lib.check(self.lib.ganb200_bmm(self._p(a), self._p(b), self._p(c), Bn, M, N, Ka, a.shape[1] * a.shape[2], sam,
                                        sak, b.shape[1] * b.shape[2], sbk, sbn, float(alpha), st))
        return c

    def softmax_rows_fwd(self, x):
        """softmax over the last dimension"""
        st = self._prep(x)
        y = torch.empty_like(x)
        n = x.shape[-1]
        _lib.check(self.lib.ganb200_softmax_rows_fwd(self._p(x), self._p(y), x.numel() // max(n, 1), n, st))
        return y

    def softmax_rows_bwd(self, gy, y):
        st = self._prep(gy, y)
        gx = torch.empty_like(y)
        n = y.shape[-1]
        _lib.check(self.lib.ganb200_softmax_rows_bwd(self._p(gy), self._p(y), self._p(gx), y.numel() // max(n, 1), n, st))
        return gx

    # -- pixelwise feature-vector normalisation (generators/generator_11.py:9-16) ---------------------
    def pixnorm_fwd(self, x, eps):
        N, C, HW = _nchw(x)
        st = self._prep(x)
        y = torch.empty_like(x)
        r = torch.empty((N, HW), dtype=torch.float32, device=x.device)
        _lib.check(self.lib.ganb200_pixnorm_fwd(self._p(x), self._p(y), self._p(r), N, C, HW, eps, st))
        return y, r

    def pixnorm_bwd(self, gy, y, r):
        N, C, HW = _nchw(y)
        st = self._prep(gy, y, r)
        gx = torch.empty_like(y)
        _lib.check(self.lib.ganb200_pixnorm_bwd(self._p(gy), self._p(y), self._p(r), self._p(gx), N, C, HW, st))
        return gx

    def pool_sum2_ok(self, t0, t1, out):
        return (t0.dim() == 4 and t0.shape == t1.shape and t0.dtype == torch.float32 and
                bool(self.lib.ganb200_pool_sum2_eligible(t0.shape[-2], t0.shape[-1], out)))

    def pool_sum2(self, t0, t1, out):
        """AdaptiveAvgPool2d(out)(t0 + t1) reading each branch once"""
        N, C, H, W = t0.shape
        st = self._prep(t0, t1)
        y = torch.empty((N, C, out, out), dtype=torch.float32, device=t0.device)
        _lib.check(self.lib.ganb200_pool_sum2(self._p(t0), self._p(t1), self._p(y), N * C, H, W, out, st))
        return y

    def sk_bwd_finish(self, g, att, gp):
        """g_i = att[:, i] * g + adjoint of AdaptiveAvgPool2d(gp.shape[-1]) applied to gp -> (g0, g1)"""
        N, C, H, W = g.shape
        st = self._prep(g, att, gp)
        g0, g1 = torch.empty_like(g), torch.empty_like(g)
        _lib.check(self.lib.ganb200_sk_bwd_finish(self._p(g), self._p(att), self._p(gp), self._p(g0), self._p(g1), N, C, H,
                                                  W, gp.shape[-1], st))
        return g0, g1

    def sk_merge2_fwd(self, t0, t1, att):
        """att[:, 0] * t0 + att[:, 1] * t1, att: [N, 2, C] contiguous"""
        N, C, HW = _nchw(t0)
        st = self._prep(t0, t1, att)
        out = torch.empty_like(t0)
        _lib.check(self.lib.ganb200_sk_merge2_fwd(self._p(t0), self._p(t1), self._p(att), self._p(out), N, C, HW, st))
        return out

    def sk_merge2_bwd(self, g, t0, t1, att, att_only=False):
        """-> (g_t0, g_t1, g_att[N, 2, C]); att_only: (None, None, g_att)"""
        N, C, HW = _nchw(t0)
        st = self._prep(g, t0, t1, att)
        g0, g1 = (None, None) if att_only else (torch.empty_like(t0), torch.empty_like(t1))
        gatt = torch.empty((N, 2, C), dtype=torch.float32, device=t0.device)
        _lib.check(self.lib.ganb200_sk_merge2_bwd(self._p(g), self._p(t0), self._p(t1), self._p(att), self._p(g0),
                                                  self._p(g1), self._p(gatt), N, C, HW, st))
        return g0, g1, gatt

    def reduce(self, x, y, rmode, scale=1.0):
        N, C, HW = _nchw(x)
        st = self._prep(x, y)
        shape = {R_HW: (N, C), R_NHW: (C,), R_CHW: (N,), R_ALL: (1,)}[rmode]
        out = torch.empty(shape, dtype=torch.float32, device=x.device)
        _lib.check(self.lib.ganb200_reduce(self._p(x), self._p(y), self._p(out), N, C, HW, rmode, scale, st))
        return out

    # -- PReLU ----------------------------------------------------------------------------------
    def prelu_fwd(self, x, a):
        N, C, HW = _nchw(x)
        st = self._prep(x, a)
        y = torch.empty_like(x)
        _lib.check(self.lib.ganb200_prelu_fwd(self._p(x), self._p(a), self._p(y), N, C, HW, a.numel(), st))
        return y

    def maskscale(self, g, x, a):
        N, C, HW = _nchw(x)
        st = self._prep(g, x, a)
        y = torch.empty_like(x)
        _lib.check(self.lib.ganb200_maskscale(self._p(g), self._p(x), self._p(a), self._p(y), N, C, HW, a.numel(), st))
        return y

    def prelu_bwd(self, gy, x, a, need_gx=True, need_ga=True):
        N, C, HW = _nchw(x)
        st = self._prep(gy, x, a)
        gx = torch.empty_like(x) if need_gx else None
        ga = torch.empty_like(a) if need_ga else None
        _lib.check(self.lib.ganb200_prelu_bwd(self._p(gy), self._p(x), self._p(a), self._p(gx), self._p(ga), N, C, HW,
                                              a.numel(), st))
        return gx, ga

    def masked_dot(self, u, v, x, a_n):
        N, C, HW = _nchw(x)
        st = self._prep(u, v, x)
        out = torch.empty((a_n,), dtype=torch.float32, device=x.device)
        _lib.check(self.lib.ganb200_masked_dot(self._p(u), self._p(v), self._p(x), self._p(out), N, C, HW, a_n, st))
        return out

    # -- BatchNorm ------------------------------------------------------------------------------
    def bn_stats(self, x):
        N, C, HW = _nchw(x)
        st = self._prep(x)
        mean = torch.empty((C,), dtype=torch.float32, device=x.device)
        var = torch.empty((C,), dtype=torch.float32, device=x.device)
        _lib.check(self.lib.ganb200_bn_stats(self._p(x), self._p(mean), self._p(var), N, C, HW, st))
        return mean, var

    def bn_apply(self, x, mean, var, gamma, beta, running_mean, running_var, eps, momentum, count_total, nbt=None):
        N, C, HW = _nchw(x)
        st = self._prep(x, mean, var, gamma, beta, running_mean, running_var)
        if nbt is not None and nbt.dtype != torch.int64:
            raise _lib.GanB200Error("num_batches_tracked must be int64")
        y = torch.empty_like(x)
        invstd = torch.empty((C,), dtype=torch.float32, device=x.device)
        _lib.check(self.lib.ganb200_bn_apply(self._p(x), self._p(mean), self._p(var), self._p(gamma), self._p(beta),
                                             self._p(y), self._p(invstd), self._p(running_mean),
                                             self._p(running_var), self._p(nbt), N, C, HW, eps, momentum,
                                             count_total, st))
        return y, invstd

    # N*HW up to which the one-launch kernels are used (the kernels take up to 8192; GANB200_BN_SMALL_MAX lowers it)
    BN_SMALL_MAX = min(8192, int(os.environ.get("GANB200_BN_SMALL_MAX", "8192")))

    def bn_small_ok(self, x):
        N, C, HW = _nchw(x)
        return N * HW <= self.BN_SMALL_MAX

    def bn_small_sync_ok(self, x, peer):
        """data parallel: can the fused launch carry the sync-BN exchange (gan_b200.peer.PeerExchange, all C blocks
        co-resident)?"""
        if peer is None or not self.bn_small_ok(x):
            return False
        if self._bn_sync_cap is None:
            self._bn_sync_cap = int(self.lib.ganb200_bn_small_sync_max_channels())
        return x.shape[1] <= self._bn_sync_cap

    _bn_sync_cap = None

    def bn_small_fwd(self, x, gamma, beta, running_mean, running_var, eps, momentum, nbt=None, slope=None, peer=None):
        """one launch: batch statistics, y (optionally through the following per-channel PReLU), running statistics
        -> (y, mean, invstd).  `peer` (gan_b200.peer.PeerExchange): statistics over the GLOBAL batch, swapped between
        the ranks inside the launch."""
        N, C, HW = _nchw(x)
        st = self._prep(x, gamma, beta, running_mean, running_var, slope)
        if slope is not None and slope.numel() != C:
            raise _lib.GanB200Error("bn_small_fwd: the fused PReLU needs one slope per channel")
        if nbt is not None and nbt.dtype != torch.int64:
            raise _lib.GanB200Error("num_batches_tracked must be int64")
        y = torch.empty_like(x)
        mean = torch.empty((C,), dtype=torch.float32, device=x.device)
        invstd = torch.empty((C,), dtype=torch.float32, device=x.device)
        if peer is not None:
            peer.count()
            _lib.check(self.lib.ganb200_bn_small_fwd_sync(self._p(x), self._p(gamma), self._p(beta), self._p(y),
                                                          self._p(mean), self._p(invstd), self._p(running_mean),
                                                          self._p(running_var), self._p(nbt), N, C, HW, eps, momentum,
                                                          self._p(slope), self._p(peer.table), peer.rank, peer.world, st))
            return y, mean, invstd
        _lib.check(self.lib.ganb200_bn_small_fwd(self._p(x), self._p(gamma), self._p(beta), self._p(y), self._p(mean),
                                                 self._p(invstd), self._p(running_mean), self._p(running_var),
                                                 self._p(nbt), N, C, HW, eps, momentum, self._p(slope), st))
        return y, mean, invstd

    def bn_small_bwd(self, gy, x, mean, invstd, gamma, beta=None, slope=None, peer=None):
        """one launch: (gx, ggamma, gbeta[, gslope when the PReLU is fused]).  `peer`: dx from the sums over the global
        batch (swapped inside the launch); ggamma / gbeta / gslope stay rank-local partial sums."""
        N, C, HW = _nchw(x)
        st = self._prep(gy, x, mean, invstd, gamma, beta, slope)
        gx = torch.empty_like(x)
        gg = torch.empty((C,), dtype=torch.float32, device=x.device)
        gb = torch.empty((C,), dtype=torch.float32, device=x.device)
        gs = torch.empty((C,), dtype=torch.float32, device=x.device) if slope is not None else None
        if peer is not None:
            peer.count()
            _lib.check(self.lib.ganb200_bn_small_bwd_sync(self._p(gy), self._p(x), self._p(mean), self._p(invstd),
                                                          self._p(gamma), self._p(gx), self._p(gg), self._p(gb), N, C, HW,
                                                          self._p(beta), self._p(slope), self._p(gs), self._p(peer.table),
                                                          peer.rank, peer.world, st))
            return (gx, gg, gb) if slope is None else (gx, gg, gb, gs)
        _lib.check(self.lib.ganb200_bn_small_bwd(self._p(gy), self._p(x), self._p(mean), self._p(invstd),
                                                 self._p(gamma), self._p(gx), self._p(gg), self._p(gb), N, C, HW,
                                                 self._p(beta), self._p(slope), self._p(gs), st))
        return (gx, gg, gb) if slope is None else (gx, gg, gb, gs)

    def bn_bwd_reduce(self, gy, x, mean, invstd):
        N, C, HW = _nchw(x)
        st = self._prep(gy, x, mean, invstd)
        s1 = torch.empty((C,), dtype=torch.float32, device=x.device)
        s2 = torch.empty((C,), dtype=torch.float32, device=x.device)
        _lib.check(self.lib.ganb200_bn_bwd_reduce(self._p(gy), self._p(x), self._p(mean), self._p(invstd),
                                                  self._p(s1), self._p(s2), N, C, HW, st))
        return s1, s2

    def bn_bwd_apply(self, gy, x, mean, invstd, gamma, sum_gy, sum_gy_xhat, count_total):
        N, C, HW = _nchw(x)
        st = self._prep(gy, x, mean, invstd, gamma, sum_gy, sum_gy_xhat)
        gx = torch.empty_like(x)
        gg = torch.empty((C,), dtype=torch.float32, device=x.device)
        gb = torch.empty((C,), dtype=torch.float32, device=x.device)
        _lib.check(self.lib.ganb200_bn_bwd_apply(self._p(gy), self._p(x), self._p(mean), self._p(invstd),
                                                 self._p(gamma), self._p(sum_gy), self._p(sum_gy_xhat), self._p(gx),
                                                 self._p(gg), self._p(gb), N, C, HW, count_total, st))
        return gx, gg, gb

    # -- softmax over branches, minibatch stddev, BCE ---------------------------------------------
    def softmax_m_fwd(self, x):
        B, M, C = x.shape
        st = self._prep(x)
        y = torch.empty_like(x)
        _lib.check(self.lib.ganb200_softmax_m_fwd(self._p(x), self._p(y), B, M, C, st))
        return y

    def softmax_m_bwd(self, gy, y):
        B, M, C = y.shape
        st = self._prep(gy, y)
        gx = torch.empty_like(y)
        _lib.check(self.lib.ganb200_softmax_m_bwd(self._p(gy), self._p(y), self._p(gx), B, M, C, st))
        return gx

    def mbstd_fwd(self, x, G, eps):
        st = self._prep(x)
        out = torch.empty((1,), dtype=torch.float32, device=x.device)
        _lib.check(self.lib.ganb200_mbstd_fwd(self._p(x), self._p(out), G, x.numel() // G, eps, st))
        return out

    def mbstd_bwd(self, g, x, G, eps):
        st = self._prep(g, x)
        gx = torch.empty_like(x)
        _lib.check(self.lib.ganb200_mbstd_bwd(self._p(g), self._p(x), self._p(gx), G, x.numel() // G, eps, st))
        return gx

    def mbstd_bwd_bwd(self, g, x, ggx, G, eps):
        st = self._prep(g, x, ggx)
        out_gg = torch.empty((1,), dtype=torch.float32, device=x.device)
        out_gx = torch.empty_like(x)
        _lib.check(self.lib.ganb200_mbstd_bwd_bwd(self._p(g), self._p(x), self._p(ggx), self._p(out_gg),
                                                  self._p(out_gx), G, x.numel() // G, eps, st))
        return out_gg, out_gx

    def bce_fwd(self, p, t, sticky=None):
        """-> (loss[1], flag[1] int32).  `sticky`: a caller-owned int32[1] that is only ever set (CUDA-graph replays)."""
        st = self._prep(p, t)
        out = torch.empty((1,), dtype=torch.float32, device=p.device)
        if sticky is not None:
            _lib.check(self.lib.ganb200_bce_fwd_sticky(self._p(p), self._p(t), self._p(out), self._p(sticky),
                                                       p.numel(), st))
            return out, sticky
        flag = torch.empty((1,), dtype=torch.int32, device=p.device)
        _lib.check(self.lib.ganb200_bce_fwd(self._p(p), self._p(t), self._p(out), self._p(flag), p.numel(), st))
        return out, flag

    def bce_bwd(self, g, p, t):
        st = self._prep(g, p, t)
        gp = torch.empty_like(p)
        _lib.check(self.lib.ganb200_bce_bwd(self._p(g), self._p(p), self._p(t), self._p(gp), p.numel(), st))
        return gp

    # -- fused StyleConv layer ----------------------------------------------------------------------
    def _prep_any(self, *tensors):
        """like _prep but accepts the int64 rng state"""
        st = self._prep(*[t for t in tensors if t is None or t.dtype != torch.int64])
        return st

    def noise_fill(self, shape, rng_state, layer, device):
        z = torch.empty(shape, dtype=torch.float32, device=device)
        st = self._prep(z)
        _lib.check(self.lib.ganb200_noise_fill(self._p(z), z.numel(), self._p(rng_state), int(layer), st))
        return z

    def demod(self, s, wsq, alpha, eps):
        B, C = s.shape
        Kout = wsq.shape[0]
        st = self._prep(s, wsq)
        sinv = torch.empty((B, Kout), dtype=torch.float32, device=s.device)
        _lib.check(self.lib.ganb200_demod(self._p(s), self._p(wsq), self._p(sinv), B, C, Kout, alpha, eps, st))
        return sinv

    def demod_bwd_s(self, gs1, s, Ad, sinv, wsq, alpha, out=None):
        B, C = s.shape
        Kout = wsq.shape[0]
        st = self._prep(gs1, s, Ad, sinv, wsq, out)
        gs = out if out is not None else torch.empty_like(s)
        _lib.check(self.lib.ganb200_demod_bwd_s(self._p(gs1), self._p(s), self._p(Ad), self._p(sinv), self._p(wsq),
                                                self._p(gs), B, C, Kout, alpha, st))
        return gs

    def demod_bwd_w_(self, gw, w, Ad, sinv, s, alpha):
        Kout, C = w.shape[:2]
        RS = w.numel() // (Kout * C)
        st = self._prep(gw, w, Ad, sinv, s)
        _lib.check(self.lib.ganb200_demod_bwd_w(self._p(gw), self._p(w), self._p(Ad), self._p(sinv), self._p(s),
                                                s.shape[0], C, Kout, RS, alpha, st))
        return gw

    def style_tail_fwd(self, y, ns, slope, rng_state, layer):
        """-> (u, a): u = y + ns*z written IN PLACE into y (u is y) when ns is given; a = PReLU(u) or None."""
        N, Kc, HW = _nchw(y)
        st = self._prep(y, ns, slope)
        a = torch.empty_like(y) if slope is not None else None
        _lib.check(self.lib.ganb200_style_tail_fwd(self._p(y), self._p(y), self._p(a), self._p(ns), self._p(slope),
                                                   slope.numel() if slope is not None else 0, N, Kc, HW,
                                                   self._p(rng_state), int(layer), st))
        return y, a

    def style_tail_bwd(self, ga, u, ns, slope, rng_state, layer, need_Ad=True):
        """-> (g_u, Ad[N,K], Z[N,K] or None, Pm[N,K] or None)"""
        N, Kc, HW = _nchw(u)
        st = self._prep(ga, u, ns, slope)
        dev = u.device
        gu = torch.empty_like(u) if slope is not None else ga
        Ad = torch.empty((N, Kc), dtype=torch.float32, device=dev) if need_Ad else None
        Z = torch.empty((N, Kc), dtype=torch.float32, device=dev) if ns is not None else None
        Pm = torch.empty((N, Kc), dtype=torch.float32, device=dev) if slope is not None else None
        _lib.check(self.lib.ganb200_style_tail_bwd(self._p(ga), self._p(u), self._p(gu) if slope is not None else None,
                                                   self._p(ns), self._p(slope),
                                                   slope.numel() if slope is not None else 0, self._p(Ad), self._p(Z),
                                                   self._p(Pm), N, Kc, HW, self._p(rng_state), int(layer), st))
        return gu, Ad, Z, Pm

    def modconv_bwd(self, x, w, s, sinv, gu, pad, alpha, need_x=True, need_s=True, need_w=True, engine=None):
        """fused backward of the modulated conv -> (gx, gs1, gw) or None when the shape / engine is not eligible"""
        N, C, H, W = x.shape
        Kout, _, R, S = w.shape
        d = self._desc(N, C, H, W, Kout, R, S, 1, pad, 1, alpha, engine)
        if d.engine == 1 or sinv is None or not self.lib.ganb200_modconv_bwd_eligible(ctypes.byref(d)):
            return None
        st = self._prep(x, w, s, sinv, gu)
        if CONV_CENSUS is not None:
            for op in (DGRAD, WGRAD):
                key = (op, N, C, H, W, Kout, R, S, 1, pad, 1, op == WGRAD, True, False, "tc")
                CONV_CENSUS[key] = CONV_CENSUS.get(key, 0) + 1
        gx = torch.empty_like(x) if need_x else None
        gs1 = torch.empty((N, C), dtype=torch.float32, device=x.device) if need_s else None
        gw = torch.empty_like(w) if need_w else None
        nb = self.lib.ganb200_modconv_bwd_workspace_bytes(ctypes.byref(d))
        ws = self._workspace(nb, x.device)
        wp = self.packed_weight(w, d, DGRAD) if (need_x or need_s) else None
        if wp is not None:
            _lib.check(self.lib.ganb200_modconv_bwd_pk(ctypes.byref(d), self._p(x), self._p(w), self._p(wp), self._p(s),
                                                       self._p(sinv), self._p(gu), self._p(gx), self._p(gs1),
                                                       self._p(gw), self._p(ws), nb, st))
            return gx, gs1, gw
        _lib.check(self.lib.ganb200_modconv_bwd(ctypes.byref(d), self._p(x), self._p(w), self._p(s), self._p(sinv),
                                                self._p(gu), self._p(gx), self._p(gs1), self._p(gw), self._p(ws), nb,
                                                st))
        return gx, gs1, gw

    def scale_reduce(self, dxs, s, x, need_gx=True, need_gs=True):
        N, C, HW = _nchw(dxs)
        st = self._prep(dxs, s, x)
        gx = torch.empty_like(dxs) if need_gx else None
        gs1 = torch.empty((N, C), dtype=torch.float32, device=dxs.device) if need_gs else None
        _lib.check(self.lib.ganb200_scale_reduce(self._p(dxs), self._p(s), self._p(x), self._p(gx), self._p(gs1), N, C,
                                                 HW, st))
        return gx, gs1

    # -- grouped launches (gan_b200/group.py, csrc/group.cu) ------------------------------------------
    def linear_group(self, plan, a_base, b_base, o_base):
        from . import group as GR
        st = self._prep(a_base, b_base, o_base)
        tab, n, blocks = GR.glin_table(plan, o_base.device)
        _lib.check(self.lib.ganb200_linear_group(self._p(tab), n, blocks, self._p(a_base), self._p(b_base),
                                                 self._p(o_base), st))
        return o_base

    def bn1d_group_stats(self, plan, x_base, N, total_c):
        from . import group as GR
        st = self._prep(x_base)
        tab, n, blocks = GR.gbn_table(plan, x_base.device)
        mean = torch.empty((total_c,), dtype=torch.float32, device=x_base.device)
        var = torch.empty((total_c,), dtype=torch.float32, device=x_base.device)
        _lib.check(self.lib.ganb200_bn1d_group_stats(self._p(tab), n, blocks, self._p(x_base), N, self._p(mean),
                                                     self._p(var), st))
        return mean, var

    def bn1d_group_apply(self, plan, x_base, N, mean, var, eps, momentum, count_total):
        from . import group as GR
        st = self._prep(x_base, mean, var)
        tab, n, blocks = GR.gbn_table(plan, x_base.device)
        y = torch.empty_like(x_base)
        invstd = torch.empty_like(mean)
        _lib.check(self.lib.ganb200_bn1d_group_apply(self._p(tab), n, blocks, self._p(x_base), self._p(y), N,
                                                     self._p(mean), self._p(var), self._p(invstd), eps, momentum,
                                                     int(count_total), st))
        return y, invstd

    def bn1d_group_bwd_reduce(self, plan, gy_base, x_base, N, mean, invstd, total_c, has_slope):
        from . import group as GR
        st = self._prep(gy_base, x_base, mean, invstd)
        tab, n, blocks = GR.gbn_table(plan, x_base.device)
        s1, s2 = torch.empty_like(mean), torch.empty_like(mean)
        gsl = torch.empty_like(mean) if has_slope else None
        _lib.check(self.lib.ganb200_bn1d_group_bwd_reduce(self._p(tab), n, blocks, self._p(gy_base), self._p(x_base), N,
                                                          self._p(mean), self._p(invstd), self._p(s1), self._p(s2),
                                                          self._p(gsl), st))
        return s1, s2, gsl

    def bn1d_group_bwd_apply(self, plan, gy_base, x_base, N, mean, invstd, s1, s2, count_total, total_c):
        from . import group as GR
        st = self._prep(gy_base, x_base, mean, invstd, s1, s2)
        tab, n, blocks = GR.gbn_table(plan, x_base.device)
        gx = torch.empty_like(x_base)
        gbias = torch.empty_like(mean)
        _lib.check(self.lib.ganb200_bn1d_group_bwd_apply(self._p(tab), n, blocks, self._p(gy_base), self._p(x_base),
                                                         self._p(gx), N, self._p(mean), self._p(invstd), self._p(s1),
                                                         self._p(s2), int(count_total), self._p(gbias), st))
        return gx, gbias

    # -- optimiser / misc -------------------------------------------------------------------------
    def adam_blocks(self, tab, block_tab, n_blocks, chunk, lr, b1, b2, eps, wd, decoupled, step, grad_scale,
                    step_dev=None):
        """tab: uint8 device tensor holding ganb200_adam_entry[]; block_tab: int32 (n_blocks,2); step_dev: int64 (1,)"""
        st = ctypes.c_void_p(torch.cuda.current_stream(tab.device).cuda_stream)
        LAUNCH_COUNT[0] += 1
        _lib.check(self.lib.ganb200_multi_tensor_adam_blocks(self._p(tab), self._p(block_tab), n_blocks, chunk, lr,
                                                             b1, b2, eps, wd, int(decoupled), step,
                                                             self._p(step_dev), grad_scale, st))

    def multi_tensor_copy(self, tab, block_tab, n_blocks, chunk, stream=None):
        """tab: device table of ganb200_adam_entry (p = dst, g = src); one launch copies every entry"""
        st = ctypes.c_void_p((stream or torch.cuda.current_stream(tab.device)).cuda_stream)
        LAUNCH_COUNT[0] += 1
        _lib.check(self.lib.ganb200_multi_tensor_copy(self._p(tab), self._p(block_tab), n_blocks, chunk, st))

    def gather_flip(self, cache, idx, flips):
        """cache:[n,C,H,W] fp32 on the device, idx: int64[B], flips: uint8/bool[B] -> [B,C,H,W] (flipped where asked)"""
        st = self._prep(cache)
        B = idx.numel()
        _, C, H, W = cache.shape
        if idx.dtype != torch.int64 or not idx.is_cuda or not flips.is_cuda:
            raise _lib.GanB200Error("gather_flip: idx must be int64 and idx / flips on the device")
        flips = flips.to(torch.uint8)
        out = torch.empty((B, C, H, W), dtype=torch.float32, device=cache.device)
        _lib.check(self.lib.ganb200_gather_flip(self._p(cache), ctypes.c_void_p(idx.data_ptr()),
                                                ctypes.c_void_p(flips.data_ptr()), self._p(out), B, C, H, W, st))
        return out

    def make_grid_u8(self, x, nrow=8, padding=2):
        """torchvision make_grid(normalize=True) + save_image quantisation on the device -> uint8 [GH, GW, 3]"""
        N, C, H, W = x.shape
        st = self._prep(x)
        LAUNCH_COUNT[0] += 1
        mm = torch.empty((2,), dtype=torch.float32, device=x.device)
        _lib.check(self.lib.ganb200_minmax(self._p(x), x.numel(), self._p(mm), st))
        xm = min(nrow, N)
        ym = (N + xm - 1) // xm
        gh, gw = (H, W) if N == 1 else ((H + padding) * ym + padding, (W + padding) * xm + padding)
        out = torch.empty((gh, gw, 3), dtype=torch.uint8, device=x.device)
        _lib.check(self.lib.ganb200_make_grid_u8(self._p(x), self._p(mm), ctypes.c_void_p(out.data_ptr()), N, C, H, W,
                                                 nrow, padding, st))
        return out

    def counter_add(self, ctr, inc):
        st = ctypes.c_void_p(torch.cuda.current_stream(ctr.device).cuda_stream)
        LAUNCH_COUNT[0] += 1
        _lib.check(self.lib.ganb200_counter_add(self._p(ctr), int(inc), st))

    def chan_copy(self, src, src_c0, dst, dst_c0, ncopy, accumulate=False):
        """dst[:, dst_c0:dst_c0+ncopy] (+)= src[:, src_c0:src_c0+ncopy] on [N,C,...] tensors."""
        N, Cs, HW = _nchw(src)
        N2, Cd, HW2 = _nchw(dst)
        assert N == N2 and HW == HW2, (src.shape, dst.shape)
        st = self._prep(src, dst)
        _lib.check(self.lib.ganb200_chan_copy(self._p(src), self._p(dst), N, Cs, src_c0, Cd, dst_c0, ncopy, HW,
                                              int(accumulate), st))
        return dst

    def mixup(self, t, f, lam):
        st = self._prep(t, f)
        out = torch.empty_like(t)
        _lib.check(self.lib.ganb200_mixup(self._p(t), self._p(f), float(lam), self._p(out), t.numel(), st))
        return out

    def mixup_wgan_f64(self, label, t, f):
        B, C, H, W = t.shape
        st = self._prep(label, t, f, allow_f64=True)   # the reference's float64 label quirk (units/mixup.py:19-27)
        out = torch.empty(t.shape, dtype=torch.float64, device=t.device)
        _lib.check(self.lib.ganb200_mixup_wgan_f64(self._p(label), self._p(t), self._p(f), self._p(out), B, C, H, W, st))
        return out

    def clamp_(self, x, lo, hi):
        bump_param_epoch()   # in-place parameter update torch does not see (wgan weight clipping)
        st = self._prep(x)
        _lib.check(self.lib.ganb200_clamp_(self._p(x), x.numel(), lo, hi, st))
        return x

    def spectral_power_iter(self, w2d, u, v, n_iter, eps):
        """u, v advance in place (torch's order: u first); returns sigma = u.(W v) as a 1-element tensor."""
        rows, cols = w2d.shape
        st = self._prep(w2d, u, v)
        sigma = torch.empty((1,), dtype=torch.float32, device=w2d.device)
        _lib.check(self.lib.ganb200_spectral_power_iter(self._p(w2d), self._p(u), self._p(v), self._p(sigma), rows,
                                                        cols, n_iter, eps, st))
        return sigma

    def spectral_power_iter_batched(self, table, n_jobs, max_smem_floats, n_iter, eps):
        """table: device uint8 tensor holding n_jobs ganb200_spectral_job records (gan_b200.spectral.JobTable)."""
        st = self._prep(table)
        _lib.check(self.lib.ganb200_spectral_power_iter_batched(self._p(table), n_jobs, max_smem_floats, n_iter, eps,
                                                                st))

    def spectral_scale(self, w, sigma):
        st = self._prep(w, sigma)
        out = torch.empty_like(w)
        _lib.check(self.lib.ganb200_spectral_scale(self._p(w), self._p(sigma), self._p(out), w.numel(), st))
        return out

    def spectral_scale_bwd(self, g, w, u, v, sigma):
        st = self._prep(g, w, u, v, sigma)
        LAUNCH_COUNT[0] += 1
        gw = torch.empty_like(w)
        scratch = torch.empty((1,), dtype=torch.float32, device=w.device)
        _lib.check(self.lib.ganb200_spectral_scale_bwd(self._p(g), self._p(w), self._p(u), self._p(v), self._p(sigma),
                                                       self._p(scratch), self._p(gw), w.shape[0],
                                                       w.numel() // w.shape[0], st))
        return gw


_backend = None


def _nvtx_wrap(obj):
    """GANB200_NVTX=1: an NVTX range per kernel-class call (range name = the backend method: conv_fprop, bn_small_fwd,
    style_tail_bwd, linear_group, ...), so that an nsys / ncu timeline groups the launches of one logical operator.
    Off by default: ~1 us of host time per call."""
    import functools
    nv = torch.cuda.nvtx
    for name in dir(obj):
        fn = getattr(obj, name)
        if name.startswith("_") or not callable(fn) or name in ("bn_small_ok",):
            continue

        def make(f, label):
            @functools.wraps(f)
            def wrapped(*a, **k):
                nv.range_push(label)
                try:
                    return f(*a, **k)
                finally:
                    nv.range_pop()
            return wrapped
        setattr(obj, name, make(fn, "ganb200." + name))
    return obj


def backend():
    global _backend
    if _backend is None:
        _backend = CudaBackend()
        if os.environ.get("GANB200_NVTX") == "1":
            _nvtx_wrap(_backend)
    return _backend


def set_backend(b):
    """tests only: install another kernel backend (returns the previous one)."""
    global _backend
    prev = _backend
    _backend = b
    return prev
