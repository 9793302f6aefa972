"""Collected last on purpose (a failure here must not hide another test).

ganb200_dwconv2d_wgrad_static — the allocation-free variant of the depthwise wgrad (include/ganb200.h: no counterpart in
the reference) — against the existing entry point
ganb200_dwconv2d_wgrad, through the C-ABI (ctypes, raw device pointers), on two calls in a row: the built-in scratch keeps
the previous call's contents, so the second call (a different, smaller shape) must not see them."""
import ctypes
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _lib():
    from gan_b200 import _lib
    return _lib.load()


def test_static_scratch_rejects_oversized_shapes_before_any_launch():
    """host-side check only (no GPU needed): 8 000 channels need 288 000 bytes of partials, the scratch holds 262 144"""
    lib = _lib()
    cap = lib.ganb200_dwconv2d_wgrad_static_max_bytes()
    assert cap == 256 * 1024
    assert lib.ganb200_dwconv2d_wgrad_workspace_bytes(1, 8000) > cap
    rc = lib.ganb200_dwconv2d_wgrad_static(None, None, None, 1, 8000, 8, 8, 3, 3, 1, 1, 1.0, None)
    assert rc == -5  # GANB200_WORKSPACE_TOO_SMALL
    msg = lib.ganb200_last_error().decode()
    assert "static scratch" in msg and "288000" in msg and "262144" in msg


@pytest.mark.gpu
def test_static_scratch_wgrad_equals_workspace_wgrad_on_two_calls_in_a_row():
    lib = _lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    assert lib.ganb200_device_check(0) == 0
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    gen = torch.Generator().manual_seed(11)
    # (N, C, H, W, stride): ShuffleNetV2 x2.0 stage shapes of discriminator_7_wgangpsn, then a smaller call
    for N, C, H, W, stride in [(16, 244, 32, 32, 2), (4, 24, 16, 16, 1)]:
        P = (H + 2 - 3) // stride + 1
        x = torch.randn(N, C, H, W, generator=gen).to(dev)
        dy = torch.randn(N, C, P, P, generator=gen).to(dev)
        want = torch.empty(C, 1, 3, 3, device=dev)
        got = torch.full((C, 1, 3, 3), float("nan"), device=dev)
        nb = lib.ganb200_dwconv2d_wgrad_workspace_bytes(N, C)
        assert 0 < nb <= lib.ganb200_dwconv2d_wgrad_static_max_bytes()
        ws = torch.empty(nb, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        assert lib.ganb200_dwconv2d_wgrad(p(x), p(dy), p(want), p(ws), nb, N, C, H, W, 3, 3, stride, 1, 0.7, stream) == 0
        before = torch.cuda.memory_allocated(dev)
        rc = lib.ganb200_dwconv2d_wgrad_static(p(x), p(dy), p(got), N, C, H, W, 3, 3, stride, 1, 0.7, stream)
        assert rc == 0, lib.ganb200_last_error()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before
        assert torch.isfinite(got).all()
        rel = float((got - want).norm() / want.norm())
        assert rel < 2e-5, rel  # the tolerance of tests/test_gpu_kernels.py for this kernel (same kernels: expect 0)
