"""CPU: the C-ABI shared library loads and exports every symbol include/ganb200.h declares, and the ctypes
table in gan_b200/_lib.py covers exactly that set.  No compute call is made (there is no GPU here)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ganb200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ganb200_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as ge
    ge.build()
    from gan_b200 import _lib
    return _lib.load()


def test_header_table_and_exports_agree(lib):
    from gan_b200 import _lib
    names = declared_functions()
    assert len(names) >= 30
    assert sorted(_lib.SIGNATURES) == names
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ganb200.h but not exported by libganb200.so"


def test_exports_are_plain_c(lib):
    from gan_b200 import _lib
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.LIB_PATH], text=True)
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    assert set(declared_functions()) <= exported
    assert not any("torch" in s or "at::" in s for s in exported)


def test_abi_version_and_error_text(lib):
    assert lib.ganb200_abi_version() >= 1
    # invalid descriptor: reported through the status code + thread-local text, no CUDA needed
    from gan_b200._lib import ConvDesc
    d = ConvDesc(1, 1, 4, 4, 1, 3, 3, 1, 1, 1, 9, 9, 1.0, 1)
    rc = lib.ganb200_conv2d_fprop(ctypes.byref(d), None, None, None, None, None, None, None, 0, None)
    assert rc == -1  # GANB200_BAD_SHAPE
    assert b"P/Q" in lib.ganb200_last_error()


def test_cuda_backend_refuses_cpu_tensors(lib):
    import torch
    from gan_b200 import kernels as KN
    from gan_b200._lib import GanB200Error
    k = KN.CudaBackend()
    with pytest.raises(GanB200Error):
        k.ew_unary(KN.U_SIGMOID, torch.zeros(4))


def test_sass_has_tcgen05_and_tma():
    """The conv engine must be real Blackwell code: UTC*MMA (tcgen05.mma), LDTM (tcgen05.ld), UTMALDG (TMA)."""
    from gan_b200 import _lib
    sass = subprocess.check_output(["cuobjdump", "-sass", _lib.LIB_PATH], text=True)
    assert "UTCHMMA" in sass or "UTCMMA" in sass
    assert "LDTM" in sass
    assert "UTMALDG" in sass
