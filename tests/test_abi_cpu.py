"""CPU tests (no GPU): the C-ABI libraries load and export every symbol their
headers declare; host-only logic (gallery, config) works; anything that needs
the device fails loudly instead of falling back."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle
from tests.util import assert_bitwise

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header, prefix):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(" + prefix + r"[0-9a-zA-Z_]*)\s*\(", txt)))


def test_libbmgx_exports_every_declared_symbol():
    import cedar_b200.abi as abi
    names = _declared("bmgx.h", "bmgx")
    assert len(names) >= 40
    missing = [n for n in names if not hasattr(abi.lib, n)]
    assert not missing, missing
    # and the ctypes table covers every compute entry point
    compute = [n for n in names if re.match(r"bmgx[23]_", n)]
    assert not [n for n in compute if n not in abi.SIGNATURES]
    assert b"sm_100a" in abi.lib.bmgx_version()


def test_libcedar_exports_every_declared_symbol():
    import cedar_b200.solver as cs
    names = _declared("cedar_capi.h", "cdr")
    missing = [n for n in names if not hasattr(cs.lib, n)]
    assert not missing, missing


def test_get_bc_mapping():
    """BMG_get_bc, src/2d/ftn/BMG_get_bc.f90:13-22"""
    import cedar_b200.abi as abi
    want = {0: 0, 1: 2, 2: 1, 3: 3, 4: 5, 5: 6, 6: 7, 7: 8}
    for mask, ibc in want.items():
        out = C.c_int(-1)
        abi.lib.bmgx_get_bc(mask, C.byref(out))
        assert out.value == ibc


@pytest.mark.parametrize("name,nx,ny,dx,dy", [("poisson", 400, 400, 1, 1), ("poisson", 31, 64, 1, 1),
                                               ("fe", 37, 37, 1, 1), ("diag_diffusion", 50, 20, 1.0, 1e-4)])
def test_gallery_matches_oracle(name, nx, ny, dx, dy):
    """host-layer gallery (src/2d/gallery.cc) is host-only code: runs without a GPU"""
    import cedar_b200.solver as cs
    op = cs.Operator2(gallery=name, nx=nx, ny=ny, dx=dx, dy=dy)
    assert_bitwise(op.array(), oracle.gallery2(name, nx, ny, dx, dy), name)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import cedar_b200.abi as abi
    import cedar_b200.solver as cs
    assert abi.lib.bmgx_device_count() == 0
    a = np.zeros((5, 6, 6))
    p = a.ctypes.data_as(C.c_void_p)
    st = abi.lib.bmgx2_residual(0, p, p, p, p, 6, 6, 0, 0, 5, 0, 0, 0, 0, None)
    assert st == 1, "a compute call without a device must return BMGX_ERR_CUDA"
    with pytest.raises(cs.CedarError):
        cs.Solver2(cs.Operator2(gallery="poisson", nx=15, ny=15))
