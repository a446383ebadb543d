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


def test_reference_c_interface_symbols_and_storage_conversion(tmp_path):
    """include/cedar/capi.h: every bmg2_* / bmg3_* / bmg_timer_save name of the reference's C interface is exported
    (include/cedar/2d/interface/c/operator.h:21-29, solver.h:13-15, topo.h:14-20 and the 3D twins), and
    bmg2_operator_set converts vertex-based signed entries to the symmetric storage as
    src/2d/interface/c/operator.cc:30-58 does: the 9-point FE stencil entered vertex by vertex must give gallery::fe"""
    import cedar_b200.solver as cs
    names = []
    for h in ("cedar/2d/interface/c/operator.h", "cedar/2d/interface/c/solver.h", "cedar/2d/interface/c/topo.h",
              "cedar/3d/interface/c/operator.h", "cedar/3d/interface/c/solver.h", "cedar/3d/interface/c/topo.h",
              "cedar/interface/c/timer.h"):
        names += _declared(h, "bmg")
    assert len(set(names)) >= 19
    assert not [n for n in names if not hasattr(cs.lib, n)]

    class Coord(C.Structure):
        _fields_ = [("i", C.c_uint), ("j", C.c_uint), ("dir", C.c_int)]

    L = cs.lib
    L.bmg2_topo_create.restype = C.c_void_p
    L.bmg2_topo_create.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.c_int, C.c_int]
    L.bmg2_operator_create.restype = C.c_void_p
    L.bmg2_operator_create.argtypes = [C.c_void_p]
    L.bmg2_operator_set.argtypes = [C.c_void_p, C.c_uint, C.POINTER(Coord), C.POINTER(C.c_double)]
    L.bmg2_operator_dump.argtypes = [C.c_void_p]
    L.bmg2_operator_destroy.argtypes = [C.c_void_p]
    L.bmg2_topo_destroy.argtypes = [C.c_void_p]
    nx, ny = 7, 5
    one_x, one_y = (C.c_uint * 1)(nx), (C.c_uint * 1)(ny)
    assert not L.bmg2_topo_create(None, nx, ny, one_x, one_y, 2, 1)          # a process grid is refused, loudly
    topo = L.bmg2_topo_create(None, nx, ny, one_x, one_y, 1, 1)
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        op = L.bmg2_operator_create(topo)
        assert op
        # BMG2_C=0, W=1, S=2, SW=3, NW=4, SE=5, N=6, NE=7, E=8 (include/cedar/2d/base_types.h)
        dirs = {0: (0, 0), 1: (-1, 0), 2: (0, -1), 3: (-1, -1), 4: (-1, 1), 5: (1, -1), 6: (0, 1), 7: (1, 1), 8: (1, 0)}
        ent = [(i, j, d) for j in range(ny) for i in range(nx) for d, (di, dj) in dirs.items()
               if 0 <= i + di < nx and 0 <= j + dj < ny]
        coords = (Coord * len(ent))(*[Coord(i, j, d) for i, j, d in ent])
        vals = (C.c_double * len(ent))(*[8.0 if d == 0 else -1.0 for _, _, d in ent])
        L.bmg2_operator_set(op, len(ent), coords, vals)
        assert all(v == (8.0 if d == 0 else 1.0) for v, (_, _, d) in zip(vals, ent))     # negated in place, as the reference does
        L.bmg2_operator_dump(op)
        got = np.zeros((5, ny + 2, nx + 2))
        for line in open(tmp_path / "op0-0.txt"):
            ij, rest = line.split(" ", 2)[:2], line.split(" ")[2:]
            i, j = int(ij[0].rstrip(",")), int(ij[1])
            got[:, j, i] = [float(v) for v in rest]
        assert_bitwise(got, oracle.gallery2("fe", nx, ny), "vertex-based input -> symmetric storage")
        L.bmg2_operator_destroy(op)
    finally:
        os.chdir(cwd)
        L.bmg2_topo_destroy(topo)
