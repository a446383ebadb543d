"""CPU oracle (TEST INFRASTRUCTURE ONLY).

ctypes loader for ``oracle/_build/liboracle.so`` -- the plain-C restatement of
the reference's serial BoxMG path (see ``oracle/bmg2.c``).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl
reference`` legs may import this package; the product (``cedar_b200``) never
does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "liboracle.so")

dp = C.POINTER(C.c_double)


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("bmg2.c", "ml2.c", "bmg3.c", "ml3.c", "oracle.h", "Makefile")]
    srcs = [s for s in srcs if os.path.exists(s)]
    if (not force and os.path.exists(_LIB)
            and all(os.path.getmtime(_LIB) >= os.path.getmtime(s) for s in srcs)):
        return _LIB
    subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        _lib = C.CDLL(_LIB)
        _lib.orc2_l2norm.restype = C.c_double
        _lib.orc2_solver_create.restype = C.c_void_p
        _lib.orc2_solver_level_ptr.restype = dp
        if hasattr(_lib, "orc3_l2norm"):
            _lib.orc3_l2norm.restype = C.c_double
            _lib.orc3_solver_create.restype = C.c_void_p
            _lib.orc3_solver_level_ptr.restype = dp
    return _lib


def P(a):
    """double* of a C-contiguous float64 numpy array (no copy)."""
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"], (a.dtype, a.flags)
    return a.ctypes.data_as(dp)


# ---------------------------------------------------------------------------
# numpy-facing helpers.  Arrays are stored as numpy arrays of shape
# (ndirs, ny+2, nx+2) / (ny+2, nx+2): C order of that shape == the reference's
# Fortran order (i fastest, then j, then dir).
# ---------------------------------------------------------------------------

def gallery2(name, nx, ny, dx=1.0, dy=1.0):
    L = lib()
    nst = 5 if name == "fe" else 3
    so = np.zeros((nst, ny + 2, nx + 2))
    if name == "poisson":
        L.orc2_gallery_poisson(P(so), nx, ny)
    elif name == "diag_diffusion":
        L.orc2_gallery_diag_diffusion(P(so), nx, ny, C.c_double(dx), C.c_double(dy))
    elif name == "fe":
        L.orc2_gallery_fe(P(so), nx, ny)
    else:
        raise ValueError(name)
    return so


def rhs_sin2(nx, ny):
    b = np.zeros((ny + 2, nx + 2))
    lib().orc2_rhs_sin(P(b), nx, ny)
    return b


def exact_sin2(nx, ny):
    q = np.zeros((ny + 2, nx + 2))
    lib().orc2_exact_sin(P(q), nx, ny)
    return q


RELAX = {"point": 0, "line-x": 1, "line-y": 2, "line-xy": 3}


class Solver2:
    """Mirror of cdr2::solver<fsten> on the CPU oracle."""

    def __init__(self, so, relaxation="point", nrelax_pre=2, nrelax_post=1,
                 num_levels=-1, min_coarse=3, cycle="v", periodic=(False, False)):
        self.so = np.ascontiguousarray(so)
        nst, jj, ii = self.so.shape
        self.nx, self.ny = ii - 2, jj - 2
        lib().orc2_solver_create_per.restype = C.c_void_p
        self.h = lib().orc2_solver_create_per(P(self.so), self.nx, self.ny, nst, RELAX[relaxation],
                                              nrelax_pre, nrelax_post, num_levels, min_coarse,
                                              int(bool(periodic[0])) + 2 * int(bool(periodic[1])))
        if not self.h:
            raise ValueError("oracle solver could not be created")
        self.h = C.c_void_p(self.h)
        self.nlevels = lib().orc2_solver_nlevels(self.h)
        if cycle == "f":       # include/cedar/cycle/fcycle.h
            lib().orc2_solver_set_cycle(self.h, 1)

    def level_shape(self, lvl):
        nx, ny = C.c_int(), C.c_int()
        lib().orc2_solver_level_shape(self.h, lvl, C.byref(nx), C.byref(ny))
        return nx.value, ny.value

    def level_array(self, lvl, which):
        names = {"A": 0, "P": 1, "x": 2, "res": 3, "b": 4, "SOR0": 5, "SOR1": 6}
        nx, ny = self.level_shape(lvl)
        nd = {"A": (3 if (lvl == 0 and self.so.shape[0] == 3) else 5), "P": 8, "SOR0": 2, "SOR1": 2}.get(which, 1)
        ptr = lib().orc2_solver_level_ptr(self.h, lvl, names[which])
        if not ptr:
            return None
        a = np.ctypeslib.as_array(ptr, shape=(nd, ny + 2, nx + 2))
        return a if nd > 1 else a[0]

    def vcycle(self, x, b):
        lib().orc2_solver_vcycle(self.h, P(x), P(b))

    def solve(self, b, x=None, maxiter=10, tol=1e-8):
        if x is None:
            x = np.zeros_like(b)
        hist = np.zeros(maxiter)
        res0 = C.c_double()
        it = lib().orc2_solver_solve(self.h, P(b), P(x), maxiter, C.c_double(tol), P(hist), C.byref(res0))
        return x, hist[:it], res0.value

    def __del__(self):
        try:
            lib().orc2_solver_destroy(self.h)
        except Exception:
            pass


# ---------------------------------------------------------------------------
# 3D: arrays of shape (ndirs, nz+2, ny+2, nx+2) / (nz+2, ny+2, nx+2)
# ---------------------------------------------------------------------------

def gallery3(name, nx, ny, nz):
    L = lib()
    nst = 14 if name == "fe" else 4
    so = np.zeros((nst, nz + 2, ny + 2, nx + 2))
    if name == "poisson":
        L.orc3_gallery_poisson(P(so), nx, ny, nz)
    elif name == "fe":
        L.orc3_gallery_fe(P(so), nx, ny, nz)
    else:
        raise ValueError(name)
    return so


def rhs_sin3(nx, ny, nz):
    b = np.zeros((nz + 2, ny + 2, nx + 2))
    lib().orc3_rhs_sin(P(b), nx, ny, nz)
    return b


def exact_sin3(nx, ny, nz):
    q = np.zeros((nz + 2, ny + 2, nx + 2))
    lib().orc3_exact_sin(P(q), nx, ny, nz)
    return q


class Solver3:
    """Mirror of cdr3::solver<fsten> (point relaxation) on the CPU oracle."""

    def __init__(self, so, nrelax_pre=2, nrelax_post=1, num_levels=-1, min_coarse=3):
        self.so = np.ascontiguousarray(so)
        nst, kk, jj, ii = self.so.shape
        self.nx, self.ny, self.nz = ii - 2, jj - 2, kk - 2
        self.h = C.c_void_p(lib().orc3_solver_create(P(self.so), self.nx, self.ny, self.nz, nst, nrelax_pre,
                                                     nrelax_post, num_levels, min_coarse))
        self.nlevels = lib().orc3_solver_nlevels(self.h)

    def level_shape(self, lvl):
        nx, ny, nz = C.c_int(), C.c_int(), C.c_int()
        lib().orc3_solver_level_shape(self.h, lvl, C.byref(nx), C.byref(ny), C.byref(nz))
        return nx.value, ny.value, nz.value

    def level_array(self, lvl, which):
        names = {"A": 0, "P": 1, "x": 2, "res": 3, "b": 4, "SOR0": 5}
        nx, ny, nz = self.level_shape(lvl)
        nd = {"A": (4 if (lvl == 0 and self.so.shape[0] == 4) else 14), "P": 26, "SOR0": 2}.get(which, 1)
        ptr = lib().orc3_solver_level_ptr(self.h, lvl, names[which])
        if not ptr:
            return None
        a = np.ctypeslib.as_array(ptr, shape=(nd, nz + 2, ny + 2, nx + 2))
        return a if nd > 1 else a[0]

    def vcycle(self, x, b):
        lib().orc3_solver_vcycle(self.h, P(x), P(b))

    def solve(self, b, x=None, maxiter=10, tol=1e-8):
        if x is None:
            x = np.zeros_like(b)
        hist = np.zeros(maxiter)
        res0 = C.c_double()
        it = lib().orc3_solver_solve(self.h, P(b), P(x), maxiter, C.c_double(tol), P(hist), C.byref(res0))
        return x, hist[:it], res0.value

    def __del__(self):
        try:
            lib().orc3_solver_destroy(self.h)
        except Exception:
            pass
