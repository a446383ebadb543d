"""ctypes bindings of the host layer's C API (``include/cedar_capi.h``): build a
BoxMG solver on the GPU from host operators and run solves.  Used by the tests
and ``bench.py``; mirrors ``cdr2::solver<fsten>`` of the C++ layer."""
import ctypes as C
import json
import os

import numpy as np

from . import abi  # loads libbmgx.so first (RTLD_GLOBAL)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libcedar_b200.so")
if not os.path.exists(LIB_PATH):
    raise ImportError(f"{LIB_PATH} is missing: run __graft_entry__.build() (there is no CPU fallback)")

lib = C.CDLL(LIB_PATH)
_dp = C.POINTER(C.c_double)
lib.cdr_last_error.restype = C.c_char_p
lib.cdr_gpu_launches.restype = C.c_ulonglong
lib.cdr2_operator_data.restype = _dp
lib.cdr2_operator_data.argtypes = [C.c_void_p]
lib.cdr2_operator_create.argtypes = [C.c_int, C.c_uint, C.c_uint, C.POINTER(C.c_void_p)]
lib.cdr2_operator_gallery.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_double, C.c_double, C.POINTER(C.c_void_p)]
lib.cdr2_operator_destroy.argtypes = [C.c_void_p]
lib.cdr2_solver_create.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p)]
lib.cdr2_solver_destroy.argtypes = [C.c_void_p]
lib.cdr2_solver_nlevels.argtypes = [C.c_void_p]
lib.cdr2_solver_level_shape.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint), C.POINTER(C.c_uint)]
lib.cdr2_solver_level_array.argtypes = [C.c_void_p, C.c_int, C.c_char_p, _dp, C.c_size_t, C.POINTER(C.c_size_t)]
lib.cdr2_solver_solve.argtypes = [C.c_void_p, _dp, _dp, _dp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double)]
lib.cdr2_solver_vcycle.argtypes = [C.c_void_p, _dp, _dp]
lib.cdr2_solver_time_vcycles.argtypes = [C.c_void_p, _dp, C.c_int, C.c_int, C.POINTER(C.c_float)]
lib.cdr2_solver_time_solves.argtypes = [C.c_void_p, _dp, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_int)]
lib.cdr2_solve_host.argtypes = [C.c_int, C.c_uint, C.c_uint, _dp, _dp, _dp, C.c_char_p, _dp, C.c_int, C.POINTER(C.c_int)]
lib.cdr2_solver_buffers.argtypes = [C.c_void_p, C.POINTER(_dp), C.POINTER(_dp)]
lib.cdr2_solver_vcycle_pinned.argtypes = [C.c_void_p]
lib.cdr2_solver_solve_pinned.argtypes = [C.c_void_p, _dp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double)]
lib.cdr2_solver_matvec.argtypes = [C.c_void_p, _dp, _dp]
lib.cdr2_solver_vcycle0_pinned.argtypes = [C.c_void_p]
lib.cdr2_solver_vcycle_batch.argtypes = [C.c_void_p, C.c_int, C.POINTER(_dp), C.POINTER(_dp)]
lib.cdr_pinned_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
lib.cdr_pinned_free.argtypes = [C.c_void_p]
lib.cdr_set_log.argtypes = [C.c_char_p]
lib.cdr2_solver_vcycle_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]


class CedarError(RuntimeError):
    pass


class PinnedArrays:
    """float64 arrays in pinned host memory (cdr_pinned_alloc); ``arrays`` are numpy views, valid while this
    object lives."""

    def __init__(self, n, shape):
        self.ptrs, self.arrays = [], []
        count = int(np.prod(shape))
        for _ in range(n):
            p = C.c_void_p()
            if lib.cdr_pinned_alloc(C.byref(p), count * 8) != 0:
                raise CedarError(lib.cdr_last_error().decode())
            self.ptrs.append(p)
            a = np.ctypeslib.as_array(C.cast(p, _dp), shape=(count,)).reshape(shape)
            a[...] = 0.0
            self.arrays.append(a)

    def __iter__(self):
        return iter(self.arrays)

    def __getitem__(self, i):
        return self.arrays[i]

    def __len__(self):
        return len(self.arrays)

    def __del__(self):
        for p in self.ptrs:
            lib.cdr_pinned_free(p)
        self.ptrs = []


def _check(rc):
    if rc != 0:
        raise CedarError(lib.cdr_last_error().decode())


def _P(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_dp)


def set_log(levels=""):
    lib.cdr_set_log(levels.encode())


def gpu_launches():
    return int(lib.cdr_gpu_launches())


def make_config(relaxation="point", nrelax_pre=2, nrelax_post=1, maxiter=10, tol=1e-8, num_levels=None,
                cycle="v", cuda_graph=True, log=("error",), coarse_tail=None, periodic=None):
    conf = {"log": list(log),
            "solver": {"relaxation": relaxation, "cycle": {"type": cycle, "nrelax-pre": nrelax_pre, "nrelax-post": nrelax_post},
                       "max-iter": maxiter, "tol": tol, "cuda-graph": bool(cuda_graph)}}
    if num_levels:
        conf["solver"]["num-levels"] = int(num_levels)
    if coarse_tail is not None:
        conf["solver"]["coarse-tail"] = int(coarse_tail)
    if periodic is not None:
        conf["grid"] = {"periodic": [bool(p) for p in periodic]}
    return json.dumps(conf)


def _host_view(owner, p, shape):
    """numpy view of library-owned host memory that keeps `owner` (whose destructor frees it) alive as long as the
    view or anything sliced from it is"""
    n = int(np.prod(shape))
    buf = (C.c_double * n).from_address(C.addressof(p.contents))
    buf._owner = owner
    return np.frombuffer(buf, dtype=np.float64).reshape(shape)


class Operator2:
    """Host-side 2D stencil operator, shape (ndirs, ny+2, nx+2)."""

    def __init__(self, so=None, gallery=None, nx=None, ny=None, dx=1.0, dy=1.0):
        self.h = C.c_void_p()
        if gallery is not None:
            _check(lib.cdr2_operator_gallery(gallery.encode(), nx, ny, dx, dy, C.byref(self.h)))
            self.nst = 5 if gallery == "fe" else 3
            self.nx, self.ny = nx, ny
        else:
            so = np.ascontiguousarray(so, dtype=np.float64)
            self.nst, jj, ii = so.shape
            self.nx, self.ny = ii - 2, jj - 2
            _check(lib.cdr2_operator_create(self.nst, self.nx, self.ny, C.byref(self.h)))
            self.array()[...] = so

    def array(self):
        p = lib.cdr2_operator_data(self.h)
        if not p:
            raise CedarError(lib.cdr_last_error().decode())
        return _host_view(self, p, (self.nst, self.ny + 2, self.nx + 2))

    def __del__(self):
        if getattr(self, "h", None):
            lib.cdr2_operator_destroy(self.h)
            self.h = None


class Solver2:
    NDIRS = {"P": 8, "SOR0": 2, "SOR1": 2, "x": 1, "res": 1, "b": 1}

    def __init__(self, op, config_json=None, **kw):
        self.op = op
        self.h = C.c_void_p()
        cfg = config_json if config_json is not None else make_config(**kw)
        _check(lib.cdr2_solver_create(op.h, cfg.encode(), C.byref(self.h)))
        self.nlevels = lib.cdr2_solver_nlevels(self.h)

    def level_shape(self, lvl):
        nx, ny = C.c_uint(), C.c_uint()
        _check(lib.cdr2_solver_level_shape(self.h, lvl, C.byref(nx), C.byref(ny)))
        return nx.value, ny.value

    def level_array(self, lvl, which):
        nx, ny = self.level_shape(lvl)
        nd = self.NDIRS.get(which, (self.op.nst if lvl == 0 else 5))
        out = np.zeros((nd, ny + 2, nx + 2))
        w = C.c_size_t()
        _check(lib.cdr2_solver_level_array(self.h, lvl, which.encode(), _P(out), out.size, C.byref(w)))
        if w.value == 0:
            return None
        assert w.value == out.size, (w.value, out.size)
        return out if nd > 1 else out[0]

    def solve(self, b, x=None, maxiter_cap=1000):
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.zeros_like(b) if x is None else np.ascontiguousarray(x, dtype=np.float64)
        hist = np.zeros(maxiter_cap)
        n = C.c_int()
        r0 = C.c_double()
        _check(lib.cdr2_solver_solve(self.h, _P(b), _P(x), _P(hist), maxiter_cap, C.byref(n), C.byref(r0)))
        return x, hist[:n.value], r0.value

    def vcycle(self, x, b):
        _check(lib.cdr2_solver_vcycle(self.h, _P(b), _P(x)))
        return x

    def buffers(self):
        """(b, x): numpy views of the solver's pinned host staging buffers."""
        pb, px = _dp(), _dp()
        _check(lib.cdr2_solver_buffers(self.h, C.byref(pb), C.byref(px)))
        shape = (self.op.ny + 2, self.op.nx + 2)
        return np.ctypeslib.as_array(pb, shape=shape), np.ctypeslib.as_array(px, shape=shape)

    def vcycle_pinned(self):
        """upload b, x from the pinned buffers; one cycle; download x"""
        _check(lib.cdr2_solver_vcycle_pinned(self.h))

    def matvec(self, x):
        """y = A x with the fine-level operator"""
        y = np.zeros_like(x)
        _check(lib.cdr2_solver_matvec(self.h, _P(np.ascontiguousarray(x)), _P(y)))
        return y

    def vcycle0_pinned(self):
        """upload b from the pinned buffer; x0 = 0 on the device; one cycle; download x"""
        _check(lib.cdr2_solver_vcycle0_pinned(self.h))

    def pinned_grid_funcs(self, n):
        """n pinned host arrays of the solver's grid-function shape (freed with the returned owner)"""
        shape = (self.op.ny + 2, self.op.nx + 2)
        return PinnedArrays(n, shape)

    def vcycle_batch(self, bs, xs):
        """One cycle from x0 = 0 for each right-hand side of ``bs``, results into ``xs`` (lists of C-contiguous
        float64 arrays, ideally from pinned_grid_funcs); copies of consecutive right-hand sides overlap the cycles."""
        n = len(bs)
        assert len(xs) == n
        pb = (_dp * n)(*[_P(a) for a in bs])
        px = (_dp * n)(*[_P(a) for a in xs])
        _check(lib.cdr2_solver_vcycle_batch(self.h, n, pb, px))
        return xs

    def solve_pinned(self, maxiter_cap=1000):
        hist = np.zeros(maxiter_cap)
        n = C.c_int()
        r0 = C.c_double()
        _check(lib.cdr2_solver_solve_pinned(self.h, _P(hist), maxiter_cap, C.byref(n), C.byref(r0)))
        return hist[:n.value], r0.value

    def time_vcycles(self, b, warmup=3, n=10):
        ms = C.c_float()
        _check(lib.cdr2_solver_time_vcycles(self.h, _P(np.ascontiguousarray(b)), warmup, n, C.byref(ms)))
        return ms.value

    def time_solves(self, b, warmup=1, n=3):
        ms = C.c_float()
        it = C.c_int()
        _check(lib.cdr2_solver_time_solves(self.h, _P(np.ascontiguousarray(b)), warmup, n, C.byref(ms), C.byref(it)))
        return ms.value, it.value

    def __del__(self):
        if getattr(self, "h", None):
            lib.cdr2_solver_destroy(self.h)
            self.h = None


def solve_host(so, b, config_json=None, **kw):
    """One-shot end-to-end solve from host arrays (upload, setup, solve, download)."""
    so = np.ascontiguousarray(so, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    nst, jj, ii = so.shape
    x = np.zeros_like(b)
    hist = np.zeros(1000)
    n = C.c_int()
    cfg = config_json if config_json is not None else make_config(**kw)
    _check(lib.cdr2_solve_host(nst, ii - 2, jj - 2, _P(so), _P(b), _P(x), cfg.encode(), _P(hist), 1000, C.byref(n)))
    return x, hist[:n.value]


# ---------------------------------------------------------------------------- 3D
_vp = C.c_void_p
lib.cdr3_operator_data.restype = _dp
lib.cdr3_operator_data.argtypes = [_vp]
lib.cdr3_operator_create.argtypes = [C.c_int, C.c_uint, C.c_uint, C.c_uint, C.POINTER(_vp)]
lib.cdr3_operator_gallery.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_uint, C.POINTER(_vp)]
lib.cdr3_operator_destroy.argtypes = [_vp]
lib.cdr3_solver_create.argtypes = [_vp, C.c_char_p, C.POINTER(_vp)]
lib.cdr3_solver_destroy.argtypes = [_vp]
lib.cdr3_solver_nlevels.argtypes = [_vp]
lib.cdr3_solver_level_shape.argtypes = [_vp, C.c_int, C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.POINTER(C.c_uint)]
lib.cdr3_solver_level_array.argtypes = [_vp, C.c_int, C.c_char_p, _dp, C.c_size_t, C.POINTER(C.c_size_t)]
lib.cdr3_solver_solve.argtypes = [_vp, _dp, _dp, _dp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double)]
lib.cdr3_solver_vcycle.argtypes = [_vp, _dp, _dp]
lib.cdr3_solver_buffers.argtypes = [_vp, C.POINTER(_dp), C.POINTER(_dp)]
lib.cdr3_solver_vcycle_pinned.argtypes = [_vp]
lib.cdr3_solver_matvec.argtypes = [_vp, _dp, _dp]
lib.cdr3_solver_vcycle0_pinned.argtypes = [_vp]
lib.cdr3_solver_vcycle_batch.argtypes = [_vp, C.c_int, C.POINTER(_dp), C.POINTER(_dp)]
lib.cdr3_solver_time_vcycles.argtypes = [_vp, _dp, C.c_int, C.c_int, C.POINTER(C.c_float)]
lib.cdr3_solver_time_solves.argtypes = [_vp, _dp, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_int)]


lib.cdr3_plane_relax.argtypes = [_vp, C.c_char_p, C.c_char_p, C.c_int, _dp, _dp]


def plane_relax3(op, direction, x, b, up=False, plane_config=None):
    """one zebra plane sweep through the kernel registry (test/3d/test_planes.cc); x is updated in place"""
    cfg = json.dumps({"log": ["error"], "plane-config": plane_config or {}})
    _check(lib.cdr3_plane_relax(op.h, cfg.encode(), direction.encode(), int(up), _P(np.ascontiguousarray(b)), _P(x)))
    return x


class Operator3:
    """Host-side 3D stencil operator, shape (ndirs, nz+2, ny+2, nx+2); ndirs = 4 or 14."""

    def __init__(self, so=None, gallery=None, nx=None, ny=None, nz=None):
        self.h = C.c_void_p()
        if gallery is not None:
            _check(lib.cdr3_operator_gallery(gallery.encode(), nx, ny, nz, C.byref(self.h)))
            self.nst = 14 if gallery == "fe" else 4
            self.nx, self.ny, self.nz = nx, ny, nz
        else:
            so = np.ascontiguousarray(so, dtype=np.float64)
            self.nst, kk, jj, ii = so.shape
            self.nx, self.ny, self.nz = ii - 2, jj - 2, kk - 2
            _check(lib.cdr3_operator_create(self.nst, self.nx, self.ny, self.nz, C.byref(self.h)))
            self.array()[...] = so

    def array(self):
        p = lib.cdr3_operator_data(self.h)
        if not p:
            raise CedarError(lib.cdr_last_error().decode())
        return _host_view(self, p, (self.nst, self.nz + 2, self.ny + 2, self.nx + 2))

    def __del__(self):
        if getattr(self, "h", None):
            lib.cdr3_operator_destroy(self.h)
            self.h = None


class Solver3:
    """cdr3::solver<fsten> on the GPU (setup at construction)."""
    NDIRS = {"P": 26, "SOR0": 2, "x": 1, "res": 1, "b": 1}

    def __init__(self, op, config_json=None, **kw):
        self.op = op
        self.h = C.c_void_p()
        cfg = config_json if config_json is not None else make_config(**kw)
        _check(lib.cdr3_solver_create(op.h, cfg.encode(), C.byref(self.h)))
        self.nlevels = lib.cdr3_solver_nlevels(self.h)

    def level_shape(self, lvl):
        nx, ny, nz = C.c_uint(), C.c_uint(), C.c_uint()
        _check(lib.cdr3_solver_level_shape(self.h, lvl, C.byref(nx), C.byref(ny), C.byref(nz)))
        return nx.value, ny.value, nz.value

    def level_array(self, lvl, which):
        nx, ny, nz = self.level_shape(lvl)
        nd = self.NDIRS.get(which, (self.op.nst if lvl == 0 else 14))
        out = np.zeros((nd, nz + 2, ny + 2, nx + 2))
        w = C.c_size_t()
        _check(lib.cdr3_solver_level_array(self.h, lvl, which.encode(), _P(out), out.size, C.byref(w)))
        if w.value == 0:
            return None
        assert w.value == out.size, (w.value, out.size)
        return out if nd > 1 else out[0]

    def solve(self, b, x=None, maxiter_cap=1000):
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.zeros_like(b) if x is None else np.ascontiguousarray(x, dtype=np.float64)
        hist = np.zeros(maxiter_cap)
        n = C.c_int()
        r0 = C.c_double()
        _check(lib.cdr3_solver_solve(self.h, _P(b), _P(x), _P(hist), maxiter_cap, C.byref(n), C.byref(r0)))
        return x, hist[:n.value], r0.value

    def vcycle(self, x, b):
        _check(lib.cdr3_solver_vcycle(self.h, _P(b), _P(x)))
        return x

    def buffers(self):
        pb, px = _dp(), _dp()
        _check(lib.cdr3_solver_buffers(self.h, C.byref(pb), C.byref(px)))
        shape = (self.op.nz + 2, self.op.ny + 2, self.op.nx + 2)
        return np.ctypeslib.as_array(pb, shape=shape), np.ctypeslib.as_array(px, shape=shape)

    def vcycle_pinned(self):
        _check(lib.cdr3_solver_vcycle_pinned(self.h))

    def matvec(self, x):
        """y = A x with the fine-level operator"""
        y = np.zeros_like(x)
        _check(lib.cdr3_solver_matvec(self.h, _P(np.ascontiguousarray(x)), _P(y)))
        return y

    def vcycle0_pinned(self):
        """upload b from the pinned buffer; x0 = 0 on the device; one cycle; download x"""
        _check(lib.cdr3_solver_vcycle0_pinned(self.h))

    def pinned_grid_funcs(self, n):
        return PinnedArrays(n, (self.op.nz + 2, self.op.ny + 2, self.op.nx + 2))

    def vcycle_batch(self, bs, xs):
        """3D counterpart of Solver2.vcycle_batch"""
        n = len(bs)
        assert len(xs) == n
        pb = (_dp * n)(*[_P(a) for a in bs])
        px = (_dp * n)(*[_P(a) for a in xs])
        _check(lib.cdr3_solver_vcycle_batch(self.h, n, pb, px))
        return xs

    def time_vcycles(self, b, warmup=3, n=10):
        ms = C.c_float()
        _check(lib.cdr3_solver_time_vcycles(self.h, _P(np.ascontiguousarray(b)), warmup, n, C.byref(ms)))
        return ms.value

    def time_solves(self, b, warmup=1, n=3):
        ms = C.c_float()
        it = C.c_int()
        _check(lib.cdr3_solver_time_solves(self.h, _P(np.ascontiguousarray(b)), warmup, n, C.byref(ms), C.byref(it)))
        return ms.value, it.value

    def __del__(self):
        if getattr(self, "h", None):
            lib.cdr3_solver_destroy(self.h)
            self.h = None
