"""ctypes binding of the slab-decomposed 3D solver (``cdr3_dist_*`` of include/cedar_capi.h,
cedar_b200/host/dist3.cc): the grid is cut into slabs of xy planes, one per GPU; ghost planes are pulled out of the
two neighbours' device memory over NVLink.  Same protocol as cedar_b200/dist2.py."""
import ctypes as C
import json

import numpy as np

from . import solver as _cs
from .dist2 import HANDLE_BYTES, TorchComm   # noqa: F401

lib = _cs.lib
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_check = _cs._check

lib.cdr3_dist_create.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_int, _ip, C.c_int, C.c_int, C.c_int, C.c_int,
                                 C.c_int, C.POINTER(C.c_void_p)]
lib.cdr3_dist_destroy.argtypes = [C.c_void_p]
lib.cdr3_dist_destroy.restype = None
lib.cdr3_dist_info.argtypes = [C.c_void_p, _ip, _ip, _ip, _ip, _ip]
lib.cdr3_dist_slab_extent.argtypes = [C.c_void_p, C.c_int, C.c_int, _ip, _ip, _ip, _ip]
lib.cdr3_dist_export_handles.argtypes = [C.c_void_p, C.c_void_p]
lib.cdr3_dist_connect.argtypes = [C.c_void_p, C.c_void_p]
lib.cdr3_dist_set_operator.argtypes = [C.c_void_p, C.c_int, _dp]
lib.cdr3_dist_setup.argtypes = [C.c_void_p]
lib.cdr3_dist_upload.argtypes = [C.c_void_p, C.c_int, C.c_char_p, _dp]
lib.cdr3_dist_download.argtypes = [C.c_void_p, C.c_int, C.c_char_p, _dp]
lib.cdr3_dist_level_array.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p, _dp]
lib.cdr3_dist_upload_level.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p, _dp]
lib.cdr3_dist_extended_array.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p, _dp]
lib.cdr3_dist_level_ptr.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p]
lib.cdr3_dist_level_ptr.restype = C.c_void_p
lib.cdr3_dist_halo_exchanger.argtypes = [C.c_void_p]
lib.cdr3_dist_halo_exchanger.restype = C.c_void_p
lib.cdr3_dist_exchange.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
lib.cdr3_dist_vcycle.argtypes = [C.c_void_p]
lib.cdr3_dist_sync.argtypes = [C.c_void_p]
lib.cdr3_dist_solve.argtypes = [C.c_void_p, C.c_int, C.c_double, _dp, C.c_int, _ip, _dp]
lib.cdr3_dist_time_vcycles.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float)]
lib.cdr3_dist_plan_json.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p,
                                    C.c_char_p, C.c_size_t, _ip]
lib.cdr3_dist_geometry.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _ip]


def _P(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_dp)


def plan(nx, ny, nz, nslabs, nst=4, pre=1, post=1, agglomerate=129, halo=5, what="cycle"):
    """(operation list, distributed levels); host only"""
    buf = C.create_string_buffer(1 << 20)
    nd = C.c_int()
    _check(lib.cdr3_dist_plan_json(nx, ny, nz, nslabs, nst, pre, post, agglomerate, halo, what.encode(), buf, len(buf), C.byref(nd)))
    return json.loads(buf.value.decode()), nd.value


def geometry(nx, ny, nz, nslabs, ndist, halo, lvl, slab):
    v = (C.c_int * 8)()
    _check(lib.cdr3_dist_geometry(nx, ny, nz, nslabs, ndist, halo, lvl, slab, v))
    g = dict(zip(("nx", "ny", "z0", "nz", "wb", "wt", "oz", "kkc"), list(v)))
    g["ii"], g["jj"], g["kk"] = g["nx"] + 2, g["ny"] + 2, g["nz"] + g["wb"] + g["wt"]
    return g


class DistSolver3:
    """Slab-decomposed BoxMG solver, point relaxation.  `slabs`: ids owned by this process (all when comm is None)."""

    def __init__(self, nx, ny, nz, nslabs, nst=4, slabs=None, nrelax_pre=1, nrelax_post=1, agglomerate=129, halo=5,
                 cuda_graph=True, comm=None):
        self.nx, self.ny, self.nz, self.np, self.nst, self.comm = nx, ny, nz, nslabs, nst, comm
        if slabs is None:
            slabs = [comm.rank] if comm is not None else list(range(nslabs))
        self.owned = list(slabs)
        ids = (C.c_int * len(self.owned))(*self.owned)
        self.h = C.c_void_p()
        _check(lib.cdr3_dist_create(nx, ny, nz, nslabs, nst, len(self.owned), ids, nrelax_pre, nrelax_post, agglomerate, halo,
                                    int(cuda_graph), C.byref(self.h)))
        v = [C.c_int() for _ in range(5)]
        _check(lib.cdr3_dist_info(self.h, *[C.byref(t) for t in v]))
        self.ndist, self.nlevels, self.halo, self.exchanges_per_cycle, self.ops_per_cycle = [t.value for t in v]
        handles = C.create_string_buffer(nslabs * HANDLE_BYTES)
        if comm is not None and comm.world > 1:
            _check(lib.cdr3_dist_export_handles(self.h, handles))
            mine = handles.raw
            allh = comm.allgather_bytes(b"".join(mine[b * HANDLE_BYTES:(b + 1) * HANDLE_BYTES] for b in self.owned))
            owners = comm.allgather_bytes(np.asarray(self.owned, dtype=np.int32).tobytes())
            merged = bytearray(nslabs * HANDLE_BYTES)
            for hb, ob in zip(allh, owners):
                for k, b in enumerate(np.frombuffer(ob, dtype=np.int32)):
                    merged[b * HANDLE_BYTES:(b + 1) * HANDLE_BYTES] = hb[k * HANDLE_BYTES:(k + 1) * HANDLE_BYTES]
            handles = C.create_string_buffer(bytes(merged), nslabs * HANDLE_BYTES)
        _check(lib.cdr3_dist_connect(self.h, handles))
        if comm is not None:
            comm.barrier()

    def extent(self, lvl, slab):
        """(nx, ny, z0, nz) of a slab on a level"""
        v = [C.c_int() for _ in range(4)]
        _check(lib.cdr3_dist_slab_extent(self.h, lvl, slab, *[C.byref(t) for t in v]))
        return tuple(t.value for t in v)

    def set_operator(self, slab, so):
        so = np.ascontiguousarray(so, dtype=np.float64)
        nx, ny, _, nz = self.extent(0, slab)
        assert so.shape == (self.nst, nz + 2, ny + 2, nx + 2), so.shape
        _check(lib.cdr3_dist_set_operator(self.h, slab, _P(so)))

    def setup(self):
        _check(lib.cdr3_dist_setup(self.h))
        if self.comm is not None:
            self.comm.barrier()

    def upload(self, slab, which, a):
        _check(lib.cdr3_dist_upload(self.h, slab, which.encode(), _P(np.ascontiguousarray(a, dtype=np.float64))))

    def download(self, slab, which="x"):
        nx, ny, _, nz = self.extent(0, slab)
        out = np.zeros((nz + 2, ny + 2, nx + 2))
        _check(lib.cdr3_dist_download(self.h, slab, which.encode(), _P(out)))
        return out

    def level_array(self, slab, lvl, which):
        """own planes of A or P of a level: (entries, nzl, ny+2, nx+2)"""
        nx, ny, _, nz = self.extent(lvl, slab)
        ne = 26 if which == "P" else (self.nst if lvl == 0 else 14)
        out = np.zeros((ne, nz, ny + 2, nx + 2))
        _check(lib.cdr3_dist_level_array(self.h, slab, lvl, which.encode(), _P(out)))
        return out

    def _entries(self, lvl, which):
        return (self.nst if lvl == 0 else 14) if which == "A" else 1

    def upload_level(self, slab, lvl, which, a):
        """own planes of a level array: (entries, nzl, ny+2, nx+2)"""
        _check(lib.cdr3_dist_upload_level(self.h, slab, lvl, which.encode(), _P(np.ascontiguousarray(a, dtype=np.float64))))

    def extended_array(self, slab, lvl, which):
        """the whole halo-extended array of a level: (entries, kk, ny+2, nx+2)"""
        g = geometry(self.nx, self.ny, self.nz, self.np, self.ndist, self.halo, lvl, slab)
        out = np.zeros((self._entries(lvl, which), g["kk"], g["jj"], g["ii"]))
        _check(lib.cdr3_dist_extended_array(self.h, slab, lvl, which.encode(), _P(out)))
        return out

    def exchange(self, lvl, which, slab=None):
        """one ghost exchange through the reference's halo_exchanger_base interface (k = ndist + 1 - lvl); collective"""
        slab = self.owned[0] if slab is None else slab
        ptr = lib.cdr3_dist_level_ptr(self.h, slab, lvl, which.encode())
        if not ptr:
            raise _cs.CedarError(lib.cdr_last_error().decode())
        _check(lib.cdr3_dist_exchange(self.h, self.ndist + 1 - lvl, 1 if which == "A" else 0, ptr))
        _check(lib.cdr3_dist_sync(self.h))

    def vcycle(self, sync=True):
        _check(lib.cdr3_dist_vcycle(self.h))
        if sync:
            _check(lib.cdr3_dist_sync(self.h))

    def solve(self, maxiter=10, tol=1e-8):
        hist = np.zeros(max(maxiter, 1))
        n, r0 = C.c_int(), C.c_double()
        _check(lib.cdr3_dist_solve(self.h, maxiter, tol, _P(hist), len(hist), C.byref(n), C.byref(r0)))
        return hist[:n.value], r0.value

    def time_vcycles(self, warmup, n):
        ms = C.c_float()
        _check(lib.cdr3_dist_time_vcycles(self.h, warmup, n, C.byref(ms)))
        return ms.value

    def close(self):
        if getattr(self, "h", None):
            lib.cdr3_dist_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


def split_slabs(a, extents):
    """cut the local arrays (own planes + one zero ghost plane below and above) out of a global array with ghost layer"""
    out = {}
    for b, (_, _, z0, nz) in extents.items():
        loc = np.array(a[..., z0:z0 + nz + 2, :, :])
        loc[..., 0, :, :] = 0
        loc[..., -1, :, :] = 0
        out[b] = loc
    return out
