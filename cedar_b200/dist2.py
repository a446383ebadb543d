"""ctypes binding of the block-decomposed 2D solver (``cdr2_dist_*`` of include/cedar_capi.h,
cedar_b200/host/dist2.cc): one process per GPU, ghost exchange by pulls out of the neighbours'
device memory over NVLink.  The library does no inter-process communication itself; this wrapper
moves the 64-byte CUDA IPC handles once with ``torch.distributed`` (any backend: it is a host
all-gather of a few bytes) -- what an MPI application would do with MPI_Allgather.

Also here: the block-wise problem generators of the bench and the tests (host side, numpy)."""
import ctypes as C
import json
import os
import time
import math

import numpy as np

from . import solver as _cs   # loads libbmgx.so + libcedar_b200.so (raises if missing: no CPU fallback)

lib = _cs.lib
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
HANDLE_BYTES = 64

lib.cdr2_dist_create.argtypes = [C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int, _ip, C.c_int, C.c_int, C.c_char_p,
                                 C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
lib.cdr2_dist_sweep_rects.argtypes = [C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                      _ip, C.c_int, _ip]
lib.cdr2_dist_destroy.argtypes = [C.c_void_p]
lib.cdr2_dist_destroy.restype = None
lib.cdr2_dist_info.argtypes = [C.c_void_p, _ip, _ip, _ip, _ip, _ip]
lib.cdr2_dist_block_extent.argtypes = [C.c_void_p, C.c_int, C.c_int, _ip, _ip, _ip, _ip]
lib.cdr2_dist_export_handles.argtypes = [C.c_void_p, C.c_void_p]
lib.cdr2_dist_connect.argtypes = [C.c_void_p, C.c_void_p]
lib.cdr2_dist_set_operator.argtypes = [C.c_void_p, C.c_int, _dp]
lib.cdr2_dist_setup.argtypes = [C.c_void_p]
lib.cdr2_dist_upload.argtypes = [C.c_void_p, C.c_int, C.c_char_p, _dp]
lib.cdr2_dist_download.argtypes = [C.c_void_p, C.c_int, C.c_char_p, _dp]
lib.cdr2_dist_level_array.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p, _dp]
lib.cdr2_dist_upload_level.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p, _dp]
lib.cdr2_dist_extended_array.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p, _dp]
lib.cdr2_dist_level_ptr.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p]
lib.cdr2_dist_level_ptr.restype = C.c_void_p
lib.cdr2_dist_halo_exchanger.argtypes = [C.c_void_p]
lib.cdr2_dist_halo_exchanger.restype = C.c_void_p
lib.cdr2_dist_exchange.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
lib.cdr2_dist_vcycle.argtypes = [C.c_void_p]
lib.cdr2_dist_sync.argtypes = [C.c_void_p]
lib.cdr2_dist_solve.argtypes = [C.c_void_p, C.c_int, C.c_double, _dp, C.c_int, _ip, _dp]
lib.cdr2_dist_time_vcycles.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float)]
lib.cdr2_dist_vcycle_stream.argtypes = [C.c_void_p, C.c_int, C.POINTER(_dp), C.POINTER(_dp)]
lib.cdr2_dist_profile_ops.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_float), C.c_int, _ip]
lib.cdr2_dist_plan_json.argtypes = [C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int, C.c_int,
                                    C.c_int, C.c_char_p, C.c_char_p, C.c_size_t, _ip]
lib.cdr2_dist_geometry.argtypes = [C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _ip]

_check = _cs._check


def _P(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_dp)


# ------------------------------------------------------------------ host-only views of the schedule
# Sweeps can be split so that the ghost exchange runs beside the interior relaxation (blocks at least this wide; 0 = never).
# Default off: with one exchange per level left (deep halos) the latency it hides (~6 us per exchange) is smaller than
# what the split costs (a frames launch + fork/join, ~15-25 us per level): 16385x8193 on 8 GPUs 1.344 ms per cycle without,
# 1.398 ms with (profiles/r02_overlap_8gpu.txt); 2 GPUs 1.280 / 1.301 ms.
SPLIT_MIN = 0


def plan(nx, ny, px, py, nst=5, pre=1, post=1, agglomerate=513, halo=15, what="cycle", split_min=SPLIT_MIN, relaxation="point"):
    """(operation list, number of distributed levels): include/cedar/2d/mpi/plan.h; needs no device"""
    buf = C.create_string_buffer(1 << 20)
    nd = C.c_int()
    _check(lib.cdr2_dist_plan_json(nx, ny, px, py, nst, pre, post, relaxation.encode(), agglomerate, halo, split_min, what.encode(),
                                   buf, len(buf), C.byref(nd)))
    return json.loads(buf.value.decode()), nd.value


def sweep_rects(nx, ny, px, py, ndist, halo, lvl, block, ncolours, part):
    """rectangles (i0, j0, ni, nj, wi0, wi1, wj0, wj1) of a split sweep: part 1 = interior, 2 = frames"""
    out = (C.c_int * 40)()
    n = C.c_int()
    _check(lib.cdr2_dist_sweep_rects(nx, ny, px, py, ndist, halo, lvl, block, ncolours, part, out, 5, C.byref(n)))
    return [tuple(out[8 * k:8 * k + 8]) for k in range(n.value)]


def geometry(nx, ny, px, py, ndist, halo, lvl, block):
    """dict: own extent, halo widths, array extents and the coarse window of a block on a level"""
    v = (C.c_int * 12)()
    _check(lib.cdr2_dist_geometry(nx, ny, px, py, ndist, halo, lvl, block, v))
    k = ("x0", "y0", "nx", "ny", "wl", "wr", "wb", "wt", "ox", "oy", "iic", "jjc")
    g = dict(zip(k, list(v)))
    g["ii"], g["jj"] = g["nx"] + g["wl"] + g["wr"], g["ny"] + g["wb"] + g["wt"]
    return g


def grid_decomp(n, nprocs):
    """include/cedar/decomp.h:59-86: prime factors, largest first, onto the currently longest dimension"""
    n, dec, f, p = list(n), [1, 1], [], 2
    m = nprocs
    while m > 1:
        while m % p == 0:
            f.append(p)
            m //= p
        p += 1
    for fac in sorted(f, reverse=True):
        ind = 1 if n[1] > n[0] else 0
        dec[ind] *= fac
        n[ind] //= fac
    return dec


# ------------------------------------------------------------------ the solver
class TorchComm:
    """moves the IPC handles between the ranks: one host all-gather over torch.distributed"""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def allgather_bytes(self, b):
        out = [None] * self.world
        self.dist.all_gather_object(out, bytes(b), group=self.group)
        return out

    def barrier(self):
        self.dist.barrier(group=self.group)


class DistSolver2:
    """Block-decomposed BoxMG solver.  `blocks`: ids owned by this process (all of them when comm is None: every block
    on this process's GPU).  relaxation: "point", "line-x", "line-y" or "line-xy" (lines that cross block cuts are
    transposed over peer memory and solved whole, include/cedar/2d/mpi/plan.h)."""

    def __init__(self, nx, ny, decomp_shape, nst=5, blocks=None, nrelax_pre=1, nrelax_post=1, agglomerate=513, halo=15,
                 cuda_graph=True, comm=None, split_min=SPLIT_MIN, relaxation="point"):
        px, py = decomp_shape
        self.nx, self.ny, self.px, self.py, self.nst, self.comm = nx, ny, px, py, nst, comm
        self.pre, self.post, self.agglomerate = nrelax_pre, nrelax_post, agglomerate
        nb = px * py
        if blocks is None:
            blocks = [comm.rank] if comm is not None else list(range(nb))
        self.owned = list(blocks)
        ids = (C.c_int * len(self.owned))(*self.owned)
        self.h = C.c_void_p()
        self.split_min = split_min
        self.relaxation = relaxation
        _t = [time.perf_counter()]
        _timing = os.environ.get("CDR_DIST_TIMING") == "1"
        _check(lib.cdr2_dist_create(nx, ny, px, py, nst, len(self.owned), ids, nrelax_pre, nrelax_post, relaxation.encode(),
                                    agglomerate, halo, split_min, int(cuda_graph), C.byref(self.h)))
        _t.append(time.perf_counter())
        v = [C.c_int() for _ in range(5)]
        _check(lib.cdr2_dist_info(self.h, *[C.byref(t) for t in v]))
        self.ndist, self.nlevels, self.halo, self.exchanges_per_cycle, self.ops_per_cycle = [t.value for t in v]
        # exchange the IPC handles of the arenas (nothing to do when every block lives here)
        handles = C.create_string_buffer(nb * HANDLE_BYTES)
        if comm is not None and comm.world > 1:
            _check(lib.cdr2_dist_export_handles(self.h, handles))
            mine = handles.raw
            allh = comm.allgather_bytes(b"".join(mine[b * HANDLE_BYTES:(b + 1) * HANDLE_BYTES] for b in self.owned))
            owners = comm.allgather_bytes(np.asarray(self.owned, dtype=np.int32).tobytes())
            merged = bytearray(nb * HANDLE_BYTES)
            for hb, ob in zip(allh, owners):
                for k, b in enumerate(np.frombuffer(ob, dtype=np.int32)):
                    merged[b * HANDLE_BYTES:(b + 1) * HANDLE_BYTES] = hb[k * HANDLE_BYTES:(k + 1) * HANDLE_BYTES]
            handles = C.create_string_buffer(bytes(merged), nb * HANDLE_BYTES)
        _t.append(time.perf_counter())
        _check(lib.cdr2_dist_connect(self.h, handles))
        _t.append(time.perf_counter())
        if comm is not None:
            comm.barrier()        # every arena is mapped (and cleared) before anyone signals into it
        _t.append(time.perf_counter())
        if _timing and (comm is None or comm.rank == 0):
            print("cdr2 dist timing (ms): create %.1f  handles %.1f  connect %.1f  barrier %.1f" %
                  tuple((b - a) * 1e3 for a, b in zip(_t[:-1], _t[1:])), flush=True)

    def extent(self, lvl, block):
        v = [C.c_int() for _ in range(4)]
        _check(lib.cdr2_dist_block_extent(self.h, lvl, block, *[C.byref(t) for t in v]))
        return tuple(t.value for t in v)

    def set_operator(self, block, so):
        so = np.ascontiguousarray(so, dtype=np.float64)
        _, _, nxl, nyl = self.extent(0, block)
        assert so.shape == (self.nst, nyl + 2, nxl + 2), (so.shape, (self.nst, nyl + 2, nxl + 2))
        _check(lib.cdr2_dist_set_operator(self.h, block, _P(so)))

    def setup(self):
        _check(lib.cdr2_dist_setup(self.h))
        if self.comm is not None:
            self.comm.barrier()

    def upload(self, block, which, a):
        _check(lib.cdr2_dist_upload(self.h, block, which.encode(), _P(np.ascontiguousarray(a, dtype=np.float64))))

    def download(self, block, which="x"):
        _, _, nxl, nyl = self.extent(0, block)
        out = np.zeros((nyl + 2, nxl + 2))
        _check(lib.cdr2_dist_download(self.h, block, which.encode(), _P(out)))
        return out

    def level_array(self, block, lvl, which):
        """own points of a level array: (planes, nyl, nxl), or (nyl, nxl) for grid functions"""
        _, _, nxl, nyl = self.extent(lvl, block)
        npl = {"A": self.nst if lvl == 0 else 5, "P": 8, "SOR": 2}.get(which, 1)
        out = np.zeros((npl, nyl, nxl))
        _check(lib.cdr2_dist_level_array(self.h, block, lvl, which.encode(), _P(out)))
        return out if npl > 1 else out[0]

    def upload_level(self, block, lvl, which, a):
        """own points of a level array: (nyl, nxl) for grid functions, (planes, nyl, nxl) for the operator (which = "A")"""
        _check(lib.cdr2_dist_upload_level(self.h, block, lvl, which.encode(), _P(np.ascontiguousarray(a, dtype=np.float64))))

    def extended_array(self, block, lvl, which):
        """the whole halo-extended array of a level: (jj, ii) or (planes, jj, ii)"""
        g = geometry(self.nx, self.ny, self.px, self.py, self.ndist, self.halo, lvl, block)
        npl = (self.nst if lvl == 0 else 5) if which == "A" else 1
        out = np.zeros((npl, g["jj"], g["ii"]))
        _check(lib.cdr2_dist_extended_array(self.h, block, lvl, which.encode(), _P(out)))
        return out if npl > 1 else out[0]

    def exchange(self, lvl, which, block=None):
        """one ghost exchange of an array through the reference's halo_exchanger_base interface (exchange_func /
        exchange_sten with the reference's level number k = ndist + 1 - lvl); collective"""
        block = self.owned[0] if block is None else block
        ptr = lib.cdr2_dist_level_ptr(self.h, block, lvl, which.encode())
        if not ptr:
            raise _cs.CedarError(lib.cdr_last_error().decode())
        _check(lib.cdr2_dist_exchange(self.h, self.ndist + 1 - lvl, 1 if which == "A" else 0, ptr))
        _check(lib.cdr2_dist_sync(self.h))

    def vcycle(self, sync=True):
        _check(lib.cdr2_dist_vcycle(self.h))
        if sync:
            _check(lib.cdr2_dist_sync(self.h))

    def sync(self):
        _check(lib.cdr2_dist_sync(self.h))

    def solve(self, maxiter=10, tol=1e-8):
        """on the resident x, b; returns (history of relative residuals, initial residual norm)"""
        hist = np.zeros(max(maxiter, 1))
        n, r0 = C.c_int(), C.c_double()
        _check(lib.cdr2_dist_solve(self.h, maxiter, tol, _P(hist), len(hist), C.byref(n), C.byref(r0)))
        return hist[:n.value], r0.value

    def time_vcycles(self, warmup, n):
        ms = C.c_float()
        _check(lib.cdr2_dist_time_vcycles(self.h, warmup, n, C.byref(ms)))
        return ms.value

    def profile_ops(self, n=5):
        """[(operation dict, ms)] for one cycle: eager, host-synchronised per operation (diagnostic)"""
        ops, _ = plan(self.nx, self.ny, self.px, self.py, self.nst, self.pre, self.post, self.agglomerate, self.halo, "cycle",
                      self.split_min, self.relaxation)
        out = (C.c_float * len(ops))()
        m = C.c_int()
        _check(lib.cdr2_dist_profile_ops(self.h, n, out, len(ops), C.byref(m)))
        assert m.value == len(ops)
        return list(zip(ops, list(out)))

    def vcycle_stream(self, bs, xs):
        n = len(bs)
        pb = (_dp * n)(*[_P(a) for a in bs])
        px = (_dp * n)(*[_P(a) for a in xs])
        _check(lib.cdr2_dist_vcycle_stream(self.h, n, pb, px))

    def close(self):
        if getattr(self, "h", None):
            lib.cdr2_dist_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


# ------------------------------------------------------------------ block-wise problem generators (host, numpy)
class Blocks:
    """extents of the blocks of a px x py decomposition on level 0 without a solver (aligned to 2**ndist)"""

    def __init__(self, nx, ny, px, py, ndist, halo=1):
        self.nx, self.ny, self.px, self.py, self.ndist, self.halo = nx, ny, px, py, ndist, halo

    def extent(self, block, lvl=0):
        g = geometry(self.nx, self.ny, self.px, self.py, self.ndist, self.halo, lvl, block)
        return g["x0"], g["y0"], g["nx"], g["ny"]


def gallery_fe_block(gn, ext):
    """own entries of gallery::fe (src/2d/gallery.cc:78-113) on a block: centre 8, every coupling between two
    interior points 1"""
    x0, y0, nxl, nyl = ext
    so = np.zeros((5, nyl + 2, nxl + 2))
    gi = x0 + np.arange(nxl)[None, :]
    gj = y0 + np.arange(nyl)[:, None]
    own = so[:, 1:-1, 1:-1]
    own[0] = 8.0
    own[1] = (gi >= 1) * np.ones((nyl, 1))
    own[2] = (gj >= 1) * np.ones((1, nxl))
    own[3] = (gi >= 1) * (gj >= 1)
    own[4] = (gi >= 1) * (gj >= 1)
    return so


def gallery_hetero_block(gn, ext, nblk=64, seed=20260411):
    """own entries of the heterogeneous 9-point operator of config 5: bilinear-FE-like stencil
    of -div(D grad u), D piecewise constant on an nblk x nblk checkerboard of log-uniform blocks D = 10**U(-3,3);
    edge couplings are the mean of D over the two cells they border, diagonal couplings the D of the cell they
    cross, the centre the sum of the eight couplings (Dirichlet: couplings to the boundary stay in the centre, as
    gallery::fe does)."""
    gnx, gny = gn
    rng = np.random.default_rng(seed)
    D = 10.0 ** rng.uniform(-3, 3, size=(nblk, nblk))
    x0, y0, nxl, nyl = ext

    def cell(ci, cj):       # coefficient of the cell whose lower-left grid node is (ci, cj), nodes 0..gn+1
        bi = np.clip(ci * nblk // (gnx + 1), 0, nblk - 1)
        bj = np.clip(cj * nblk // (gny + 1), 0, nblk - 1)
        return D[bj, bi]

    gi = (x0 + 1 + np.arange(nxl))[None, :] + np.zeros((nyl, 1), dtype=np.int64)     # node indices 1..gn of own points
    gj = (y0 + 1 + np.arange(nyl))[:, None] + np.zeros((1, nxl), dtype=np.int64)
    sw, se, nw, ne = cell(gi - 1, gj - 1), cell(gi, gj - 1), cell(gi - 1, gj), cell(gi, gj)
    cw, cs_, ce, cn = 0.5 * (sw + nw), 0.5 * (sw + se), 0.5 * (se + ne), 0.5 * (nw + ne)
    so = np.zeros((5, nyl + 2, nxl + 2))
    own = so[:, 1:-1, 1:-1]
    own[0] = cw + cs_ + ce + cn + sw + se + nw + ne
    own[1] = cw * (gi >= 2)
    own[2] = cs_ * (gj >= 2)
    own[3] = sw * ((gi >= 2) & (gj >= 2))          # KSW at S couples S and S-x-y: crosses the cell south-west of S
    own[4] = sw * ((gi >= 2) & (gj >= 2))          # KNW at S couples S-x and S-y: crosses the same cell
    return so


def rhs_sin_block(gn, ext):
    """examples/basic-2d-ser/poisson.cc:15-37 on a block"""
    gnx, gny = gn
    x0, y0, nxl, nyl = ext
    hx, hy = 1.0 / (gnx + 1), 1.0 / (gny + 1)
    x = ((x0 + 1 + np.arange(nxl)) * hx)[None, :]
    y = ((y0 + 1 + np.arange(nyl)) * hy)[:, None]
    out = np.zeros((nyl + 2, nxl + 2))
    pi = math.pi
    out[1:-1, 1:-1] = 8 * (pi * pi) * np.sin(2 * pi * x) * np.sin(2 * pi * y) * (hx * hy)
    return out


def split_blocks(a, extents):
    """cut the local arrays (own points + a zero ghost ring) out of a global array with ghost ring"""
    out = {}
    for b, (x0, y0, nxl, nyl) in extents.items():
        loc = np.array(a[..., y0:y0 + nyl + 2, x0:x0 + nxl + 2])
        loc[..., 0, :] = 0; loc[..., -1, :] = 0; loc[..., :, 0] = 0; loc[..., :, -1] = 0
        out[b] = loc
    return out
