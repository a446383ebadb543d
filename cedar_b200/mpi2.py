"""Block-decomposed 2D BoxMG across GPUs -- the B200 counterpart of the reference's
``cdr2::mpi::solver`` (include/cedar/2d/mpi/solver.h) with its MPI halo exchange
(src/2d/mpi/halo.cc, MSG/Tausch) replaced by NCCL send/recv over NVLink.

Layout.  The global (nx, ny) grid is cut into px x py blocks following the reference's
``grid_decomp`` (include/cedar/decomp.h:59-86: prime factors, largest first, onto the
currently longest dimension) and ``block_partition::low(id) = id*n/p``
(include/cedar/mpi/block_partition.h:14-17), with every block boundary rounded down
to a multiple of 2**ndist so that each block starts on a coarse point of all ``ndist``
distributed levels (for the 2**k+1 grids of the headline configs the rounding is a
no-op and the blocks are exactly the reference's).  A block then keeps the index
parity of the serial grid on every level, so the SERIAL kernels of ``include/bmgx.h``
run unchanged on the local arrays ((nxl+2) x (nyl+2), one ghost layer) and the
decomposed solve is arithmetically identical, point by point, to the serial one --
the partition-independent "serial hierarchy" of SURVEY.md section 8(e).  Ghost layers
are exchanged where the reference's MPI kernels exchange them: after every colour
pass of a sweep, after the residual, after the coarse-grid correction, and between
the stages of the setup.  Levels whose global size drops to ``agglomerate`` or below
are gathered onto rank 0 and handled by the serial, CUDA-graph-captured
``cdr2::solver`` (the reference's redist / coarse-level agglomeration).

One process per GPU (``torch.distributed``, backend nccl); the same code also runs
all blocks inside one process (``LocalTransport``) -- used by the single-GPU parity
tests, which compare it bitwise with the serial solver.

The compute backend is the C ABI (``CudaKernels``); there is no CPU fallback.  The
CPU tests of the host logic inject ``tests/oracle_backend.py`` explicitly.
"""
import ctypes as C
import math
import types

import numpy as np
import torch


# ------------------------------------------------------------------ decomposition
def _factor(n):
    f, p = [], 2
    while n > 1:
        while n % p == 0:
            f.append(p)
            n //= p
        p += 1
    return sorted(f, reverse=True)


def grid_decomp(n, nprocs):
    """include/cedar/decomp.h:59-86"""
    n = list(n)
    dec = [1] * len(n)
    for fac in _factor(nprocs):
        ind = max(range(len(n)), key=lambda j: (n[j], -j))   # first of the longest
        dec[ind] *= fac
        n[ind] //= fac
    return dec


def coarse_n(n):
    return int((n - 1) / 2.0 + 1)     # include/cedar/2d/solver.h:86-87


class Decomp2:
    """Block extents on every distributed level.  lo/n are 0-based first global
    interior index and count, per level, per block coordinate."""

    def __init__(self, nx, ny, px, py, ndist):
        self.nx, self.ny, self.px, self.py, self.ndist = nx, ny, px, py, ndist
        self.gn = [(nx, ny)]
        for _ in range(ndist):
            self.gn.append((coarse_n(self.gn[-1][0]), coarse_n(self.gn[-1][1])))
        al = 1 << ndist
        self.lo = []      # lo[l][dim][coord]
        for lvl in range(ndist + 1):
            per_dim = []
            for dim, (n0, p) in enumerate(((nx, px), (ny, py))):
                lo0 = [((c * n0 // p) // al) * al for c in range(p)]
                per_dim.append([v >> lvl for v in lo0] + [self.gn[lvl][dim]])
            self.lo.append(per_dim)
        for dim in range(2):
            for c in range((px, py)[dim]):
                if self.lo[ndist][dim][c + 1] - self.lo[ndist][dim][c] < 2:
                    raise ValueError("blocks too small for %d distributed levels" % ndist)

    @property
    def nblocks(self):
        return self.px * self.py

    def coords(self, b):
        return b % self.px, b // self.px        # x fastest, like the reference's topology

    def block(self, cx, cy):
        return cy * self.px + cx

    def extent(self, lvl, b):
        cx, cy = self.coords(b)
        lx, ly = self.lo[lvl][0], self.lo[lvl][1]
        return lx[cx], ly[cy], lx[cx + 1] - lx[cx], ly[cy + 1] - ly[cy]

    def neighbour(self, b, dx, dy):
        cx, cy = self.coords(b)
        cx, cy = cx + dx, cy + dy
        if cx < 0 or cx >= self.px or cy < 0 or cy >= self.py:
            return None
        return self.block(cx, cy)


# ------------------------------------------------------------------ compute backend
class CudaKernels:
    """The C-ABI kernels (include/bmgx.h) on torch CUDA tensors, enqueued on `stream`."""

    def __init__(self, stream):
        from . import abi          # raises if libbmgx.so is missing: no CPU fallback
        if not torch.cuda.is_available():
            raise RuntimeError("CudaKernels needs a CUDA device (there is no CPU fallback)")
        self.abi, self.L = abi, abi.lib
        self.stream = stream
        self.s = C.c_void_p(stream.cuda_stream)
        self._norm_work = None
        self.launches = 0

    @staticmethod
    def _p(t):
        return C.c_void_p(t.data_ptr())

    def _ck(self, rc, n=1):
        self.abi.check(rc)
        self.launches += n

    def recip(self, A, sor):
        nst, jj, ii = A.shape
        self._ck(self.L.bmgx2_SETUP_recip(self._p(A), self._p(sor), ii, jj, nst, 2, self.s))

    def relax_pass(self, A, b, x, sor, updown, p):
        nst, jj, ii = A.shape
        self._ck(self.L.bmgx2_relax_GS_pass(1, self._p(A), self._p(b), self._p(x), self._p(sor), ii, jj, 1,
                                            1 if nst == 3 else 0, nst, 2, 1, updown, 0, p, self.s))

    def residual(self, A, b, x, r):
        nst, jj, ii = A.shape
        self._ck(self.L.bmgx2_residual(0, self._p(A), self._p(b), self._p(x), self._p(r), ii, jj, 0,
                                       1 if nst == 3 else 0, nst, 0, 0, 0, 0, self.s))

    def restrict(self, r, bc, P):
        jj, ii = r.shape
        jjc, iic = bc.shape
        self._ck(self.L.bmgx2_restrict(self._p(r), self._p(bc), self._p(P), ii, jj, iic, jjc, 0, self.s))

    def interp_add(self, x, xc, r, A, P):
        nst, jj, ii = A.shape
        jjc, iic = xc.shape
        self._ck(self.L.bmgx2_interp_add(self._p(x), self._p(xc), self._p(r), self._p(A), self._p(P), iic, jjc, ii, jj,
                                         nst, 0, self.s))

    def setup_interp_stage(self, A, P, stage):
        nst, jj, ii = A.shape
        _, jjc, iic = P.shape
        self._ck(self.L.bmgx2_SETUP_interp_OI_stage(self._p(A), None, self._p(P), ii, jj, iic, jjc,
                                                    1 if nst == 3 else 0, nst, 0, 0, stage, self.s))

    def relax_oop(self, A, b, x, xout, sor, updown):
        """one whole sweep, all colours in a single pass over memory (out of place)"""
        nst, jj, ii = A.shape
        self._ck(self.L.bmgx2_relax_GS_oop(1, self._p(A), self._p(b), self._p(x), self._p(xout), self._p(sor), ii, jj, 1,
                                           1 if nst == 3 else 0, nst, 2, 1, updown, 0, self.s))

    def galerkin(self, A, Ac, P):
        nst, jj, ii = A.shape
        _, jjc, iic = P.shape
        self._ck(self.L.bmgx2_SETUP_ITLI_ex(self._p(A), self._p(Ac), self._p(P), ii, jj, iic, jjc,
                                            1 if nst == 3 else 0, nst, 0, self.s))

    def line_copy(self, a, mask, direction, index, buf, to_buf):
        if a.dim() == 2:
            npl, (jj, ii) = 1, a.shape
        else:
            npl, jj, ii = a.shape
        self._ck(self.L.bmgx2_line_copy(self._p(a), ii, jj, npl, mask, direction, index, self._p(buf), int(to_buf),
                                        self.s))

    def sumsq(self, r, out):
        """sum of squares of the interior of r -> out (1-element device tensor)"""
        if self._norm_work is None:
            self._norm_work = torch.empty(self.L.bmgx_norm_work_len(), dtype=torch.float64, device=r.device)
        jj, ii = r.shape
        self._ck(self.L.bmgx2_sumsq(self._p(r), ii, jj, self._p(self._norm_work), self._p(out), self.s), 2)


# ------------------------------------------------------------------ transports
class LocalTransport:
    """All blocks live in this process: a message is a device-to-device copy."""
    rank, world = 0, 1

    def __init__(self, decomp):
        self.owned = list(range(decomp.nblocks))

    def owner(self, b):
        return 0

    def run(self, msgs):
        # msgs: list of (src_block, dst_block, tag, sendbuf | None, recvbuf | None)
        box = {(s, d, tag): sb for s, d, tag, sb, rb in msgs if sb is not None}
        for s, d, tag, sb, rb in msgs:
            if rb is not None:
                rb.copy_(box[(s, d, tag)])

    def allreduce_sum(self, t):
        return t


class DistTransport:
    """One block per rank over torch.distributed (nccl between GPUs; gloo in the CPU
    tests of the host logic).  Ghost lines travel as batched isend/irecv pairs."""

    def __init__(self, decomp, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world != decomp.nblocks:
            raise ValueError("need one rank per block: %d ranks, %d blocks" % (self.world, decomp.nblocks))
        self.owned = [self.rank]

    def owner(self, b):
        return b

    def run(self, msgs):
        dist = self.dist
        ops = []
        for s, d, tag, sb, rb in msgs:
            if sb is not None and rb is not None:      # cannot happen with one block per rank
                rb.copy_(sb)
            elif sb is not None:
                ops.append(dist.P2POp(dist.isend, sb, d, self.group))
            elif rb is not None:
                ops.append(dist.P2POp(dist.irecv, rb, s, self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()

    def allreduce_sum(self, t):
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return t


# ------------------------------------------------------------------ the solver
LL, LR, LA, LB, LSW, LNW, LNE, LSE = (1 << k for k in range(8))       # planes of CI as bit masks
DOWN, UP = 0, 1


def compute_num_levels(nx, ny, min_coarse=3):
    """include/cedar/2d/solver.h:57-73"""
    ng = 0
    while True:
        ng += 1
        if min((nx - 1) // (1 << ng) + 1, (ny - 1) // (1 << ng) + 1) < min_coarse:
            return ng


class _Level:
    def __init__(self):
        self.A, self.P, self.x, self.b, self.res, self.sor = {}, {}, {}, {}, {}, {}


class DistSolver2:
    """Block-decomposed BoxMG solver with point relaxation (see the module docstring).

    so_local: {block id: array (nst, nyl+2, nxl+2)} for the blocks this process owns --
    the block's own stencil entries; ghost entries are filled by the solver."""

    def __init__(self, so_local, nx, ny, decomp_shape, nrelax_pre=2, nrelax_post=1, agglomerate=513,
                 min_coarse=3, ndist=None, transport=None, kernels=None, device=None, coarse_factory=None,
                 fused_sweep=True, fuse_min=64):
        self.nx, self.ny = nx, ny
        self.npre, self.npost = nrelax_pre, nrelax_post
        self.nlevels = compute_num_levels(nx, ny, min_coarse)
        px, py = decomp_shape
        if ndist is None:
            ndist, n = 0, (nx, ny)
            while ndist < self.nlevels - 1 and min(n) > agglomerate:
                ndist += 1
                n = (coarse_n(n[0]), coarse_n(n[1]))
        self.ndist = ndist
        self.dec = Decomp2(nx, ny, px, py, ndist)
        self.T = transport if transport is not None else LocalTransport(self.dec)
        self.owned = self.T.owned
        if kernels is None:
            self.device = torch.device(device if device is not None else "cuda")
            self.stream = torch.cuda.Stream(device=self.device)
            kernels = CudaKernels(self.stream)
        else:
            self.device = torch.device(device if device is not None else "cpu")
            self.stream = None
        self.K = kernels
        self.nst = next(iter(so_local.values())).shape[0]
        self.coarse_factory = coarse_factory
        self.exchanges = 0
        self.fused_sweep, self.fuse_min = fused_sweep, fuse_min
        self.ext = {}
        with self._on_stream():
            self._allocate(so_local)
            self._setup()
            if fused_sweep:
                for lvl in range(self.ndist):
                    self._build_ext(lvl)

    # -- helpers -------------------------------------------------------------
    def _on_stream(self):
        import contextlib
        return torch.cuda.stream(self.stream) if self.stream is not None else contextlib.nullcontext()

    def _zeros(self, *shape):
        return torch.zeros(shape, dtype=torch.float64, device=self.device)

    def _allocate(self, so_local):
        self.lv = [_Level() for _ in range(self.ndist + 1)]
        for lvl, L in enumerate(self.lv):
            for b in self.owned:
                _, _, nxl, nyl = self.dec.extent(lvl, b)
                jj, ii = nyl + 2, nxl + 2
                if lvl == 0:
                    a = torch.as_tensor(np.ascontiguousarray(so_local[b]), dtype=torch.float64).to(self.device)
                    assert a.shape == (self.nst, jj, ii), (a.shape, (self.nst, jj, ii))
                    L.A[b] = a.contiguous()
                else:
                    L.A[b] = self._zeros(5, jj, ii)
                    L.P[b] = self._zeros(8, jj, ii)
                L.x[b], L.b[b], L.res[b] = self._zeros(jj, ii), self._zeros(jj, ii), self._zeros(jj, ii)
                L.sor[b] = self._zeros(2, jj, ii)
        self._ss = self._zeros(1)

    # -- ghost exchange ------------------------------------------------------
    def _lines(self, lvl, arrs, phase, pick, masks):
        """move one line per (block, side): pick(side, n_own) -> (send index, recv index)"""
        K, dec = self.K, self.dec
        msgs, unpack = [], []
        for b in self.owned:
            a = arrs[b]
            npl = 1 if a.dim() == 2 else a.shape[0]
            mask = masks if masks is not None else (1 << npl) - 1
            nsel = bin(mask & ((1 << npl) - 1)).count("1")
            _, _, nxl, nyl = dec.extent(lvl, b)
            n_own, length = (nxl, nyl + 2) if phase == 0 else (nyl, nxl + 2)
            for sgn in (-1, 1):
                nb = dec.neighbour(b, sgn if phase == 0 else 0, sgn if phase == 1 else 0)
                if nb is None:
                    continue
                send_idx, recv_idx = pick(sgn, n_own)
                if send_idx is not None:
                    sbuf = torch.empty(nsel * length, dtype=torch.float64, device=self.device)
                    K.line_copy(a, mask, phase, send_idx, sbuf, True)
                    msgs.append((b, nb, (phase, sgn), sbuf, None))
                if recv_idx is not None:
                    rbuf = torch.empty(nsel * length, dtype=torch.float64, device=self.device)
                    msgs.append((nb, b, (phase, -sgn), None, rbuf))
                    unpack.append((a, mask, recv_idx, rbuf))
        self.T.run(msgs)
        for a, mask, recv_idx, rbuf in unpack:
            K.line_copy(a, mask, phase, recv_idx, rbuf, False)
        self.exchanges += 1

    def exchange(self, lvl, arrs, xmask=None, ymask=None):
        """standard ghost exchange: first/last own line -> the neighbour's ghost line;
        columns first, then rows including the fresh ghost columns (corners)"""
        std = lambda sgn, n: (2, 1) if sgn < 0 else (n + 1, n + 2)
        self._lines(lvl, arrs, 0, std, xmask)
        self._lines(lvl, arrs, 1, std, ymask)

    def push_weights(self, lvl):
        """Interpolation weights of the F-points just left of / below a block's first
        coarse line are computed by the neighbour that owns those F-points, in ITS ghost
        line, but stored (reference layout) at THIS block's first line: move them over."""
        push = lambda sgn, n: (None, 2) if sgn < 0 else (n + 2, None)
        P = self.lv[lvl].P
        self._lines(lvl, P, 0, push, LL | LR | LSW | LNW | LNE | LSE)
        self._lines(lvl, P, 1, push, LA | LB | LSW | LNW | LNE | LSE)

    # -- setup phase (multilevel::setup, include/cedar/multilevel.h:243-265) ---
    def _setup(self):
        K = self.K
        for lvl in range(self.ndist):
            F, Cc = self.lv[lvl], self.lv[lvl + 1]
            self.exchange(lvl, F.A)                                   # operator entries stored at the neighbours
            for b in self.owned:
                K.setup_interp_stage(F.A[b], Cc.P[b], 1)
            # edge weights of the ghost lines: own values of the neighbour across that line
            self.exchange(lvl + 1, Cc.P, xmask=LA | LB, ymask=LL | LR)
            for b in self.owned:
                K.setup_interp_stage(F.A[b], Cc.P[b], 2)
            self.push_weights(lvl + 1)
            self.exchange(lvl + 1, Cc.P)
            for b in self.owned:
                K.galerkin(F.A[b], Cc.A[b], Cc.P[b])
                K.recip(F.A[b], F.sor[b])
        self.exchange(self.ndist, self.lv[self.ndist].A)
        self._setup_coarse()

    def _own(self, t):
        return t[..., 1:-1, 1:-1]

    def _gather(self, lvl, arrs, planes):
        """own interiors of every block -> global array (with zero ghost ring) on block 0's owner"""
        dec = self.dec
        gx, gy = dec.gn[lvl]
        root = self.T.owner(0) == self.T.rank
        msgs, place = [], []
        for b in range(dec.nblocks):
            mine, x0, y0, nxl, nyl = b in self.owned, *dec.extent(lvl, b)
            sbuf = self._own(arrs[b]).reshape(planes, nyl, nxl).contiguous() if mine else None
            rbuf = torch.empty((planes, nyl, nxl), dtype=torch.float64, device=self.device) if root else None
            if mine or root:
                msgs.append((b, 0, ("g", b), sbuf, rbuf))
            if root:
                place.append((rbuf, x0, y0, nxl, nyl))
        self.T.run(msgs)
        if not root:
            return None
        g = self._zeros(planes, gy + 2, gx + 2)
        for rbuf, x0, y0, nxl, nyl in place:
            g[:, 1 + y0:1 + y0 + nyl, 1 + x0:1 + x0 + nxl] = rbuf
        return g

    def _scatter(self, lvl, g, arrs):
        """global array on block 0's owner -> own interiors of every block"""
        dec = self.dec
        root = self.T.owner(0) == self.T.rank
        msgs, place = [], []
        for b in range(dec.nblocks):
            mine, x0, y0, nxl, nyl = b in self.owned, *dec.extent(lvl, b)
            sbuf = g[1 + y0:1 + y0 + nyl, 1 + x0:1 + x0 + nxl].contiguous() if root else None
            rbuf = torch.empty((nyl, nxl), dtype=torch.float64, device=self.device) if mine else None
            if mine or root:
                msgs.append((0, b, ("s", b), sbuf, rbuf))
            if mine:
                place.append((arrs[b], rbuf))
        self.T.run(msgs)
        for a, rbuf in place:
            self._own(a).copy_(rbuf)

    def _setup_coarse(self):
        """agglomeration: the operator of level `ndist` goes to rank 0, which owns the
        remaining levels through the serial solver (CUDA-graph captured)"""
        lvl = self.ndist
        nst = self.nst if lvl == 0 else 5
        g = self._gather(lvl, self.lv[lvl].A, nst)
        self.coarse = None
        if g is not None:
            nl = self.nlevels - lvl
            if self.coarse_factory is not None:
                self.coarse = self.coarse_factory(g, self.npre, self.npost, nl)
            else:
                from . import solver as cs
                if self.stream is not None:
                    self.stream.synchronize()
                op = cs.Operator2(g.cpu().numpy())
                self.coarse = _SerialCoarse(cs, op, self.npre, self.npost, nl, self.stream)
            gx, gy = self.dec.gn[lvl]
            self._gb, self._gx = self._zeros(gy + 2, gx + 2), self._zeros(gy + 2, gx + 2)

    def _coarse_solve(self, initial_guess=False):
        lvl = self.ndist
        L = self.lv[lvl]
        g = self._gather(lvl, L.b, 1)
        gx0 = self._gather(lvl, L.x, 1) if initial_guess else None
        if g is not None:
            self._gb.copy_(g[0])
            if initial_guess:
                self._gx.copy_(gx0[0])
            else:
                self._gx.zero_()
            self.coarse.vcycle(self._gx, self._gb)
        self._scatter(lvl, self._gx if g is not None else None, L.x)

    # -- wide halos: one exchange per sweep instead of one per colour -----------
    # A whole multicolour sweep changes a point only through values at most 4 lines away (9-point,
    # 4 colours; fewer for red-black).  A block therefore keeps, for relaxation only, copies of its
    # arrays extended by WIDE = 5 lines towards every neighbour (4 recomputed redundantly + the
    # outermost as frozen ghost line): fill them with ONE exchange, run the single-pass fused sweep
    # (bmgx2_relax_GS_oop) on the extended arrays, and copy the block and its ghost ring back -- the
    # ghost ring then already holds the neighbours' new values, bit for bit, so no exchange follows.
    # WIDE is odd so that the extended array keeps the index parity (colouring) of the global grid.
    WIDE = 5

    def _build_ext(self, lvl):
        dec, W = self.dec, self.WIDE
        for b in self.owned:
            _, _, nxl, nyl = dec.extent(lvl, b)
            if min(nxl, nyl) < max(self.fuse_min, W):
                return                                   # small level: per-colour passes
        E = _Level()
        E.x2, E.off, E.b_current = {}, {}, False
        L = self.lv[lvl]
        for b in self.owned:
            _, _, nxl, nyl = dec.extent(lvl, b)
            e = [(W - 1) if dec.neighbour(b, dx, dy) is not None else 0 for dx, dy in ((-1, 0), (1, 0), (0, -1), (0, 1))]
            jj, ii = nyl + 2 + e[2] + e[3], nxl + 2 + e[0] + e[1]
            E.off[b] = (e[0], e[2], nxl, nyl)
            E.A[b] = self._zeros(L.A[b].shape[0], jj, ii)
            E.sor[b] = self._zeros(2, jj, ii)
            E.x[b], E.x2[b], E.b[b] = self._zeros(jj, ii), self._zeros(jj, ii), self._zeros(jj, ii)
            self._centre(E, b, E.A[b]).copy_(L.A[b])
            self._centre(E, b, E.sor[b]).copy_(L.sor[b])
        self.ext[lvl] = E
        self.exchange_wide(lvl, [E.A, E.sor])

    @staticmethod
    def _centre(E, b, t):
        ox, oy, nxl, nyl = E.off[b]
        return t[..., oy:oy + nyl + 2, ox:ox + nxl + 2]

    def _halo(self, lvl, arrs_list, geom, W):
        """one-phase halo exchange with all (up to 8) neighbours: the W own lines next to a side (a W x W
        corner for the diagonal neighbours) go into the neighbour's W halo lines.  geom(b) = (ox, oy): the
        block's own interior starts at [oy + 1, ox + 1] of its arrays."""
        dec = self.dec
        msgs, unpack = [], []

        def span(o, n, d, recv):
            if d == 0:
                return slice(o + 1, o + 1 + n)
            if recv:
                return slice(o + 1 - W, o + 1) if d < 0 else slice(o + 1 + n, o + 1 + n + W)
            return slice(o + 1, o + 1 + W) if d < 0 else slice(o + 1 + n - W, o + 1 + n)

        for b in self.owned:
            ox, oy = geom(b)
            _, _, nxl, nyl = dec.extent(lvl, b)
            for dy in (-1, 0, 1):
                for dx in (-1, 0, 1):
                    nb = dec.neighbour(b, dx, dy) if (dx or dy) else None
                    if nb is None:
                        continue
                    rs, cs = span(oy, nyl, dy, False), span(ox, nxl, dx, False)
                    sbuf = torch.cat([a[b][..., rs, cs].reshape(-1) for a in arrs_list])
                    rbuf = torch.empty_like(sbuf)
                    msgs.append((b, nb, ("h", dx, dy), sbuf, None))
                    msgs.append((nb, b, ("h", -dx, -dy), None, rbuf))
                    rr, cr = span(oy, nyl, dy, True), span(ox, nxl, dx, True)
                    unpack.append(([a[b][..., rr, cr] for a in arrs_list], rbuf))
        self.T.run(msgs)
        for views, rbuf in unpack:
            o = 0
            for v in views:
                v.copy_(rbuf[o:o + v.numel()].reshape(v.shape))
                o += v.numel()
        self.exchanges += 1

    def exchange_wide(self, lvl, arrs_list):
        """fill the WIDE lines beyond every neighbour side of the extended arrays with the neighbours' own values"""
        E = self.ext[lvl]
        self._halo(lvl, arrs_list, lambda b: E.off[b][:2], self.WIDE)

    def exchange1(self, lvl, arrs):
        """the standard one-line ghost exchange of grid functions, all neighbours in one phase"""
        self._halo(lvl, [arrs], lambda b: (0, 0), 1)

    def _sweep_fused(self, lvl, x, b, updown):
        E, K = self.ext[lvl], self.K
        arrs = [E.x]
        for blk in self.owned:
            self._centre(E, blk, E.x[blk]).copy_(x[blk])
        if not E.b_current:          # the right-hand side of a level changes once per visit: it rides along once
            for blk in self.owned:
                self._centre(E, blk, E.b[blk]).copy_(b[blk])
            arrs.append(E.b)
            E.b_current = True
        self.exchange_wide(lvl, arrs)
        for blk in self.owned:
            K.relax_oop(E.A[blk], E.b[blk], E.x[blk], E.x2[blk], E.sor[blk], updown)
            x[blk].copy_(self._centre(E, blk, E.x2[blk]))

    # -- the cycle (include/cedar/cycle/vcycle.h:57-115) ------------------------
    def _sweep(self, lvl, x, b, updown):
        if lvl in self.ext:
            return self._sweep_fused(lvl, x, b, updown)
        L, K = self.lv[lvl], self.K
        npass = 2 if L.A[self.owned[0]].shape[0] == 3 else 4
        for p in range(npass):
            for blk in self.owned:
                K.relax_pass(L.A[blk], b[blk], x[blk], L.sor[blk], updown, p)
            self.exchange1(lvl, x)

    def _ncycle(self, lvl, x, b):
        L, Cc, K = self.lv[lvl], self.lv[lvl + 1], self.K
        if lvl in self.ext:
            self.ext[lvl].b_current = False
        for _ in range(self.npre):
            self._sweep(lvl, x, b, DOWN)
        for blk in self.owned:
            K.residual(L.A[blk], b[blk], x[blk], L.res[blk])
        self.exchange1(lvl, L.res)
        for blk in self.owned:
            K.restrict(L.res[blk], Cc.b[blk], Cc.P[blk])
            Cc.x[blk].zero_()
        if lvl + 1 == self.ndist:
            self._coarse_solve()
        else:
            self._ncycle(lvl + 1, Cc.x, Cc.b)
        self.exchange1(lvl + 1, Cc.x)
        for blk in self.owned:
            K.interp_add(x[blk], Cc.x[blk], L.res[blk], L.A[blk], Cc.P[blk])
        self.exchange1(lvl, x)
        for _ in range(self.npost):
            self._sweep(lvl, x, b, UP)

    def vcycle(self, x, b):
        """one V-cycle on {block: device array (nyl+2, nxl+2)}; x's ghosts must be current"""
        with self._on_stream():
            if self.ndist == 0:
                self.lv[0].b, self.lv[0].x = b, x
                self._coarse_solve(initial_guess=True)
                self.exchange(0, x)
            else:
                self._ncycle(0, x, b)

    def vcycle_graphed(self, x, b, warm=4):
        """vcycle() replayed from one CUDA graph per (x, b) pair: the kernels, the NCCL ghost
        exchanges and the agglomerated serial cycle (a child graph) are captured together, so a
        replay costs one launch instead of several hundred.  All ranks must call it together."""
        key = tuple(sorted((k, x[k].data_ptr(), b[k].data_ptr()) for k in self.owned))
        g = self._graphs.get(key) if hasattr(self, "_graphs") else None
        if g is None:
            if not hasattr(self, "_graphs"):
                self._graphs = {}
            saved = {k: x[k].clone() for k in self.owned}
            for _ in range(warm):          # lets the serial coarse solver reach its replay state
                self.vcycle(x, b)
            self.stream.synchronize()
            for k in self.owned:
                x[k].copy_(saved[k])
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            l0, e0 = self.K.launches, self.exchanges
            with torch.cuda.graph(g, stream=self.stream):
                self._ncycle(0, x, b) if self.ndist > 0 else self.vcycle(x, b)
            if not hasattr(self, "_graph_costs"):
                self._graph_costs = {}
            self._graph_costs[key] = (self.K.launches - l0, self.exchanges - e0)
            self._graphs[key] = g
            for k in self.owned:
                x[k].copy_(saved[k])
        with self._on_stream():
            g.replay()
        self.K.launches += self._graph_costs[key][0]
        self.exchanges += self._graph_costs[key][1]

    def vcycle_stream(self, x, b, bhs, xhs, graphed=True):
        """A stream of right-hand sides from and to pinned host memory (this rank's block of each): for every
        pair (bh, xh) one cycle from x0 = 0 on bh, the result in xh.  The upload of the next right-hand side and
        the download of the previous solution run on their own streams while the cycle (kernels + NCCL ghost
        exchanges) runs on the solver's; device staging is double-buffered each way.  Host-API counterpart of
        cdr2_solver_vcycle_batch (include/cedar_capi.h) for the block-decomposed solver; all ranks call it with
        the same number of pairs."""
        (r,) = self.owned
        st = getattr(self, "_stage", None)
        if st is None:
            st = self._stage = types.SimpleNamespace(
                up=torch.cuda.Stream(), down=torch.cuda.Stream(),
                db=[torch.empty_like(b[r]) for _ in range(2)], dx=[torch.empty_like(x[r]) for _ in range(2)],
                b_ready=[torch.cuda.Event() for _ in range(2)], b_free=[torch.cuda.Event() for _ in range(2)],
                x_ready=[torch.cuda.Event() for _ in range(2)], x_free=[torch.cuda.Event() for _ in range(2)])
        cyc = self.vcycle_graphed if graphed else self.vcycle
        for k, (bh, xh) in enumerate(zip(bhs, xhs)):
            q = k & 1
            with torch.cuda.stream(st.up):
                if k >= 2:
                    st.up.wait_event(st.b_free[q])
                st.db[q].copy_(bh, non_blocking=True)
                st.b_ready[q].record()
            with torch.cuda.stream(self.stream):
                self.stream.wait_event(st.b_ready[q])
                b[r].copy_(st.db[q])
                st.b_free[q].record()
                x[r].zero_()
            cyc(x, b)
            with torch.cuda.stream(self.stream):
                if k >= 2:
                    self.stream.wait_event(st.x_free[q])
                st.dx[q].copy_(x[r])
                st.x_ready[q].record()
            with torch.cuda.stream(st.down):
                st.down.wait_event(st.x_ready[q])
                xh.copy_(st.dx[q], non_blocking=True)
                st.x_free[q].record()
        st.down.synchronize()
        self.stream.synchronize()

    def _norm(self, arrs):
        tot = self._zeros(1)
        for blk in self.owned:
            self.K.sumsq(arrs[blk], self._ss)
            tot += self._ss
        self.T.allreduce_sum(tot)
        return math.sqrt(float(tot.item()))

    def solve(self, b, x=None, maxiter=10, tol=1e-8):
        """multilevel::solve (include/cedar/multilevel.h:278-298).  Returns (x, history, res0)."""
        L, K = self.lv[0], self.K
        with self._on_stream():
            b = {k: torch.as_tensor(v, dtype=torch.float64).to(self.device).contiguous() for k, v in b.items()}
            if x is None:
                x = {k: torch.zeros_like(v) for k, v in b.items()}
            else:
                x = {k: torch.as_tensor(v, dtype=torch.float64).to(self.device).contiguous() for k, v in x.items()}
            self.exchange(0, x)
            for blk in self.owned:
                K.residual(L.A[blk], b[blk], x[blk], L.res[blk])
            res0 = self._norm(L.res)
            hist = []
            for _ in range(maxiter):
                self.vcycle(x, b)
                for blk in self.owned:
                    K.residual(L.A[blk], b[blk], x[blk], L.res[blk])
                hist.append(self._norm(L.res) / res0)
                if hist[-1] < tol:
                    break
        return x, np.array(hist), res0


class _SerialCoarse:
    """the agglomerated levels: cdr2::solver on rank 0, device buffers in and out"""

    def __init__(self, cs, op, npre, npost, nlevels, stream):
        self.cs, self.op, self.stream = cs, op, stream
        cfg = cs.make_config(relaxation="point", nrelax_pre=npre, nrelax_post=npost, num_levels=nlevels)
        self.sol = cs.Solver2(op, config_json=cfg)
        assert self.sol.nlevels == nlevels, (self.sol.nlevels, nlevels)

    def vcycle(self, x, b):
        self.cs._check(self.cs.lib.cdr2_solver_vcycle_dev(self.sol.h, C.c_void_p(b.data_ptr()), C.c_void_p(x.data_ptr()),
                                                         C.c_void_p(self.stream.cuda_stream)))


def split_blocks(a, decomp, lvl=0, blocks=None, keep_ghosts=False):
    """cut the local arrays (with ghost ring) of the given blocks out of a global array"""
    out = {}
    for b in (range(decomp.nblocks) if blocks is None else blocks):
        x0, y0, nxl, nyl = decomp.extent(lvl, b)
        loc = np.array(a[..., y0:y0 + nyl + 2, x0:x0 + nxl + 2])
        if not keep_ghosts:
            loc[..., 0, :] = 0; loc[..., -1, :] = 0; loc[..., :, 0] = 0; loc[..., :, -1] = 0
        out[b] = loc
    return out


def join_blocks(parts, decomp, lvl=0):
    """inverse of split_blocks for the own interiors (ghost ring of the result is zero)"""
    gx, gy = decomp.gn[lvl]
    first = next(iter(parts.values()))
    lead = tuple(first.shape[:-2])
    g = np.zeros(lead + (gy + 2, gx + 2))
    for b, t in parts.items():
        x0, y0, nxl, nyl = decomp.extent(lvl, b)
        t = t.detach().cpu().numpy() if hasattr(t, "detach") else np.asarray(t)
        g[..., 1 + y0:1 + y0 + nyl, 1 + x0:1 + x0 + nxl] = t[..., 1:-1, 1:-1]
    return g


# ------------------------------------------------------------------ block-wise problem generators
def gallery_fe_block(dec, b):
    """own entries of gallery::fe (src/2d/gallery.cc:78-113) for block b: centre 8, every
    coupling between two interior points 1"""
    x0, y0, nxl, nyl = dec.extent(0, b)
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


def gallery_hetero_block(dec, b, nblk=64, seed=20260411):
    """own entries of the heterogeneous 9-point operator of config 5:
    bilinear-FE-like stencil of -div(D grad u), D piecewise constant on an nblk x nblk
    checkerboard of log-uniform blocks D = 10**U(-3,3); every coupling is the mean of D over
    the cells it crosses, the centre is the sum of the eight couplings (Dirichlet boundary
    couplings kept in the centre, as gallery::fe does)."""
    gnx, gny = dec.gn[0]
    rng = np.random.default_rng(seed)
    D = 10.0 ** rng.uniform(-3, 3, size=(nblk, nblk))
    x0, y0, nxl, nyl = dec.extent(0, b)

    def cell(ci, cj):       # coefficient of the cell whose lower-left grid node is (ci, cj), nodes 0..gn+1
        bi = np.clip(ci * nblk // (gnx + 1), 0, nblk - 1)
        bj = np.clip(cj * nblk // (gny + 1), 0, nblk - 1)
        return D[bj, bi]

    gi = (x0 + 1 + np.arange(nxl))[None, :] + np.zeros((nyl, 1), dtype=int)     # node indices 1..gn of own points
    gj = (y0 + 1 + np.arange(nyl))[:, None] + np.zeros((1, nxl), dtype=int)
    sw, se, nw, ne = cell(gi - 1, gj - 1), cell(gi, gj - 1), cell(gi - 1, gj), cell(gi, gj)
    cw, cs_, ce, cn = 0.5 * (sw + nw), 0.5 * (sw + se), 0.5 * (se + ne), 0.5 * (nw + ne)
    so = np.zeros((5, nyl + 2, nxl + 2))
    own = so[:, 1:-1, 1:-1]
    own[0] = cw + cs_ + ce + cn + sw + se + nw + ne
    own[1] = cw * (gi >= 2)
    own[2] = cs_ * (gj >= 2)
    own[3] = sw * ((gi >= 2) & (gj >= 2))
    # KNW stored at S couples (S-y) and (S-x): it crosses the cell south-west of S
    own[4] = sw * ((gi >= 2) & (gj >= 2))
    return so


def rhs_sin_block(dec, b):
    """examples/basic-2d-ser/poisson.cc:15-37 on block b"""
    gnx, gny = dec.gn[0]
    x0, y0, nxl, nyl = dec.extent(0, b)
    hx, hy = 1.0 / (gnx + 1), 1.0 / (gny + 1)
    x = ((x0 + 1 + np.arange(nxl)) * hx)[None, :]
    y = ((y0 + 1 + np.arange(nyl)) * hy)[:, None]
    out = np.zeros((nyl + 2, nxl + 2))
    pi = math.pi
    out[1:-1, 1:-1] = 8 * (pi * pi) * np.sin(2 * pi * x) * np.sin(2 * pi * y) * (hx * hy)
    return out
