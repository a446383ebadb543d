"""TEST INFRASTRUCTURE: runs the operation lists of include/cedar/2d/mpi/plan.h (the schedule the CUDA
executor cedar_b200/host/dist2.cc runs on the GPUs) with the CPU oracle's kernels on numpy arrays, so that
the planner -- halo widths, validity tracking, where the exchanges go, coarse windows, agglomeration -- is
checked on CPU: all blocks in one process, or one block per rank over torch.distributed/gloo.  The plan and
the block geometry come from the product library (host-only entry points cdr2_dist_plan_json /
cdr2_dist_geometry); the arithmetic is the oracle's.  Never imported by the product."""
import numpy as np

import oracle
from oracle import P
import cedar_b200.dist2 as d2

X0, X1, B, RES, A, SOR, PI = range(7)
A_SLOT = A
NPL = {X0: 1, X1: 1, B: 1, RES: 1, SOR: 2, PI: 8}


class PlanSim:
    def __init__(self, nx, ny, px, py, so_blocks, pre=1, post=1, agglomerate=33, halo=15, dist=None, poison=True, split_min=0,
                 relaxation="point"):
        """so_blocks: {block: (nst, nyl+2, nxl+2)} own entries of the blocks this process owns"""
        self.L = oracle.lib()
        self.nx, self.ny, self.px, self.py, self.W = nx, ny, px, py, halo
        self.pre, self.post, self.dist = pre, post, dist
        self.owned = sorted(so_blocks)
        self.nst = next(iter(so_blocks.values())).shape[0]
        self.relaxation = relaxation
        self.lfac, self.lrhs = {}, {}
        kw = dict(nst=self.nst, pre=pre, post=post, agglomerate=agglomerate, halo=halo, split_min=split_min, relaxation=relaxation)
        self.pending = {}
        self.p_setup, self.nd = d2.plan(nx, ny, px, py, what="setup", **kw)
        self.p_cycle, _ = d2.plan(nx, ny, px, py, what="cycle", **kw)
        self.p_norm, _ = d2.plan(nx, ny, px, py, what="norm", **kw)
        self.nb = px * py
        self.g = {(l, b): d2.geometry(nx, ny, px, py, self.nd, halo, l, b) for l in range(self.nd + 1) for b in range(self.nb)}
        self.gn = [(nx, ny)]
        for _ in range(self.nd):
            self.gn.append(((self.gn[-1][0] - 1) // 2 + 1, (self.gn[-1][1] - 1) // 2 + 1))
        self.nlevels = 1
        while min((nx - 1) // (1 << self.nlevels) + 1, (ny - 1) // (1 << self.nlevels) + 1) >= 3:
            self.nlevels += 1
        # arrays; halos start as NaN so that any use of a never-exchanged halo value shows up in the own points
        fill = np.nan if poison else 0.0
        self.a = {}
        for l in range(self.nd + 1):
            for b in self.owned:
                g = self.g[l, b]
                for s in range(7):
                    if s == PI and l == 0:
                        continue
                    npl = (self.nst if l == 0 else 5) if s == A else NPL[s]
                    arr = np.full((npl, g["jj"], g["ii"]), fill)
                    self.a[l, b, s] = arr
        for b in self.owned:
            g = self.g[0, b]
            self._own(0, b, A)[...] = so_blocks[b][:, 1:-1, 1:-1]
            for s in (X0, B):
                self._own(0, b, s)[...] = 0.0
        self._clean_boundary()
        self.red = {b: 0.0 for b in self.owned}
        self.tmpP, self.tmpA = {}, {}
        self.coarse = {}
        self.exchanges = 0
        self.run(self.p_setup)

    # -- helpers ----------------------------------------------------------------
    def _own(self, l, b, s):
        g = self.g[l, b]
        return self.a[l, b, s][:, g["wb"]:g["wb"] + g["ny"], g["wl"]:g["wl"] + g["nx"]]

    def _clean_boundary(self):
        """the ghost ring at the physical boundary is zero, as in the serial arrays"""
        for (l, b, s), arr in self.a.items():
            g = self.g[l, b]
            if g["wl"] == 1: arr[:, :, 0] = 0.0
            if g["wr"] == 1: arr[:, :, -1] = 0.0
            if g["wb"] == 1: arr[:, 0, :] = 0.0
            if g["wt"] == 1: arr[:, -1, :] = 0.0

    def _neighbour(self, b, dx, dy):
        cx, cy = b % self.px + dx, b // self.px + dy
        return None if (cx < 0 or cx >= self.px or cy < 0 or cy >= self.py) else cy * self.px + cx

    def _rects(self, l, b):
        """(neighbour, my destination slices, its source slices) of a ghost exchange"""
        g, W = self.g[l, b], self.W
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                nb = self._neighbour(b, dx, dy) if (dx or dy) else None
                if nb is None:
                    continue
                h = self.g[l, nb]
                if dx < 0: dc, sc = slice(0, W), slice(h["wl"] + h["nx"] - W, h["wl"] + h["nx"])
                elif dx > 0: dc, sc = slice(g["wl"] + g["nx"], g["wl"] + g["nx"] + W), slice(h["wl"], h["wl"] + W)
                else: dc, sc = slice(g["wl"], g["wl"] + g["nx"]), slice(h["wl"], h["wl"] + h["nx"])
                if dy < 0: dr, sr = slice(0, W), slice(h["wb"] + h["ny"] - W, h["wb"] + h["ny"])
                elif dy > 0: dr, sr = slice(g["wb"] + g["ny"], g["wb"] + g["ny"] + W), slice(h["wb"], h["wb"] + W)
                else: dr, sr = slice(g["wb"], g["wb"] + g["ny"]), slice(h["wb"], h["wb"] + h["ny"])
                yield nb, (dr, dc), (sr, sc), (dx, dy)

    def _exchange(self, l, slots):
        self.exchanges += 1
        if self.dist is None:
            snap = {(b, s): self.a[l, b, s].copy() for b in self.owned for s in slots}
            for b in self.owned:
                for nb, (dr, dc), (sr, sc), _ in self._rects(l, b):
                    for s in slots:
                        self.a[l, b, s][:, dr, dc] = snap[nb, s][:, sr, sc]
            return
        import torch
        dist = self.dist
        (b,) = self.owned
        ops, recvs = [], []
        # what I send to nb is what nb's _rects asks of me: the mirror rectangle
        for nb, (dr, dc), _, (dx, dy) in self._rects(l, b):
            for (nb2, _, (sr, sc), (dx2, dy2)) in self._rects(l, nb):
                if nb2 == b and (dx2, dy2) == (-dx, -dy):
                    sbuf = torch.from_numpy(np.concatenate([self.a[l, b, s][:, sr, sc].reshape(-1) for s in slots]))
                    ops.append(dist.P2POp(dist.isend, sbuf, nb))
            n = sum(self.a[l, b, s][:, dr, dc].size for s in slots)
            rbuf = torch.empty(n, dtype=torch.float64)
            ops.append(dist.P2POp(dist.irecv, rbuf, nb))
            recvs.append((rbuf, dr, dc))
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        for rbuf, dr, dc in recvs:
            o = 0
            r = rbuf.numpy()
            for s in slots:
                v = self.a[l, b, s][:, dr, dc]
                v[...] = r[o:o + v.size].reshape(v.shape)
                o += v.size

    def _gather(self, l, s):
        """own points of every block -> a global array with zero ghost ring, on every rank"""
        gx, gy = self.gn[l]
        npl = self.a[l, self.owned[0], s].shape[0]
        out = np.zeros((npl, gy + 2, gx + 2))
        parts = {b: np.ascontiguousarray(self._own(l, b, s)) for b in self.owned}
        if self.dist is not None:
            allp = [None] * self.dist.get_world_size()
            self.dist.all_gather_object(allp, parts)
            parts = {k: v for p in allp for k, v in p.items()}
        for b, v in parts.items():
            g = self.g[l, b]
            out[:, 1 + g["y0"]:1 + g["y0"] + g["ny"], 1 + g["x0"]:1 + g["x0"] + g["nx"]] = v
        return out

    # -- the executor -------------------------------------------------------------
    def run(self, plan):
        for o in plan:
            getattr(self, "op_" + o["op"])(o)

    def op_exchange(self, o):
        if o["async"]:
            # the pull runs beside the following operations: deliver it as LATE as allowed (at the join), so that an
            # operation in between that reads a halo line sees the stale value and the comparison with the serial solve fails
            self.pending[o["id"]] = (o["lvl"], o["arrays"])
            return
        self._exchange(o["lvl"], o["arrays"])

    def op_join(self, o):
        self._exchange(*self.pending.pop(o["id"]))

    def op_sync(self, o):
        self.exchanges += 1

    def _each(self, o):
        for b in self.owned:
            yield b, self.g[o["lvl"], b]

    def _ifd(self, l):
        return 1 if (l == 0 and self.nst == 3) else 0

    def _nstl(self, l):
        return self.nst if l == 0 else 5

    def op_sweep(self, o):
        l = o["lvl"]
        if o["part"]:
            return self._sweep_rects(o)
        for b, g in self._each(o):
            src, dst = self.a[l, b, o["a"]], self.a[l, b, o["b"]]
            dst[...] = src
            assert self.L.orc2_relax_gs(1, P(self.a[l, b, A]), P(self.a[l, b, B]), P(dst), P(self.a[l, b, SOR]), g["ii"], g["jj"], 1,
                                        self._ifd(l), self._nstl(l), 2, 1, o["dir"], 0) == 0

    def _sweep_rects(self, o):
        """bmgx2_relax_GS_oop_rects: every rectangle is swept as a grid of its own (outer lines = ghost ring from the
        input iterate), only its window is stored"""
        l = o["lvl"]
        nc = 2 if self._ifd(l) else 4
        for b, g in self._each(o):
            src, dst = self.a[l, b, o["a"]], self.a[l, b, o["b"]]
            for (i0, j0, ni, nj, wi0, wi1, wj0, wj1) in d2.sweep_rects(self.nx, self.ny, self.px, self.py, self.nd, self.W, l, b, nc,
                                                                       o["part"]):
                assert i0 % 2 == 0 and j0 % 2 == 0
                cut = lambda a: np.array(a[:, j0:j0 + nj, i0:i0 + ni], order="C", copy=True)   # a copy even when the slice is contiguous
                A, bb, q, sor = cut(self.a[l, b, A_SLOT]), cut(self.a[l, b, B]), cut(src), cut(self.a[l, b, SOR])
                assert self.L.orc2_relax_gs(1, P(A), P(bb), P(q), P(sor), ni, nj, 1, self._ifd(l), self._nstl(l), 2, 1, o["dir"], 0) == 0
                dst[0, j0 + wj0 - 1:j0 + wj1, i0 + wi0 - 1:i0 + wi1] = q[0, wj0 - 1:wj1, wi0 - 1:wi1]

    def op_residual(self, o):
        l = o["lvl"]
        for b, g in self._each(o):
            self.L.orc2_residual(0, P(self.a[l, b, A]), P(self.a[l, b, B]), P(self.a[l, b, o["a"]]), P(self.a[l, b, RES]), g["ii"],
                                 g["jj"], 0, self._ifd(l), self._nstl(l))

    def _window(self, l, b, s):
        g = self.g[l, b]
        return self.a[l + 1, b, s][:, g["oy"]:g["oy"] + g["jjc"], g["ox"]:g["ox"] + g["iic"]]

    def op_restrict(self, o):
        l = o["lvl"]
        for b, g in self._each(o):
            bc = np.ascontiguousarray(self._window(l, b, B))
            pw = np.ascontiguousarray(self._window(l, b, PI))
            assert self.L.orc2_restrict(P(self.a[l, b, RES]), P(bc), P(pw), g["ii"], g["jj"], g["iic"], g["jjc"], 0) == 0
            self._window(l, b, B)[:, 1:-1, 1:-1] = bc[:, 1:-1, 1:-1]

    def op_interp(self, o):
        l = o["lvl"]
        for b, g in self._each(o):
            xc = np.ascontiguousarray(self._window(l, b, o["b"]))
            pw = np.ascontiguousarray(self._window(l, b, PI))
            assert self.L.orc2_interp_add(P(self.a[l, b, o["a"]]), P(xc), P(self.a[l, b, RES]), P(self.a[l, b, A]), P(pw), g["iic"],
                                          g["jjc"], g["ii"], g["jj"], self._nstl(l), 0) == 0

    def op_zero(self, o):
        for b, _ in self._each(o):
            self.a[o["lvl"], b, o["a"]][...] = 0.0

    def op_copy(self, o):
        for b, _ in self._each(o):
            self.a[o["lvl"], b, o["b"]][...] = self.a[o["lvl"], b, o["a"]]

    def op_gather(self, o):
        self.exchanges += 1
        self.GB = self._gather(self.nd, B)[0]

    def op_coarse(self, o):
        self.GX = np.zeros_like(self.GB)
        self.csol.vcycle(self.GX, self.GB)

    def op_scatter(self, o):
        l = o["lvl"]
        for b, g in self._each(o):
            r0, c0 = g["y0"] + 1 - g["wb"], g["x0"] + 1 - g["wl"]
            self.a[l, b, o["a"]][0] = self.GX[r0:r0 + g["jj"], c0:c0 + g["ii"]]

    def op_recip(self, o):
        l = o["lvl"]
        for b, g in self._each(o):
            self.L.orc2_setup_recip(P(self.a[l, b, A]), P(self.a[l, b, SOR]), g["ii"], g["jj"], self._nstl(l), 2)

    def op_setup_interp(self, o):
        l = o["lvl"]
        for b, g in self._each(o):
            pw = np.zeros((8, g["jjc"], g["iic"]))
            for stage in (1, 2):
                assert self.L.orc2_setup_interp_oi_stage(P(self.a[l, b, A]), None, P(pw), g["ii"], g["jj"], g["iic"], g["jjc"],
                                                         self._ifd(l), self._nstl(l), 0, 0, stage) == 0
            self.tmpP[b] = pw
            self._window(l, b, PI)[...] = pw

    def op_galerkin(self, o):
        l = o["lvl"]
        for b, g in self._each(o):
            ac = np.zeros((5, g["jjc"], g["iic"]))
            assert self.L.orc2_setup_itli_ex(P(self.a[l, b, A]), P(ac), P(self.tmpP[b]), g["ii"], g["jj"], g["iic"], g["jjc"],
                                             self._ifd(l), self._nstl(l), 0) == 0
            self._window(l, b, A)[...] = ac

    def op_gather_op(self, o):
        self.exchanges += 1
        self.GA = self._gather(self.nd, A)

    def op_setup_coarse(self, o):
        self.csol = oracle.Solver2(self.GA, self.relaxation, self.pre, self.post, num_levels=self.nlevels - self.nd)
        assert self.csol.nlevels == self.nlevels - self.nd

    # -- line relaxation across block cuts: the blocks transpose their lines (plan.h line_group).  Here: the serial
    # factor of the gathered operator, right-hand sides of the own points in the serial kernels' term order, DPTTS2 on
    # whole lines, own segments back
    def _allgather(self, parts):
        if self.dist is None:
            return parts
        allp = [None] * self.dist.get_world_size()
        self.dist.all_gather_object(allp, parts)
        return {k: v for p in allp for k, v in p.items()}

    def op_setup_lines(self, o):
        self.exchanges += 1
        l, dirn = o["lvl"], o["dir"]
        GA = self._gather(l, A)
        gx, gy = self.gn[l]
        if dirn == 0:
            sor = np.zeros((2, gy + 2, gx + 2))
            assert self.L.orc2_setup_lines_x(P(GA), P(sor), gx + 2, gy + 2, self._nstl(l), 0) == 0
        else:
            sor = np.zeros((2, gx + 2, gy + 2))
            assert self.L.orc2_setup_lines_y(P(GA), P(sor), gx + 2, gy + 2, self._nstl(l), 0) == 0
        self.lfac[l, dirn] = sor

    def op_lines_rhs(self, o):
        l, dirn, t0 = o["lvl"], o["dir"], o["part"] & 1
        for b, g in self._each(o):
            so, qf, q = self.a[l, b, A], self.a[l, b, B][0], self.a[l, b, o["a"]][0]
            assert g["y0"] % 2 == 0 and g["x0"] % 2 == 0
            if dirn == 0:
                rs, cs = slice(g["wb"] + t0, g["wb"] + g["ny"], 2), slice(g["wl"], g["wl"] + g["nx"])
            else:
                rs, cs = slice(g["wb"], g["wb"] + g["ny"]), slice(g["wl"] + t0, g["wl"] + g["nx"], 2)
            sh = lambda a, dj, di: a[slice(rs.start + dj, rs.stop + dj, rs.step), slice(cs.start + di, cs.stop + di, cs.step)]
            if dirn == 0:
                r = sh(qf, 0, 0) + sh(so[2], 0, 0) * sh(q, -1, 0) + sh(so[2], 1, 0) * sh(q, 1, 0)
            else:
                r = sh(qf, 0, 0) + sh(so[1], 0, 0) * sh(q, 0, -1) + sh(so[1], 0, 1) * sh(q, 0, 1)
            if not self._ifd(l):
                r = r + sh(so[3], 0, 0) * sh(q, -1, -1) + sh(so[4], 0, 1) * sh(q, -1, 1) + sh(so[4], 1, 0) * sh(q, 1, -1) \
                    + sh(so[3], 1, 1) * sh(q, 1, 1)
            self.lrhs[b] = (rs, cs, r)

    def op_lines_solve(self, o):
        l, dirn, t0 = o["lvl"], o["dir"], o["part"] & 1
        gx, gy = self.gn[l]
        parts = {}
        for b, g in self._each(o):
            rs, cs, r = self.lrhs[b]
            parts[b] = r
        parts = self._allgather(parts)
        R = np.zeros((gy, gx))      # global interior; only the lines of the colour are filled
        for b, r in parts.items():
            g = self.g[l, b]
            if dirn == 0:
                R[g["y0"] + t0:g["y0"] + g["ny"]:2, g["x0"]:g["x0"] + g["nx"]] = r
            else:
                R[g["y0"]:g["y0"] + g["ny"], g["x0"] + t0:g["x0"] + g["nx"]:2] = r
        sor = self.lfac[l, dirn]
        if dirn == 0:
            lines = np.arange(t0, gy, 2)
            bb, d, e = R[lines, :].copy(), sor[0, 1 + lines, 1:-1], sor[1, 1 + lines, 2:-1]
        else:
            lines = np.arange(t0, gx, 2)
            bb, d, e = R[:, lines].T.copy(), sor[0, 1 + lines, 1:-1], sor[1, 1 + lines, 2:-1]
        n = bb.shape[1]
        if n == 1:
            bb[:, 0] = bb[:, 0] * (1.0 / d[:, 0])
        else:
            for i in range(1, n):
                bb[:, i] = bb[:, i] - bb[:, i - 1] * e[:, i - 1]
            bb[:, n - 1] = bb[:, n - 1] / d[:, n - 1]
            for i in range(n - 2, -1, -1):
                bb[:, i] = bb[:, i] / d[:, i] - bb[:, i + 1] * e[:, i]
        if dirn == 0:
            R[lines, :] = bb
        else:
            R[:, lines] = bb.T
        for b, g in self._each(o):
            x = self.a[l, b, o["a"]][0]
            if dirn == 0:
                x[g["wb"] + t0:g["wb"] + g["ny"]:2, g["wl"]:g["wl"] + g["nx"]] = R[g["y0"] + t0:g["y0"] + g["ny"]:2, g["x0"]:g["x0"] + g["nx"]]
            else:
                x[g["wb"]:g["wb"] + g["ny"], g["wl"] + t0:g["wl"] + g["nx"]:2] = R[g["y0"]:g["y0"] + g["ny"], g["x0"] + t0:g["x0"] + g["nx"]:2]

    def op_sumsq(self, o):
        for b, _ in self._each(o):
            self.red[b] = float((self._own(o["lvl"], b, o["a"]) ** 2).sum())

    def op_norm(self, o):
        self.exchanges += 1
        parts = dict(self.red)
        if self.dist is not None:
            allp = [None] * self.dist.get_world_size()
            self.dist.all_gather_object(allp, parts)
            parts = {k: v for p in allp for k, v in p.items()}
        self.norm2 = sum(parts[b] for b in sorted(parts))

    # -- user level ---------------------------------------------------------------
    def set(self, b, slot, local):
        self._own(0, b, slot)[0] = local[1:-1, 1:-1]

    def get(self, b, slot=X0):
        g = self.g[0, b]
        out = np.zeros((g["ny"] + 2, g["nx"] + 2))
        out[1:-1, 1:-1] = self._own(0, b, slot)[0]
        return out

    def vcycle(self):
        self.run(self.p_cycle)

    def norm(self):
        self.run(self.p_norm)
        return float(np.sqrt(self.norm2))
