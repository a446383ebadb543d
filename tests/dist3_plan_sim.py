"""TEST INFRASTRUCTURE: the operation lists of the slab-decomposed 3D solver (cedar_b200/host/dist3.cc; planner of
include/cedar/2d/mpi/plan.h on the slabs of include/cedar/3d/mpi/plan.h) run with the CPU oracle's 3D kernels on numpy
arrays, all slabs in one process.  Halos start as NaN.  Never imported by the product."""
import numpy as np

import oracle
from oracle import P
import cedar_b200.dist3 as d3

X0, X1, B, RES, A, SOR, PI = range(7)


class PlanSim3:
    def __init__(self, nx, ny, nz, nslabs, so_slabs, pre=1, post=1, agglomerate=9, halo=5):
        self.L = oracle.lib()
        self.n, self.np_, self.W, self.pre, self.post = (nx, ny, nz), nslabs, halo, pre, post
        self.owned = sorted(so_slabs)
        self.nst = next(iter(so_slabs.values())).shape[0]
        kw = dict(nst=self.nst, pre=pre, post=post, agglomerate=agglomerate, halo=halo)
        self.p_setup, self.nd = d3.plan(nx, ny, nz, nslabs, what="setup", **kw)
        self.p_cycle, _ = d3.plan(nx, ny, nz, nslabs, what="cycle", **kw)
        self.g = {(l, b): d3.geometry(nx, ny, nz, nslabs, self.nd, halo, l, b) for l in range(self.nd + 1) for b in range(nslabs)}
        self.gn = [(nx, ny, nz)]
        for _ in range(self.nd):
            self.gn.append(tuple((v - 1) // 2 + 1 for v in self.gn[-1]))
        self.nlevels = 1
        while min((v - 1) // (1 << self.nlevels) + 1 for v in (nx, ny, nz)) >= 3:
            self.nlevels += 1
        self.a = {}
        for l in range(self.nd + 1):
            for b in self.owned:
                g = self.g[l, b]
                for s in range(7):
                    if s == PI and l == 0:
                        continue
                    ne = (self.nst if l == 0 else 14) if s == A else {SOR: 2, PI: 26}.get(s, 1)
                    self.a[l, b, s] = np.full((ne, g["kk"], g["jj"], g["ii"]), np.nan)
        for b in self.owned:
            self._own(0, b, A)[...] = so_slabs[b][:, 1:-1]
            for s in (X0, B):
                self._own(0, b, s)[...] = 0.0
        for (l, b, s), arr in self.a.items():      # ghost entries at the physical boundary are zero
            g = self.g[l, b]
            arr[:, :, 0, :] = 0.0; arr[:, :, -1, :] = 0.0; arr[:, :, :, 0] = 0.0; arr[:, :, :, -1] = 0.0
            if g["wb"] == 1: arr[:, 0] = 0.0
            if g["wt"] == 1: arr[:, -1] = 0.0
        self.tmpP = {}
        self.exchanges = 0
        self.run(self.p_setup)

    def _own(self, l, b, s):
        g = self.g[l, b]
        return self.a[l, b, s][:, g["wb"]:g["wb"] + g["nz"]]

    def run(self, plan):
        for o in plan:
            getattr(self, "op_" + o["op"])(o)

    def op_exchange(self, o):
        self.exchanges += 1
        l, W = o["lvl"], self.W
        snap = {(b, s): self.a[l, b, s].copy() for b in self.owned for s in o["arrays"]}
        for b in self.owned:
            g = self.g[l, b]
            for dz in (-1, 1):
                nb = b + dz
                if nb < 0 or nb >= self.np_:
                    continue
                h = self.g[l, nb]
                for s in o["arrays"]:
                    if dz < 0:
                        self.a[l, b, s][:, 0:W] = snap[nb, s][:, h["wb"] + h["nz"] - W:h["wb"] + h["nz"]]
                    else:
                        self.a[l, b, s][:, g["wb"] + g["nz"]:g["wb"] + g["nz"] + W] = snap[nb, s][:, h["wb"]:h["wb"] + W]

    def op_sync(self, o):
        self.exchanges += 1

    def op_join(self, o):
        pass

    def _ifd(self, l):
        return 1 if (l == 0 and self.nst == 4) else 0

    def _nstl(self, l):
        return self.nst if l == 0 else 14

    def op_sweep(self, o):
        l = o["lvl"]
        for b in self.owned:
            g = self.g[l, b]
            src, dst = self.a[l, b, o["a"]], self.a[l, b, o["b"]]
            if o["a"] != o["b"]:
                dst[...] = src
            assert self.L.orc3_relax_gs(1, P(self.a[l, b, A]), P(self.a[l, b, B]), P(dst), P(self.a[l, b, SOR]), g["ii"], g["jj"],
                                        g["kk"], self._ifd(l), self._nstl(l), 2, 1, o["dir"], 0) == 0

    def op_residual(self, o):
        l = o["lvl"]
        for b in self.owned:
            g = self.g[l, b]
            self.L.orc3_residual(1, 1, self._ifd(l), P(self.a[l, b, o["a"]]), P(self.a[l, b, B]), P(self.a[l, b, A]),
                                 P(self.a[l, b, RES]), g["ii"], g["jj"], g["kk"], self._nstl(l))

    def _window(self, l, b, s):
        g = self.g[l, b]
        return self.a[l + 1, b, s][:, g["oz"]:g["oz"] + g["kkc"]]

    def op_restrict(self, o):
        l = o["lvl"]
        for b in self.owned:
            g, c = self.g[l, b], self.g[l + 1, b]
            bc = np.array(self._window(l, b, B), copy=True)
            pw = np.array(self._window(l, b, PI), copy=True)
            assert self.L.orc3_restrict(P(self.a[l, b, RES]), P(bc), P(pw), g["ii"], g["jj"], g["kk"], c["ii"], c["jj"], g["kkc"], 0) == 0
            self._window(l, b, B)[:, 1:-1, 1:-1, 1:-1] = bc[:, 1:-1, 1:-1, 1:-1]

    def op_interp(self, o):
        l = o["lvl"]
        for b in self.owned:
            g, c = self.g[l, b], self.g[l + 1, b]
            xc = np.array(self._window(l, b, o["b"]), copy=True)
            pw = np.array(self._window(l, b, PI), copy=True)
            assert self.L.orc3_interp_add(P(self.a[l, b, o["a"]]), P(xc), P(self.a[l, b, A]), P(self.a[l, b, RES]), P(pw), c["ii"],
                                          c["jj"], g["kkc"], g["ii"], g["jj"], g["kk"], self._nstl(l), 0) == 0

    def op_zero(self, o):
        for b in self.owned:
            self.a[o["lvl"], b, o["a"]][...] = 0.0

    def op_copy(self, o):
        for b in self.owned:
            self.a[o["lvl"], b, o["b"]][...] = self.a[o["lvl"], b, o["a"]]

    def _gather(self, s):
        gx, gy, gz = self.gn[self.nd]
        ne = self.a[self.nd, self.owned[0], s].shape[0]
        out = np.zeros((ne, gz + 2, gy + 2, gx + 2))
        for b in self.owned:
            g = self.g[self.nd, b]
            out[:, 1 + g["z0"]:1 + g["z0"] + g["nz"]] = self._own(self.nd, b, s)
        return out

    def op_gather(self, o):
        self.exchanges += 1
        self.GB = self._gather(B)[0]

    def op_coarse(self, o):
        self.GX = np.zeros_like(self.GB)
        self.csol.vcycle(self.GX, self.GB)

    def op_scatter(self, o):
        for b in self.owned:
            g = self.g[o["lvl"], b]
            p0 = g["z0"] + 1 - g["wb"]
            self.a[o["lvl"], b, o["a"]][0] = self.GX[p0:p0 + g["kk"]]

    def op_recip(self, o):
        l = o["lvl"]
        for b in self.owned:
            g = self.g[l, b]
            self.L.orc3_setup_recip(P(self.a[l, b, A]), P(self.a[l, b, SOR]), g["ii"], g["jj"], g["kk"], self._nstl(l), 2)

    def op_setup_interp(self, o):
        l = o["lvl"]
        for b in self.owned:
            g, c = self.g[l, b], self.g[l + 1, b]
            pw = np.zeros((26, g["kkc"], c["jj"], c["ii"]))
            assert self.L.orc3_setup_interp_oi(P(self.a[l, b, A]), None, P(pw), g["ii"], g["jj"], g["kk"], c["ii"], c["jj"], g["kkc"],
                                               self._ifd(l), self._nstl(l), 0, 0, None) == 0
            self.tmpP[b] = pw
            self._window(l, b, PI)[...] = pw

    def op_galerkin(self, o):
        l = o["lvl"]
        for b in self.owned:
            g, c = self.g[l, b], self.g[l + 1, b]
            ac = np.zeros((14, g["kkc"], c["jj"], c["ii"]))
            assert self.L.orc3_setup_itli_ex(P(self.a[l, b, A]), P(ac), P(self.tmpP[b]), g["ii"], g["jj"], g["kk"], c["ii"], c["jj"],
                                             g["kkc"], self._ifd(l), self._nstl(l), 0) == 0
            self._window(l, b, A)[...] = ac

    def op_gather_op(self, o):
        self.exchanges += 1
        self.GA = self._gather(A)

    def op_setup_coarse(self, o):
        self.csol = oracle.Solver3(self.GA, self.pre, self.post, num_levels=self.nlevels - self.nd)
        assert self.csol.nlevels == self.nlevels - self.nd

    def set(self, b, slot, local):
        self._own(0, b, slot)[0] = local[1:-1]

    def get(self, b, slot=X0):
        g = self.g[0, b]
        out = np.zeros((g["nz"] + 2, g["jj"], g["ii"]))
        out[1:-1] = self._own(0, b, slot)[0]
        return out

    def vcycle(self):
        self.run(self.p_cycle)
