"""TEST INFRASTRUCTURE: CPU restatement of the reference's 3D solver with PLANE relaxation
(include/cedar/3d/relax_planes.h:35-261, src/3d/relax_planes.cc:26-238, multilevel.h:165-222),
driven from Python over the C oracle kernels (oracle/bmg3.c for the 3D grid transfer and
setup kernels, oracle.Solver2 for the in-plane 2D solves).

The reference quirk of SURVEY.md 8(a) P3 is restated as it is: copy_coeff is called
without a plane index, so every plane solver holds the coefficients of the LAST interior
plane; one oracle.Solver2 per direction and level therefore stands for all of them.
"""
import ctypes as C

import numpy as np

from . import P, Solver2, lib

# off-plane neighbour offsets (dx,dy,dz) in the reference's term order, per direction
_RHS27 = {
    "xy": [(0, 0, -1), (-1, 0, -1), (-1, 1, -1), (0, 1, -1), (1, 1, -1), (1, 0, -1), (1, -1, -1), (0, -1, -1), (-1, -1, -1),
           (-1, 0, 1), (-1, 1, 1), (0, 1, 1), (1, 1, 1), (1, 0, 1), (0, 0, 1), (1, -1, 1), (0, -1, 1), (-1, -1, 1)],
    "xz": [(-1, 1, 0), (0, 1, 0), (1, 1, 0), (-1, 1, -1), (0, 1, -1), (1, 1, -1), (-1, 1, 1), (0, 1, 1), (1, 1, 1),
           (-1, -1, 0), (0, -1, 0), (1, -1, 0), (-1, -1, -1), (0, -1, -1), (1, -1, -1), (-1, -1, 1), (0, -1, 1), (1, -1, 1)],
    "yz": [(-1, 1, 0), (-1, 0, 0), (-1, -1, 0), (-1, 1, -1), (-1, 0, -1), (-1, -1, -1), (-1, 1, 1), (-1, 0, 1), (-1, -1, 1),
           (1, 1, 0), (1, 0, 0), (1, -1, 0), (1, 1, -1), (1, 0, -1), (1, -1, -1), (1, 1, 1), (1, 0, 1), (1, -1, 1)],
}
_RHS7 = {"xy": [(0, 0, -1), (0, 0, 1)], "xz": [(0, -1, 0), (0, 1, 0)], "yz": [(-1, 0, 0), (1, 0, 0)]}

# (entry index 0-based, storage shift) of the coupling p ~ p+d (BMG3_SymStd_residual.f90:74-100)
_CPL = {(-1, 0, 0): (1, (0, 0, 0)), (-1, 1, 0): (5, (0, 1, 0)), (0, 1, 0): (2, (0, 1, 0)), (1, 1, 0): (4, (1, 1, 0)),
        (1, 0, 0): (1, (1, 0, 0)), (1, -1, 0): (5, (1, 0, 0)), (0, -1, 0): (2, (0, 0, 0)), (-1, -1, 0): (4, (0, 0, 0)),
        (0, 0, -1): (3, (0, 0, 0)), (-1, 0, -1): (6, (0, 0, 0)), (-1, 1, -1): (7, (0, 1, 0)), (0, 1, -1): (8, (0, 1, 0)),
        (1, 1, -1): (9, (1, 1, 0)), (1, 0, -1): (10, (1, 0, 0)), (1, -1, -1): (11, (1, 0, 0)), (0, -1, -1): (12, (0, 0, 0)),
        (-1, -1, -1): (13, (0, 0, 0)),
        (0, 0, 1): (3, (0, 0, 1)), (-1, 0, 1): (10, (0, 0, 1)), (-1, 1, 1): (11, (0, 1, 1)), (0, 1, 1): (12, (0, 1, 1)),
        (1, 1, 1): (13, (1, 1, 1)), (1, 0, 1): (6, (1, 0, 1)), (1, -1, 1): (7, (1, 0, 1)), (0, -1, 1): (8, (0, 0, 1)),
        (-1, -1, 1): (9, (0, 0, 1))}
# in-plane 2D slots c,w,s,sw,nw <- 3D entries (relax_planes.h:84-161)
_COEFF = {"xy": [0, 1, 2, 4, 5], "xz": [0, 1, 3, 6, 10], "yz": [0, 2, 3, 12, 8]}
_AXIS = {"xy": 0, "xz": 1, "yz": 2}      # numpy axis (k,j,i order) of the plane index


def _plane(a, d, ipl):
    """view of plane ipl (array index) of a (.., nz+2, ny+2, nx+2) array, as a 2D (v, u) array"""
    if d == "xy":
        return a[..., ipl, :, :]
    if d == "xz":
        return a[..., :, ipl, :]
    return a[..., :, :, ipl]


def _shift(a, off):
    """a[k+dz, j+dy, i+dx] on the interior points"""
    dx, dy, dz = off
    nz, ny, nx = a.shape[0] - 2, a.shape[1] - 2, a.shape[2] - 2
    return a[1 + dz:1 + dz + nz, 1 + dy:1 + dy + ny, 1 + dx:1 + dx + nx]


class _Planes:
    """the plane solvers of one direction on one level"""

    def __init__(self, so, d, plane_kw, iters=1):
        """iters: cycles per plane visit (plane-config "max-iter", default 1; the reference's own plane test uses 50
        with tol 1e-30, test/3d/test-planes-ser.json)"""
        self.iters = iters
        nst = so.shape[0]
        n = so.shape[1 + _AXIS[d]] - 2
        slots = _COEFF[d][:5 if nst == 14 else 3]
        so2 = np.ascontiguousarray(np.stack([_plane(so[e], d, n) for e in slots]))    # last interior plane
        self.d, self.n, self.so2 = d, n, so2
        self.slv = Solver2(so2, **plane_kw)

    def relax(self, so, x, b, down):
        d, nst = self.d, so.shape[0]
        offs = (_RHS27 if nst == 14 else _RHS7)[d]
        # b + off-plane couplings * x for ALL planes would use values of the wrong colour:
        # build the right-hand side plane by plane, like copy_rhs
        for beg in ((1, 2) if down else (2, 1)):
            for ipl in range(beg, self.n + 1, 2):
                rhs3 = None
                x2 = np.ascontiguousarray(_plane(x, d, ipl))
                b2 = np.zeros_like(x2)
                acc = _plane(b, d, ipl)[1:-1, 1:-1].copy()
                for off in offs:
                    e, s = _CPL[off]
                    cp = _plane_int(_shift(so[e], s), d, ipl)
                    xv = _plane_int(_shift(x, off), d, ipl)
                    acc = acc + cp * xv
                b2[1:-1, 1:-1] = acc
                for _ in range(self.iters):
                    self.slv.vcycle(x2, b2)
                _plane(x, d, ipl)[...] = x2


def plane_rhs(so, x, b, d, ipl):
    """copy_rhs (src/3d/relax_planes.cc:26-178): right-hand side of plane ipl's 2D problem, interior points"""
    offs = (_RHS27 if so.shape[0] == 14 else _RHS7)[d]
    acc = _plane(b, d, ipl)[1:-1, 1:-1].copy()
    for off in offs:
        e, s = _CPL[off]
        acc = acc + _plane_int(_shift(so[e], s), d, ipl) * _plane_int(_shift(x, off), d, ipl)
    return acc


def _plane_int(a_int, d, ipl):
    """plane ipl (array index in the full array) of an interior-only (nz,ny,nx) array"""
    if d == "xy":
        return a_int[ipl - 1, :, :]
    if d == "xz":
        return a_int[:, ipl - 1, :]
    return a_int[:, :, ipl - 1]


class Solver3Planes:
    """cdr3::solver<fsten> with relaxation plane-xy / plane-xz / plane-yz / plane-xyz"""

    def __init__(self, so, relaxation="plane-xyz", nrelax_pre=2, nrelax_post=1, min_coarse=3, plane_kw=None):
        L = self.L = lib()
        self.relaxation, self.npre, self.npost = relaxation, nrelax_pre, nrelax_post
        # default plane_config: line-xy, one cycle per visit, default V(2,1) (src/kernel_params.cc:72-78)
        self.plane_kw = plane_kw or dict(relaxation="line-xy", nrelax_pre=2, nrelax_post=1)
        so = np.ascontiguousarray(so)
        nst, kk, jj, ii = so.shape
        nx, ny, nz = ii - 2, jj - 2, kk - 2
        ng = 0
        while True:        # include/cedar/3d/solver.h:54-72
            ng += 1
            if min((nx - 1) // (1 << ng) + 1, (ny - 1) // (1 << ng) + 1, (nz - 1) // (1 << ng) + 1) < min_coarse:
                break
        self.nlevels = ng
        self.A, self.Pw, self.x, self.b, self.res, self.planes = [so], [None], [], [], [], []
        for l in range(ng):
            nst_l, kk, jj, ii = self.A[l].shape
            self.x.append(np.zeros((kk, jj, ii))); self.b.append(np.zeros((kk, jj, ii))); self.res.append(np.zeros((kk, jj, ii)))
            if l == ng - 1:
                break
            nxc, nyc, nzc = [int((n - 3) / 2.0 + 1) + 2 for n in (ii, jj, kk)]
            ci = np.zeros((26, nzc, nyc, nxc)); soc = np.zeros((14, nzc, nyc, nxc))
            ifd = 1 if nst_l == 4 else 0
            L.orc3_setup_interp_oi(P(self.A[l]), P(soc), P(ci), ii, jj, kk, nxc, nyc, nzc, ifd, nst_l, 0, 0, None)
            L.orc3_setup_itli_ex(P(self.A[l]), P(soc), P(ci), ii, jj, kk, nxc, nyc, nzc, ifd, nst_l, 0)
            self.A.append(soc); self.Pw.append(ci)
            # multilevel.h:199-222 setup order for plane-xyz: xy, xz, yz
            dirs = {"plane-xyz": ["xy", "xz", "yz"]}.get(relaxation, [relaxation[-2:]])
            self.planes.append({d: _Planes(self.A[l], d, self.plane_kw) for d in dirs})
        C_ = self.A[-1]
        nst_c, kk, jj, ii = C_.shape
        self.n1, self.n2 = (ii - 2) * (jj - 1) + 2, (ii - 2) * (jj - 2) * (kk - 2)
        self.abd = np.zeros((self.n2, self.n1)); self.bbd = np.zeros(self.n2)
        assert L.orc3_setup_cg_lu(P(C_), ii, jj, kk, nst_c, P(self.abd), self.n1, self.n2, 0) == 0

    def _sweep(self, l, x, b, down):
        pl = self.planes[l]
        if self.relaxation == "plane-xyz":       # multilevel.h:179-183 / 208-212
            order = ["xy", "yz", "xz"] if down else ["xz", "yz", "xy"]
        else:
            order = [self.relaxation[-2:]]
        for d in order:
            pl[d].relax(self.A[l], x, b, down)

    def _ncycle(self, l, x, b):
        L = self.L
        A = self.A[l]
        nst, kk, jj, ii = A.shape
        ifd = 1 if nst == 4 else 0
        kkc, jjc, iic = self.x[l + 1].shape
        for _ in range(self.npre):
            self._sweep(l, x, b, True)
        L.orc3_residual(1, 1, ifd, P(x), P(b), P(A), P(self.res[l]), ii, jj, kk, nst)
        L.orc3_restrict(P(self.res[l]), P(self.b[l + 1]), P(self.Pw[l + 1]), ii, jj, kk, iic, jjc, kkc, 0)
        self.x[l + 1][...] = 0
        if l + 1 == self.nlevels - 1:
            L.orc3_solve_cg(P(self.x[l + 1]), P(self.b[l + 1]), iic, jjc, kkc, P(self.abd), P(self.bbd), self.n1, self.n2, 0)
        else:
            self._ncycle(l + 1, self.x[l + 1], self.b[l + 1])
        L.orc3_interp_add(P(x), P(self.x[l + 1]), P(A), P(self.res[l]), P(self.Pw[l + 1]), iic, jjc, kkc, ii, jj, kk, nst, 0)
        for _ in range(self.npost):
            self._sweep(l, x, b, False)

    def vcycle(self, x, b):
        self._ncycle(0, x, b)

    def solve(self, b, x=None, maxiter=10, tol=1e-8):
        L = self.L
        x = np.zeros_like(b) if x is None else x
        A = self.A[0]
        nst, kk, jj, ii = A.shape
        ifd = 1 if nst == 4 else 0
        L.orc3_l2norm.restype = C.c_double
        L.orc3_residual(1, 1, ifd, P(x), P(b), P(A), P(self.res[0]), ii, jj, kk, nst)
        r0 = L.orc3_l2norm(P(self.res[0]), ii, jj, kk)
        hist = []
        for _ in range(maxiter):
            self.vcycle(x, b)
            L.orc3_residual(1, 1, ifd, P(x), P(b), P(A), P(self.res[0]), ii, jj, kk, nst)
            hist.append(L.orc3_l2norm(P(self.res[0]), ii, jj, kk) / r0)
            if hist[-1] < tol:
                break
        return x, np.array(hist), r0
