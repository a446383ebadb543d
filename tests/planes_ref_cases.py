"""The reference's own plane-relaxation test, restated (test/3d/test_planes.cc:13-265 with its checker
test/3d/pyplanes.pyx): six cases XY7 ... YZ27.  One zebra plane sweep (DOWN: odd planes, then even planes) with the
plane problems solved to convergence (test/3d/test-planes-ser.json: point relaxation, max-iter 50, tol 1e-30); then
every EVEN plane -- relaxed last, so its neighbours no longer change -- must equal the sparse direct solve of the
in-plane matrix for the right-hand side copy_rhs builds from the final iterate, to 1e-8 absolute.  pyamg's
stencil_grid is restated with scipy.sparse (pyamg is not in the image); the in-plane stencils are pyplanes.pyx's."""
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as sla

from oracle import planes3

# name, direction, (nx, ny, nz) -- test_planes.cc:17-19, 59-61, 105-107, 147-149, 185-187, 227-229
CASES = [("poisson", "xy", (31, 35, 5)), ("fe", "xy", (31, 35, 5)), ("poisson", "xz", (31, 5, 35)), ("fe", "xz", (31, 5, 35)),
         ("poisson", "yz", (5, 31, 35)), ("fe", "yz", (5, 31, 35))]
PLANE_CONFIG = {"log-planes": False, "solver": {"relaxation": "point", "max-iter": 50, "tol": 1e-30}}
TOL = 1e-8


def index_field(nx, ny, nz):
    """x(i,j,k) = b(i,j,k) = k*nx*ny + j*nx + i on the interior (array indices), zero ghosts (test_planes.cc:30-37)"""
    a = np.zeros((nz + 2, ny + 2, nx + 2))
    k, j, i = np.meshgrid(np.arange(1, nz + 1), np.arange(1, ny + 1), np.arange(1, nx + 1), indexing="ij")
    a[1:-1, 1:-1, 1:-1] = k * nx * ny + j * nx + i
    return a


def stencil_grid(stencil, shape):
    """pyamg.gallery.stencil_grid for a 3x3 stencil on an (n2, n1) grid, Dirichlet: CSR matrix"""
    n2, n1 = shape
    A = sp.lil_matrix((n1 * n2, n1 * n2))
    idx = np.arange(n1 * n2).reshape(n2, n1)
    for dv in (-1, 0, 1):
        for du in (-1, 0, 1):
            w = stencil[dv + 1][du + 1]
            if w == 0:
                continue
            src = idx[max(0, -dv):n2 - max(0, dv), max(0, -du):n1 - max(0, du)]
            dst = idx[max(0, dv):n2 - max(0, -dv), max(0, du):n1 - max(0, -du)]
            A[src.ravel(), dst.ravel()] = w
    return A.tocsr()


def inplane_matrix(name, d, n):
    """the in-plane matrices of pyplanes.pyx (pyplane, pyplane27, pyplane_xz, ..., pyplane_yz27)"""
    nx, ny, nz = n
    if name == "fe":
        st = [[-1.0, -1.0, -1.0], [-1.0, 26.0, -1.0], [-1.0, -1.0, -1.0]]
    else:
        hx, hy, hz = 1.0 / (nx + 1), 1.0 / (ny + 1), 1.0 / (nz + 1)
        xh, yh, zh = hy * hz / hx, hx * hz / hy, hx * hy / hz
        c = 2 * xh + 2 * yh + 2 * zh
        u, v = {"xy": (xh, yh), "xz": (xh, zh), "yz": (yh, zh)}[d]
        st = [[0, -v, 0], [-u, c, -u], [0, -v, 0]]
    shape = {"xy": (ny, nx), "xz": (nz, nx), "yz": (nz, ny)}[d]
    return stencil_grid(st, shape)


def check_even_planes(name, d, n, so, x_after, b):
    """max over the even planes of |x - spsolve(A_plane, copy_rhs(x))|"""
    A = inplane_matrix(name, d, n)
    npl = {"xy": n[2], "xz": n[1], "yz": n[0]}[d]
    worst = 0.0
    for ipl in range(2, npl + 1, 2):
        rhs = planes3.plane_rhs(so, x_after, b, d, ipl)
        want = sla.spsolve(A, rhs.ravel()).reshape(rhs.shape)
        got = planes3._plane(x_after, d, ipl)[1:-1, 1:-1]
        worst = max(worst, float(np.abs(got - want).max()))
    return worst
