"""Host-side operator generators that the reference does not have: the 5-point
finite-volume discretisation of -div(D grad u) on the unit square with a heterogeneous coefficient field.  The
same arrays feed the CPU oracle and the CUDA path, so parity runs compare the solvers, not two generators.

Layout = the reference's symmetric stencil storage (include/cedar/2d/stencil_op.h:9-13): shape (3, ny+2, nx+2),
slots c=0, w=1, s=2, off-diagonals stored positive, ghost ring zero; Dirichlet boundary edges are dropped from the
W/S slots and kept in the centre, as gallery::poisson does (src/2d/gallery.cc:20-36)."""
import numpy as np


def checkerboard_field(nx, ny, nblk=64, seed=20260411, lo=-3.0, hi=3.0):
    """D on the nx-by-ny interior points: nblk x nblk blocks, D = 10**U(lo, hi) per block."""
    rng = np.random.default_rng(seed)
    blocks = 10.0 ** rng.uniform(lo, hi, size=(nblk, nblk))
    jb = (np.arange(ny) * nblk) // ny
    ib = (np.arange(nx) * nblk) // nx
    return blocks[np.ix_(jb, ib)]


def fv_diffusion5(Dx, Dy=None):
    """5-point operator for -d/dx(Dx du/dx) - d/dy(Dy du/dy); Dx, Dy of shape (ny, nx) (cell values at the unknowns).
    Edge coefficients are harmonic means of the two cells; boundary edges use the cell's own value."""
    Dx = np.asarray(Dx, dtype=np.float64)
    Dy = Dx if Dy is None else np.asarray(Dy, dtype=np.float64)
    ny, nx = Dx.shape
    hx, hy = 1.0 / (nx + 1), 1.0 / (ny + 1)
    xh, yh = hy / hx, hx / hy

    def harm(a, b):
        return 2.0 * a * b / (a + b)
    so = np.zeros((3, ny + 2, nx + 2))
    ew = np.empty((ny, nx + 1))          # x-edges: ew[:, i] between unknowns i-1 and i (0 and nx: boundary)
    ew[:, 1:nx] = harm(Dx[:, :-1], Dx[:, 1:])
    ew[:, 0], ew[:, nx] = Dx[:, 0], Dx[:, -1]
    es = np.empty((ny + 1, nx))
    es[1:ny, :] = harm(Dy[:-1, :], Dy[1:, :])
    es[0, :], es[ny, :] = Dy[0, :], Dy[-1, :]
    ew *= xh
    es *= yh
    so[1, 1:-1, 2:nx + 1] = ew[:, 1:nx]
    so[2, 2:ny + 1, 1:-1] = es[1:ny, :]
    so[0, 1:-1, 1:-1] = ew[:, :-1] + ew[:, 1:] + es[:-1, :] + es[1:, :]
    return so


def hetero5(nx, ny, nblk=64, seed=20260411):
    """config C3(b): isotropic checkerboard coefficient, jumps of up to 10**6"""
    D = checkerboard_field(nx, ny, nblk, seed)
    return fv_diffusion5(D, D)
