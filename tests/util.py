"""Shared helpers for the parity tests (problem generators, device round trips)."""
import ctypes as C

import numpy as np


def random_op2(nx, ny, nine, rng, hetero=True):
    """Heterogeneous symmetric M-matrix-like 5/9-pt operator in Cedar storage:
    array (ndirs, ny+2, nx+2), dirs c,w,s[,sw,nw]; couplings across the boundary
    are absent from w/s/sw/nw (as in src/2d/gallery.cc) but kept in the centre."""
    nst = 5 if nine else 3
    so = np.zeros((nst, ny + 2, nx + 2))

    def coef(shape):
        return 10.0 ** rng.uniform(-2, 2, size=shape) if hetero else np.ones(shape)

    # C++ 0-based indices as in gallery.cc
    so[2, 2:ny + 1, 1:nx + 1] = coef((ny - 1, nx)) if ny > 1 else 0
    so[1, 1:ny + 1, 2:nx + 1] = coef((ny, nx - 1)) if nx > 1 else 0
    if nine and nx > 1 and ny > 1:
        so[3, 2:ny + 1, 2:nx + 1] = coef((ny - 1, nx - 1))
        so[4, 2:ny + 1, 2:nx + 1] = coef((ny - 1, nx - 1))
    full = np.zeros((ny + 3, nx + 3))
    w = np.zeros((ny + 3, nx + 3)); w[:ny + 2, :nx + 2] = so[1]
    s = np.zeros((ny + 3, nx + 3)); s[:ny + 2, :nx + 2] = so[2]
    full[:-1, :-1] = w[:-1, :-1] + w[:-1, 1:] + s[:-1, :-1] + s[1:, :-1]
    if nine:
        sw = np.zeros((ny + 3, nx + 3)); sw[:ny + 2, :nx + 2] = so[3]
        nw = np.zeros((ny + 3, nx + 3)); nw[:ny + 2, :nx + 2] = so[4]
        full[:-1, :-1] += sw[:-1, :-1] + sw[1:, 1:] + nw[:-1, 1:] + nw[1:, :-1]
    c = full[:ny + 2, :nx + 2] + coef((ny + 2, nx + 2)) * 0.1 + 1e-3
    so[0, 1:ny + 1, 1:nx + 1] = c[1:ny + 1, 1:nx + 1]
    return so


def random_gf(nx, ny, rng, ghosts_zero=True):
    q = rng.standard_normal((ny + 2, nx + 2))
    if ghosts_zero:
        q[0, :] = 0; q[-1, :] = 0; q[:, 0] = 0; q[:, -1] = 0
    return q


def coarse_n(n):
    return int((n - 1) / 2.0 + 1)


def dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def host(t):
    return t.detach().cpu().numpy()


def ptr(t):
    return C.c_void_p(t.data_ptr())


def assert_bitwise(a, b, what=""):
    a = np.asarray(a); b = np.asarray(b)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    same = (a == b) | (np.isnan(a) & np.isnan(b))
    if not same.all():
        idx = np.argwhere(~same)
        k = tuple(idx[0])
        raise AssertionError(f"{what}: {len(idx)} of {a.size} entries differ; first at {k}: {a[k]!r} vs {b[k]!r}; "
                             f"max abs diff {np.nanmax(np.abs(a - b))}")


# 3D entry order (src/3d/ftn/BMG_stencils_f90.h:43-48) and, per entry, the two
# grid points it couples relative to its storage point S: (a, b) = offsets of the
# two end points (read off BMG3_SymStd_residual.f90:74-100)
ENTRIES3 = ["p", "pw", "ps", "b", "psw", "pnw", "bw", "bnw", "bn", "bne", "be", "bse", "bs", "bsw"]
ENDS3 = {1: ((0, 0, 0), (-1, 0, 0)), 2: ((0, 0, 0), (0, -1, 0)), 3: ((0, 0, 0), (0, 0, -1)),
         4: ((0, 0, 0), (-1, -1, 0)), 5: ((0, -1, 0), (-1, 0, 0)), 6: ((0, 0, 0), (-1, 0, -1)),
         7: ((0, -1, 0), (-1, 0, -1)), 8: ((0, -1, 0), (0, 0, -1)), 9: ((-1, -1, 0), (0, 0, -1)),
         10: ((-1, 0, 0), (0, 0, -1)), 11: ((-1, 0, 0), (0, -1, -1)), 12: ((0, 0, 0), (0, -1, -1)),
         13: ((0, 0, 0), (-1, -1, -1))}
SEVEN3 = (1, 2, 3)


def random_op3(nx, ny, nz, full, rng, hetero=True):
    """Heterogeneous symmetric diagonally dominant 7/27-point operator in Cedar
    storage: (ndirs, nz+2, ny+2, nx+2); a coupling is present iff both end points
    are interior (src/3d/gallery.cc); the diagonal is the row sum of all couplings
    plus a positive shift (so boundary couplings are "kept in the centre")."""
    nst = 14 if full else 4
    so = np.zeros((nst, nz + 2, ny + 2, nx + 2))
    rowsum = np.zeros((nz + 2, ny + 2, nx + 2))
    n = (nx, ny, nz)
    for e in (range(1, 14) if full else SEVEN3):
        (a, b) = ENDS3[e]
        # storage points S (0-based array index) for which S+a and S+b are interior (1..n)
        lo = [max(1 - a[d], 1 - b[d], 1) for d in range(3)]
        hi = [min(n[d] - a[d], n[d] - b[d], n[d] + 1) for d in range(3)]
        if any(lo[d] > hi[d] for d in range(3)):
            continue
        shape = (hi[2] - lo[2] + 1, hi[1] - lo[1] + 1, hi[0] - lo[0] + 1)
        v = 10.0 ** rng.uniform(-2, 2, size=shape) if hetero else np.ones(shape)
        so[e, lo[2]:hi[2] + 1, lo[1]:hi[1] + 1, lo[0]:hi[0] + 1] = v
        for end in (a, b):
            rowsum[lo[2] + end[2]:hi[2] + 1 + end[2], lo[1] + end[1]:hi[1] + 1 + end[1],
                   lo[0] + end[0]:hi[0] + 1 + end[0]] += v
    shift = (10.0 ** rng.uniform(-2, 1, size=rowsum.shape) if hetero else 1.0)
    so[0, 1:-1, 1:-1, 1:-1] = (rowsum * (1.0 + 0.05 * rng.uniform(size=rowsum.shape)) + shift)[1:-1, 1:-1, 1:-1]
    return so


def random_gf3(nx, ny, nz, rng):
    q = np.zeros((nz + 2, ny + 2, nx + 2))
    q[1:-1, 1:-1, 1:-1] = rng.standard_normal((nz, ny, nx))
    return q


def periodic_problem2(nx, ny, periodic):
    """operator, right-hand side and exact solution of the reference's examples/basic-2d-ser/periodic.cc:16-170
    (5-point Laplacian with periodic wrap in the chosen directions), in Cedar storage with the periodic ghost copies"""
    px, py = periodic
    so = np.zeros((3, ny + 2, nx + 2))
    n1, n2 = nx - (1 if px else 0), ny - (1 if py else 0)
    hx, hy = 1.0 / (n1 + 1), 1.0 / (n2 + 1)
    xh, yh = hy / hx, hx / hy
    ibeg, jbeg = (1 if px else 2), (1 if py else 2)          # C++ 0-based indices with ghost: interior 1..n
    so[2, jbeg:ny + 1, 1:nx + 1] = 1.0 * yh
    so[1, 1:ny + 1, ibeg:nx + 1] = 1.0 * xh
    so[0, 1:ny + 1, 1:nx + 1] = 2 * xh + 2 * yh
    if px:
        so[:, 1:ny + 1, ibeg - 1] = so[:, 1:ny + 1, nx]
        so[:, 1:ny + 1, nx + 1] = so[:, 1:ny + 1, ibeg]
    if py:
        so[:, jbeg - 1, 1:nx + 1] = so[:, ny, 1:nx + 1]
        so[:, ny + 1, 1:nx + 1] = so[:, jbeg, 1:nx + 1]
    jj, ii = np.meshgrid(np.arange(ny + 2), np.arange(nx + 2), indexing="ij")
    b = np.zeros((ny + 2, nx + 2))
    rhs = 8 * (np.pi * np.pi) * np.sin(2 * np.pi * (ii * hx)) * np.sin(2 * np.pi * (jj * hy)) * (hx * hy)
    b[1:-1, 1:-1] = rhs[1:-1, 1:-1]
    if px:
        b[:, 0] = b[:, nx]; b[:, nx + 1] = b[:, 1]
    if py:
        b[0, :] = b[ny, :]; b[ny + 1, :] = b[1, :]
    exact = np.zeros((ny + 2, nx + 2))
    exact[1:-1, 1:-1] = (np.sin(2 * np.pi * (ii * hx)) * np.sin(2 * np.pi * (jj * hy)))[1:-1, 1:-1]
    return so, b, exact
