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
