"""CPU tests: the hand-written 3D oracle (oracle/bmg3.c) against the REFERENCE's own
3D Fortran kernels, transliterated and compiled at build time into
oracle/_ref/libbmgref.so by oracle/f2c_ref.py.  Bitwise on the kernels restated in
the reference's expression order; to rounding (1e-12 of the diagonal scale) on the
interpolation setup and the Galerkin product, restated in neighbour-offset form."""
import ctypes as C

import numpy as np
import pytest

import oracle
from oracle import P, ref
from tests.util import coarse_n, random_gf3, random_op3

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libbmgref.so not built (no /root/reference)")

SIZES = [(3, 3, 3), (5, 7, 4), (8, 8, 8), (9, 6, 13), (17, 12, 10)]


def ref_setup(so, nxc, nyc, nzc, full):
    """interpolation + Galerkin operator from the transliterated reference"""
    R = ref.lib()
    nst, KK, JJ, II = so.shape
    ifd = 0 if full else 1
    ci = np.zeros((26, nzc + 2, nyc + 2, nxc + 2))
    soc = np.zeros((14, nzc + 2, nyc + 2, nxc + 2))
    yo = np.zeros((14, 2, JJ, II))
    R.BMG3_SymStd_SETUP_interp_OI(P(so), P(soc), P(ci), II, JJ, KK, nxc + 2, nyc + 2, nzc + 2, ifd, nst, 0, 0, P(yo))
    f = R.BMG3_SymStd_SETUP_ITLI27_ex if full else R.BMG3_SymStd_SETUP_ITLI07_ex
    f(P(so), P(soc), P(ci), II, JJ, KK, nxc + 2, nyc + 2, nzc + 2, 0)
    return ci, soc


@pytest.mark.parametrize("full", [False, True])
@pytest.mark.parametrize("nx,ny,nz", SIZES)
def test_3d_kernels_vs_reference(nx, ny, nz, full):
    R, L = ref.lib(), oracle.lib()
    rng = np.random.default_rng(nx * 131 + ny * 17 + nz + full)
    so = random_op3(nx, ny, nz, full, rng)
    nst, ifd = so.shape[0], (0 if full else 1)
    II, JJ, KK = nx + 2, ny + 2, nz + 2
    q, b = random_gf3(nx, ny, nz, rng), random_gf3(nx, ny, nz, rng)
    r1, r2 = np.zeros_like(q), np.zeros_like(q)
    L.orc3_residual(1, 1, ifd, P(q), P(b), P(so), P(r1), II, JJ, KK, nst)
    R.BMG3_SymStd_residual(1, 1, ifd, P(q), P(b), P(so), P(r2), II, JJ, KK, nst)
    assert np.array_equal(r1, r2)
    s1, s2 = np.zeros((2, KK, JJ, II)), np.zeros((2, KK, JJ, II))
    L.orc3_setup_recip(P(so), P(s1), II, JJ, KK, nst, 2)
    R.BMG3_SymStd_SETUP_recip(P(so), P(s2), II, JJ, KK, nst, 2)
    assert np.array_equal(s1, s2)
    for ud in (0, 1):
        q1, q2 = q.copy(), q.copy()
        L.orc3_relax_gs(1, P(so), P(b), P(q1), P(s1), II, JJ, KK, ifd, nst, 2, 1, ud, 0)
        R.BMG3_SymStd_relax_GS(1, P(so), P(b), P(q2), P(s2), II, JJ, KK, ifd, nst, 2, 1, ud, 0)
        assert np.array_equal(q1, q2)
    # setup kernels: to rounding
    nxc, nyc, nzc = coarse_n(nx), coarse_n(ny), coarse_n(nz)
    ci2, c2 = ref_setup(so, nxc, nyc, nzc, full)
    ci1 = np.zeros_like(ci2); c1 = np.zeros_like(c2)
    L.orc3_setup_interp_oi(P(so), P(c1), P(ci1), II, JJ, KK, nxc + 2, nyc + 2, nzc + 2, ifd, nst, 0, 0, None)
    assert np.max(np.abs(ci1 - ci2)) <= 1e-12, np.max(np.abs(ci1 - ci2))
    # Galerkin on the REFERENCE's weights, so that only the product is compared
    L.orc3_setup_itli_ex(P(so), P(c1), P(ci2), II, JJ, KK, nxc + 2, nyc + 2, nzc + 2, ifd, nst, 0)
    scale = np.max(np.abs(c2[0]))
    assert np.max(np.abs(c1 - c2)) <= 1e-12 * scale, (np.max(np.abs(c1 - c2)), scale)
    # transfer kernels: bitwise (on the reference's weights)
    v = random_gf3(nxc, nyc, nzc, rng)
    rc1, rc2 = np.zeros_like(v), np.zeros_like(v)
    L.orc3_restrict(P(q), P(rc1), P(ci2), II, JJ, KK, nxc + 2, nyc + 2, nzc + 2, 0)
    R.BMG3_SymStd_restrict(P(q), P(rc2), P(ci2), II, JJ, KK, nxc + 2, nyc + 2, nzc + 2, 0)
    assert np.array_equal(rc1, rc2)
    qa, qb, ra, rb = q.copy(), q.copy(), b.copy(), b.copy()
    L.orc3_interp_add(P(qa), P(v), P(so), P(ra), P(ci2), nxc + 2, nyc + 2, nzc + 2, II, JJ, KK, nst, 0)
    R.BMG3_SymStd_interp_add(P(qb), P(v), P(so), P(rb), P(ci2), nxc + 2, nyc + 2, nzc + 2, II, JJ, KK, nst, 0)
    assert np.array_equal(ra, rb)
    assert np.array_equal(qa, qb)


@pytest.mark.parametrize("full", [False, True])
@pytest.mark.parametrize("nx,ny,nz", [(3, 3, 3), (4, 3, 5)])
def test_3d_coarse_solve_vs_reference(nx, ny, nz, full):
    R, L = ref.lib(), oracle.lib()
    rng = np.random.default_rng(7 + nx + full)
    so = random_op3(nx, ny, nz, full, rng)
    nst = so.shape[0]
    II, JJ, KK = nx + 2, ny + 2, nz + 2
    n1, n2 = nx * (ny + 1) + 2, nx * ny * nz
    a1, a2 = np.zeros((n2, n1)), np.zeros((n2, n1))
    assert L.orc3_setup_cg_lu(P(so), II, JJ, KK, nst, P(a1), n1, n2, 0) == 0
    R.BMG3_SymStd_SETUP_cg_LU(P(so), II, JJ, KK, nst, P(a2), n1, n2, 0)
    assert np.array_equal(a1, a2)
    b = random_gf3(nx, ny, nz, rng)
    x1, x2 = np.zeros_like(b), np.zeros_like(b)
    bb1, bb2 = np.zeros(n2), np.zeros(n2)
    L.orc3_solve_cg(P(x1), P(b), II, JJ, KK, P(a1), P(bb1), n1, n2, 0)
    R.BMG3_SymStd_SOLVE_cg(P(x2), P(b), II, JJ, KK, P(a2), P(bb2), n1, n2, 0)
    assert np.array_equal(x1, x2)
    # and it solves the system
    r = np.zeros_like(b)
    L.orc3_residual(1, 1, 0 if full else 1, P(x1), P(b), P(so), P(r), II, JJ, KK, nst)
    assert np.max(np.abs(r)) < 1e-10 * np.max(np.abs(b))
