"""CPU tests: the hand-written oracle (oracle/bmg2.c) against the REFERENCE's own
Fortran kernels, transliterated and compiled at build time into
oracle/_ref/libbmgref.so by oracle/f2c_ref.py (built where /root/reference
exists; the prebuilt library travels).  Bitwise on every 2D kernel."""
import numpy as np
import pytest

import oracle
from oracle import P, ref
from tests.util import coarse_n, random_gf, random_op2

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libbmgref.so not built (no /root/reference)")

SIZES = [(3, 3), (7, 5), (16, 16), (37, 64), (130, 97)]


@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", SIZES)
def test_2d_kernels_bitwise_vs_reference(nx, ny, nine):
    R, L = ref.lib(), oracle.lib()
    rng = np.random.default_rng(nx * 31 + ny + nine)
    so = random_op2(nx, ny, nine, rng)
    nst, ifd = so.shape[0], (0 if nine else 1)
    II, JJ = nx + 2, ny + 2
    q, b = random_gf(nx, ny, rng), random_gf(nx, ny, rng)
    # residual (by-reference scalars in the reference)
    r1, r2 = np.zeros_like(q), np.zeros_like(q)
    L.orc2_residual(0, P(so), P(b), P(q), P(r1), II, JJ, 0, ifd, nst)
    R.BMG2_SymStd_residual(ref.ip(0), P(so), P(b), P(q), P(r2), ref.ip(II), ref.ip(JJ), ref.ip(0), ref.ip(ifd),
                           ref.ip(nst), ref.ip(0), ref.ip(0), ref.ip(1), ref.ip(0))
    assert np.array_equal(r1, r2)
    # point relaxation
    s1, s2 = np.zeros((2, JJ, II)), np.zeros((2, JJ, II))
    L.orc2_setup_recip(P(so), P(s1), II, JJ, nst, 2)
    R.BMG2_SymStd_SETUP_recip(P(so), P(s2), II, JJ, nst, 2)
    assert np.array_equal(s1, s2)
    for ud in (0, 1):
        q1, q2 = q.copy(), q.copy()
        L.orc2_relax_gs(1, P(so), P(b), P(q1), P(s1), II, JJ, 1, ifd, nst, 2, 1, ud, 0)
        R.BMG2_SymStd_relax_GS(1, P(so), P(b), P(q2), P(s2), II, JJ, 1, ifd, nst, 2, 1, ud, 0)
        assert np.array_equal(q1, q2)
    # interpolation setup, Galerkin, restriction, interpolate-and-add
    nxc, nyc = coarse_n(nx), coarse_n(ny)
    ci1 = np.zeros((8, nyc + 2, nxc + 2)); ci2 = ci1.copy()
    c1 = np.zeros((5, nyc + 2, nxc + 2)); c2 = c1.copy()
    L.orc2_setup_interp_oi(P(so), P(c1), P(ci1), II, JJ, nxc + 2, nyc + 2, ifd, nst, 0, 0)
    R.BMG2_SymStd_SETUP_interp_OI(P(so), P(c2), P(ci2), II, JJ, nxc + 2, nyc + 2, ifd, nst, 0, 0)
    assert np.array_equal(ci1, ci2)
    L.orc2_setup_itli_ex(P(so), P(c1), P(ci1), II, JJ, nxc + 2, nyc + 2, ifd, nst, 0)
    R.BMG2_SymStd_SETUP_ITLI_ex(P(so), P(c2), P(ci2), II, JJ, nxc + 2, nyc + 2, ifd, nst, 0)
    assert np.array_equal(c1, c2)
    v = random_gf(nxc, nyc, rng)
    rc1, rc2 = np.zeros_like(v), np.zeros_like(v)
    L.orc2_restrict(P(q), P(rc1), P(ci1), II, JJ, nxc + 2, nyc + 2, 0)
    R.BMG2_SymStd_restrict(P(q), P(rc2), P(ci2), II, JJ, nxc + 2, nyc + 2, 0)
    assert np.array_equal(rc1, rc2)
    qa, qb, ra, rb = q.copy(), q.copy(), b.copy(), b.copy()
    L.orc2_interp_add(P(qa), P(v), P(ra), P(so), P(ci1), nxc + 2, nyc + 2, II, JJ, nst, 0)
    R.BMG2_SymStd_interp_add(P(qb), P(v), P(rb), P(so), P(ci2), nxc + 2, nyc + 2, II, JJ, nst, 0)
    assert np.array_equal(qa, qb) and np.array_equal(ra, rb)
    # line relaxation (LAPACK bound to the restated DPTTRF / DPTTS2)
    x1, x2 = np.zeros((2, JJ, II)), np.zeros((2, JJ, II))
    y1, y2 = np.zeros((2, II, JJ)), np.zeros((2, II, JJ))
    L.orc2_setup_lines_x(P(so), P(x1), II, JJ, nst, 0); R.BMG2_SymStd_SETUP_lines_x(P(so), P(x2), II, JJ, nst, 0)
    L.orc2_setup_lines_y(P(so), P(y1), II, JJ, nst, 0); R.BMG2_SymStd_SETUP_lines_y(P(so), P(y2), II, JJ, nst, 0)
    assert np.array_equal(x1[:, 1:-1, 1:-1], x2[:, 1:-1, 1:-1]) and np.array_equal(y1[:, 1:-1, 1:-1], y2[:, 1:-1, 1:-1])
    w = np.zeros(4 * max(II, JJ))
    for ud in (0, 1):
        q1, q2 = q.copy(), q.copy()
        L.orc2_relax_lines_x(1, P(so), P(b), P(q1), P(x1), P(w), II, JJ, 1, ifd, nst, 1, ud, 0)
        R.BMG2_SymStd_relax_lines_x(1, P(so), P(b), P(q2), P(x2), P(w), II, JJ, 1, ifd, nst, 1, ud, 0)
        assert np.array_equal(q1, q2)
        L.orc2_relax_lines_y(1, P(so), P(b), P(q1), P(y1), P(w), II, JJ, 1, ifd, nst, 1, ud, 0)
        R.BMG2_SymStd_relax_lines_y(1, P(so), P(b), P(q2), P(y2), P(w), II, JJ, 1, ifd, nst, 1, ud, 0)
        assert np.array_equal(q1, q2)


@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(3, 3), (4, 5), (7, 4)])
def test_2d_coarse_solve_bitwise_vs_reference(nx, ny, nine):
    R, L = ref.lib(), oracle.lib()
    rng = np.random.default_rng(nx * 7 + ny + nine)
    so = random_op2(nx, ny, nine, rng)
    nst = so.shape[0]
    II, JJ = nx + 2, ny + 2
    n1, n2 = nx + 2, nx * ny
    a1, a2 = np.zeros((n2, n1)), np.zeros((n2, n1))
    assert L.orc2_setup_cg_lu(P(so), II, JJ, nst, P(a1), n1, n2, 0) == 0
    R.BMG2_SymStd_SETUP_cg_LU(P(so), ref.ip(II), ref.ip(JJ), ref.ip(nst), P(a2), ref.ip(n1), ref.ip(n2), ref.ip(0))
    assert np.array_equal(a1, a2)
    b = random_gf(nx, ny, rng)
    x1, x2, w1, w2 = np.zeros_like(b), np.zeros_like(b), np.zeros(n2), np.zeros(n2)
    L.orc2_solve_cg(P(x1), P(b), II, JJ, P(a1), P(w1), n1, n2, 0)
    R.BMG2_SymStd_SOLVE_cg(P(x2), P(b), II, JJ, P(a2), P(w2), n1, n2, 0)
    assert np.array_equal(x1, x2)


@pytest.mark.parametrize("jpn", [1, 2, 3, -1, -2, -3])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(3, 3), (4, 4), (8, 6), (7, 5), (16, 17), (37, 64)])
def test_periodic_point_relaxation_bitwise_vs_reference(nx, ny, nine, jpn):
    """Periodic branch of BMG2_SymStd_relax_GS (f90:139-224): x ghosts refreshed per row and colour, y ghost rows once
    per sweep.  Ghosts deliberately start inconsistent with the interior: the reference does not fix them up front."""
    R, L = ref.lib(), oracle.lib()
    rng = np.random.default_rng(nx * 131 + ny * 7 + nine + 10 * abs(jpn))
    II, JJ = nx + 2, ny + 2
    nst, ifd = (5, 0) if nine else (3, 1)
    # periodic operators carry couplings in the ghost layer too
    so = 10.0 ** rng.uniform(-1, 1, size=(nst, JJ, II))
    so[0] = so[1:].sum(axis=0) * 2.5 + 1.0
    b = rng.standard_normal((JJ, II))
    q = rng.standard_normal((JJ, II))
    s1, s2 = np.zeros((2, JJ, II)), np.zeros((2, JJ, II))
    L.orc2_setup_recip(P(so), P(s1), II, JJ, nst, 2)
    R.BMG2_SymStd_SETUP_recip(P(so), P(s2), II, JJ, nst, 2)
    assert np.array_equal(s1, s2)
    q1, q2 = q.copy(), q.copy()
    for ud in (0, 1, 1, 0):
        assert L.orc2_relax_gs(1, P(so), P(b), P(q1), P(s1), II, JJ, 1, ifd, nst, 2, 1, ud, jpn) == 0
        R.BMG2_SymStd_relax_GS(1, P(so), P(b), P(q2), P(s2), II, JJ, 1, ifd, nst, 2, 1, ud, jpn)
        assert np.array_equal(q1, q2)
    assert not np.array_equal(q1, q)


@pytest.mark.parametrize("jpn", [1, 2, 3, -1, -2, -3])
@pytest.mark.parametrize("nx,ny", [(4, 4), (8, 6), (7, 5), (16, 17), (37, 64), (64, 38)])
def test_periodic_restrict_bitwise_vs_reference(nx, ny, jpn):
    """Periodic branch of BMG2_SymStd_restrict (f90:96-128): refreshes the fine vector's ghost layer (mutating its
    argument) when the coarse grid is the N/2+1 coarsening, then applies the same stencil."""
    R, L = ref.lib(), oracle.lib()
    rng = np.random.default_rng(nx * 17 + ny + 100 * abs(jpn))
    II, JJ = nx + 2, ny + 2
    for iic, jjc in [(II // 2 + 1, JJ // 2 + 1), (coarse_n(nx) + 2, coarse_n(ny) + 2)]:
        if 2 * (iic - 1) > II or 2 * (jjc - 1) > JJ:
            continue            # the stencil must stay inside the fine array
        ci = rng.uniform(0, 1, size=(8, jjc, iic))
        q = rng.standard_normal((JJ, II))
        q1, q2 = q.copy(), q.copy()
        c1, c2 = np.zeros((jjc, iic)), np.zeros((jjc, iic))
        assert L.orc2_restrict(P(q1), P(c1), P(ci), II, JJ, iic, jjc, jpn) == 0
        R.BMG2_SymStd_restrict(P(q2), P(c2), P(ci), II, JJ, iic, jjc, jpn)
        assert np.array_equal(c1, c2) and np.array_equal(q1, q2)


@pytest.mark.parametrize("jpn", [1, 2, 3, -1, -2, -3])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(4, 4), (8, 6), (7, 5), (16, 17), (37, 64)])
def test_periodic_interp_add_bitwise_vs_reference(nx, ny, nine, jpn):
    """Periodic branch of BMG2_SymStd_interp_add (f90:143-165): the same correction, then the ghost layer of the
    fine vector refreshed (y rows, then x columns)."""
    R, L = ref.lib(), oracle.lib()
    rng = np.random.default_rng(nx * 19 + ny + nine + 100 * abs(jpn))
    II, JJ = nx + 2, ny + 2
    iic, jjc = (II - 2) // 2 + 3, (JJ - 2) // 2 + 3          # IICF, JJCF: what the loops of the kernel span
    nst = 5 if nine else 3
    so = 10.0 ** rng.uniform(-1, 1, size=(nst, JJ, II))
    so[0] = so[1:].sum(axis=0) * 2.5 + 1.0
    ci = rng.uniform(0, 1, size=(8, jjc, iic))
    v = rng.standard_normal((jjc, iic))
    q, r = rng.standard_normal((JJ, II)), rng.standard_normal((JJ, II))
    qa, qb, ra, rb = q.copy(), q.copy(), r.copy(), r.copy()
    assert L.orc2_interp_add(P(qa), P(v), P(ra), P(so), P(ci), iic, jjc, II, JJ, nst, jpn) == 0
    R.BMG2_SymStd_interp_add(P(qb), P(v), P(rb), P(so), P(ci), iic, jjc, II, JJ, nst, jpn)
    assert np.array_equal(qa, qb) and np.array_equal(ra, rb)


@pytest.mark.parametrize("jpn", [1, 2, 3, -1, -2, -3])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(4, 4), (8, 6), (7, 5), (16, 17), (37, 64)])
def test_periodic_galerkin_bitwise_vs_reference(nx, ny, nine, jpn):
    """Periodic branch of BMG2_SymStd_SETUP_ITLI_ex (f90:216-247, 333-364): the same Galerkin products, then the ghost
    layer of all five planes of the coarse operator refreshed."""
    R, L = ref.lib(), oracle.lib()
    rng = np.random.default_rng(nx * 23 + ny + nine + 100 * abs(jpn))
    II, JJ = nx + 2, ny + 2
    iic, jjc = II // 2 + 1 + II % 2, JJ // 2 + 1 + JJ % 2
    nst, ifd = (5, 0) if nine else (3, 1)
    so = 10.0 ** rng.uniform(-1, 1, size=(nst, JJ, II))
    so[0] = so[1:].sum(axis=0) * 2.5 + 1.0
    ci = rng.uniform(0, 1, size=(8, jjc, iic))
    c1 = rng.standard_normal((5, jjc, iic))
    c2 = c1.copy()
    assert L.orc2_setup_itli_ex(P(so), P(c1), P(ci), II, JJ, iic, jjc, ifd, nst, jpn) == 0
    R.BMG2_SymStd_SETUP_ITLI_ex(P(so), P(c2), P(ci), II, JJ, iic, jjc, ifd, nst, jpn)
    assert np.array_equal(c1, c2)


@pytest.mark.parametrize("ibc", [1, 2, 3])
@pytest.mark.parametrize("nine", [False, True])
def test_residual_has_no_periodic_branch(nine, ibc):
    """BMG2_SymStd_residual ignores its boundary code (f90:87-119 is the whole kernel): with a periodic code and
    non-zero ghost values the reference gives what the restatement gives with code 0 -- the periodic cycle relies
    on the ghost refreshes of relaxation / restriction / interpolation alone."""
    R, L = ref.lib(), oracle.lib()
    nx, ny = 9, 6
    rng = np.random.default_rng(5 + nine + 10 * ibc)
    II, JJ = nx + 2, ny + 2
    nst, ifd = (5, 0) if nine else (3, 1)
    so = 10.0 ** rng.uniform(-1, 1, size=(nst, JJ, II))
    q, b = rng.standard_normal((JJ, II)), rng.standard_normal((JJ, II))
    r1, r2 = np.zeros_like(q), np.zeros_like(q)
    L.orc2_residual(0, P(so), P(b), P(q), P(r1), II, JJ, 0, ifd, nst)
    R.BMG2_SymStd_residual(ref.ip(0), P(so), P(b), P(q), P(r2), ref.ip(II), ref.ip(JJ), ref.ip(0), ref.ip(ifd),
                           ref.ip(nst), ref.ip(ibc), ref.ip(0), ref.ip(1), ref.ip(0))
    assert np.array_equal(r1, r2)


@pytest.mark.parametrize("jpn", [1, 2, 3, -1, -2, -3])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(4, 4), (8, 6), (7, 5), (6, 9), (16, 17), (37, 64), (40, 31)])
def test_periodic_interp_setup_bitwise_vs_reference(nx, ny, nine, jpn):
    """Periodic branch of BMG2_SymStd_SETUP_interp_OI (f90:258-618): wrapped fine index sequences, the first and last
    coarse lines from the wrapped fine lines; even and odd extents (the IIC == IICF cases of l.296-317)."""
    R, L = ref.lib(), oracle.lib()
    rng = np.random.default_rng(nx * 29 + ny + nine + 100 * abs(jpn))
    II, JJ = nx + 2, ny + 2
    iic, jjc = II // 2 + 1 + II % 2, JJ // 2 + 1 + JJ % 2
    nst, ifd = (5, 0) if nine else (3, 1)
    so = 10.0 ** rng.uniform(-1, 1, size=(nst, JJ, II))       # periodic operators carry couplings in the ghost layer too
    so[0] = so[1:].sum(axis=0) * 2.5 + 1.0
    soc = np.zeros((5, jjc, iic))
    c1 = rng.standard_normal((8, jjc, iic))
    c2 = c1.copy()
    assert L.orc2_setup_interp_oi(P(so), P(soc), P(c1), II, JJ, iic, jjc, ifd, nst, jpn, 0) == 0
    R.BMG2_SymStd_SETUP_interp_OI(P(so), P(soc), P(c2), II, JJ, iic, jjc, ifd, nst, jpn, 0)
    assert np.array_equal(c1, c2)


@pytest.mark.parametrize("ibc", [1, 2, 3, -1, -2, -3])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(2, 2), (3, 3), (4, 5), (7, 4), (5, 6)])
def test_periodic_coarse_solve_bitwise_vs_reference(nx, ny, nine, ibc):
    """Periodic branches of BMG2_SymStd_SETUP_cg_LU (f90:121-217, 258-331: dense upper triangle with the wrap-around
    couplings, DPOTRF) and BMG2_SymStd_SOLVE_cg (f90:104-166: DPOTRS, mean subtraction for every code but 0, ghost
    layer of the solution); DPOTRF / DPOTRS bound to the restatements in both libraries."""
    R, L = ref.lib(), oracle.lib()
    rng = np.random.default_rng(nx * 7 + ny + nine + 50 * abs(ibc) + (ibc < 0))
    nst = 5 if nine else 3
    II, JJ = nx + 2, ny + 2
    so = 10.0 ** rng.uniform(-1, 0, size=(nst, JJ, II))
    so[0] = 12.0 + rng.uniform(0, 1, size=(JJ, II))             # diagonally dominant whatever wraps around
    n = nx * ny
    a1, a2 = np.zeros((n, n)), np.zeros((n, n))
    rc = L.orc2_setup_cg_lu(P(so), II, JJ, nst, P(a1), n, n, ibc)
    R.BMG2_SymStd_SETUP_cg_LU(P(so), ref.ip(II), ref.ip(JJ), ref.ip(nst), P(a2), ref.ip(n), ref.ip(n), ref.ip(ibc))
    assert rc == 0 and np.array_equal(a1, a2)
    b = rng.standard_normal((JJ, II))
    x1 = rng.standard_normal((JJ, II))
    x2 = x1.copy()
    w1, w2 = np.zeros(n), np.zeros(n)
    assert L.orc2_solve_cg(P(x1), P(b), II, JJ, P(a1), P(w1), n, n, ibc) == 0
    R.BMG2_SymStd_SOLVE_cg(P(x2), P(b), II, JJ, P(a2), P(w2), n, n, ibc)
    assert np.array_equal(x1, x2)


def test_dense_cholesky_restatement_against_scipy():
    """orc_dpotf2_upper / orc_dpotrs_upper (what DPOTRF / DPOTRS are bound to) against scipy's LAPACK"""
    import scipy.linalg as sl
    L = oracle.lib()
    rng = np.random.default_rng(3)
    for n in (1, 2, 5, 12, 40):
        M = rng.standard_normal((n, n))
        A = M @ M.T + n * np.eye(n)
        a = np.array(A.T, order="C")           # column-major upper triangle == row-major of the transpose
        assert L.orc_dpotf2_upper(n, P(a), n) == 0
        U = np.triu(a.T)
        assert np.allclose(U, sl.cholesky(A, lower=False), rtol=1e-12, atol=1e-12)
        b = rng.standard_normal(n)
        x = b.copy()
        L.orc_dpotrs_upper(n, P(a), n, P(x))
        assert np.allclose(x, np.linalg.solve(A, b), rtol=1e-10, atol=1e-12)


@pytest.mark.parametrize("jpn", [1, 2, 3, -1, -2, -3, -4])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(4, 4), (8, 6), (7, 5), (16, 17), (37, 64), (64, 38)])
def test_periodic_line_relaxation_bitwise_vs_reference(nx, ny, nine, jpn):
    """Periodic halves of SETUP_lines_x/_y (f90:76-83) and relax_lines_x/_y (x: f90:170-306, y: f90:178-296): wrap
    couplings added to the end diagonals, Sherman-Morrison correction with a second solve per line, ghost refresh after
    every colour -- and, for lines NOT periodic in their own direction, the other direction's ghosts refreshed only once
    at the end (x: f90:162-167).  Ghosts start inconsistent with the interior, as the reference does not fix them."""
    R, L = ref.lib(), oracle.lib()
    rng = np.random.default_rng(nx * 17 + ny * 3 + nine + 10 * abs(jpn))
    II, JJ = nx + 2, ny + 2
    nst, ifd = (5, 0) if nine else (3, 1)
    so = 10.0 ** rng.uniform(-1, 1, size=(nst, JJ, II))     # periodic operators carry couplings in the ghost layer too
    so[0] = so[1:].sum(axis=0) * 2.5 + 1.0
    b = rng.standard_normal((JJ, II))
    q = rng.standard_normal((JJ, II))
    x1, x2 = np.zeros((2, JJ, II)), np.zeros((2, JJ, II))
    y1, y2 = np.zeros((2, II, JJ)), np.zeros((2, II, JJ))
    assert L.orc2_setup_lines_x(P(so), P(x1), II, JJ, nst, jpn) == 0
    R.BMG2_SymStd_SETUP_lines_x(P(so), P(x2), II, JJ, nst, jpn)
    assert L.orc2_setup_lines_y(P(so), P(y1), II, JJ, nst, jpn) == 0
    R.BMG2_SymStd_SETUP_lines_y(P(so), P(y2), II, JJ, nst, jpn)
    assert np.array_equal(x1[:, 1:-1, 1:-1], x2[:, 1:-1, 1:-1]) and np.array_equal(y1[:, 1:-1, 1:-1], y2[:, 1:-1, 1:-1])
    w1, w2 = np.zeros(4 * max(II, JJ)), np.zeros(4 * max(II, JJ))
    q1, q2 = q.copy(), q.copy()
    for ud in (0, 1, 1, 0):
        assert L.orc2_relax_lines_x(1, P(so), P(b), P(q1), P(x1), P(w1), II, JJ, 1, ifd, nst, 1, ud, jpn) == 0
        R.BMG2_SymStd_relax_lines_x(1, P(so), P(b), P(q2), P(x2), P(w2), II, JJ, 1, ifd, nst, 1, ud, jpn)
        assert np.array_equal(q1, q2), f"x-lines updown={ud}"
        assert L.orc2_relax_lines_y(1, P(so), P(b), P(q1), P(y1), P(w1), II, JJ, 1, ifd, nst, 1, ud, jpn) == 0
        R.BMG2_SymStd_relax_lines_y(1, P(so), P(b), P(q2), P(y2), P(w2), II, JJ, 1, ifd, nst, 1, ud, jpn)
        assert np.array_equal(q1, q2), f"y-lines updown={ud}"
    assert not np.array_equal(q1, q)
