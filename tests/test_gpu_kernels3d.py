"""GPU parity tests (through the C ABI) of the 3D kernels against the CPU oracle.

Bitwise for point relaxation, residual, restriction, interpolate-and-add and the
coarse solve (same expression order, -fmad=false vs -ffp-contract=off); to
rounding for the interpolation setup and the Galerkin product, which oracle and
kernels both write in neighbour-offset form with different summation orders
(|diff| <= 1e-12 x diagonal scale; north_star: coarse operators to 1e-12)."""
import numpy as np
import pytest

import oracle
from oracle import P
from tests.util import assert_bitwise, coarse_n, dev, host, ptr, random_gf3, random_op3

pytestmark = pytest.mark.gpu

SIZES = [(1, 1, 1), (3, 3, 3), (2, 5, 3), (5, 7, 4), (8, 8, 8), (9, 6, 13), (17, 12, 10), (33, 33, 33), (70, 35, 18),
         (129, 20, 9)]


@pytest.fixture(scope="module")
def abi(cuda):
    import cedar_b200.abi as a
    return a


@pytest.mark.parametrize("full", [False, True])
@pytest.mark.parametrize("nx,ny,nz", SIZES)
def test_residual_relax(cuda, abi, nx, ny, nz, full):
    rng = np.random.default_rng(nx * 1000 + ny * 10 + nz + full)
    so = random_op3(nx, ny, nz, full, rng)
    nst, ifd = so.shape[0], (0 if full else 1)
    II, JJ, KK = nx + 2, ny + 2, nz + 2
    q, b = random_gf3(nx, ny, nz, rng), random_gf3(nx, ny, nz, rng)
    L = oracle.lib()
    ref = np.zeros_like(q)
    L.orc3_residual(1, 1, ifd, P(q), P(b), P(so), P(ref), II, JJ, KK, nst)
    d_so, d_q, d_b = dev(cuda, so), dev(cuda, q), dev(cuda, b)
    d_r = cuda.zeros_like(d_q)
    abi.check(abi.lib.bmgx3_residual(1, 1, ifd, ptr(d_q), ptr(d_b), ptr(d_so), ptr(d_r), II, JJ, KK, nst, None))
    assert_bitwise(host(d_r), ref, "residual3")
    sor = np.zeros((2, KK, JJ, II))
    L.orc3_setup_recip(P(so), P(sor), II, JJ, KK, nst, 2)
    d_sor = cuda.zeros(sor.shape, dtype=cuda.float64, device="cuda")
    abi.check(abi.lib.bmgx3_SETUP_recip(ptr(d_so), ptr(d_sor), II, JJ, KK, nst, 2, None))
    assert_bitwise(host(d_sor), sor, "SETUP_recip3")
    d_q2 = cuda.full_like(d_q, 7.0)
    for updown in (0, 1, 0):
        assert L.orc3_relax_gs(1, P(so), P(b), P(q), P(sor), II, JJ, KK, ifd, nst, 2, 1, updown, 0) == 0
        if not full:
            # single-pass out-of-place form first (reads d_q), then the in-place one
            abi.check(abi.lib.bmgx3_relax_GS_oop(1, ptr(d_so), ptr(d_b), ptr(d_q), ptr(d_q2), ptr(d_sor), II, JJ, KK, ifd,
                                                 nst, 2, 1, updown, 0, None))
        abi.check(abi.lib.bmgx3_relax_GS(1, ptr(d_so), ptr(d_b), ptr(d_q), ptr(d_sor), II, JJ, KK, ifd, nst, 2, 1,
                                         updown, 0, None))
        assert_bitwise(host(d_q), q, f"relax_GS3 updown={updown}")
        if not full:
            assert_bitwise(host(d_q2), q, f"relax_GS3_oop updown={updown}")


@pytest.mark.parametrize("full", [False, True])
@pytest.mark.parametrize("nx,ny,nz", [s for s in SIZES if min(s) >= 3])
def test_setup_and_transfer(cuda, abi, nx, ny, nz, full):
    rng = np.random.default_rng(nx * 31 + ny * 7 + nz + full)
    so = random_op3(nx, ny, nz, full, rng)
    nst, ifd = so.shape[0], (0 if full else 1)
    II, JJ, KK = nx + 2, ny + 2, nz + 2
    nxc, nyc, nzc = coarse_n(nx), coarse_n(ny), coarse_n(nz)
    IIC, JJC, KKC = nxc + 2, nyc + 2, nzc + 2
    L = oracle.lib()
    ci = np.zeros((26, KKC, JJC, IIC)); soc = np.zeros((14, KKC, JJC, IIC))
    L.orc3_setup_interp_oi(P(so), P(soc), P(ci), II, JJ, KK, IIC, JJC, KKC, ifd, nst, 0, 0, None)
    L.orc3_setup_itli_ex(P(so), P(soc), P(ci), II, JJ, KK, IIC, JJC, KKC, ifd, nst, 0)
    d_so = dev(cuda, so)
    d_ci = cuda.zeros(ci.shape, dtype=cuda.float64, device="cuda")
    d_soc = cuda.zeros(soc.shape, dtype=cuda.float64, device="cuda")
    abi.check(abi.lib.bmgx3_SETUP_interp_OI(ptr(d_so), ptr(d_soc), ptr(d_ci), II, JJ, KK, IIC, JJC, KKC, ifd, nst, 0, 0,
                                            None, None))
    g_ci = host(d_ci)
    assert np.max(np.abs(g_ci - ci)) <= 1e-12, np.max(np.abs(g_ci - ci))
    # Galerkin on the oracle's weights so that only the product is compared
    d_ci.copy_(dev(cuda, ci))
    f = abi.lib.bmgx3_SETUP_ITLI27_ex if full else abi.lib.bmgx3_SETUP_ITLI07_ex
    abi.check(f(ptr(d_so), ptr(d_soc), ptr(d_ci), II, JJ, KK, IIC, JJC, KKC, 0, None))
    g_soc = host(d_soc)
    scale = np.max(np.abs(soc[0]))
    assert np.max(np.abs(g_soc - soc)) <= 1e-12 * scale, (np.max(np.abs(g_soc - soc)), scale)
    # invariant: the coarse operator is symmetric by construction and P^T A P of an
    # operator with zero row sums has (nearly) zero row sums in the interior -- covered
    # by the oracle comparison; restriction / interpolation are bitwise
    q, b = random_gf3(nx, ny, nz, rng), random_gf3(nx, ny, nz, rng)
    v = random_gf3(nxc, nyc, nzc, rng)
    rc = np.zeros_like(v)
    L.orc3_restrict(P(q), P(rc), P(ci), II, JJ, KK, IIC, JJC, KKC, 0)
    d_q, d_b, d_v = dev(cuda, q), dev(cuda, b), dev(cuda, v)
    d_rc = cuda.zeros_like(d_v)
    abi.check(abi.lib.bmgx3_restrict(ptr(d_q), ptr(d_rc), ptr(d_ci), II, JJ, KK, IIC, JJC, KKC, 0, None))
    assert_bitwise(host(d_rc), rc, "restrict3")
    L.orc3_interp_add(P(q), P(v), P(so), P(b), P(ci), IIC, JJC, KKC, II, JJ, KK, nst, 0)
    abi.check(abi.lib.bmgx3_interp_add(ptr(d_q), ptr(d_v), ptr(d_so), ptr(d_b), ptr(d_ci), IIC, JJC, KKC, II, JJ, KK,
                                       nst, 0, None))
    assert_bitwise(host(d_b), b, "interp_add3 res")
    assert_bitwise(host(d_q), q, "interp_add3 q")


@pytest.mark.parametrize("full", [False, True])
@pytest.mark.parametrize("nx,ny,nz", [(2, 2, 2), (3, 3, 3), (4, 3, 5), (6, 6, 4)])
def test_coarse_solve(cuda, abi, nx, ny, nz, full):
    rng = np.random.default_rng(11 + nx + full)
    so = random_op3(nx, ny, nz, full, rng)
    nst = so.shape[0]
    II, JJ, KK = nx + 2, ny + 2, nz + 2
    n1, n2 = nx * (ny + 1) + 2, nx * ny * nz
    L = oracle.lib()
    abd = np.zeros((n2, n1))
    assert L.orc3_setup_cg_lu(P(so), II, JJ, KK, nst, P(abd), n1, n2, 0) == 0
    b = random_gf3(nx, ny, nz, rng)
    x = np.zeros_like(b); bbd = np.zeros(n2)
    L.orc3_solve_cg(P(x), P(b), II, JJ, KK, P(abd), P(bbd), n1, n2, 0)
    d_so, d_b = dev(cuda, so), dev(cuda, b)
    d_abd = cuda.full(abd.shape, 3.0, dtype=cuda.float64, device="cuda")
    d_bbd = cuda.zeros(n2, dtype=cuda.float64, device="cuda")
    d_x = cuda.zeros_like(d_b)
    abi.check(abi.lib.bmgx3_SETUP_cg_LU(ptr(d_so), II, JJ, KK, nst, ptr(d_abd), n1, n2, 0, None))
    assert_bitwise(host(d_abd), abd, "SETUP_cg_LU3")
    abi.check(abi.lib.bmgx3_SOLVE_cg(ptr(d_x), ptr(d_b), II, JJ, KK, ptr(d_abd), ptr(d_bbd), n1, n2, 0, None))
    assert_bitwise(host(d_x), x, "SOLVE_cg3")


@pytest.mark.parametrize("full", [False, True])
@pytest.mark.parametrize("nx,ny,nz", [(1, 1, 1), (3, 4, 5), (17, 9, 12), (130, 33, 7)])
def test_matvec3(cuda, abi, nx, ny, nz, full):
    rng = np.random.default_rng(nx * 100 + ny * 10 + nz + full)
    so = random_op3(nx, ny, nz, full, rng)
    nst, ifd = so.shape[0], (0 if full else 1)
    II, JJ, KK = nx + 2, ny + 2, nz + 2
    q = random_gf3(nx, ny, nz, rng)
    ref = np.zeros_like(q)
    oracle.lib().orc3_matvec(1, 1, ifd, P(so), P(ref), P(q), II, JJ, KK, nst)
    d_so, d_q = dev(cuda, so), dev(cuda, q)
    d_y = cuda.zeros_like(d_q)
    abi.check(abi.lib.bmgx3_matvec(1, ptr(d_so), ptr(d_y), ptr(d_q), II, JJ, KK, 1, ifd, nst, None))
    assert_bitwise(host(d_y), ref, "matvec3")


def test_relax27_residual_at_513_cubed(cuda, abi):
    """BASELINE configs[3] at its full size: one DOWN and one UP 8-colour sweep and the residual of a 27-point operator
    on 513^3 against the CPU oracle (all host cores), bitwise.  The operator is a random 2D field per stencil entry
    times a random factor per plane (15 GB of independent random numbers would only slow the test down)."""
    n = 513
    II = JJ = KK = n + 2
    rng = np.random.default_rng(513)
    a2 = rng.uniform(0.5, 1.5, size=(14, 1, JJ, II))
    ck = rng.uniform(0.5, 1.5, size=(14, KK, 1, 1))
    so = np.empty((14, KK, JJ, II))
    np.multiply(a2, ck, out=so)
    so[0] *= 30.0                                            # a dominant diagonal keeps the sweep's values O(1)
    q, b = random_gf3(n, n, n, rng), random_gf3(n, n, n, rng)
    L = oracle.lib()
    sor = np.zeros((2, KK, JJ, II))
    L.orc3_setup_recip(P(so), P(sor), II, JJ, KK, 14, 2)
    d_so, d_q, d_b, d_sor = dev(cuda, so), dev(cuda, q), dev(cuda, b), dev(cuda, sor)
    for updown in (0, 1):
        assert L.orc3_relax_gs(1, P(so), P(b), P(q), P(sor), II, JJ, KK, 0, 14, 2, 1, updown, 0) == 0
        abi.check(abi.lib.bmgx3_relax_GS(1, ptr(d_so), ptr(d_b), ptr(d_q), ptr(d_sor), II, JJ, KK, 0, 14, 2, 1, updown, 0, None))
        assert_bitwise(host(d_q), q, f"27-point sweep at 513^3, updown={updown}")
    ref = np.zeros_like(q)
    L.orc3_residual(1, 1, 0, P(q), P(b), P(so), P(ref), II, JJ, KK, 14)
    d_r = cuda.zeros_like(d_q)
    abi.check(abi.lib.bmgx3_residual(1, 1, 0, ptr(d_q), ptr(d_b), ptr(d_so), ptr(d_r), II, JJ, KK, 14, None))
    assert_bitwise(host(d_r), ref, "27-point residual at 513^3")
