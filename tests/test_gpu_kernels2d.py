"""GPU parity tests (through the C ABI) of the 2D kernels against the CPU oracle.

Bitwise for every stencil kernel (same expression order, -fmad=false vs
-ffp-contract=off); 1e-12 relative for the line solves (warp-scan evaluation
order differs from DPTTS2's sequential loop)."""
import ctypes as C

import numpy as np
import pytest

import oracle
from oracle import P
from tests.util import (assert_bitwise, coarse_n, dev, host, ptr, random_gf, random_op2)

pytestmark = pytest.mark.gpu

SIZES = [(1, 1), (2, 3), (3, 3), (7, 5), (16, 16), (31, 31), (37, 64), (58, 58), (58, 59), (130, 97), (400, 257), (1030, 301), (245, 30), (246, 33), (247, 18)]


@pytest.fixture(scope="module")
def abi(cuda):
    import cedar_b200.abi as a
    return a


def ifd_of(nine):
    return 0 if nine else 1


@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", SIZES)
def test_residual(cuda, abi, nx, ny, nine):
    rng = np.random.default_rng(nx * 1000 + ny + nine)
    so = random_op2(nx, ny, nine, rng)
    q = random_gf(nx, ny, rng); b = random_gf(nx, ny, rng)
    ref = np.zeros_like(q)
    oracle.lib().orc2_residual(0, P(so), P(b), P(q), P(ref), nx + 2, ny + 2, 0, ifd_of(nine), so.shape[0])
    d_so, d_q, d_b = dev(cuda, so), dev(cuda, q), dev(cuda, b)
    d_r = cuda.zeros_like(d_q)
    abi.check(abi.lib.bmgx2_residual(0, ptr(d_so), ptr(d_b), ptr(d_q), ptr(d_r), nx + 2, ny + 2, 0, ifd_of(nine),
                                     so.shape[0], 0, 0, 0, 0, None))
    assert_bitwise(host(d_r), ref, "residual")


@pytest.mark.parametrize("updown", [0, 1])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", SIZES)
def test_relax_gs(cuda, abi, nx, ny, nine, updown):
    rng = np.random.default_rng(nx * 77 + ny + 5 * nine + updown)
    so = random_op2(nx, ny, nine, rng)
    q = random_gf(nx, ny, rng); b = random_gf(nx, ny, rng)
    sor = np.zeros((2, ny + 2, nx + 2))
    L = oracle.lib()
    L.orc2_setup_recip(P(so), P(sor), nx + 2, ny + 2, so.shape[0], 2)
    d_so, d_q, d_b = dev(cuda, so), dev(cuda, q), dev(cuda, b)
    d_sor = cuda.zeros(sor.shape, dtype=cuda.float64, device="cuda")
    abi.check(abi.lib.bmgx2_SETUP_recip(ptr(d_so), ptr(d_sor), nx + 2, ny + 2, so.shape[0], 2, None))
    assert_bitwise(host(d_sor), sor, "SETUP_recip")
    d_q2 = cuda.full_like(d_q, 7.0)
    for sweep in range(3):
        assert L.orc2_relax_gs(1, P(so), P(b), P(q), P(sor), nx + 2, ny + 2, 1, ifd_of(nine), so.shape[0], 2, 1, updown, 0) == 0
        # out-of-place form first (reads d_q), then the in-place one
        abi.check(abi.lib.bmgx2_relax_GS_oop(1, ptr(d_so), ptr(d_b), ptr(d_q), ptr(d_q2), ptr(d_sor), nx + 2, ny + 2,
                                             1, ifd_of(nine), so.shape[0], 2, 1, updown, 0, None))
        abi.check(abi.lib.bmgx2_relax_GS(1, ptr(d_so), ptr(d_b), ptr(d_q), ptr(d_sor), nx + 2, ny + 2, 1, ifd_of(nine),
                                         so.shape[0], 2, 1, updown, 0, None))
        assert_bitwise(host(d_q), q, f"relax_GS sweep {sweep}")
        assert_bitwise(host(d_q2), q, f"relax_GS_oop sweep {sweep}")


def _setup_pair(nx, ny, nine, rng):
    so = random_op2(nx, ny, nine, rng)
    nxc, nyc = coarse_n(nx), coarse_n(ny)
    ci = np.zeros((8, nyc + 2, nxc + 2))
    soc = np.zeros((5, nyc + 2, nxc + 2))
    return so, nxc, nyc, ci, soc


@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [s for s in SIZES if min(s) >= 3])
def test_setup_interp_and_galerkin(cuda, abi, nx, ny, nine):
    rng = np.random.default_rng(nx * 13 + ny + nine)
    so, nxc, nyc, ci, soc = _setup_pair(nx, ny, nine, rng)
    L = oracle.lib()
    nst = so.shape[0]
    assert L.orc2_setup_interp_oi(P(so), P(soc), P(ci), nx + 2, ny + 2, nxc + 2, nyc + 2, ifd_of(nine), nst, 0, 0) == 0
    assert L.orc2_setup_itli_ex(P(so), P(soc), P(ci), nx + 2, ny + 2, nxc + 2, nyc + 2, ifd_of(nine), nst, 0) == 0
    d_so = dev(cuda, so)
    d_ci = cuda.zeros(ci.shape, dtype=cuda.float64, device="cuda")
    d_soc = cuda.zeros(soc.shape, dtype=cuda.float64, device="cuda")
    abi.check(abi.lib.bmgx2_SETUP_interp_OI(ptr(d_so), ptr(d_soc), ptr(d_ci), nx + 2, ny + 2, nxc + 2, nyc + 2,
                                            ifd_of(nine), nst, 0, 0, None))
    assert_bitwise(host(d_ci), ci, "SETUP_interp_OI")
    abi.check(abi.lib.bmgx2_SETUP_ITLI_ex(ptr(d_so), ptr(d_soc), ptr(d_ci), nx + 2, ny + 2, nxc + 2, nyc + 2,
                                          ifd_of(nine), nst, 0, None))
    assert_bitwise(host(d_soc), soc, "SETUP_ITLI_ex")
    # structural invariants: R = P^T  <Rr, v> == <r, Pv>
    r = random_gf(nx, ny, rng); v = random_gf(nxc, nyc, rng)
    rc = np.zeros_like(v)
    assert L.orc2_restrict(P(r), P(rc), P(ci), nx + 2, ny + 2, nxc + 2, nyc + 2, 0) == 0
    d_r, d_rc = dev(cuda, r), cuda.zeros(v.shape, dtype=cuda.float64, device="cuda")
    abi.check(abi.lib.bmgx2_restrict(ptr(d_r), ptr(d_rc), ptr(d_ci), nx + 2, ny + 2, nxc + 2, nyc + 2, 0, None))
    assert_bitwise(host(d_rc), rc, "restrict")
    # interp_add: q += P v + res/diag at F points
    q = random_gf(nx, ny, rng); res = random_gf(nx, ny, rng)
    d_q, d_res, d_v = dev(cuda, q), dev(cuda, res), dev(cuda, v)
    assert L.orc2_interp_add(P(q), P(v), P(res), P(so), P(ci), nxc + 2, nyc + 2, nx + 2, ny + 2, nst, 0) == 0
    abi.check(abi.lib.bmgx2_interp_add(ptr(d_q), ptr(d_v), ptr(d_res), ptr(d_so), ptr(d_ci), nxc + 2, nyc + 2,
                                       nx + 2, ny + 2, nst, 0, None))
    assert_bitwise(host(d_q)[1:-1, 1:-1], q[1:-1, 1:-1], "interp_add q")
    assert_bitwise(host(d_res), res, "interp_add res (in-place scaling)")
    # adjointness with res = 0: <R r, v> == <r, P v>
    q0 = np.zeros((ny + 2, nx + 2)); z = np.zeros_like(q0)
    L.orc2_interp_add(P(q0), P(v), P(z), P(so), P(ci), nxc + 2, nyc + 2, nx + 2, ny + 2, nst, 0)
    lhs = float((rc[1:-1, 1:-1] * v[1:-1, 1:-1]).sum()); rhs = float((r[1:-1, 1:-1] * q0[1:-1, 1:-1]).sum())
    assert abs(lhs - rhs) <= 1e-10 * max(abs(lhs), abs(rhs), 1.0)


@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(3, 3), (4, 4), (3, 5), (7, 4), (12, 9)])
def test_coarse_solve(cuda, abi, nx, ny, nine):
    rng = np.random.default_rng(nx * 3 + ny + nine)
    so = random_op2(nx, ny, nine, rng)
    nst = so.shape[0]
    nabd1, nabd2 = nx + 2, nx * ny
    abd = np.zeros((nabd2, nabd1)); bbd = np.zeros(nabd2)
    L = oracle.lib()
    assert L.orc2_setup_cg_lu(P(so), nx + 2, ny + 2, nst, P(abd), nabd1, nabd2, 0) == 0
    b = random_gf(nx, ny, rng); x = np.zeros_like(b)
    assert L.orc2_solve_cg(P(x), P(b), nx + 2, ny + 2, P(abd), P(bbd), nabd1, nabd2, 0) == 0
    d_so, d_b = dev(cuda, so), dev(cuda, b)
    d_abd = cuda.full(abd.shape, 3.0, dtype=cuda.float64, device="cuda")
    d_bbd = cuda.zeros(nabd2, dtype=cuda.float64, device="cuda")
    d_x = cuda.zeros_like(d_b)
    abi.check(abi.lib.bmgx2_SETUP_cg_LU(ptr(d_so), nx + 2, ny + 2, nst, ptr(d_abd), nabd1, nabd2, 0, None))
    assert_bitwise(host(d_abd), abd, "SETUP_cg_LU")
    abi.check(abi.lib.bmgx2_SOLVE_cg(ptr(d_x), ptr(d_b), nx + 2, ny + 2, ptr(d_abd), ptr(d_bbd), nabd1, nabd2, 0, None))
    assert_bitwise(host(d_x), x, "SOLVE_cg")
    # and it really solves A x = b
    r = np.zeros_like(b)
    L.orc2_residual(0, P(so), P(b), P(x), P(r), nx + 2, ny + 2, 0, ifd_of(nine), nst)
    assert np.abs(r).max() <= 1e-9 * np.abs(b).max()


@pytest.mark.parametrize("updown", [0, 1])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(3, 3), (7, 5), (31, 33), (64, 40), (130, 97), (300, 1000)])
def test_lines(cuda, abi, nx, ny, nine, updown):
    rng = np.random.default_rng(nx * 5 + ny + nine + 2 * updown)
    so = random_op2(nx, ny, nine, rng)
    nst = so.shape[0]
    L = oracle.lib()
    II, JJ = nx + 2, ny + 2
    sorx = np.zeros((2, JJ, II)); sory = np.zeros((2, II, JJ))
    assert L.orc2_setup_lines_x(P(so), P(sorx), II, JJ, nst, 0) == 0
    assert L.orc2_setup_lines_y(P(so), P(sory), II, JJ, nst, 0) == 0
    d_so = dev(cuda, so)
    d_sorx = cuda.zeros(sorx.shape, dtype=cuda.float64, device="cuda")
    d_sory = cuda.zeros(sory.shape, dtype=cuda.float64, device="cuda")
    abi.check(abi.lib.bmgx2_SETUP_lines_x(ptr(d_so), ptr(d_sorx), II, JJ, nst, 0, None))
    abi.check(abi.lib.bmgx2_SETUP_lines_y(ptr(d_so), ptr(d_sory), II, JJ, nst, 0, None))
    assert_bitwise(host(d_sorx)[:, 1:-1, 1:-1], sorx[:, 1:-1, 1:-1], "SETUP_lines_x")
    assert_bitwise(host(d_sory)[:, 1:-1, 1:-1], sory[:, 1:-1, 1:-1], "SETUP_lines_y")
    q = random_gf(nx, ny, rng); b = random_gf(nx, ny, rng)
    d_q, d_b = dev(cuda, q), dev(cuda, b)
    d_w = cuda.zeros_like(d_q)
    work = np.zeros(max(II, JJ) + 2)
    for sweep in range(2):
        assert L.orc2_relax_lines_x(1, P(so), P(b), P(q), P(sorx), P(work), II, JJ, 1, ifd_of(nine), nst, 1, updown, 0) == 0
        abi.check(abi.lib.bmgx2_relax_lines_x(1, ptr(d_so), ptr(d_b), ptr(d_q), ptr(d_sorx), ptr(d_w), II, JJ, 1,
                                              ifd_of(nine), nst, 1, updown, 0, None))
        scale = np.abs(q).max()
        assert np.abs(host(d_q) - q).max() <= 1e-12 * scale, "relax_lines_x"
        assert L.orc2_relax_lines_y(1, P(so), P(b), P(q), P(sory), P(work), II, JJ, 1, ifd_of(nine), nst, 1, updown, 0) == 0
        abi.check(abi.lib.bmgx2_relax_lines_y(1, ptr(d_so), ptr(d_b), ptr(d_q), ptr(d_sory), ptr(d_w), II, JJ, 1,
                                              ifd_of(nine), nst, 1, updown, 0, None))
        scale = np.abs(q).max()
        assert np.abs(host(d_q) - q).max() <= 1e-12 * scale, "relax_lines_y"


@pytest.mark.parametrize("nx,ny", [(1, 1), (5, 9), (300, 1000), (2049, 1025)])
def test_sumsq(cuda, abi, nx, ny):
    rng = np.random.default_rng(nx + ny)
    q = rng.standard_normal((ny + 2, nx + 2))  # ghosts non-zero on purpose: must be ignored
    d_q = dev(cuda, q)
    d_w = cuda.zeros(abi.lib.bmgx_norm_work_len(), dtype=cuda.float64, device="cuda")
    d_o = cuda.zeros(1, dtype=cuda.float64, device="cuda")
    abi.check(abi.lib.bmgx2_sumsq(ptr(d_q), nx + 2, ny + 2, ptr(d_w), ptr(d_o), None))
    got = float(host(d_o)[0])
    ref = oracle.lib().orc2_l2norm(P(q), nx + 2, ny + 2) ** 2
    assert abs(got - ref) <= 1e-13 * ref
    a2 = d_o.clone()
    abi.check(abi.lib.bmgx2_sumsq(ptr(d_q), nx + 2, ny + 2, ptr(d_w), ptr(d_o), None))
    assert float(host(a2)[0]) == float(host(d_o)[0])  # run-to-run deterministic


def test_periodic_is_refused(cuda, abi):
    """entry points whose periodic branch is not built yet refuse the boundary code instead of guessing
    (in-place point relaxation has it: test_relax_gs_periodic)"""
    d = cuda.zeros((5, 6, 6), dtype=cuda.float64, device="cuda")
    st = abi.lib.bmgx2_relax_GS_oop(1, ptr(d), ptr(d[0]), ptr(d[1]), ptr(d[2]), ptr(d), 6, 6, 1, 0, 5, 2, 1, 0, 2, None)
    assert st == 2 and b"periodic" in abi.lib.bmgx_last_error()


@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(7, 5), (37, 64), (130, 97)])
def test_batch_context(cuda, abi, nx, ny, nine):
    """bmgx_set_batch(nb): the batch-aware entry points process nb grid functions stored back
    to back with one shared operator -- bitwise what nb separate calls give."""
    rng = np.random.default_rng(nx * 13 + ny + nine)
    nb = 3
    so = random_op2(nx, ny, nine, rng)
    nst, ifd = so.shape[0], ifd_of(nine)
    II, JJ = nx + 2, ny + 2
    nxc, nyc = coarse_n(nx), coarse_n(ny)
    IIC, JJC = nxc + 2, nyc + 2
    d_so = dev(cuda, so)
    d_sorx = cuda.zeros((2, JJ, II), dtype=cuda.float64, device="cuda")
    d_sory = cuda.zeros((2, II, JJ), dtype=cuda.float64, device="cuda")
    abi.check(abi.lib.bmgx2_SETUP_lines_x(ptr(d_so), ptr(d_sorx), II, JJ, nst, 0, None))
    abi.check(abi.lib.bmgx2_SETUP_lines_y(ptr(d_so), ptr(d_sory), II, JJ, nst, 0, None))
    ci = dev(cuda, rng.uniform(0.1, 0.5, size=(8, JJC, IIC)))
    q = np.stack([random_gf(nx, ny, rng) for _ in range(nb)]); b = np.stack([random_gf(nx, ny, rng) for _ in range(nb)])
    qc = np.stack([random_gf(nxc, nyc, rng) for _ in range(nb)])

    def run(batched):
        d_q, d_b, d_qc = dev(cuda, q), dev(cuda, b), dev(cuda, qc)
        d_r = cuda.zeros_like(d_q); d_w = cuda.zeros_like(d_q); d_rc = cuda.zeros_like(d_qc)
        groups = [(slice(None), nb)] if batched else [(slice(k, k + 1), 1) for k in range(nb)]
        for sl, n in groups:
            abi.check(abi.lib.bmgx_set_batch(n))
            try:
                Q, B, R, W, QC, RC = d_q[sl], d_b[sl], d_r[sl], d_w[sl], d_qc[sl], d_rc[sl]
                abi.check(abi.lib.bmgx2_relax_lines_x(1, ptr(d_so), ptr(B), ptr(Q), ptr(d_sorx), ptr(W), II, JJ, 1, ifd, nst, 1, 0, 0, None))
                abi.check(abi.lib.bmgx2_relax_lines_y(1, ptr(d_so), ptr(B), ptr(Q), ptr(d_sory), ptr(W), II, JJ, 1, ifd, nst, 1, 0, 0, None))
                abi.check(abi.lib.bmgx2_residual(0, ptr(d_so), ptr(B), ptr(Q), ptr(R), II, JJ, 0, ifd, nst, 0, 0, 0, 0, None))
                abi.check(abi.lib.bmgx2_restrict(ptr(R), ptr(RC), ptr(ci), II, JJ, IIC, JJC, 0, None))
                abi.check(abi.lib.bmgx2_interp_add(ptr(Q), ptr(QC), ptr(R), ptr(d_so), ptr(ci), IIC, JJC, II, JJ, nst, 0, None))
            finally:
                abi.check(abi.lib.bmgx_set_batch(1))
        return host(d_q), host(d_r), host(d_rc)

    one = run(False)
    many = run(True)
    for a, c, what in zip(one, many, ("q", "res", "coarse rhs")):
        assert_bitwise(c, a, f"batch context: {what}")
    assert abi.lib.bmgx_get_batch() == 1


@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(1, 1), (3, 3), (7, 5), (37, 64), (130, 97), (1030, 301)])
def test_matvec(cuda, abi, nx, ny, nine):
    rng = np.random.default_rng(nx * 31 + ny + nine)
    so = random_op2(nx, ny, nine, rng)
    q = random_gf(nx, ny, rng)
    ref = np.zeros_like(q)
    oracle.lib().orc2_matvec(0, P(so), P(ref), P(q), nx + 2, ny + 2, 0, ifd_of(nine), so.shape[0])
    d_so, d_q = dev(cuda, so), dev(cuda, q)
    d_y = cuda.zeros_like(d_q)
    abi.check(abi.lib.bmgx2_matvec(0, ptr(d_so), ptr(d_y), ptr(d_q), nx + 2, ny + 2, 0, ifd_of(nine), so.shape[0], None))
    assert_bitwise(host(d_y), ref, "matvec")


@pytest.mark.parametrize("jpn", [1, 2, 3, -3])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(2, 3), (4, 4), (7, 5), (37, 64), (400, 257)])
def test_relax_gs_periodic(cuda, abi, nx, ny, nine, jpn):
    """Periodic branch of point relaxation (BMG2_SymStd_relax_GS.f90:139-224; oracle pinned against the reference's
    own kernel in tests/test_ref_parity_cpu.py): x ghosts per colour pass, y ghost rows once per sweep; bitwise."""
    rng = np.random.default_rng(nx * 131 + ny * 7 + nine + 10 * abs(jpn))
    II, JJ = nx + 2, ny + 2
    nst = 5 if nine else 3
    so = 10.0 ** rng.uniform(-1, 1, size=(nst, JJ, II))     # periodic operators carry couplings in the ghost layer
    so[0] = so[1:].sum(axis=0) * 2.5 + 1.0
    b, q = rng.standard_normal((JJ, II)), rng.standard_normal((JJ, II))
    sor = np.zeros((2, JJ, II))
    L = oracle.lib()
    L.orc2_setup_recip(P(so), P(sor), II, JJ, nst, 2)
    d_so, d_q, d_b, d_sor = dev(cuda, so), dev(cuda, q), dev(cuda, b), dev(cuda, sor)
    for sweep, updown in enumerate((0, 1, 1, 0)):
        assert L.orc2_relax_gs(1, P(so), P(b), P(q), P(sor), II, JJ, 1, ifd_of(nine), nst, 2, 1, updown, jpn) == 0
        abi.check(abi.lib.bmgx2_relax_GS(1, ptr(d_so), ptr(d_b), ptr(d_q), ptr(d_sor), II, JJ, 1, ifd_of(nine), nst, 2,
                                         1, updown, jpn, None))
        assert_bitwise(host(d_q), q, f"periodic relax_GS sweep {sweep}")


@pytest.mark.parametrize("jpn", [1, 2, 3, -3])
@pytest.mark.parametrize("nx,ny", [(4, 4), (8, 6), (7, 5), (64, 38), (400, 258)])
def test_restrict_periodic(cuda, abi, nx, ny, jpn):
    """Periodic branch of restriction (BMG2_SymStd_restrict.f90:96-128): ghost refresh of the fine vector (argument
    mutated) for the N/2+1 coarsening, then the same stencil; coarse result and fine vector bitwise."""
    rng = np.random.default_rng(nx * 17 + ny + 100 * abs(jpn))
    II, JJ = nx + 2, ny + 2
    iic, jjc = II // 2 + 1, JJ // 2 + 1
    ci = rng.uniform(0, 1, size=(8, jjc, iic))
    q = rng.standard_normal((JJ, II))
    ref = np.zeros((jjc, iic))
    d_q, d_ci = dev(cuda, q), dev(cuda, ci)
    d_c = cuda.zeros((jjc, iic), dtype=cuda.float64, device="cuda")
    assert oracle.lib().orc2_restrict(P(q), P(ref), P(ci), II, JJ, iic, jjc, jpn) == 0
    abi.check(abi.lib.bmgx2_restrict(ptr(d_q), ptr(d_c), ptr(d_ci), II, JJ, iic, jjc, jpn, None))
    assert_bitwise(host(d_c), ref, "periodic restrict")
    assert_bitwise(host(d_q), q, "fine vector after the ghost refresh")


@pytest.mark.parametrize("jpn", [1, 2, 3, -3])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(4, 4), (8, 6), (7, 5), (37, 64), (400, 258)])
def test_interp_add_periodic(cuda, abi, nx, ny, nine, jpn):
    """Periodic branch of interpolate-and-add (BMG2_SymStd_interp_add.f90:143-165): same correction, then the ghost
    layer of the fine vector refreshed; fine vector and scaled residual bitwise."""
    rng = np.random.default_rng(nx * 19 + ny + nine + 100 * abs(jpn))
    II, JJ = nx + 2, ny + 2
    iic, jjc = II // 2 + 1 + II % 2, JJ // 2 + 1 + JJ % 2        # N/2+1 coarsening (even N), (N+1)/2+1 (odd N)
    nst = 5 if nine else 3
    so = 10.0 ** rng.uniform(-1, 1, size=(nst, JJ, II))
    so[0] = so[1:].sum(axis=0) * 2.5 + 1.0
    ci = rng.uniform(0, 1, size=(8, jjc, iic))
    v = rng.standard_normal((jjc, iic))
    q, r = rng.standard_normal((JJ, II)), rng.standard_normal((JJ, II))
    # a direction that is not periodic keeps the zero ghost layer every Cedar array has (include/cedar/array.h:65);
    # the reference's loops run over those ghost cells too on even grids, adding zeros
    if abs(jpn) == 1:
        for a in (q, r, v):
            a[:, 0] = a[:, -1] = 0.0
    if abs(jpn) == 2:
        for a in (q, r, v):
            a[0, :] = a[-1, :] = 0.0
    d_q, d_r, d_so, d_ci, d_v = dev(cuda, q), dev(cuda, r), dev(cuda, so), dev(cuda, ci), dev(cuda, v)
    assert oracle.lib().orc2_interp_add(P(q), P(v), P(r), P(so), P(ci), iic, jjc, II, JJ, nst, jpn) == 0
    abi.check(abi.lib.bmgx2_interp_add(ptr(d_q), ptr(d_v), ptr(d_r), ptr(d_so), ptr(d_ci), iic, jjc, II, JJ, nst, jpn, None))
    assert_bitwise(host(d_q), q, "periodic interp_add")
    assert_bitwise(host(d_r), r, "scaled residual")


@pytest.mark.parametrize("jpn", [1, 2, 3, -3])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(4, 4), (8, 6), (7, 5), (37, 64), (400, 258)])
def test_galerkin_periodic(cuda, abi, nx, ny, nine, jpn):
    """Periodic branch of the Galerkin coarse operator (BMG2_SymStd_SETUP_ITLI_ex.f90:216-247, 333-364): same
    products, then the ghost layer of the five coarse planes refreshed; bitwise."""
    rng = np.random.default_rng(nx * 23 + ny + nine + 100 * abs(jpn))
    II, JJ = nx + 2, ny + 2
    iic, jjc = II // 2 + 1 + II % 2, JJ // 2 + 1 + JJ % 2
    nst = 5 if nine else 3
    so = 10.0 ** rng.uniform(-1, 1, size=(nst, JJ, II))
    so[0] = so[1:].sum(axis=0) * 2.5 + 1.0
    ci = rng.uniform(0, 1, size=(8, jjc, iic))
    soc = np.zeros((5, jjc, iic))
    d_so, d_ci, d_soc = dev(cuda, so), dev(cuda, ci), dev(cuda, soc)
    assert oracle.lib().orc2_setup_itli_ex(P(so), P(soc), P(ci), II, JJ, iic, jjc, ifd_of(nine), nst, jpn) == 0
    abi.check(abi.lib.bmgx2_SETUP_ITLI_ex(ptr(d_so), ptr(d_soc), ptr(d_ci), II, JJ, iic, jjc, ifd_of(nine), nst, jpn, None))
    assert_bitwise(host(d_soc), soc, "periodic SETUP_ITLI_ex")


# ---- periodic interpolation setup and coarsest-grid solve ---------------------------------------------------------
@pytest.mark.parametrize("jpn", [1, 2, 3, -1, -2, -3])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(4, 4), (8, 6), (7, 5), (16, 17), (37, 64), (300, 301), (129, 256)])
def test_periodic_interp_setup_bitwise(cuda, abi, nx, ny, nine, jpn):
    """bmgx2_SETUP_interp_OI, periodic codes (BMG2_SymStd_SETUP_interp_OI.f90:258-618), against the oracle restatement
    (pinned bitwise to the reference's kernel, tests/test_ref_parity_cpu.py)"""
    L = oracle.lib()
    rng = np.random.default_rng(nx * 29 + ny + nine + 100 * abs(jpn))
    II, JJ = nx + 2, ny + 2
    iic, jjc = II // 2 + 1 + II % 2, JJ // 2 + 1 + JJ % 2
    nst, ifd = (5, 0) if nine else (3, 1)
    so = 10.0 ** rng.uniform(-1, 1, size=(nst, JJ, II))
    so[0] = so[1:].sum(axis=0) * 2.5 + 1.0
    c0 = rng.standard_normal((8, jjc, iic))
    want = c0.copy()
    assert L.orc2_setup_interp_oi(P(so), None, P(want), II, JJ, iic, jjc, ifd, nst, jpn, 0) == 0
    d_ci = dev(cuda, c0)
    abi.check(abi.lib.bmgx2_SETUP_interp_OI(ptr(dev(cuda, so)), None, ptr(d_ci), II, JJ, iic, jjc, ifd, nst, jpn, 0, None))
    assert_bitwise(host(d_ci), want, "periodic SETUP_interp_OI")


@pytest.mark.parametrize("ibc", [1, 2, 3, -1, -2, -3])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(2, 2), (3, 3), (4, 5), (7, 4), (5, 6)])
def test_periodic_coarse_solve_bitwise(cuda, abi, nx, ny, nine, ibc):
    """bmgx2_SETUP_cg_LU / bmgx2_SOLVE_cg with periodic codes: dense upper triangle + DPOTF2 order, DPOTRS, mean
    subtraction, ghost layer of the solution (SETUP_cg_LU.f90:121-217, 258-331; SOLVE_cg.f90:104-166)"""
    import ctypes as C
    L = oracle.lib()
    rng = np.random.default_rng(nx * 7 + ny + nine + 50 * abs(ibc) + (ibc < 0))
    nst = 5 if nine else 3
    II, JJ = nx + 2, ny + 2
    so = 10.0 ** rng.uniform(-1, 0, size=(nst, JJ, II))
    so[0] = 12.0 + rng.uniform(0, 1, size=(JJ, II))
    n = nx * ny
    a = np.zeros((n, n))
    assert L.orc2_setup_cg_lu(P(so), II, JJ, nst, P(a), n, n, ibc) == 0
    d_abd = dev(cuda, rng.standard_normal((n, n)))                    # the entry point clears it
    abi.check(abi.lib.bmgx2_SETUP_cg_LU(ptr(dev(cuda, so)), II, JJ, nst, ptr(d_abd), n, n, ibc, None))
    info = C.c_int(-1)
    abi.check(abi.lib.bmgx_cg_LU_info(C.byref(info), None))
    assert info.value == 0
    assert_bitwise(host(d_abd), a, "periodic ABD")
    b = rng.standard_normal((JJ, II))
    x0 = rng.standard_normal((JJ, II))
    want = x0.copy()
    w = np.zeros(n)
    assert L.orc2_solve_cg(P(want), P(b), II, JJ, P(a), P(w), n, n, ibc) == 0
    d_x, d_w = dev(cuda, x0), dev(cuda, np.zeros(n))
    abi.check(abi.lib.bmgx2_SOLVE_cg(ptr(d_x), ptr(dev(cuda, b)), II, JJ, ptr(d_abd), ptr(d_w), n, n, ibc, None))
    assert_bitwise(host(d_x), want, "periodic SOLVE_cg")


@pytest.mark.parametrize("jpn", [1, 2, 3, -3, -4])
@pytest.mark.parametrize("nine", [False, True])
@pytest.mark.parametrize("nx,ny", [(2, 3), (4, 4), (7, 5), (9, 5), (17, 9), (37, 64), (257, 130), (400, 257), (1030, 70)])
def test_lines_periodic(cuda, abi, nx, ny, nine, jpn):
    """Periodic halves of the line setup and the line sweeps (SETUP_lines_x.f90:76-83, relax_lines_x.f90:170-306,
    relax_lines_y.f90:178-296; oracle pinned bitwise against the reference's own Fortran in
    tests/test_ref_parity_cpu.py): the factor bitwise, the sweeps -- Sherman-Morrison second solve per line, ghost
    refresh per colour -- to the tolerance of the non-periodic line sweeps (the block-scan solve rounds the chunk-entry
    values differently from DPTTS2), ghost layer included."""
    rng = np.random.default_rng(nx * 131 + ny * 7 + nine + 10 * abs(jpn))
    II, JJ = nx + 2, ny + 2
    nst = 5 if nine else 3
    so = 10.0 ** rng.uniform(-1, 1, size=(nst, JJ, II))     # periodic operators carry couplings in the ghost layer
    so[0] = so[1:].sum(axis=0) * 2.5 + 25.0                 # every line's tridiagonal matrix positive definite
    if jpn == -4:                                           # not periodic: no couplings to or inside the ghost layer
        so[:, 0, :] = 0; so[:, -1, :] = 0; so[:, :, 0] = 0; so[:, :, -1] = 0
        so[1, :, 1] = 0; so[2, 1, :] = 0
        if nine:
            so[3, :, 1] = 0; so[3, 1, :] = 0; so[4, :, 1] = 0; so[4, -2, :] = 0
    b, q = rng.standard_normal((JJ, II)), rng.standard_normal((JJ, II))
    L = oracle.lib()
    sorx, sory = np.zeros((2, JJ, II)), np.zeros((2, II, JJ))
    assert L.orc2_setup_lines_x(P(so), P(sorx), II, JJ, nst, jpn) == 0
    assert L.orc2_setup_lines_y(P(so), P(sory), II, JJ, nst, jpn) == 0
    d_so, d_q, d_b = dev(cuda, so), dev(cuda, q), dev(cuda, b)
    d_sorx, d_sory = dev(cuda, np.zeros_like(sorx)), dev(cuda, np.zeros_like(sory))
    abi.check(abi.lib.bmgx2_SETUP_lines_x(ptr(d_so), ptr(d_sorx), II, JJ, nst, jpn, None))
    abi.check(abi.lib.bmgx2_SETUP_lines_y(ptr(d_so), ptr(d_sory), II, JJ, nst, jpn, None))
    assert_bitwise(host(d_sorx)[:, 1:-1, 1:-1], sorx[:, 1:-1, 1:-1], "periodic SETUP_lines_x")
    assert_bitwise(host(d_sory)[:, 1:-1, 1:-1], sory[:, 1:-1, 1:-1], "periodic SETUP_lines_y")
    work = np.zeros(4 * max(II, JJ))
    d_w = cuda.zeros((JJ, II), dtype=cuda.float64, device="cuda")
    for sweep, updown in enumerate((0, 1, 1, 0)):
        assert L.orc2_relax_lines_x(1, P(so), P(b), P(q), P(sorx), P(work), II, JJ, 1, ifd_of(nine), nst, 1, updown, jpn) == 0
        abi.check(abi.lib.bmgx2_relax_lines_x(1, ptr(d_so), ptr(d_b), ptr(d_q), ptr(d_sorx), ptr(d_w), II, JJ, 1,
                                              ifd_of(nine), nst, 1, updown, jpn, None))
        scale = max(1.0, np.abs(q).max())
        assert np.abs(host(d_q) - q).max() <= 1e-12 * scale, f"periodic relax_lines_x sweep {sweep}"
        assert L.orc2_relax_lines_y(1, P(so), P(b), P(q), P(sory), P(work), II, JJ, 1, ifd_of(nine), nst, 1, updown, jpn) == 0
        abi.check(abi.lib.bmgx2_relax_lines_y(1, ptr(d_so), ptr(d_b), ptr(d_q), ptr(d_sory), ptr(d_w), II, JJ, 1,
                                              ifd_of(nine), nst, 1, updown, jpn, None))
        scale = max(1.0, np.abs(q).max())
        assert np.abs(host(d_q) - q).max() <= 1e-12 * scale, f"periodic relax_lines_y sweep {sweep}"
