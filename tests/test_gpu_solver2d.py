"""GPU parity of the full 2D BoxMG setup + V-cycle + solve (C++ host layer over
the C ABI) against the CPU oracle and the reference's golden residual history."""
import numpy as np
import pytest

import oracle
from tests.util import assert_bitwise, random_op2, random_gf

pytestmark = pytest.mark.gpu

# README.md:52-63 of the reference (7-level hierarchy, see tests/test_oracle_golden.py)
README_HISTORY = [0.388629, 0.0443548, 0.00494131, 0.000513399, 5.44908e-05, 5.60612e-06,
                  5.86933e-07, 6.04942e-08, 6.30975e-09, 6.52713e-10]


@pytest.fixture(scope="module")
def cs(cuda):
    import cedar_b200.solver as s
    return s


def test_readme_history(cs):
    nx = ny = 400
    op = cs.Operator2(gallery="poisson", nx=nx, ny=ny)
    assert_bitwise(op.array(), oracle.gallery2("poisson", nx, ny), "gallery::poisson")
    b = oracle.rhs_sin2(nx, ny)
    sol = cs.Solver2(op, relaxation="point", nrelax_pre=1, nrelax_post=1, maxiter=10, tol=1e-10, num_levels=7)
    assert sol.nlevels == 7
    x, hist, _ = sol.solve(b)
    assert len(hist) == 10
    for got, want in zip(hist, README_HISTORY):
        # the README prints 6 significant digits: agree to one unit of the last one
        assert abs(got - want) <= 2e-6 * want, (got, want)
    err = np.abs((oracle.exact_sin2(nx, ny) - x)[1:-1, 1:-1]).max()
    assert "%.6g" % err == "2.04592e-05"


@pytest.mark.parametrize("graph", [False, True])
@pytest.mark.parametrize("case", [
    ("poisson", 400, 400, "point", 1, 1), ("fe", 257, 129, "point", 2, 1), ("fe", 100, 37, "point", 1, 1),
    ("poisson", 63, 200, "point", 2, 2)])
def test_point_solver_matches_oracle(cs, case, graph):
    name, nx, ny, relax, pre, post = case
    so = oracle.gallery2(name, nx, ny)
    b = oracle.rhs_sin2(nx, ny)
    ref = oracle.Solver2(so, relax, pre, post)
    xr, hr, r0r = ref.solve(b, maxiter=8, tol=1e-10)
    op = cs.Operator2(so)
    sol = cs.Solver2(op, relaxation=relax, nrelax_pre=pre, nrelax_post=post, maxiter=8, tol=1e-10, cuda_graph=graph)
    assert sol.nlevels == ref.nlevels
    for lvl in range(1, sol.nlevels):
        assert sol.level_shape(lvl) == ref.level_shape(lvl)
        assert_bitwise(sol.level_array(lvl, "P"), ref.level_array(lvl, "P"), f"P level {lvl}")
        assert_bitwise(sol.level_array(lvl, "A"), ref.level_array(lvl, "A"), f"A level {lvl}")
    x, h, r0 = sol.solve(b)
    assert len(h) == len(hr)                      # identical iteration count
    assert abs(r0 - r0r) <= 1e-13 * r0r
    assert np.allclose(h, hr, rtol=1e-10, atol=0)  # north-star tolerance; in practice ~1e-15
    assert_bitwise(x, xr, "solution")              # every kernel is bitwise => so is x


@pytest.mark.parametrize("nine", [False, True])
def test_heterogeneous_vcycle(cs, nine):
    rng = np.random.default_rng(7 + nine)
    nx, ny = 150, 93
    so = random_op2(nx, ny, nine, rng)
    b = random_gf(nx, ny, rng)
    ref = oracle.Solver2(so, "point", 2, 1)
    op = cs.Operator2(so)
    sol = cs.Solver2(op, relaxation="point", nrelax_pre=2, nrelax_post=1, maxiter=5, tol=1e-30)
    for lvl in range(1, sol.nlevels):
        assert_bitwise(sol.level_array(lvl, "A"), ref.level_array(lvl, "A"), f"A level {lvl}")
    x = random_gf(nx, ny, rng); xr = x.copy()
    for it in range(3):   # eager, capture, replay
        ref.vcycle(xr, b)
        sol.vcycle(x, b)
        assert_bitwise(x, xr, f"vcycle {it}")


@pytest.mark.parametrize("relax", ["line-x", "line-y", "line-xy"])
@pytest.mark.parametrize("case", [("diag_diffusion", 200, 120, 1.0, 1e-4), ("diag_diffusion", 97, 130, 1e-4, 1.0),
                                  ("fe", 65, 65, 1, 1)])
def test_line_solver_matches_oracle(cs, case, relax):
    name, nx, ny, dx, dy = case
    so = oracle.gallery2(name, nx, ny, dx, dy)
    b = oracle.rhs_sin2(nx, ny)
    ref = oracle.Solver2(so, relax, 2, 1)
    xr, hr, _ = ref.solve(b, maxiter=6, tol=1e-10)
    sol = cs.Solver2(cs.Operator2(so), relaxation=relax, nrelax_pre=2, nrelax_post=1, maxiter=6, tol=1e-10)
    x, h, _ = sol.solve(b)
    assert len(h) == len(hr)
    # the warp-scan line solve rounds differently from DPTTS2's sequential loop, and these
    # problems converge to the rounding floor (relative residual ~1e-13) in three cycles
    assert np.allclose(h, hr, rtol=1e-8, atol=2e-12), (h, hr)
    assert np.abs(x - xr).max() <= 1e-10 * np.abs(xr).max()


def test_solve_host_end_to_end(cs):
    nx, ny = 300, 200
    so = oracle.gallery2("fe", nx, ny)
    b = oracle.rhs_sin2(nx, ny)
    x, h = cs.solve_host(so, b, relaxation="point", nrelax_pre=1, nrelax_post=1, maxiter=20, tol=1e-10)
    assert h[-1] < 1e-10
    r = np.zeros_like(b)
    oracle.lib().orc2_residual(0, oracle.P(so), oracle.P(b), oracle.P(x), oracle.P(r), nx + 2, ny + 2, 0, 0, 5)
    assert np.linalg.norm(r[1:-1, 1:-1]) <= 1.01e-10 * np.linalg.norm(b[1:-1, 1:-1])


@pytest.mark.parametrize("case", [("poisson", 129, 129, 1, 1), ("fe", 100, 77, 2, 1), ("fe", 257, 257, 1, 1)])
def test_fcycle_matches_oracle(cs, case):
    """cycle type "f" (include/cedar/cycle/fcycle.h:49-83) against the oracle's F-cycle (oracle/ml2.c fmg_cycle,
    itself checked against a Python composition of the pinned kernels): iterate bitwise, history to 1e-10"""
    name, nx, ny, pre, post = case
    so = oracle.gallery2(name, nx, ny)
    b = oracle.rhs_sin2(nx, ny)
    ref = oracle.Solver2(so, "point", pre, post, cycle="f")
    xr, hr, r0r = ref.solve(b, maxiter=3, tol=1e-30)
    sol = cs.Solver2(cs.Operator2(so), relaxation="point", cycle="f", nrelax_pre=pre, nrelax_post=post, maxiter=3, tol=1e-30)
    x, h, r0 = sol.solve(b)
    assert_bitwise(x, xr, "F-cycle iterate")
    assert len(h) == len(hr) and np.allclose(h, hr, rtol=1e-10, atol=0) and abs(r0 - r0r) <= 1e-12 * r0r


def test_fcycle_runs(cs):
    nx, ny = 129, 129
    so = oracle.gallery2("poisson", nx, ny)
    b = oracle.rhs_sin2(nx, ny)
    sol = cs.Solver2(cs.Operator2(so), relaxation="point", cycle="f", maxiter=3, tol=1e-9)
    x, h, _ = sol.solve(b)
    # the reference's F-cycle restarts from x = 0 on every call (fcycle.h:77-78), so
    # repeating it is idempotent: one FMG pass reaches a few percent and stays there
    assert h[0] < 0.05 and abs(h[-1] - h[0]) <= 1e-12


@pytest.mark.parametrize("graph", [True, False])
def test_vcycle_batch_matches_single_calls(cs, graph):
    """A stream of right-hand sides through cdr2_solver_vcycle_batch (copies overlapped with the cycles on
    separate streams) gives, for every right-hand side, exactly the V-cycle of the oracle from x0 = 0."""
    nx, ny = 257, 200
    so = oracle.gallery2("fe", nx, ny)
    ref = oracle.Solver2(so, "point", 1, 1)
    sol = cs.Solver2(cs.Operator2(so), relaxation="point", nrelax_pre=1, nrelax_post=1, cuda_graph=graph)
    nrhs = 7
    rng = np.random.default_rng(77)
    bs = sol.pinned_grid_funcs(nrhs)
    xs = sol.pinned_grid_funcs(nrhs)
    for k in range(nrhs):
        bs[k][1:-1, 1:-1] = rng.standard_normal((ny, nx))
        xs[k][...] = 123.0            # must be overwritten, not used as an initial guess
    sol.vcycle_batch(list(bs), list(xs))
    hb, hx = sol.buffers()
    for k in range(nrhs):
        xr = np.zeros_like(bs[k])
        ref.vcycle(xr, bs[k].copy())
        assert_bitwise(xs[k], xr, f"rhs {k}")
        hb[...] = bs[k]
        hx[...] = 55.0
        sol.vcycle0_pinned()
        assert_bitwise(hx, xr, f"vcycle0 rhs {k}")
    # pageable host arrays are accepted too
    b2 = [np.array(bs[k]) for k in range(3)]
    x2 = [np.empty_like(b2[k]) for k in range(3)]
    sol.vcycle_batch(b2, x2)
    for k in range(3):
        assert_bitwise(x2[k], xs[k], f"pageable rhs {k}")


@pytest.mark.parametrize("nine", [False, True])
def test_solver_matvec(cs, nine):
    """cdr2_solver_matvec (registry kernel `matvec`) == the oracle's restatement of BMG2_SymStd_UTILS_matvec."""
    nx, ny = 70, 45
    rng = np.random.default_rng(8 + nine)
    so = random_op2(nx, ny, nine, rng)
    x = random_gf(nx, ny, rng)
    ref = np.zeros_like(x)
    oracle.lib().orc2_matvec(1, oracle.P(so), oracle.P(ref), oracle.P(x), nx + 2, ny + 2, 1, 0 if nine else 1, so.shape[0])
    sol = cs.Solver2(cs.Operator2(so), relaxation="point")
    assert_bitwise(sol.matvec(x), ref, "matvec")


@pytest.mark.parametrize("relax", ["point", "line-xy"])
def test_hetero_c3b_matches_oracle(cs, relax):
    """config C3(b): 5-point finite-volume operator on the checkerboard coefficient field (jumps up to 1e6), the
    same array into the oracle and the CUDA solver: identical iteration count, history to 1e-10 relative."""
    from cedar_b200 import gallery
    nx, ny = 300, 260
    so = gallery.hetero5(nx, ny, nblk=8)
    b = oracle.rhs_sin2(nx, ny)
    ref = oracle.Solver2(so, relax, 1, 1)
    xr, hr, r0r = ref.solve(b, maxiter=12, tol=1e-10)
    sol = cs.Solver2(cs.Operator2(so), relaxation=relax, nrelax_pre=1, nrelax_post=1, maxiter=12, tol=1e-10)
    x, h, r0 = sol.solve(b)
    assert len(h) == len(hr)
    if relax == "point":       # every point kernel is bitwise
        assert np.allclose(h, hr, rtol=1e-10, atol=0)
        assert_bitwise(x, xr, "solution")
    else:
        # the line solves are chunked scans (rounding differs from DPTTS2's sequential order) and the operator's
        # condition number carries the 1e6 jumps: the histories agree to 1e-10 relative down to a floor of 1e-12
        assert np.all(np.abs(h - hr) <= 1e-10 * hr + 1e-12), (h, hr)
        assert np.max(np.abs(x - xr)) <= 1e-6 * np.max(np.abs(xr))


def test_indefinite_coarse_operator_is_reported(cs):
    """DPBTRF INFO (BMG2_SymStd_SETUP_cg_LU.f90:124-131 logs 'Coarse grid Cholesky decomposition failed'): a
    coarsest-grid operator that is not positive definite must fail the setup loudly, not leave a half-factored ABD"""
    nx = ny = 5                      # one level: the fine operator IS the coarse-solve operator
    so = oracle.gallery2("fe", nx, ny)
    so[0, 1:-1, 1:-1] = -8.0         # negative diagonal: first pivot <= 0
    with pytest.raises(cs.CedarError, match="Cholesky"):
        cs.Solver2(cs.Operator2(so), relaxation="point", nrelax_pre=1, nrelax_post=1, num_levels=1)
    # a singular one (zero row sums everywhere, pure Neumann-like): the last pivot collapses
    so = oracle.gallery2("fe", nx, ny)
    so[0, 1:-1, 1:-1] = 0.0
    with pytest.raises(cs.CedarError, match="Cholesky"):
        cs.Solver2(cs.Operator2(so), relaxation="point", nrelax_pre=1, nrelax_post=1, num_levels=1)
    # and a good operator still sets up after the failures (the status word is reset by the next factorisation)
    sol = cs.Solver2(cs.Operator2(oracle.gallery2("fe", nx, ny)), relaxation="point", nrelax_pre=1, nrelax_post=1, num_levels=1)
    assert sol.nlevels == 1


@pytest.mark.parametrize("case", [("fe", 257, 257, 1, 1), ("poisson", 400, 300, 2, 1), ("fe", 1025, 513, 1, 2), ("fe", 130, 67, 0, 1)])
def test_coarse_tail_kernel_is_bitwise_and_saves_launches(cs, case):
    """solver.coarse-tail (default 33): the levels up to that size run their whole sub-cycle in ONE kernel
    (bmgx2_vcycle_tail).  Same bits as the level-by-level cycle (coarse-tail 0) and as the oracle; fewer launches."""
    name, nx, ny, pre, post = case
    so = oracle.gallery2(name, nx, ny)
    rng = np.random.default_rng(7)
    so[0, 1:-1, 1:-1] += rng.uniform(0, 0.3, size=(ny, nx))
    b = random_gf(nx, ny, rng)
    ref = oracle.Solver2(so, "point", pre, post)
    xr, hr, _ = ref.solve(b, maxiter=4, tol=1e-30)
    out = {}
    for tail in (0, 33, 129):
        sol = cs.Solver2(cs.Operator2(so), relaxation="point", nrelax_pre=pre, nrelax_post=post, maxiter=4, tol=1e-30,
                         coarse_tail=tail, cuda_graph=False)
        l0 = cs.gpu_launches()
        x, h, _ = sol.solve(b)
        out[tail] = (x, h, cs.gpu_launches() - l0)
        assert_bitwise(x, xr, f"coarse-tail {tail} vs oracle")
        assert np.allclose(h, hr, rtol=1e-10, atol=0)
    assert out[33][2] < out[0][2] and out[129][2] <= out[33][2], {k: v[2] for k, v in out.items()}


@pytest.mark.parametrize("case", [((True, False), 300, 300, 1, 1), ((False, True), 300, 300, 1, 1), ((True, False), 129, 200, 2, 1),
                                  ((True, True), 64, 96, 1, 1), ((False, True), 37, 50, 1, 2)])
def test_periodic_solver_matches_oracle(cs, case):
    """grid.periodic (reference examples/basic-2d-ser/periodic.cc, periodic-config.json): every kernel of the cycle takes
    its periodic branch -- interpolation setup, Galerkin operator, in-place point relaxation, restriction,
    interpolate-and-add, dense coarse solve with mean subtraction.  Hierarchy and iterate bitwise the oracle's (itself
    pinned kernel by kernel to the reference's Fortran), history to 1e-10."""
    from tests.util import periodic_problem2
    per, nx, ny, pre, post = case
    so, b, exact = periodic_problem2(nx, ny, per)
    if all(per):
        so[0] += 0.05 * so[0, 1, 1]          # the doubly periodic Laplacian is singular: make it definite
        so[0, 0, :] = so[0, ny, :]; so[0, ny + 1, :] = so[0, 1, :]; so[0, :, 0] = so[0, :, nx]; so[0, :, nx + 1] = so[0, :, 1]
    ref = oracle.Solver2(so, "point", pre, post, periodic=per)
    xr, hr, r0r = ref.solve(b, maxiter=8, tol=1e-30)
    assert hr[-1] < hr[0]
    sol = cs.Solver2(cs.Operator2(so), relaxation="point", nrelax_pre=pre, nrelax_post=post, maxiter=8, tol=1e-30, periodic=per)
    assert sol.nlevels == ref.nlevels
    for lvl in range(1, sol.nlevels):
        assert_bitwise(sol.level_array(lvl, "P"), ref.level_array(lvl, "P"), f"P level {lvl}")
        assert_bitwise(sol.level_array(lvl, "A"), ref.level_array(lvl, "A"), f"A level {lvl}")
    x, h, r0 = sol.solve(b)
    assert_bitwise(x, xr, "periodic solve")
    assert len(h) == len(hr) and np.allclose(h, hr, rtol=1e-10, atol=0) and abs(r0 - r0r) <= 1e-12 * r0r


@pytest.mark.parametrize("relaxation", ["line-x", "line-y", "line-xy"])
@pytest.mark.parametrize("case", [((True, False), 300, 300, 1, 1), ((False, True), 129, 200, 2, 1), ((True, True), 64, 96, 1, 1),
                                  ((True, True), 257, 130, 1, 2)])
def test_periodic_line_relaxation_solver_matches_oracle(cs, case, relaxation):
    """grid.periodic with zebra line relaxation (relax_lines_x.f90:170-306, relax_lines_y.f90:178-296: Sherman-Morrison
    second solve per periodic line, ghost refresh per colour).  Hierarchy bitwise the oracle's, the residual history to the
    tolerance of the non-periodic line cycles (the GPU line solve agrees with DPTTS2 to rounding), down to a floor."""
    from tests.util import periodic_problem2
    per, nx, ny, pre, post = case
    so, b, exact = periodic_problem2(nx, ny, per)
    if all(per):
        so[0] += 0.05 * so[0, 1, 1]          # the doubly periodic Laplacian is singular: make it definite
        so[0, 0, :] = so[0, ny, :]; so[0, ny + 1, :] = so[0, 1, :]; so[0, :, 0] = so[0, :, nx]; so[0, :, nx + 1] = so[0, :, 1]
    ref = oracle.Solver2(so, relaxation, pre, post, periodic=per)
    xr, hr, r0r = ref.solve(b, maxiter=8, tol=1e-30)
    assert hr[-1] < hr[0]
    sol = cs.Solver2(cs.Operator2(so), relaxation=relaxation, nrelax_pre=pre, nrelax_post=post, maxiter=8, tol=1e-30, periodic=per)
    assert sol.nlevels == ref.nlevels
    for lvl in range(1, sol.nlevels):
        assert_bitwise(sol.level_array(lvl, "P"), ref.level_array(lvl, "P"), f"P level {lvl}")
        assert_bitwise(sol.level_array(lvl, "A"), ref.level_array(lvl, "A"), f"A level {lvl}")
    x, h, r0 = sol.solve(b)
    assert len(h) == len(hr) and abs(r0 - r0r) <= 1e-12 * r0r
    assert np.allclose(h, hr, rtol=1e-8, atol=1e-12), (h, hr)
    assert np.abs(x - xr).max() <= 1e-9 * np.abs(xr).max()
