"""GPU parity of 3D plane relaxation (include/cedar/3d/relax_planes.h over the C ABI):
staging kernels bitwise against numpy restatements of copy32/copy23/copy_coeff/copy_rhs
(src/3d/relax_planes.cc), and the full solver with plane-xy/xz/yz/xyz relaxation against
the CPU oracle (oracle/planes3.py) to 1e-10 (the in-plane line solves are warp scans on
the GPU, 1e-12 relative to the sequential LAPACK order; everything else is bitwise)."""
import numpy as np
import pytest

import oracle
from oracle import planes3
from tests.util import assert_bitwise, dev, host, ptr, random_gf3, random_op3

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def abi(cuda):
    import cedar_b200.abi as a
    return a


@pytest.fixture(scope="module")
def cs(cuda):
    import cedar_b200.solver as s
    return s


@pytest.mark.parametrize("full", [False, True])
@pytest.mark.parametrize("d", ["xy", "xz", "yz"])
def test_staging_kernels(cuda, abi, d, full):
    nx, ny, nz = 13, 9, 11
    rng = np.random.default_rng(3 + full)
    so = random_op3(nx, ny, nz, full, rng)
    x, b = random_gf3(nx, ny, nz, rng), random_gf3(nx, ny, nz, rng)
    II, JJ, KK = nx + 2, ny + 2, nz + 2
    code = {"xy": 0, "xz": 1, "yz": 2}[d]
    npl = {"xy": nz, "xz": ny, "yz": nx}[d]
    d_so, d_x, d_b = dev(cuda, so), dev(cuda, x), dev(cuda, b)
    shape2 = planes3._plane(x, d, 1).shape
    for ipl in (1, 2, npl):
        # copy32
        d_x2 = cuda.zeros(shape2, dtype=cuda.float64, device="cuda")
        abi.check(abi.lib.bmgx3_plane_copy(ptr(d_x), ptr(d_x2), II, JJ, KK, code, ipl + 1, 1, None))
        assert_bitwise(host(d_x2), planes3._plane(x, d, ipl), "copy32")
        # copy_rhs
        offs = (planes3._RHS27 if full else planes3._RHS7)[d]
        acc = planes3._plane(b, d, ipl)[1:-1, 1:-1].copy()
        for off in offs:
            e, s = planes3._CPL[off]
            acc = acc + planes3._plane_int(planes3._shift(so[e], s), d, ipl) * planes3._plane_int(planes3._shift(x, off), d, ipl)
        d_b2 = cuda.zeros(shape2, dtype=cuda.float64, device="cuda")
        abi.check(abi.lib.bmgx3_plane_rhs(ptr(d_so), ptr(d_x), ptr(d_b), ptr(d_b2), II, JJ, KK, so.shape[0], code, ipl + 1, None))
        assert_bitwise(host(d_b2)[1:-1, 1:-1], acc, "copy_rhs")
        # copy_coeff
        slots = planes3._COEFF[d][:5 if full else 3]
        want = np.stack([planes3._plane(so[e], d, ipl) for e in slots])
        d_so2 = cuda.zeros(want.shape, dtype=cuda.float64, device="cuda")
        abi.check(abi.lib.bmgx3_plane_coeff(ptr(d_so), ptr(d_so2), II, JJ, KK, so.shape[0], code, ipl + 1, None))
        assert_bitwise(host(d_so2), want, "copy_coeff")
        # copy23
        y2 = rng.standard_normal(shape2)
        abi.check(abi.lib.bmgx3_plane_copy(ptr(d_x), ptr(dev(cuda, y2)), II, JJ, KK, code, ipl + 1, 0, None))
        xw = x.copy()
        planes3._plane(xw, d, ipl)[...] = y2
        assert_bitwise(host(d_x), xw, "copy23")
        d_x.copy_(dev(cuda, x))


@pytest.mark.parametrize("case", [("poisson", 17, 17, 17, "plane-xyz"), ("fe", 17, 17, 17, "plane-xyz"),
                                  ("poisson", 20, 13, 17, "plane-xy"), ("fe", 12, 21, 17, "plane-xz"),
                                  ("poisson", 17, 12, 20, "plane-yz"), ("fe", 33, 33, 33, "plane-xyz")])
def test_plane_solver_matches_oracle(cs, case):
    name, nx, ny, nz, relax = case
    so = oracle.gallery3(name, nx, ny, nz)
    b = oracle.rhs_sin3(nx, ny, nz)
    ref = planes3.Solver3Planes(so, relax, 1, 1)
    xr, hr, r0r = ref.solve(b, maxiter=6, tol=1e-10)
    sol = cs.Solver3(cs.Operator3(so), relaxation=relax, nrelax_pre=1, nrelax_post=1, maxiter=6, tol=1e-10)
    assert sol.nlevels == ref.nlevels
    x, h, r0 = sol.solve(b)
    assert len(h) == len(hr)
    assert abs(r0 - r0r) <= 1e-13 * r0r
    assert np.all(np.abs(h - hr) <= 1e-8 * hr + 2e-15), (h, hr)
    # thousands of line-relaxed plane cycles, each 1e-12 away from the sequential LAPACK order
    assert np.max(np.abs(x - xr)) <= 1e-7 * np.max(np.abs(xr))
    # it converges.  How fast depends on the reference quirk kept on purpose (relax_planes.h:77-161: every
    # plane solver gets the coefficients of the LAST interior plane): harmless on the fine level of the
    # constant-coefficient galleries, but the Galerkin coarse operators are not plane-invariant near the
    # boundary, so deeper hierarchies (33^3: 5 levels) contract by only ~0.7 per cycle, here and in the oracle
    assert h[-1] < h[0] * (1e-2 if max(nx, ny, nz) <= 20 else 0.5)


# ---- the reference's own plane test (test/3d/test_planes.cc:13-265) on the GPU path ---------------------------------
from tests import planes_ref_cases as ref_planes  # noqa: E402


@pytest.mark.parametrize("name,d,n", ref_planes.CASES)
def test_reference_plane_cases(cs, name, d, n):
    """SerialPlanes.XY7 ... YZ27: build_kernel_manager, setup<plane_relax<dir>>, run(DOWN) through the registry
    (cdr3_plane_relax); even planes against the sparse direct solve to the reference's 1e-8, and the whole sweep
    against the pinned CPU oracle (point-relaxed plane solves are bitwise kernels: 1e-12 relative)"""
    nx, ny, nz = n
    so = oracle.gallery3(name, nx, ny, nz)
    x = ref_planes.index_field(nx, ny, nz)
    b = x.copy()
    xg = cs.plane_relax3(cs.Operator3(so), d, x.copy(), b, plane_config=ref_planes.PLANE_CONFIG)
    assert ref_planes.check_even_planes(name, d, n, so, xg, b) < ref_planes.TOL
    kw = ref_planes.PLANE_CONFIG["solver"]
    pl = planes3._Planes(so, d, dict(relaxation="point", nrelax_pre=2, nrelax_post=1), iters=kw["max-iter"])
    xo = x.copy()
    pl.relax(so, xo, b, True)
    assert np.max(np.abs(xg - xo)) <= 1e-12 * np.max(np.abs(xo))


@pytest.mark.parametrize("case", [("poisson", 17, 17, 17, "plane-xyz"), ("fe", 20, 13, 17, "plane-xy"), ("fe", 17, 12, 20, "plane-yz")])
def test_plane_solver_point_planes_tight(cs, case):
    """with POINT relaxation inside the planes every kernel of the plane cycles is bitwise the oracle's; what is left
    is the 3D hierarchy itself: the re-derived 3D interpolation and Galerkin kernels agree with the oracle to 1e-12
    of the diagonal, not bitwise (tests/test_gpu_kernels3d.py), and that difference in the coarse operators is what
    the history sees -- so the first residuals agree to 1e-10 (north_star's tolerance) and the bound is allowed to
    grow by the contraction factor lost per cycle (the residual drops ~25x per cycle, an operator perturbation
    does not)"""
    name, nx, ny, nz, relax = case
    so = oracle.gallery3(name, nx, ny, nz)
    b = oracle.rhs_sin3(nx, ny, nz)
    pk = dict(relaxation="point", nrelax_pre=2, nrelax_post=1)
    ref = planes3.Solver3Planes(so, relax, 1, 1, plane_kw=pk)
    xr, hr, r0r = ref.solve(b, maxiter=5, tol=1e-30)
    import json
    cfg = json.loads(cs.make_config(relaxation=relax, nrelax_pre=1, nrelax_post=1, maxiter=5, tol=1e-30))
    cfg["plane-config"] = {"solver": {"relaxation": "point", "max-iter": 1, "cycle": {"nrelax-pre": 2, "nrelax-post": 1}}}
    sol = cs.Solver3(cs.Operator3(so), config_json=json.dumps(cfg))
    x, h, r0 = sol.solve(b)
    assert len(h) == len(hr)
    dev = np.abs(h - hr) / hr
    print("relative deviation of the history:", dev, "iterate:", np.max(np.abs(x - xr)) / np.max(np.abs(xr)))
    assert dev[0] <= 1e-10 and np.all(dev <= 1e-10 * (hr[0] / hr)), dev
    assert np.max(np.abs(x - xr)) <= 1e-10 * np.max(np.abs(xr))
