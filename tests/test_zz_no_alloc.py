"""GPU test of bmgx2_relax_lines_y_static (include/bmgx.h): the y-line relaxation whose scratch is a device array built
into the library, so that the call allocates nothing.  Through the C ABI, against the existing entry point
bmgx2_relax_lines_y with caller scratch: same kernels, so the results must agree BITWISE (the tolerance of the other
kernel tests of this repo) -- on two calls in a row that share the array (the second on a smaller grid, so it runs over
the first call's leftovers), and sizes beyond the built-in limit must be refused.  Collected last on purpose."""
import numpy as np
import pytest

from tests.util import assert_bitwise, dev, host, ptr, random_gf, random_op2

pytestmark = pytest.mark.gpu

STATIC_MAX = 1 << 20


def _case(torch, abi, nx, ny, nine, seed):
    L = abi.lib
    rng = np.random.default_rng(seed)
    so = random_op2(nx, ny, nine, rng)
    nst, ifd = so.shape[0], (0 if nine else 1)
    II, JJ = nx + 2, ny + 2
    q, b = random_gf(nx, ny, rng), random_gf(nx, ny, rng)
    d_so, d_b = dev(torch, so), dev(torch, b)
    d_sor = torch.zeros((2, II, JJ), dtype=torch.float64, device="cuda")
    abi.check(L.bmgx2_SETUP_lines_y(ptr(d_so), ptr(d_sor), II, JJ, nst, 0, None))
    d_q_ref, d_q_new = dev(torch, q), dev(torch, q)
    d_work = torch.full((JJ, II), 3.0, dtype=torch.float64, device="cuda")
    return L, d_so, d_b, d_sor, d_q_ref, d_q_new, d_work, II, JJ, ifd, nst


def test_relax_lines_y_static_equals_caller_scratch_twice(cuda):
    import cedar_b200.abi as abi
    torch = cuda
    # two calls in a row on the shared array: a 9-point grid close to the limit, then a smaller 5-point grid
    for (nx, ny, nine, seed) in [(1000, 1020, True, 1), (300, 411, False, 2)]:
        L, d_so, d_b, d_sor, d_q_ref, d_q_new, d_work, II, JJ, ifd, nst = _case(torch, abi, nx, ny, nine, seed)
        assert II * JJ <= STATIC_MAX
        for updown in (0, 1):
            abi.check(L.bmgx2_relax_lines_y(1, ptr(d_so), ptr(d_b), ptr(d_q_ref), ptr(d_sor), ptr(d_work), II, JJ, 1, ifd, nst,
                                            1, updown, 0, None))
            mem = torch.cuda.memory_allocated()
            abi.check(L.bmgx2_relax_lines_y_static(1, ptr(d_so), ptr(d_b), ptr(d_q_new), ptr(d_sor), II, JJ, 1, ifd, nst, 1,
                                                   updown, 0, None))
            assert torch.cuda.memory_allocated() == mem
            torch.cuda.synchronize()
            assert_bitwise(host(d_q_new), host(d_q_ref), f"relax_lines_y_static {nx}x{ny} updown={updown}")
        assert not np.array_equal(host(d_q_new), np.zeros((JJ, II)))


def test_relax_lines_y_static_rejects_sizes_beyond_the_limit(cuda):
    import cedar_b200.abi as abi
    torch = cuda
    L = abi.lib
    II, JJ = 1025, 1025                     # 1050625 > 1048576
    assert II * JJ > STATIC_MAX
    z = torch.zeros(8, dtype=torch.float64, device="cuda")
    st = L.bmgx2_relax_lines_y_static(1, ptr(z), ptr(z), ptr(z), ptr(z), II, JJ, 1, 1, 3, 1, 0, 0, None)
    assert st != 0
    msg = L.bmgx_last_error().decode()
    assert "BMGX_LINES_Y_STATIC_MAX" in msg and "1048576" in msg
    # the largest square that fits is still accepted
    L2, d_so, d_b, d_sor, d_q_ref, d_q_new, d_work, II, JJ, ifd, nst = _case(torch, abi, 1022, 1022, False, 3)
    assert II * JJ == STATIC_MAX
    abi.check(L2.bmgx2_relax_lines_y(1, ptr(d_so), ptr(d_b), ptr(d_q_ref), ptr(d_sor), ptr(d_work), II, JJ, 1, ifd, nst, 1, 0, 0,
                                     None))
    abi.check(L2.bmgx2_relax_lines_y_static(1, ptr(d_so), ptr(d_b), ptr(d_q_new), ptr(d_sor), II, JJ, 1, ifd, nst, 1, 0, 0, None))
    torch.cuda.synchronize()
    assert_bitwise(host(d_q_new), host(d_q_ref), "relax_lines_y_static at the limit")
