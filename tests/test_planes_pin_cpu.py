"""Pins the plane-relaxation oracle (oracle/planes3.py) to the reference's own plane test
(test/3d/test_planes.cc:13-265, restated in tests/planes_ref_cases.py): six cases, one zebra sweep with converged
plane solves, even planes against a sparse direct solve, 1e-8 absolute -- the reference's tolerance."""
import numpy as np
import pytest

import oracle
from oracle import planes3
from tests import planes_ref_cases as ref


@pytest.mark.parametrize("name,d,n", ref.CASES)
def test_oracle_plane_sweep_against_direct_solve(name, d, n):
    nx, ny, nz = n
    so = oracle.gallery3(name, nx, ny, nz)
    x = ref.index_field(nx, ny, nz)
    b = x.copy()
    kw = ref.PLANE_CONFIG["solver"]
    pl = planes3._Planes(so, d, dict(relaxation=kw["relaxation"], nrelax_pre=2, nrelax_post=1), iters=kw["max-iter"])
    pl.relax(so, x, b, True)
    assert ref.check_even_planes(name, d, n, so, x, b) < ref.TOL


def test_inplane_matrix_is_the_operators_plane():
    """the matrices pyplanes.pyx hard-codes are the in-plane part of the gallery operators (so copy_coeff's 'last
    plane' shortcut is exact for them): A_plane x == what copy_coeff + the 2D residual would apply"""
    for name, d, n in ref.CASES:
        nx, ny, nz = n
        so = oracle.gallery3(name, nx, ny, nz)
        pl = planes3._Planes(so, d, dict(relaxation="point", nrelax_pre=1, nrelax_post=1))
        so2 = pl.so2
        rng = np.random.default_rng(1)
        q = np.zeros(so2.shape[1:]); q[1:-1, 1:-1] = rng.standard_normal((so2.shape[1] - 2, so2.shape[2] - 2))
        r = np.zeros_like(q)
        zero = np.zeros_like(q)
        nst2 = so2.shape[0]
        oracle.lib().orc2_residual(0, oracle.P(so2), oracle.P(zero), oracle.P(q), oracle.P(r), so2.shape[2], so2.shape[1], 0,
                                   1 if nst2 == 3 else 0, nst2)                      # r = 0 - A q
        A = ref.inplane_matrix(name, d, n)
        assert np.allclose(-r[1:-1, 1:-1].ravel(), A @ q[1:-1, 1:-1].ravel(), rtol=1e-12, atol=1e-12), (name, d)
