"""GPU tests of the slab-decomposed 3D solver (cedar_b200/host/dist3.cc, cdr3_dist_*): several slabs on ONE GPU must
reproduce the serial GPU solver BITWISE (same kernels, same operands per point: hierarchy, cycles eager and replayed,
solve) and the CPU oracle to the 3D setup kernels' tolerance; with >= 2 GPUs the same check runs one rank per GPU
(tests/run_dist3_ranks.py under torchrun)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle
from tests.util import assert_bitwise, random_gf3, random_op3

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CASES = [(33, 29, 129, 4, False, 33, 5, 1, 1), (33, 33, 129, 4, True, 33, 5, 1, 1), (40, 25, 193, 3, True, 33, 7, 2, 1),
         (65, 65, 257, 8, False, 65, 5, 1, 1), (21, 21, 129, 2, True, 17, 3, 1, 2)]


@pytest.mark.parametrize("nx,ny,nz,np_,full,agg,halo,pre,post", CASES)
def test_slabs_on_one_gpu_equal_serial(cuda, nx, ny, nz, np_, full, agg, halo, pre, post):
    import cedar_b200.dist3 as d3
    import cedar_b200.solver as cs
    rng = np.random.default_rng(nx + ny + nz + full)
    so = random_op3(nx, ny, nz, full, rng)
    b = random_gf3(nx, ny, nz, rng)
    ser = cs.Solver3(cs.Operator3(so), relaxation="point", nrelax_pre=pre, nrelax_post=post, maxiter=5, tol=1e-30)
    sol = d3.DistSolver3(nx, ny, nz, np_, nst=so.shape[0], nrelax_pre=pre, nrelax_post=post, agglomerate=agg, halo=halo)
    ext = {k: sol.extent(0, k) for k in range(np_)}
    for k, v in d3.split_slabs(so, ext).items():
        sol.set_operator(k, v)
    sol.setup()
    assert sol.nlevels == ser.nlevels and sol.ndist >= 1
    for lvl in range(1, sol.ndist + 1):
        for name in ("A", "P"):
            want = ser.level_array(lvl, name)
            for k in range(np_):
                _, _, z0, nzl = sol.extent(lvl, k)
                got = sol.level_array(k, lvl, name)
                assert_bitwise(got[:, :, 1:-1, 1:-1], want[:, 1 + z0:1 + z0 + nzl, 1:-1, 1:-1], f"{name} level {lvl} slab {k}")
    for k, v in d3.split_slabs(b, ext).items():
        sol.upload(k, "b", v)
        sol.upload(k, "x", np.zeros_like(v))
    h, r0 = sol.solve(maxiter=5, tol=1e-30)             # 2 eager cycles, capture, replays
    xs, hs, r0s = ser.solve(b)
    for k in range(np_):
        _, _, z0, nzl = ext[k]
        assert_bitwise(sol.download(k)[1:-1], xs[1 + z0:1 + z0 + nzl], f"x, slab {k}")
    assert len(h) == len(hs) and np.allclose(h, hs, rtol=1e-10, atol=0) and abs(r0 - r0s) <= 1e-12 * r0s
    xo, ho, _ = oracle.Solver3(so, pre, post).solve(b, maxiter=5, tol=1e-30)
    assert np.allclose(h, ho, rtol=1e-8, atol=0)         # the 3D setup kernels agree with the oracle to 1e-12, not bitwise
    sol.close()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_ranks_equal_serial_3d(cuda, world):
    if cuda.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29551", os.path.join(ROOT, "tests", "run_dist3_ranks.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "DIST3 RANKS OK" in r.stdout


@pytest.mark.parametrize("full", [False, True])
def test_halo_exchange_contract_3d(cuda, full):
    """the reference's ghost-exchange contract (test/3d/mpi/test_halo.cc: global-index pattern in, every ghost checked,
    grid functions and stencils, every level) through halo_exchanger_base over the slab exchange"""
    import cedar_b200.dist3 as d3
    n, nzg, slabs, W = 33, 130, 4, 3
    nst = 14 if full else 4
    sol = d3.DistSolver3(n, n, nzg, slabs, nst=nst, agglomerate=17, halo=W)
    assert sol.ndist >= 2
    for lvl in range(sol.ndist + 1):
        ext = {k: sol.extent(lvl, k) for k in range(slabs)}
        gnz = max(z0 + nz for (_, _, z0, nz) in ext.values())
        ne = nst if lvl == 0 else 14
        for k, (nx, ny, z0, nz) in ext.items():
            kk, jj, ii = np.meshgrid(np.arange(z0, z0 + nz), np.arange(ny + 2), np.arange(nx + 2), indexing="ij")
            f = ((kk * (ny + 2) + jj) * (nx + 2) + ii + 1).astype(np.float64)
            sol.upload_level(k, lvl, "x", f[None])
            sol.upload_level(k, lvl, "A", np.stack([f * ne + p for p in range(ne)]))
        sol.exchange(lvl, "x")
        sol.exchange(lvl, "A")
        for k, (nx, ny, z0, nz) in ext.items():
            g = d3.geometry(n, n, nzg, slabs, sol.ndist, W, lvl, k)
            kk, jj, ii = np.meshgrid(np.arange(g["kk"]) - g["wb"] + z0, np.arange(ny + 2), np.arange(nx + 2), indexing="ij")
            inside = (kk >= 0) & (kk < gnz)
            want = np.where(inside, (kk * (ny + 2) + jj) * (nx + 2) + ii + 1, 0).astype(np.float64)
            assert_bitwise(sol.extended_array(k, lvl, "x")[0], want, f"grid function, level {lvl}, slab {k}")
            a = sol.extended_array(k, lvl, "A")
            for p in range(ne):
                assert_bitwise(a[p], np.where(inside, want * ne + p, 0.0), f"stencil entry {p}, level {lvl}, slab {k}")
    sol.close()


@pytest.mark.parametrize("np_", [1, 3, 8])
def test_reference_mpi_poisson3_case(cuda, np_):
    """The reference's distributed acceptance test MPIPoisson3.Isotropic (test/3d/mpi/test_poisson.cc:100-134): 200^3
    7-point Poisson, point relaxation, default settings (V(2,1), at most 10 cycles, tol 1e-8): final residual norm < 1e-8
    and |x - sin(2 pi x) sin(2 pi y) sin(2 pi z)|_inf < 1e-4 -- on 1 (serial solver), 3 and 8 slabs."""
    import cedar_b200.dist3 as d3
    import cedar_b200.solver as cs
    n = 200
    op = cs.Operator3(gallery="poisson", nx=n, ny=n, nz=n)
    so = np.array(op.array())
    h = 1.0 / (n + 1)
    s1 = np.sin(2 * np.pi * h * np.arange(n + 2))
    exact = s1[:, None, None] * s1[None, :, None] * s1[None, None, :]
    b = np.zeros((n + 2,) * 3)
    b[1:-1, 1:-1, 1:-1] = (12 * np.pi ** 2 * h ** 3 * exact)[1:-1, 1:-1, 1:-1]      # examples/basic-3d-ser/poisson.cc:16-38
    if np_ == 1:
        x, hist, r0 = cs.Solver3(cs.Operator3(so), relaxation="point", nrelax_pre=2, nrelax_post=1, maxiter=10, tol=1e-8).solve(b)
    else:
        sol = d3.DistSolver3(n, n, n, np_, nst=4, nrelax_pre=2, nrelax_post=1, agglomerate=25, halo=5)
        ext = {k: sol.extent(0, k) for k in range(np_)}
        for k, v in d3.split_slabs(so, ext).items():
            sol.set_operator(k, v)
        sol.setup()
        for k, v in d3.split_slabs(b, ext).items():
            sol.upload(k, "b", v)
            sol.upload(k, "x", np.zeros_like(v))
        hist, r0 = sol.solve(maxiter=10, tol=1e-8)
        x = np.zeros_like(b)
        for k in range(np_):
            _, _, z0, nzl = ext[k]
            x[1 + z0:1 + z0 + nzl] = sol.download(k)[1:-1]
        sol.close()
    assert hist[-1] * r0 < 1e-8, (hist, r0)
    assert np.abs(exact[1:-1, 1:-1, 1:-1] - x[1:-1, 1:-1, 1:-1]).max() < 1e-4
