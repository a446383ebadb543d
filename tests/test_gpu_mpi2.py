"""GPU tests of the block-decomposed solver (cedar_b200/mpi2.py) over the C-ABI kernels:
all blocks on one GPU (LocalTransport) must reproduce the serial GPU solver and the CPU
oracle BITWISE; with >= 2 GPUs the same check runs one rank per GPU over NCCL
(tests/run_mpi2_nccl.py under torchrun)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle
from tests.util import assert_bitwise, random_gf, random_op2

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("nx,ny,px,py,nine,agg", [(257, 257, 2, 2, True, 33), (513, 257, 4, 2, True, 65),
                                                  (300, 210, 2, 2, False, 33), (1025, 1025, 2, 1, True, 129),
                                                  (129, 129, 2, 2, True, 1000)])
def test_blocks_on_one_gpu_equal_serial(cuda, nx, ny, px, py, nine, agg):
    import cedar_b200.mpi2 as m
    import cedar_b200.solver as cs
    rng = np.random.default_rng(nx + ny + nine)
    so = random_op2(nx, ny, nine, rng)
    b = random_gf(nx, ny, rng)
    ref = oracle.Solver2(so, "point", 2, 1)
    xr, hr, r0r = ref.solve(b, maxiter=4, tol=1e-12)
    ser = cs.Solver2(cs.Operator2(so), relaxation="point", nrelax_pre=2, nrelax_post=1, maxiter=4, tol=1e-12)
    xs, hs, _ = ser.solve(b)
    assert_bitwise(xs, xr, "serial GPU vs oracle")
    sol = m.DistSolver2(m.split_blocks(so, _dec(m, nx, ny, px, py, agg)), nx, ny, (px, py), 2, 1, agglomerate=agg)
    dec = sol.dec
    for lvl in range(1, sol.ndist + 1):
        assert_bitwise(m.join_blocks(sol.lv[lvl].A, dec, lvl), ref.level_array(lvl, "A"), f"A level {lvl}")
    x, h, r0 = sol.solve(m.split_blocks(b, dec), maxiter=4, tol=1e-12)
    assert_bitwise(m.join_blocks(x, dec), xr, "decomposed GPU vs oracle")
    assert len(h) == len(hr) and np.allclose(h, hr, rtol=1e-10, atol=0) and abs(r0 - r0r) <= 1e-12 * r0r
    assert sol.K.launches > 0


def _dec(m, nx, ny, px, py, agg):
    nl = m.compute_num_levels(nx, ny)
    nd, n = 0, (nx, ny)
    while nd < nl - 1 and min(n) > agg:
        nd += 1
        n = (m.coarse_n(n[0]), m.coarse_n(n[1]))
    return m.Decomp2(nx, ny, px, py, nd)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_nccl_ranks_equal_serial(cuda, world):
    if cuda.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29531", os.path.join(ROOT, "tests", "run_mpi2_nccl.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "MPI2 NCCL OK" in r.stdout
