"""CPU tests of the block-decomposed solver's host logic (cedar_b200/mpi2.py): the
reference's decomposition rule, the ghost-exchange schedule (test/2d/mpi/test_halo.h
ghost-value formula), agglomeration and the residual reduction -- with the CPU oracle
injected as the compute backend (tests/oracle_backend.py), all blocks in one process
and one block per rank over gloo (world_size 2 and 4).  The decomposed solve must be
BITWISE the serial one (same arithmetic per point)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import oracle
import cedar_b200.mpi2 as m
from tests.oracle_backend import OracleCoarse, OracleKernels
from tests.util import random_gf, random_op2


def test_grid_decomp_rule():
    # include/cedar/decomp.h:59-86
    assert m.grid_decomp((16385, 16385), 8) == [4, 2]
    assert m.grid_decomp((16385, 16385), 4) == [2, 2]
    assert m.grid_decomp((16385, 16385), 2) == [2, 1]
    assert m.grid_decomp((100, 400), 6) == [2, 3] or m.grid_decomp((100, 400), 6) == [1, 6]
    # block_partition::low for 2^k+1 grids is already aligned: blocks are the reference's
    d = m.Decomp2(16385, 16385, 4, 2, 5)
    assert [d.extent(0, b)[0] for b in range(4)] == [i * 16385 // 4 for i in range(4)]
    assert [d.extent(0, b)[2] for b in range(4)] == [4096, 4096, 4096, 4097]
    # every block starts on a coarse point of every distributed level and the pieces tile the grid
    for nx, ny, px, py, nd in [(100, 77, 2, 2, 3), (200, 131, 3, 2, 3), (65, 65, 2, 2, 3)]:
        d = m.Decomp2(nx, ny, px, py, nd)
        for lvl in range(nd + 1):
            covered = np.zeros((d.gn[lvl][1], d.gn[lvl][0]), dtype=int)
            for b in range(d.nblocks):
                x0, y0, nxl, nyl = d.extent(lvl, b)
                if lvl < nd:
                    assert x0 % 2 == 0 and y0 % 2 == 0
                covered[y0:y0 + nyl, x0:x0 + nxl] += 1
            assert (covered == 1).all()


def ndist_for(nx, ny, agg):
    nl = m.compute_num_levels(nx, ny)
    nd, n = 0, (nx, ny)
    while nd < nl - 1 and min(n) > agg:
        nd += 1
        n = (m.coarse_n(n[0]), m.coarse_n(n[1]))
    return nd


def test_halo_exchange_ghost_values():
    """after an exchange every ghost entry equals the owner's value (test_halo.h:18-114)"""
    nx, ny, px, py = 37, 29, 3, 2
    dec = m.Decomp2(nx, ny, px, py, 1)
    g = np.zeros((2, ny + 2, nx + 2))
    jj, ii = np.meshgrid(np.arange(ny + 2), np.arange(nx + 2), indexing="ij")
    g[0] = ii * 1000 + jj
    g[1] = -(ii * 1000 + jj)
    g[:, 0, :] = 0; g[:, -1, :] = 0; g[:, :, 0] = 0; g[:, :, -1] = 0
    so = {b: np.zeros((3,) + v.shape[1:]) for b, v in m.split_blocks(g, dec).items()}
    sol = m.DistSolver2.__new__(m.DistSolver2)
    sol.dec, sol.T, sol.K, sol.device, sol.exchanges = dec, m.LocalTransport(dec), OracleKernels(), torch.device("cpu"), 0
    sol.owned = sol.T.owned
    arrs = {b: torch.from_numpy(v.copy()) for b, v in m.split_blocks(g, dec).items()}     # ghosts zeroed
    sol.exchange(0, arrs)
    want = m.split_blocks(g, dec, keep_ghosts=True)
    for b in range(dec.nblocks):
        assert np.array_equal(arrs[b].numpy(), want[b]), b


CASES = [(65, 65, 2, 1, True, 9), (65, 65, 2, 2, True, 9), (129, 65, 4, 2, True, 9), (100, 77, 2, 2, True, 9),
         (65, 65, 2, 2, False, 9), (200, 131, 3, 2, False, 17), (33, 33, 2, 2, True, 100)]


@pytest.mark.parametrize("fuse_min", [8, 10 ** 9])    # wide-halo fused sweeps on every big enough level / never
@pytest.mark.parametrize("nx,ny,px,py,nine,agg", CASES)
def test_multiblock_equals_serial(nx, ny, px, py, nine, agg, fuse_min):
    rng = np.random.default_rng(nx + ny + nine)
    so = random_op2(nx, ny, nine, rng)
    b = random_gf(nx, ny, rng)
    ref = oracle.Solver2(so, "point", 2, 1)
    xr, hr, r0r = ref.solve(b, maxiter=5, tol=1e-12)
    dec = m.Decomp2(nx, ny, px, py, ndist_for(nx, ny, agg))
    sol = m.DistSolver2(m.split_blocks(so, dec), nx, ny, (px, py), 2, 1, agglomerate=agg, kernels=OracleKernels(),
                        coarse_factory=OracleCoarse, fuse_min=fuse_min)
    assert sol.nlevels == ref.nlevels and sol.ndist == dec.ndist
    if fuse_min == 8 and sol.ndist > 0 and min(nx // px, ny // py) >= 16:
        assert 0 in sol.ext            # the fused path is what runs on the fine level
    for lvl in range(1, sol.ndist + 1):
        assert np.array_equal(m.join_blocks(sol.lv[lvl].A, dec, lvl), ref.level_array(lvl, "A")), lvl
        Pr = ref.level_array(lvl, "P").copy()
        Pr[:, 0, :] = 0; Pr[:, -1, :] = 0; Pr[:, :, 0] = 0; Pr[:, :, -1] = 0
        assert np.array_equal(m.join_blocks(sol.lv[lvl].P, dec, lvl), Pr), lvl
    x, h, r0 = sol.solve(m.split_blocks(b, dec), maxiter=5, tol=1e-12)
    assert np.array_equal(m.join_blocks(x, dec), xr)
    assert len(h) == len(hr) and np.allclose(h, hr, rtol=1e-10, atol=0) and abs(r0 - r0r) <= 1e-12 * r0r


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nx, ny, px, py, nine, agg, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        oracle.lib().orc_set_threads(1)
        rng = np.random.default_rng(nx + ny + nine)
        so = random_op2(nx, ny, nine, rng)
        b = random_gf(nx, ny, rng)
        dec = m.Decomp2(nx, ny, px, py, ndist_for(nx, ny, agg))
        T = m.DistTransport(dec)
        sol = m.DistSolver2(m.split_blocks(so, dec, blocks=[rank]), nx, ny, (px, py), 2, 1, agglomerate=agg,
                            transport=T, kernels=OracleKernels(), coarse_factory=OracleCoarse, fuse_min=8)
        x, h, r0 = sol.solve(m.split_blocks(b, dec, blocks=[rank]), maxiter=5, tol=1e-12)
        np.save(os.path.join(out, f"x{rank}.npy"), x[rank].numpy())
        if rank == 0:
            np.save(os.path.join(out, "h.npy"), np.concatenate([[r0], h]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("nx,ny,px,py,nine,agg", [(65, 65, 2, 1, True, 9), (100, 77, 2, 2, False, 9)])
def test_gloo_ranks_equal_serial(tmp_path, nx, ny, px, py, nine, agg):
    world = px * py
    mp.spawn(_worker, args=(world, _free_port(), nx, ny, px, py, nine, agg, str(tmp_path)), nprocs=world, join=True)
    rng = np.random.default_rng(nx + ny + nine)
    so = random_op2(nx, ny, nine, rng)
    b = random_gf(nx, ny, rng)
    xr, hr, r0r = oracle.Solver2(so, "point", 2, 1).solve(b, maxiter=5, tol=1e-12)
    dec = m.Decomp2(nx, ny, px, py, ndist_for(nx, ny, agg))
    x = m.join_blocks({r: np.load(tmp_path / f"x{r}.npy") for r in range(world)}, dec)
    assert np.array_equal(x, xr)
    h = np.load(tmp_path / "h.npy")
    assert np.allclose(h[1:], hr, rtol=1e-10, atol=0) and abs(h[0] - r0r) <= 1e-12 * r0r
