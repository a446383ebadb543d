"""CPU tests of the multi-GPU 2D path's host logic (include/cedar/2d/mpi/plan.h: decomposition, deep-halo
validity tracking, where ghost exchanges go, coarse windows, replicated agglomerated levels).  The operation
lists come from the product library; tests/dist_plan_sim.py runs them with the CPU oracle's kernels, all blocks
in one process and one block per rank over gloo (world_size 2 and 4).  The decomposed result must be BITWISE the
serial oracle's (same arithmetic per point).  Halos start as NaN in the simulator, so a kernel
that reads a halo value the planner believes valid but nobody delivered poisons the own points."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

import oracle
import cedar_b200.dist2 as d2
from tests.dist_plan_sim import PlanSim, X0, B, A, PI
from tests.util import random_gf, random_op2


def test_grid_decomp_rule():
    # include/cedar/decomp.h:59-86
    assert d2.grid_decomp((16385, 16385), 8) == [4, 2]
    assert d2.grid_decomp((16385, 16385), 4) == [2, 2]
    assert d2.grid_decomp((16385, 16385), 2) == [2, 1]
    # block_partition::low (include/cedar/mpi/block_partition.h:14-17) on 2^k+1 grids is already aligned
    ext = [d2.geometry(16385, 16385, 4, 2, 5, 15, 0, b) for b in range(4)]
    assert [e["x0"] for e in ext] == [i * 16385 // 4 for i in range(4)]
    assert [e["nx"] for e in ext] == [4096, 4096, 4096, 4097]
    # the pieces tile the grid on every distributed level and start on coarse points
    for nx, ny, px, py, nd in [(100, 77, 2, 2, 2), (200, 131, 3, 2, 2), (129, 129, 2, 2, 2)]:
        for lvl in range(nd + 1):
            gx, gy = nx, ny
            for _ in range(lvl):
                gx, gy = (gx - 1) // 2 + 1, (gy - 1) // 2 + 1
            cov = np.zeros((gy, gx), dtype=int)
            for b in range(px * py):
                g = d2.geometry(nx, ny, px, py, nd, 5, lvl, b)
                if lvl < nd:
                    assert g["x0"] % 2 == 0 and g["y0"] % 2 == 0
                cov[g["y0"]:g["y0"] + g["ny"], g["x0"]:g["x0"] + g["nx"]] += 1
            assert (cov == 1).all()


def count(plan, *names):
    return sum(o["op"] in names for o in plan)


def test_plan_one_exchange_per_level():
    """W = 15: a V(1,1) cycle of the C5 grid needs the start exchange, one per coarser distributed level and the gather"""
    plan, nd = d2.plan(16385, 16385, 4, 2, pre=1, post=1, agglomerate=513, halo=15, split_min=0)
    assert nd == 5
    assert count(plan, "exchange") == 5 and count(plan, "gather") == 1 and count(plan, "sync") == 0
    assert count(plan, "sweep") == 10 and count(plan, "residual") == 5
    # split_min 512: on the levels with blocks >= 512 wide the exchange runs beside the interior of the sweep that needs it
    ov, _ = d2.plan(16385, 16385, 4, 2, pre=1, post=1, agglomerate=513, halo=15, split_min=512)
    assert count(ov, "exchange") == 5 and count(ov, "join") == 4 and sum(o["async"] for o in ov) == 4
    assert count(ov, "sweep") == 14 and sum(o["part"] == 1 for o in ov) == 4 and sum(o["part"] == 2 for o in ov) == 4
    for k, o in enumerate(ov):
        if o["op"] == "exchange" and o["async"]:
            assert ov[k + 1]["op"] == "sweep" and ov[k + 1]["part"] == 1 and ov[k + 2]["op"] == "join" and ov[k + 3]["part"] == 2
    # narrower halos cost more exchanges, never a wrong result (the simulator tests below)
    p7, _ = d2.plan(16385, 16385, 4, 2, pre=1, post=1, agglomerate=513, halo=7, split_min=0)
    p5, _ = d2.plan(16385, 16385, 4, 2, pre=1, post=1, agglomerate=513, halo=5, split_min=0)
    assert count(plan, "exchange") < count(p7, "exchange") <= count(p5, "exchange")
    # the first operation of every list waits for all ranks
    for what in ("setup", "cycle", "norm"):
        p, _ = d2.plan(16385, 16385, 4, 2, what=what)
        assert p[0]["op"] == "exchange" and p[0]["wait_all"]


def test_plan_never_overwrites_a_pulled_array_too_early():
    """an array read by the neighbours in exchange s is written again only after a later exchange"""
    for halo in (5, 7, 9, 15):
        for pre, post in ((1, 1), (2, 1), (0, 1), (2, 2), (3, 0)):
          for split in (0, 100):
            plan, nd = d2.plan(1025, 1025, 2, 2, pre=pre, post=post, agglomerate=65, halo=halo, split_min=split)
            step, pulled = 0, {}
            for o in plan:
                if o["op"] == "join":
                    continue
                if o["op"] in ("exchange", "sync", "gather", "gather_op", "norm"):
                    step += 1
                    for a in o["arrays"]:
                        pulled[o["lvl"], a] = step
                    if o["op"] == "gather":
                        pulled[o["lvl"], B] = step
                    continue
                writes = {"sweep": [(o["lvl"], o["b"])], "residual": [(o["lvl"], 3)], "restrict": [(o["lvl"] + 1, B)],
                          "zero": [(o["lvl"], o["a"])], "scatter": [(o["lvl"], o["a"])],
                          "interp": [(o["lvl"], o["a"]), (o["lvl"], 3)], "copy": [(o["lvl"], o["b"])]}.get(o["op"], [])
                for w in writes:
                    assert pulled.get(w, -1) < step, (halo, pre, post, o, w)


def test_halo_width_is_checked():
    import cedar_b200.solver as cs
    for halo in (3, 4, 6):
        with pytest.raises(cs.CedarError):
            d2.plan(1025, 1025, 2, 2, halo=halo)


def blocks_of(nx, ny, px, py, nd, halo):
    return {b: tuple(d2.geometry(nx, ny, px, py, nd, halo, 0, b)[k] for k in ("x0", "y0", "nx", "ny")) for b in range(px * py)}


def join(parts, ext, nx, ny):
    first = next(iter(parts.values()))
    g = np.zeros(first.shape[:-2] + (ny + 2, nx + 2))
    for b, t in parts.items():
        x0, y0, nxl, nyl = ext[b]
        g[..., 1 + y0:1 + y0 + nyl, 1 + x0:1 + x0 + nxl] = t[..., 1:-1, 1:-1]
    return g


CASES = [
    # nx, ny, px, py, nine, agglomerate, halo, pre, post, split_min (0 = sweeps never split)
    (129, 129, 2, 2, True, 17, 15, 1, 1, 0),
    (129, 129, 2, 2, True, 17, 5, 1, 1, 0),
    (129, 97, 2, 1, True, 17, 7, 2, 1, 0),
    (257, 129, 4, 2, True, 33, 15, 1, 1, 0),
    (200, 131, 3, 2, True, 20, 9, 1, 2, 0),
    (129, 129, 2, 2, False, 17, 15, 1, 1, 0),
    (150, 100, 2, 2, False, 20, 5, 2, 2, 0),
    (129, 129, 1, 2, True, 33, 5, 3, 2, 0),
    (161, 161, 2, 2, True, 17, 11, 0, 2, 0),
    # the exchange beside the interior of the sweep that needs it (delivered as late as allowed in the simulator)
    (129, 129, 2, 2, True, 17, 15, 1, 1, 40),
    (257, 257, 2, 2, True, 17, 5, 1, 1, 20),
    (257, 129, 4, 2, True, 33, 7, 2, 1, 20),
    (300, 200, 3, 2, False, 20, 5, 1, 2, 20),
    (257, 193, 2, 1, True, 17, 9, 0, 2, 30),
    (257, 193, 1, 2, True, 17, 9, 2, 2, 30),
]


@pytest.mark.parametrize("nx,ny,px,py,nine,agg,halo,pre,post,split", CASES)
def test_plan_blocks_equal_serial(nx, ny, px, py, nine, agg, halo, pre, post, split):
    rng = np.random.default_rng(nx + 3 * ny + nine + halo)
    so = random_op2(nx, ny, nine, rng)
    b = random_gf(nx, ny, rng)
    x0 = random_gf(nx, ny, rng)
    ref = oracle.Solver2(so, "point", pre, post)
    _, nd = d2.plan(nx, ny, px, py, nst=so.shape[0], pre=pre, post=post, agglomerate=agg, halo=halo)
    ext = blocks_of(nx, ny, px, py, nd, halo)
    sim = PlanSim(nx, ny, px, py, d2.split_blocks(so, ext), pre, post, agg, halo, split_min=split)
    assert sim.nlevels == ref.nlevels and sim.nd == nd and nd >= 1
    joins = sum(o["op"] == "join" for o in sim.p_cycle)
    assert joins == 0 if split == 0 else (joins > 0 or pre == 0)
    # hierarchy: own points of every distributed level's operator and interpolation
    for lvl in range(1, nd + 1):
        gx, gy = sim.gn[lvl]
        extl = {bb: tuple(sim.g[lvl, bb][k] for k in ("x0", "y0", "nx", "ny")) for bb in range(px * py)}
        for slot, name in ((A, "A"), (PI, "P")):
            mine = {bb: np.pad(sim._own(lvl, bb, slot), ((0, 0), (1, 1), (1, 1))) for bb in range(px * py)}
            want = ref.level_array(lvl, name).copy()
            want[:, 0, :] = 0; want[:, -1, :] = 0; want[:, :, 0] = 0; want[:, :, -1] = 0
            assert np.array_equal(join(mine, extl, gx, gy), want), (lvl, name)
    # cycles from a non-zero initial guess
    xr = x0.copy()
    for bb, v in d2.split_blocks(b, ext).items():
        sim.set(bb, B, v)
    for bb, v in d2.split_blocks(x0, ext).items():
        sim.set(bb, X0, v)
    for it in range(3):
        ref.vcycle(xr, b)
        sim.vcycle()
        x = join({bb: sim.get(bb) for bb in range(px * py)}, ext, nx, ny)
        assert np.array_equal(x, xr), f"cycle {it}"
    # residual norm: rank-ordered sum of the blocks' sums of squares
    r = np.zeros_like(b)
    oracle.lib().orc2_residual(0, oracle.P(so), oracle.P(b), oracle.P(xr), oracle.P(r), nx + 2, ny + 2, 0, 0 if nine else 1,
                               so.shape[0])
    assert abs(sim.norm() - np.sqrt((r[1:-1, 1:-1] ** 2).sum())) <= 1e-12 * np.sqrt((r ** 2).sum())


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nx, ny, px, py, nine, agg, halo, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        oracle.lib().orc_set_threads(1)
        rng = np.random.default_rng(nx + ny + nine)
        so = random_op2(nx, ny, nine, rng)
        b = random_gf(nx, ny, rng)
        _, nd = d2.plan(nx, ny, px, py, nst=so.shape[0], pre=2, post=1, agglomerate=agg, halo=halo)
        ext = blocks_of(nx, ny, px, py, nd, halo)
        mine = {rank: ext[rank]}
        sim = PlanSim(nx, ny, px, py, d2.split_blocks(so, mine), 2, 1, agg, halo, dist=dist)
        sim.set(rank, B, d2.split_blocks(b, mine)[rank])
        norms = [sim.norm()]
        for _ in range(3):
            sim.vcycle()
            norms.append(sim.norm())
        np.save(os.path.join(out, f"x{rank}.npy"), sim.get(rank))
        np.save(os.path.join(out, f"h{rank}.npy"), np.array(norms))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("nx,ny,px,py,nine,agg,halo", [(129, 97, 2, 1, True, 17, 15), (129, 129, 2, 2, False, 17, 7)])
def test_plan_gloo_ranks_equal_serial(tmp_path, nx, ny, px, py, nine, agg, halo):
    world = px * py
    mp.spawn(_worker, args=(world, _free_port(), nx, ny, px, py, nine, agg, halo, str(tmp_path)), nprocs=world, join=True)
    rng = np.random.default_rng(nx + ny + nine)
    so = random_op2(nx, ny, nine, rng)
    b = random_gf(nx, ny, rng)
    ref = oracle.Solver2(so, "point", 2, 1)
    xr, hr, r0r = ref.solve(b, maxiter=3, tol=0.0)
    _, nd = d2.plan(nx, ny, px, py, nst=so.shape[0], pre=2, post=1, agglomerate=agg, halo=halo)
    ext = blocks_of(nx, ny, px, py, nd, halo)
    x = join({r: np.load(tmp_path / f"x{r}.npy") for r in range(world)}, ext, nx, ny)
    assert np.array_equal(x, xr)
    hs = [np.load(tmp_path / f"h{r}.npy") for r in range(world)]
    for h in hs[1:]:
        assert np.array_equal(h, hs[0])          # every rank sees the same bits
    assert abs(hs[0][0] - r0r) <= 1e-12 * r0r and np.allclose(hs[0][1:] / hs[0][0], hr, rtol=1e-10, atol=0)


# ---- line relaxation across block cuts (plan.h line_group: lines transposed between the blocks that share them) ----
LINE_CASES = [
    # nx, ny, px, py, nine, agglomerate, halo, pre, post, relaxation
    (129, 129, 2, 2, True, 17, 5, 1, 1, "line-x"),
    (129, 129, 2, 2, True, 17, 5, 1, 1, "line-y"),
    (129, 129, 2, 2, True, 17, 5, 1, 1, "line-xy"),
    (129, 97, 2, 1, False, 17, 7, 2, 1, "line-xy"),
    (200, 131, 3, 2, True, 20, 9, 1, 2, "line-x"),
    (161, 129, 1, 3, False, 17, 5, 1, 1, "line-y"),
    (257, 129, 4, 2, True, 33, 15, 0, 2, "line-xy"),
]


@pytest.mark.parametrize("nx,ny,px,py,nine,agg,halo,pre,post,relaxation", LINE_CASES)
def test_plan_line_relaxation_equals_serial(nx, ny, px, py, nine, agg, halo, pre, post, relaxation):
    rng = np.random.default_rng(nx + 5 * ny + nine + halo)
    so = random_op2(nx, ny, nine, rng)
    b = random_gf(nx, ny, rng)
    x0 = random_gf(nx, ny, rng)
    ref = oracle.Solver2(so, relaxation, pre, post)
    cyc, nd = d2.plan(nx, ny, px, py, nst=so.shape[0], pre=pre, post=post, agglomerate=agg, halo=halo, relaxation=relaxation)
    ext = blocks_of(nx, ny, px, py, nd, halo)
    sim = PlanSim(nx, ny, px, py, d2.split_blocks(so, ext), pre, post, agg, halo, relaxation=relaxation)
    assert sim.nlevels == ref.nlevels and sim.nd == nd and nd >= 1
    # per colour: right-hand sides, a synchronisation of all ranks, the solve, an exchange that waits for all ranks
    kinds = [o["op"] for o in cyc]
    ndir = 2 if relaxation == "line-xy" else 1
    assert kinds.count("lines_rhs") == kinds.count("lines_solve") == nd * (pre + post) * 2 * ndir
    for k, o in enumerate(cyc):
        if o["op"] == "lines_solve":
            assert cyc[k - 1]["op"] == "sync" and cyc[k - 2]["op"] == "lines_rhs"
            assert cyc[k + 1]["op"] == "exchange" and cyc[k + 1]["wait_all"] and cyc[k + 1]["arrays"] == [o["a"]]
    assert "sweep" not in kinds
    xr = x0.copy()
    for bb, v in d2.split_blocks(b, ext).items():
        sim.set(bb, B, v)
    for bb, v in d2.split_blocks(x0, ext).items():
        sim.set(bb, X0, v)
    for it in range(3):
        ref.vcycle(xr, b)
        sim.vcycle()
        x = join({bb: sim.get(bb) for bb in range(px * py)}, ext, nx, ny)
        assert np.array_equal(x, xr), f"cycle {it}"


def _line_worker(rank, world, port, nx, ny, px, py, nine, agg, halo, relaxation, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        oracle.lib().orc_set_threads(1)
        rng = np.random.default_rng(nx + ny + nine)
        so = random_op2(nx, ny, nine, rng)
        b = random_gf(nx, ny, rng)
        _, nd = d2.plan(nx, ny, px, py, nst=so.shape[0], agglomerate=agg, halo=halo, relaxation=relaxation)
        ext = blocks_of(nx, ny, px, py, nd, halo)
        mine = {rank: ext[rank]}
        sim = PlanSim(nx, ny, px, py, d2.split_blocks(so, mine), 1, 1, agg, halo, dist=dist, relaxation=relaxation)
        sim.set(rank, B, d2.split_blocks(b, mine)[rank])
        for _ in range(2):
            sim.vcycle()
        np.save(os.path.join(out, f"x{rank}.npy"), sim.get(rank))
    finally:
        dist.destroy_process_group()


def test_plan_line_relaxation_gloo_ranks(tmp_path):
    nx, ny, px, py, nine, agg, halo, relaxation = 129, 97, 2, 1, True, 17, 5, "line-xy"
    world = px * py
    mp.spawn(_line_worker, args=(world, _free_port(), nx, ny, px, py, nine, agg, halo, relaxation, str(tmp_path)), nprocs=world,
             join=True)
    rng = np.random.default_rng(nx + ny + nine)
    so = random_op2(nx, ny, nine, rng)
    b = random_gf(nx, ny, rng)
    ref = oracle.Solver2(so, relaxation, 1, 1)
    xr = np.zeros_like(b)
    for _ in range(2):
        ref.vcycle(xr, b)
    _, nd = d2.plan(nx, ny, px, py, nst=so.shape[0], agglomerate=agg, halo=halo, relaxation=relaxation)
    ext = blocks_of(nx, ny, px, py, nd, halo)
    x = join({r: np.load(tmp_path / f"x{r}.npy") for r in range(world)}, ext, nx, ny)
    assert np.array_equal(x, xr)
