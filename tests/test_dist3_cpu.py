"""CPU tests of the slab-decomposed 3D path's host logic: the product's operation lists
(cdr3_dist_plan_json: planner of include/cedar/2d/mpi/plan.h with colour depth 2 across planes, slabs of
include/cedar/3d/mpi/plan.h) run with the CPU oracle's 3D kernels (tests/dist3_plan_sim.py) must reproduce the serial
oracle solver BITWISE -- hierarchy and cycles; halos start as NaN."""
import numpy as np
import pytest

import oracle
import cedar_b200.dist3 as d3
from tests.dist3_plan_sim import PlanSim3, X0, B, A, PI
from tests.util import random_gf3, random_op3


def count(plan, *names):
    return sum(o["op"] in names for o in plan)


def test_plan_of_the_c4_grid():
    """513^3 on 8 slabs: levels 513 and 257 distributed, <= 129 planes replicated; V(1,1) with 5 halo planes"""
    for nst in (4, 14):
        plan, nd = d3.plan(513, 513, 513, 8, nst=nst, pre=1, post=1, agglomerate=129, halo=5)
        assert nd == 2
        assert count(plan, "gather") == 1 and count(plan, "sweep") == 4 and count(plan, "residual") == 2
        assert count(plan, "exchange") <= 6
        assert plan[0]["op"] == "exchange" and plan[0]["wait_all"]
        # the 27-point sweeps work in place, the 7-point fine sweep from one array into the other
        fine = [o for o in plan if o["op"] == "sweep" and o["lvl"] == 0]
        assert all((o["a"] == o["b"]) == (nst == 14) for o in fine)
        assert all(o["a"] == o["b"] for o in plan if o["op"] == "sweep" and o["lvl"] > 0)
    g = [d3.geometry(513, 513, 513, 8, 2, 5, 0, b) for b in range(8)]
    assert [v["z0"] for v in g] == [64 * b for b in range(8)] and [v["nz"] for v in g] == [64] * 7 + [65]


CASES = [
    # nx, ny, nz, slabs, 27-point?, agglomerate, halo, pre, post
    (17, 13, 65, 2, False, 17, 5, 1, 1),
    (17, 13, 65, 2, True, 17, 5, 1, 1),
    (12, 21, 129, 4, False, 17, 5, 2, 1),
    (9, 9, 129, 4, True, 17, 3, 1, 1),
    (13, 10, 97, 3, True, 20, 7, 1, 2),
    (9, 17, 65, 2, False, 9, 5, 0, 2),
]


@pytest.mark.parametrize("nx,ny,nz,np_,full,agg,halo,pre,post", CASES)
def test_plan_slabs_equal_serial(nx, ny, nz, np_, full, agg, halo, pre, post):
    rng = np.random.default_rng(nx + 3 * ny + 7 * nz + full)
    so = random_op3(nx, ny, nz, full, rng)
    b = random_gf3(nx, ny, nz, rng)
    x0 = random_gf3(nx, ny, nz, rng)
    ref = oracle.Solver3(so, pre, post)
    _, nd = d3.plan(nx, ny, nz, np_, nst=so.shape[0], pre=pre, post=post, agglomerate=agg, halo=halo)
    ext = {k: tuple(d3.geometry(nx, ny, nz, np_, nd, halo, 0, k)[v] for v in ("nx", "ny", "z0", "nz")) for k in range(np_)}
    sim = PlanSim3(nx, ny, nz, np_, d3.split_slabs(so, ext), pre, post, agg, halo)
    assert sim.nlevels == ref.nlevels and sim.nd == nd >= 1
    for lvl in range(1, nd + 1):
        for slot, name in ((A, "A"), (PI, "P")):
            want = ref.level_array(lvl, name)
            for k in range(np_):
                g = sim.g[lvl, k]
                got = sim._own(lvl, k, slot)[:, :, 1:-1, 1:-1]
                assert np.array_equal(got, want[:, 1 + g["z0"]:1 + g["z0"] + g["nz"], 1:-1, 1:-1]), (lvl, name, k)
    xr = x0.copy()
    for k, v in d3.split_slabs(b, ext).items():
        sim.set(k, B, v)
    for k, v in d3.split_slabs(x0, ext).items():
        sim.set(k, X0, v)
    for it in range(3):
        ref.vcycle(xr, b)
        sim.vcycle()
        for k in range(np_):
            _, _, z0, nzl = ext[k]
            assert np.array_equal(sim.get(k)[1:-1], xr[1 + z0:1 + z0 + nzl]), f"cycle {it}, slab {k}"
