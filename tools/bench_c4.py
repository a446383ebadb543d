#!/usr/bin/env python
"""BASELINE configs[3]: 3D 7-point / 27-point Poisson 513^3, point relaxation, V(1,1), on 1 GPU (serial cdr3::solver)
or slab-decomposed over the N GPUs of a torchrun launch (cdr3_dist_*): ms per cycle, MDOF/s, fraction of the per-GPU
whole-cycle roofline (335 / 575 algorithmic B per point), solve to 1e-10.  One JSON line per operator."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rhs_slab(n, z0, nzl):
    """examples/basic-3d-ser/poisson.cc:16-38 on the planes z0 .. z0+nzl-1 (one ghost plane below and above)"""
    h = 1.0 / (n + 1)
    c = np.sin(2 * np.pi * h * np.arange(1, n + 1))
    z = np.sin(2 * np.pi * h * (z0 + 1 + np.arange(nzl)))
    out = np.zeros((nzl + 2, n + 2, n + 2))
    out[1:-1, 1:-1, 1:-1] = 12 * np.pi ** 2 * h ** 3 * z[:, None, None] * c[None, :, None] * c[None, None, :]
    return out


def op_slab(name, n, z0, nzl):
    """own planes of gallery::poisson (7-point, src/3d/gallery.cc:7-56) / gallery::fe (27-point, :111-190)"""
    import cedar_b200.solver as cs
    if name == "poisson":
        # the same expressions as gallery::poisson (cedar_b200/host/grid3.cc:92-109) with hx = hy = hz = 1/(n+1)
        h = 1.0 / (n + 1)
        xh = yh = zh = h * h / h
        so = np.zeros((4, nzl + 2, n + 2, n + 2))
        so[0, 1:-1, 1:-1, 1:-1] = 2.0 * xh + 2.0 * yh + 2.0 * zh
        so[1, 1:-1, 1:-1, 2:-1] = 1.0 * xh
        so[2, 1:-1, 2:-1, 1:-1] = 1.0 * yh
        so[3, (2 if z0 == 0 else 1):-1, 1:-1, 1:-1] = 1.0 * zh
        return so
    op = cs.Operator3(gallery=name, nx=n, ny=n, nz=3)     # fe: constant entries; interior planes are identical ...
    a = np.array(op.array())
    so = np.zeros((a.shape[0], nzl + 2, n + 2, n + 2))
    so[:, 1:-1] = a[:, 2][:, None]
    if z0 == 0:                                           # ... except that the first plane has no couplings downwards
        so[:, 1] = a[:, 1]
    return so


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=513)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--halo", type=int, default=5)
    ap.add_argument("--agglomerate", type=int, default=129)
    a = ap.parse_args()
    import torch
    import cedar_b200.solver as cs
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", peak))
    except OSError:
        pass
    n = a.n
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("gloo")
        import cedar_b200.dist3 as d3
    for name, bpp in (("poisson", 335.0), ("fe", 575.0)):
        if world == 1:
            op = cs.Operator3(gallery=name, nx=n, ny=n, nz=n)
            t0 = time.perf_counter()
            sol = cs.Solver3(op, relaxation="point", nrelax_pre=1, nrelax_post=1, maxiter=30, tol=1e-10)
            torch.cuda.synchronize()
            setup = time.perf_counter() - t0
            # the same again with the operator already resident on the device: what the first number spends on the
            # host -> device copy of the fine operator (pageable memory; 4 / 14 planes of (n+2)^3 doubles)
            del sol
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            sol = cs.Solver3(op, relaxation="point", nrelax_pre=1, nrelax_post=1, maxiter=30, tol=1e-10)
            torch.cuda.synchronize()
            setup_resident = time.perf_counter() - t0
            b = rhs_slab(n, 0, n)
            ms = sol.time_vcycles(b, 3, a.steps)
            ms_solve, it = sol.time_solves(b, 0, 1)
            ex = 0
            del sol, op
        else:
            t0 = time.perf_counter()
            sol = d3.DistSolver3(n, n, n, world, nst=4 if name == "poisson" else 14, agglomerate=a.agglomerate, halo=a.halo,
                                 comm=d3.TorchComm())
            _, _, z0, nzl = sol.extent(0, rank)
            t1 = time.perf_counter()
            so = op_slab(name, n, z0, nzl)          # host-side generation of the slab's operator: not part of the setup
            t_gen = time.perf_counter() - t1
            sol.set_operator(rank, so)
            del so
            sol.setup()
            torch.cuda.synchronize()
            setup = time.perf_counter() - t0 - t_gen
            bl = rhs_slab(n, z0, nzl)
            sol.upload(rank, "b", bl)
            sol.time_vcycles(3, 1)
            dist.barrier()
            ms = sol.time_vcycles(0, a.steps)
            sol.upload(rank, "x", np.zeros_like(bl))
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            h, _ = sol.solve(maxiter=30, tol=1e-10)
            ms_solve, it = (time.perf_counter() - t0) * 1e3, len(h)
            ex = sol.exchanges_per_cycle
            t = torch.tensor([ms, ms_solve, setup], dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, ms_solve, setup = [float(v) for v in t.tolist()]
            dist.barrier()
            sol.close()
        if rank == 0:
            print(json.dumps({"workload": f"3D {'7' if name == 'poisson' else '27'}-point {name} {n}^3, point relaxation, V(1,1)",
                              "n_gpus": world, "decomposition": "serial" if world == 1 else f"{world} slabs of xy planes, halo {a.halo}",
                              "ms_per_cycle": ms, "mdof_per_s": n ** 3 / (ms * 1e-3) / 1e6,
                              "vcycle_roofline_frac_per_gpu": bpp * n ** 3 / (ms * 1e-3) / 1e9 / world / peak,
                              "ghost_exchanges_per_cycle": ex, "setup_s": setup,
                              **({"setup_operator_resident_s": setup_resident} if world == 1 else {}),
                              "solve": {"tol": 1e-10, "iterations": it, "ms": ms_solve}}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
