#!/usr/bin/env python
"""BASELINE configs[2]: 2D 5-point anisotropic diffusion (gallery::diag_diffusion, dx = 1, dy = 1e-4) 8193^2 with
zebra line relaxation, V(1,1): the serial cdr2::solver on one GPU, or block-decomposed over the N GPUs of a torchrun
launch with the lines that cross the block cuts transposed over peer memory (include/cedar/2d/mpi/plan.h line_group).
One JSON line per relaxation: ms per cycle (CUDA events, max over ranks), cycles to 1e-10, and -- decomposed -- whether
every rank's block equals the serial GPU solver's bit for bit (checked on a grid of a quarter the size)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rhs(n):
    rng = np.random.default_rng(1)
    b = np.zeros((n + 2, n + 2))
    b[1:-1, 1:-1] = rng.standard_normal((n, n)) * 1e-3
    return b


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=8193)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--halo", type=int, default=5)
    ap.add_argument("--agglomerate", type=int, default=513)
    ap.add_argument("--relaxations", default="line-x,line-y,line-xy")
    ap.add_argument("--decomp", default="", help="px,py (default: the reference's grid_decomp rule)")
    a = ap.parse_args()
    import torch
    import cedar_b200.solver as cs
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    n = a.n
    mk = lambda m: cs.Operator2(gallery="diag_diffusion", nx=m, ny=m, dx=1.0, dy=1e-4)
    if world == 1:
        for relax in a.relaxations.split(","):
            op = mk(n)
            torch.cuda.synchronize(); t0 = time.perf_counter()
            sol = cs.Solver2(op, relaxation=relax, nrelax_pre=1, nrelax_post=1, maxiter=30, tol=1e-10)
            torch.cuda.synchronize(); setup = time.perf_counter() - t0
            b = rhs(n)
            ms = sol.time_vcycles(b, warmup=3, n=a.steps)
            _, h, _ = sol.solve(b)
            print(json.dumps({"cfg": f"C3 diag_diffusion {n}^2 {relax} V(1,1)", "n_gpus": 1, "ms_per_cycle": ms,
                              "mdof_s": n * n / ms / 1e3, "setup_s": setup, "iters_1e-10": len(h), "final_rel_residual": float(h[-1])}),
                  flush=True)
            del sol, op
        return
    import torch.distributed as dist
    import cedar_b200.dist2 as d2
    dist.init_process_group("gloo")
    comm = d2.TorchComm()
    px, py = [int(v) for v in a.decomp.split(",")] if a.decomp else d2.grid_decomp((n, n), world)

    def problem(m, relax, agg):
        op = mk(m)                        # keep the operator alive while its host array is copied
        so = np.array(op.array())
        del op
        sol = d2.DistSolver2(m, m, (px, py), nst=3, nrelax_pre=1, nrelax_post=1, agglomerate=agg, halo=a.halo, comm=comm, relaxation=relax)
        ext = {rank: sol.extent(0, rank)}
        torch.cuda.synchronize(); t0 = time.perf_counter()
        sol.set_operator(rank, d2.split_blocks(so, ext)[rank])
        sol.setup()
        torch.cuda.synchronize()
        return sol, ext, so, time.perf_counter() - t0

    for relax in a.relaxations.split(","):
        # parity on a grid the serial solver finishes quickly
        m = (n - 1) // 4 + 1
        sol, ext, so, _ = problem(m, relax, 129)
        b = rhs(m)
        bl = d2.split_blocks(b, ext)[rank]
        sol.upload(rank, "b", bl); sol.upload(rank, "x", np.zeros_like(bl))
        h, _ = sol.solve(maxiter=6, tol=1e-30)
        x = sol.download(rank)
        xs, hs, _ = cs.Solver2(cs.Operator2(so), relaxation=relax, nrelax_pre=1, nrelax_post=1, maxiter=6, tol=1e-30).solve(b)
        x0, y0, nxl, nyl = ext[rank]
        flag = torch.tensor([int(np.array_equal(x[1:-1, 1:-1], xs[1 + y0:1 + y0 + nyl, 1 + x0:1 + x0 + nxl]))])
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        sol.close()
        # the timed problem
        sol, ext, so, setup = problem(n, relax, a.agglomerate)
        del so
        b = rhs(n)
        bl = d2.split_blocks(b, ext)[rank]
        sol.upload(rank, "b", bl); sol.upload(rank, "x", np.zeros_like(bl))
        comm.barrier()
        ms = torch.tensor([sol.time_vcycles(3, a.steps)])
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        sol.upload(rank, "x", np.zeros_like(bl))
        h, _ = sol.solve(maxiter=30, tol=1e-10)
        if rank == 0:
            print(json.dumps({"cfg": f"C3 diag_diffusion {n}^2 {relax} V(1,1)", "n_gpus": world, "decomposition": f"{px}x{py}",
                              "halo": a.halo, "ms_per_cycle": float(ms.item()), "mdof_s": n * n / float(ms.item()) / 1e3,
                              "setup_s": setup, "iters_1e-10": len(h), "final_rel_residual": float(h[-1]),
                              "exchanges_per_cycle": sol.exchanges_per_cycle,
                              "parity_vs_serial": "bitwise" if int(flag.item()) else "DIFFERS", "parity_grid": f"{m}^2"}), flush=True)
        comm.barrier()
        sol.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
