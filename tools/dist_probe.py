#!/usr/bin/env python
"""Where the time of a block-decomposed V-cycle goes: per-operation device times (eager, host-synchronised per
operation) next to the graph-replayed cycle time.  One process (all blocks on one GPU; --blocks 1x1 = the executor
alone against the serial solver) or one rank per GPU under torchrun."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=4097, help="own unknowns per block and dimension (+1 on the last block)")
    ap.add_argument("--blocks", default="1x1")
    ap.add_argument("--halo", type=int, default=15)
    ap.add_argument("--agglomerate", type=int, default=513)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--relaxation", default="point")
    ap.add_argument("--global-n", type=int, default=0, help="global grid global-n^2 instead of n-1 own unknowns per block")
    ap.add_argument("--decomp", default="", help="px,py under torchrun (default: the grid_decomp rule)")
    a = ap.parse_args()
    import torch
    import cedar_b200.dist2 as d2
    import cedar_b200.solver as cs
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    comm = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group("gloo")
        comm = d2.TorchComm()
        px, py = [int(v) for v in a.decomp.split(",")] if a.decomp else d2.grid_decomp((a.n - 1, a.n - 1), world)
    else:
        px, py = [int(v) for v in a.blocks.split("x")]
    nb = a.n - 1
    gnx, gny = (px * nb + 1, py * nb + 1) if px * py > 1 else (a.n, a.n)
    if a.global_n:
        gnx = gny = a.global_n
    sol = d2.DistSolver2(gnx, gny, (px, py), agglomerate=a.agglomerate, halo=a.halo, comm=comm, relaxation=a.relaxation)
    for k in sol.owned:
        ext = sol.extent(0, k)
        sol.set_operator(k, d2.gallery_fe_block((gnx, gny), ext))
        sol.upload(k, "b", d2.rhs_sin_block((gnx, gny), ext))
    sol.setup()
    ms = sol.time_vcycles(4, a.reps)
    prof = sol.profile_ops(5)
    if rank == 0:
        agg = {}
        for o, t in prof:
            key = (o["op"] + (":xy"[1 + o["dir"]] if o["op"].startswith("lines") else ""), o["lvl"])
            agg.setdefault(key, 0.0)
            agg[key] += t
        print(json.dumps({"grid": f"{gnx}x{gny}", "blocks": f"{px}x{py}", "ranks": world, "halo": a.halo, "relaxation": a.relaxation,
                          "ms_per_cycle_graph": ms,
                          "sum_of_ops_ms": sum(t for _, t in prof), "exchanges": sol.exchanges_per_cycle,
                          "ops": [{"op": k[0], "lvl": k[1], "ms": round(v, 4)} for k, v in agg.items()]}))
    if world == 1 and px * py == 1:
        ser = cs.Solver2(cs.Operator2(gallery="fe", nx=a.n, ny=a.n), relaxation="point", nrelax_pre=1, nrelax_post=1)
        print(json.dumps({"serial_solver_ms_per_cycle": ser.time_vcycles(d2.rhs_sin_block((a.n, a.n), (0, 0, a.n, a.n)), 4, a.reps)}))
    sol.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
