"""One rank per GPU (torchrun): the slab-decomposed 3D solver must reproduce the serial GPU solver BITWISE on a grid
every rank can also solve alone.  python -m torch.distributed.run --nproc-per-node 2 tests/run_dist3_ranks.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    import cedar_b200.dist3 as d3
    import cedar_b200.solver as cs
    from tests.util import random_gf3, random_op3
    ok = True
    for (nx, ny, nz, full, agg, halo, pre, post) in [(65, 49, 32 * world + 1, False, 33, 5, 1, 1), (49, 65, 32 * world + 1, True, 33, 5, 1, 1)]:
        rng = np.random.default_rng(nx + ny + nz)
        so = random_op3(nx, ny, nz, full, rng)
        b = random_gf3(nx, ny, nz, rng)
        sol = d3.DistSolver3(nx, ny, nz, world, nst=so.shape[0], nrelax_pre=pre, nrelax_post=post, agglomerate=agg, halo=halo,
                             comm=d3.TorchComm())
        ext = {rank: sol.extent(0, rank)}
        sol.set_operator(rank, d3.split_slabs(so, ext)[rank])
        sol.setup()
        bl = d3.split_slabs(b, ext)[rank]
        sol.upload(rank, "b", bl)
        sol.upload(rank, "x", np.zeros_like(bl))
        h, r0 = sol.solve(maxiter=6, tol=1e-30)
        x = sol.download(rank)
        ser = cs.Solver3(cs.Operator3(so), relaxation="point", nrelax_pre=pre, nrelax_post=post, maxiter=6, tol=1e-30)
        xs, hs, r0s = ser.solve(b)
        _, _, z0, nzl = ext[rank]
        same = np.array_equal(x[1:-1], xs[1 + z0:1 + z0 + nzl])
        hist = len(h) == len(hs) and np.allclose(h, hs, rtol=1e-10, atol=0)
        print(f"rank {rank}: {nx}x{ny}x{nz} {'27' if full else '7'}-point, {world} slabs, halo {halo}: bitwise {same}, history {hist}, "
              f"{sol.exchanges_per_cycle} exchanges/cycle", flush=True)
        ok = ok and same and hist
        dist.barrier()
        sol.close()
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0 and int(flag.item()) == 1:
        print("DIST3 RANKS OK", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
