"""One rank per GPU (torchrun): the peer-memory block-decomposed solver must reproduce the serial GPU solver
BITWISE on a grid rank 0 can also solve alone.  Launched by tests/test_gpu_dist2.py when several GPUs are
visible, and by hand:  python -m torch.distributed.run --nproc-per-node 2 tests/run_dist2_ranks.py"""
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
    dist.init_process_group("gloo")          # host all-gather of the IPC handles only; the data path is peer memory
    import cedar_b200.dist2 as d2
    import cedar_b200.solver as cs
    from tests.util import random_gf, random_op2
    ok = True
    for (nx, ny, nine, agg, halo, pre, post, split, relaxation) in [
            (1025, 769, True, 65, 15, 1, 1, 0, "point"), (700, 900, False, 65, 7, 2, 1, 0, "point"),
            (2049, 1537, True, 129, 15, 1, 1, 128, "point"), (1400, 1800, True, 65, 9, 2, 2, 100, "point"),
            # lines across the block cuts: transposed over peer memory, solved whole (plan.h line_group)
            (1025, 769, True, 65, 5, 1, 1, 0, "line-xy"), (1100, 900, False, 65, 5, 2, 1, 0, "line-x"),
            (2049, 1537, True, 129, 7, 1, 1, 0, "line-y")]:
        px, py = d2.grid_decomp((nx, ny), world)
        rng = np.random.default_rng(nx + ny)
        so = random_op2(nx, ny, nine, rng)
        b = random_gf(nx, ny, rng)
        sol = d2.DistSolver2(nx, ny, (px, py), nst=so.shape[0], nrelax_pre=pre, nrelax_post=post, agglomerate=agg, halo=halo,
                             comm=d2.TorchComm(), split_min=split, relaxation=relaxation)
        ext = {rank: sol.extent(0, rank)}
        sol.set_operator(rank, d2.split_blocks(so, ext)[rank])
        sol.setup()
        bl = d2.split_blocks(b, ext)[rank]
        sol.upload(rank, "b", bl)
        sol.upload(rank, "x", np.zeros_like(bl))
        h, r0 = sol.solve(maxiter=8, tol=1e-30)       # 2 eager cycles, capture, replays
        x = sol.download(rank)
        ser = cs.Solver2(cs.Operator2(so), relaxation=relaxation, nrelax_pre=pre, nrelax_post=post, maxiter=8, tol=1e-30)
        xs, hs, r0s = ser.solve(b)
        x0, y0, nxl, nyl = ext[rank]
        same = np.array_equal(x[1:-1, 1:-1], xs[1 + y0:1 + y0 + nyl, 1 + x0:1 + x0 + nxl])
        hist = len(h) == len(hs) and np.allclose(h, hs, rtol=1e-10, atol=0)
        print(f"rank {rank}: {nx}x{ny} {px}x{py} halo {halo} V({pre},{post}) {relaxation} split {split}: bitwise {same}, history {hist}, "
              f"{sol.exchanges_per_cycle} exchanges/cycle", flush=True)
        ok = ok and same and hist
        dist.barrier()
        sol.close()
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0 and int(flag.item()) == 1:
        print("DIST2 RANKS OK", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
