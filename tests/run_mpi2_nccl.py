"""torchrun entry (one rank per GPU, NCCL): the block-decomposed solver against the
serial GPU solver on the same problem -- solution bitwise, history to 1e-10.
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tests/run_mpi2_nccl.py"""
import os
import threading
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    import cedar_b200.mpi2 as m
    import cedar_b200.solver as cs
    from tests.util import random_gf, random_op2
    ok = True
    for nx, ny, nine, agg in [(1025, 513, True, 65), (600, 700, False, 65)]:
        rng = np.random.default_rng(nx + ny)
        so = random_op2(nx, ny, nine, rng)
        b = random_gf(nx, ny, rng)
        px, py = m.grid_decomp((nx, ny), world)
        nl = m.compute_num_levels(nx, ny)
        nd, n = 0, (nx, ny)
        while nd < nl - 1 and min(n) > agg:
            nd += 1
            n = (m.coarse_n(n[0]), m.coarse_n(n[1]))
        dec = m.Decomp2(nx, ny, px, py, nd)
        sol = m.DistSolver2(m.split_blocks(so, dec, blocks=[rank]), nx, ny, (px, py), 2, 1, agglomerate=agg,
                            transport=m.DistTransport(dec))
        x, h, r0 = sol.solve(m.split_blocks(b, dec, blocks=[rank]), maxiter=5, tol=1e-12)
        torch.cuda.synchronize()
        # every rank checks its own block against the serial GPU solve
        ser = cs.Solver2(cs.Operator2(so), relaxation="point", nrelax_pre=2, nrelax_post=1, maxiter=5, tol=1e-12)
        xs, hs, r0s = ser.solve(b)
        x0, y0, nxl, nyl = dec.extent(0, rank)
        mine = x[rank].cpu().numpy()[1:-1, 1:-1]
        same = np.array_equal(mine, xs[1 + y0:1 + y0 + nyl, 1 + x0:1 + x0 + nxl])
        hist_ok = len(h) == len(hs) and np.allclose(h, hs, rtol=1e-10, atol=0)
        flag = torch.tensor([int(same and hist_ok)], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            print(f"{nx}x{ny} {px}x{py} ndist={nd}: bitwise={bool(flag.item())} hist={h}", flush=True)
        ok = ok and bool(flag.item())
        # a stream of right-hand sides through vcycle_stream == the same cycles one call at a time
        bl = {rank: torch.from_numpy(np.ascontiguousarray(m.split_blocks(b, dec, blocks=[rank])[rank])).cuda()}
        xl = {rank: torch.zeros_like(bl[rank])}
        bhs = [(bl[rank] * (k + 1)).cpu().pin_memory() for k in range(5)]
        xhs = [torch.full_like(h_, 7.0).pin_memory() for h_ in bhs]
        for graphed in (False, True):
            sol.vcycle_stream(xl, bl, bhs, xhs, graphed=graphed)
            good = True
            for k in range(5):
                with torch.cuda.stream(sol.stream):
                    bl[rank].copy_(bhs[k], non_blocking=True)
                    xl[rank].zero_()
                sol.vcycle(xl, bl)
                torch.cuda.synchronize()
                good = good and torch.equal(xl[rank].cpu(), xhs[k])
            flag = torch.tensor([int(good)], device="cuda")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if rank == 0:
                print(f"  vcycle_stream graphed={graphed}: bitwise={bool(flag.item())}", flush=True)
            ok = ok and bool(flag.item())
        # captured graphs hold NCCL work: release them before the process group goes away
        if hasattr(sol, "_graphs"):
            sol._graphs.clear()
        torch.cuda.synchronize()
        dist.barrier()
    if rank == 0:
        print("MPI2 NCCL OK" if ok else "MPI2 NCCL FAILED", flush=True)
    killer = threading.Timer(20.0, lambda: os._exit(0 if ok else 1))   # never let teardown hang the run
    killer.daemon = True
    killer.start()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
