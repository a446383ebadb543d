#!/usr/bin/env python
"""bench.py -- BoxMG V-cycle throughput on B200 (BASELINE.json metric: MDOF/s).

A "step" is ONE V-cycle of the configured BoxMG hierarchy over one right-hand
side (all levels: relax, residual, restrict, coarse solve, interp_add, relax),
run through the C++ host layer (`cdr2::solver<>::vcycle`) over the C-ABI CUDA
kernels.  Default workload = BASELINE.json configs[1]: 2D 9-point (gallery::fe)
4097^2, 4-colour point relaxation, V(1,1).

  value   MDOF/s, device-resident (b, x and the hierarchy already in HBM)
  e2e     MDOF/s through the host-buffer API (include/cedar_capi.h): a stream
          of right-hand sides in pinned host memory; every timed step uploads
          its b and downloads its x (x0 = 0), copies overlapped with the
          cycles of the neighbouring steps; e2e.single_call = one synchronous
          call at a time (no overlap)
  setup   one-off Galerkin setup time (interpolation + RAP on every level)
  solve   full solve to `tol` (iterations, ms)
  roofline   the dominant kernel (fine-level point-GS sweep) timed alone
  cpu_baseline   the CPU oracle (restatement of the reference's serial path,
          OpenMP over rows) on the host cores, same workload

N > 1 (torchrun, one rank per GPU): ONE block-decomposed problem with (n-1)^2
unknowns per GPU (weak scaling) through cdr2_dist_* (ghost exchange by pulls out
of the neighbours' device memory over NVLink); the line also carries
`parity_vs_serial` (the decomposed solve against the serial GPU solver on a grid
one GPU holds, outside the timed region) and `c5` (BASELINE configs[4] as
specified: 9-point heterogeneous 16385^2 across the N GPUs, strong scaling).

`--impl reference` times the CPU implementation as the reference arm (the
reference itself -- Fortran+MPI+Boost+LAPACK -- cannot be built on this image;
DESIGN.md "Oracle") on the same global grid, with every host thread.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "MDOF/s per V-cycle (fine-grid unknowns / V-cycle time)"
BYTES_PER_PT = {"relax9": 64, "relax5": 48}  # algorithmic bytes per fine point per sweep
NCU_SUMMARY = "profiles/r02_fine2d_ncu_full_summary.csv"


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of `kernel` at 4097^2, read from the committed summary
    of this round's `ncu --set full` capture (tools/ncu_summary.py of gpurun_out/r02_fine2d_full.ncu-rep); None if absent"""
    import csv
    try:
        for row in csv.DictReader(open(os.path.join(ROOT, NCU_SUMMARY))):
            if row["kernel"].startswith(kernel):
                return float(row["dram__bytes_read.sum"]) + float(row["dram__bytes_write.sum"])
    except (OSError, KeyError, ValueError):
        pass
    return None


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=4097, help="interior unknowns per dimension")
    ap.add_argument("--gallery", default="fe", choices=["fe", "poisson", "diag_diffusion", "hetero"])
    ap.add_argument("--relaxation", default="point")
    ap.add_argument("--pre", type=int, default=1)
    ap.add_argument("--post", type=int, default=1)
    ap.add_argument("--tol", type=float, default=1e-10)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--coarse-tail", type=int, default=None, help="solver.coarse-tail (levels of at most this size run in one kernel)")
    ap.add_argument("--replicas", action="store_true", help="N > 1: independent replicas instead of one decomposed problem")
    ap.add_argument("--agglomerate", type=int, default=513, help="N > 1: levels of this size or smaller are solved replicated")
    ap.add_argument("--halo", type=int, default=15, help="N > 1: lines of the neighbours every array carries (odd)")
    ap.add_argument("--split-min", type=int, default=0, help="N > 1: split sweeps (ghost exchange beside the interior relaxation) where blocks are at least this wide; 0 = never (default: measured slower, DESIGN.md section 6)")
    ap.add_argument("--global-n", type=int, default=0, help="N > 1: strong scaling on a global-n^2 grid instead of (n-1)^2 per GPU")
    ap.add_argument("--no-c5", action="store_true", help="N > 1: skip the extra 16385^2 heterogeneous (configs[4]) measurement")
    ap.add_argument("--c5-n", type=int, default=16385)
    ap.add_argument("--transport", default="p2p", choices=["p2p", "nccl"], help="N > 1: peer-memory pulls (default) or the NCCL send/recv path")
    return ap.parse_args()


def global_grid(a, world):
    """grid the N-GPU arm solves: (n-1)^2 unknowns per GPU on the reference's block grid, or --global-n squared"""
    import cedar_b200.dist2 as d2
    if world == 1:
        return a.n, a.n, 1, 1
    if a.global_n:
        px, py = d2.grid_decomp((a.global_n, a.global_n), world)
        return a.global_n, a.global_n, px, py
    nb = a.n - 1
    px, py = d2.grid_decomp((nb, nb), world)
    return px * nb + 1, py * nb + 1, px, py


def workload(a, world=1):
    """the `config` object: identical in the repo arm and the reference arm"""
    gnx, gny, px, py = global_grid(a, world)
    pts = 9 if a.gallery in ("fe", "hetero") else 5
    return {"workload": f"2D {pts}-point {a.gallery} {gnx}x{gny}, {a.relaxation} relaxation, V({a.pre},{a.post}), "
                        f"BoxMG Galerkin hierarchy",
            "nx": gnx, "ny": gny, "relaxation": a.relaxation, "cycle": f"V({a.pre},{a.post})",
            "parallelism": "1 GPU" if world == 1 else f"{px}x{py} blocks, one per GPU",
            "l2": "inputs larger than L2 (fine level alone is > 1 GB per GPU)"}


def fine_operator(a, nx, ny):
    """fine-grid operator of the workload as a host array (product-side generators only)"""
    import cedar_b200.dist2 as d2
    import cedar_b200.solver as cs
    if a.gallery == "hetero":
        return d2.gallery_hetero_block((nx, ny), (0, 0, nx, ny))
    op = cs.Operator2(gallery=a.gallery, nx=nx, ny=ny, dx=1.0, dy=1e-4 if a.gallery == "diag_diffusion" else 1.0)
    return np.array(op.array())


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows = []
        self.p = None
        self.index = index

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        self.t.join(timeout=2)
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) >= 9 and r[5 + k].lower() == "active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def cpu_vcycles(a, nx, ny, steps, warmup, threads=None):
    """Time `steps` V-cycles of the CPU oracle on the host cores.  Returns (MDOF/s, ms/step, cores, setup_s)."""
    import oracle
    import cedar_b200.dist2 as d2
    L = oracle.lib()
    L.orc_set_threads(threads or os.cpu_count() or 1)     # torchrun exports OMP_NUM_THREADS=1: set it explicitly
    cores = L.orc_get_max_threads()
    so = fine_operator(a, nx, ny)
    b = d2.rhs_sin_block((nx, ny), (0, 0, nx, ny))
    t0 = time.perf_counter()
    s = oracle.Solver2(so, a.relaxation, a.pre, a.post)
    setup_s = time.perf_counter() - t0
    x = np.zeros_like(b)
    dt = 0.0
    for i in range(warmup + steps):
        x[...] = 0.0                     # x0 = 0 before every cycle, outside the timed region (as in the GPU arm)
        t0 = time.perf_counter()
        s.vcycle(x, b)
        if i >= warmup:
            dt += time.perf_counter() - t0
    dt /= max(steps, 1)
    return nx * ny / dt / 1e6, dt * 1e3, cores, setup_s


def run_reference(a):
    """the reference arm: the CPU implementation on the SAME global grid the N-GPU arm solves, every host thread;
    rank 0 alone works under torchrun"""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1")) if "RANK" in os.environ else a.gpus
    if rank != 0:
        return
    gnx, gny, _, _ = global_grid(a, world)
    val, ms, cores, setup_s = cpu_vcycles(a, gnx, gny, a.steps, a.warmup)
    sample = (f"{a.steps} full V-cycles of the same workload ({gnx}x{gny}) after {a.warmup} warm-ups, oracle port with "
              f"OpenMP over {cores} threads")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "MDOF/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong" if a.global_n else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload(a, world),
            "cpu_baseline": {"value": val, "unit": "MDOF/s", "cores": cores, "kind": "port", "sample": sample,
                             "setup_s": setup_s},
            "e2e": {"value": val, "unit": "MDOF/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def time_fine_sweep(a, torch, reps=20):
    """Dominant kernel alone: one fine-level point-GS sweep through the C ABI,
    CUDA events on the launching stream (the legacy default stream)."""
    import ctypes as C
    import cedar_b200.abi as abi
    import cedar_b200.dist2 as d2
    n = a.n
    nine = a.gallery in ("fe", "hetero")
    so = torch.from_numpy(fine_operator(a, n, n)).cuda()
    II = n + 2
    q = torch.zeros((II, II), dtype=torch.float64, device="cuda")
    b = torch.from_numpy(d2.rhs_sin_block((n, n), (0, 0, n, n))).cuda()
    sor = torch.zeros((2, II, II), dtype=torch.float64, device="cuda")
    P = lambda t: C.c_void_p(t.data_ptr())
    nst = so.shape[0]
    abi.check(abi.lib.bmgx2_SETUP_recip(P(so), P(sor), II, II, nst, 2, None))
    e0, e1 = C.c_void_p(), C.c_void_p()
    abi.lib.bmgx_event_create(C.byref(e0)); abi.lib.bmgx_event_create(C.byref(e1))
    q2 = torch.zeros_like(q)
    bufs = [q, q2]

    def call(i):  # fused single-pass sweep, ping-pong between the two iterates (what rbgs::run does)
        abi.check(abi.lib.bmgx2_relax_GS_oop(1, P(so), P(b), P(bufs[i & 1]), P(bufs[(i + 1) & 1]), P(sor), II, II, 1,
                                             0 if nine else 1, nst, 2, 1, i & 1, 0, None))
    for i in range(4):
        call(i)
    torch.cuda.synchronize()
    l0 = abi.lib.bmgx_launch_count()
    abi.lib.bmgx_event_record(e0, None)
    for i in range(reps):
        call(i)
    abi.lib.bmgx_event_record(e1, None)
    abi.lib.bmgx_event_sync(e1)
    launches = (abi.lib.bmgx_launch_count() - l0) / reps
    ms = C.c_float()
    abi.lib.bmgx_event_elapsed_ms(e0, e1, C.byref(ms))
    abi.lib.bmgx_event_destroy(e0); abi.lib.bmgx_event_destroy(e1)
    per = ms.value / reps
    alg = BYTES_PER_PT["relax9" if nine else "relax5"] * n * n
    del so, q, q2, bufs, b, sor
    torch.cuda.empty_cache()
    return per, alg, launches


SETUP_PARTS = {}


def _dist_problem(a, d2, cs, comm, rank, gnx, gny, px, py, gallery, pre, post, agglomerate, halo, graph=True):
    """block-decomposed solver for this rank's block + its pinned right-hand side; returns (solver, b block, setup s)"""
    import torch
    gen = d2.gallery_hetero_block if gallery == "hetero" else d2.gallery_fe_block
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sol = d2.DistSolver2(gnx, gny, (px, py), nst=5, nrelax_pre=pre, nrelax_post=post, agglomerate=agglomerate, halo=halo,
                         cuda_graph=graph, comm=comm, split_min=a.split_min)
    ext = sol.extent(0, rank)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    so = gen((gnx, gny), ext)
    t2 = time.perf_counter()
    t_gen = t2 - t1
    sol.set_operator(rank, so)
    torch.cuda.synchronize()
    t3 = time.perf_counter()
    del so
    sol.setup()
    torch.cuda.synchronize()
    t4 = time.perf_counter()
    setup_s = t4 - t0 - t_gen
    # where the setup goes (this rank): arena + CUDA IPC + handle all-gather | H2D of the block's operator from pageable
    # host memory | exchange + interpolation + Galerkin per level, replicated coarse solver
    SETUP_PARTS.update({"create_connect_ms": (t1 - t0) * 1e3, "operator_upload_ms": (t3 - t2) * 1e3,
                        "device_setup_ms": (t4 - t3) * 1e3})
    return sol, d2.rhs_sin_block((gnx, gny), ext), setup_s


def _parity_vs_serial(a, d2, cs, comm, dist, torch, rank, px, py):
    """outside the timed region: the decomposed solve (same block grid, same kernels, same exchanges) against the
    serial cdr2::solver on a grid every GPU can also solve alone; bitwise on every rank's block"""
    nb = 512
    gnx, gny = px * nb + 1, py * nb + 1
    sol, bl, _ = _dist_problem(a, d2, cs, comm, rank, gnx, gny, px, py, "hetero", a.pre, a.post, 129, a.halo)
    sol.upload(rank, "b", bl)
    sol.upload(rank, "x", np.zeros_like(bl))
    h, r0 = sol.solve(maxiter=6, tol=1e-30)
    x = sol.download(rank)
    x0, y0, nxl, nyl = sol.extent(0, rank)
    so = d2.gallery_hetero_block((gnx, gny), (0, 0, gnx, gny))
    ser = cs.Solver2(cs.Operator2(so), relaxation="point", nrelax_pre=a.pre, nrelax_post=a.post, maxiter=6, tol=1e-30)
    xs, hs, _ = ser.solve(d2.rhs_sin_block((gnx, gny), (0, 0, gnx, gny)))
    same = bool(np.array_equal(x[1:-1, 1:-1], xs[1 + y0:1 + y0 + nyl, 1 + x0:1 + x0 + nxl]))
    hist = bool(len(h) == len(hs) and np.allclose(h, hs, rtol=1e-10, atol=0))
    flag = torch.tensor([int(same), int(hist)], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    comm.barrier()
    sol.close()
    del ser
    same, hist = [bool(v) for v in flag.tolist()]
    return {"result": "bitwise" if same else "DIFFERS", "history_rel_1e-10": hist, "grid": f"{gnx}x{gny} hetero 9-point",
            "cycles": 6, "what": "every rank's block of x after 6 V-cycles (eager, captured and replayed) == the serial "
                                 "cdr2::solver's on the same GPU; relative residual history to 1e-10"}


def run_multi(a, torch, dist, world, rank, local):
    """N > 1: ONE block-decomposed problem across the N GPUs (cedar_b200/host/dist2.cc through cedar_b200/dist2.py):
    blocks of (n-1)^2 unknowns per GPU (weak scaling: the global grid is px*(n-1)+1 by py*(n-1)+1) or --global-n
    squared (strong scaling); ghost exchange by pulls out of the neighbours' memory over NVLink, levels <=
    --agglomerate solved replicated on every rank."""
    import cedar_b200.dist2 as d2
    import cedar_b200.solver as cs
    comm = d2.TorchComm()
    gnx, gny, px, py = global_grid(a, world)
    dofs = gnx * gny

    def barrier():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    # CUDA IPC between the ranks' allocations is what the exchange rests on: if any rank cannot map its peers
    # (a box that forbids it), every rank falls back to the NCCL send/recv path together
    probe = None
    try:
        probe = d2.DistSolver2(1025, 1025, (px, py), agglomerate=65, halo=a.halo, comm=comm)
        failed = 0
    except Exception as e:                       # noqa: BLE001
        failed = 1
        print(f"rank {rank}: peer-memory transport unavailable ({e}); falling back to NCCL", file=sys.stderr, flush=True)
    flag = torch.tensor([failed], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MAX)
    if probe is not None:
        probe.close()
    if int(flag.item()):
        return run_multi_nccl(a, torch, dist, world, rank, local)
    sol, bl, setup_s = _dist_problem(a, d2, cs, comm, rank, gnx, gny, px, py, a.gallery, a.pre, a.post, a.agglomerate, a.halo,
                                     graph=not a.no_graph)
    setup_parts = {k: round(v, 1) for k, v in SETUP_PARTS.items()}
    x0, y0, nxl, nyl = sol.extent(0, rank)
    sol.upload(rank, "b", bl)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    sol.time_vcycles(max(a.warmup, 3), 1)       # eager passes + graph capture
    t_pre = time.perf_counter()
    keep = torch.tensor([1], device="cuda")
    while True:                                  # keep the GPUs busy until nvidia-smi delivers its first sample (all ranks agree)
        keep[0] = 1 if (rank == 0 and sampler.p and not sampler.rows and time.perf_counter() - t_pre < 3.0) else 0
        dist.broadcast(keep, 0)
        if not int(keep.item()):
            break
        sol.time_vcycles(0, 20)
    barrier()
    l0 = cs.gpu_launches()
    ms_dev = sol.time_vcycles(0, a.steps)        # CUDA events on the launching stream around each cycle, mean
    launches = cs.gpu_launches() - l0
    barrier()
    # end to end (a): a stream of right-hand sides, this rank's block of each from pinned host memory and its block of
    # the solution back, every step; copies overlapped with the neighbouring steps' cycles (x0 = 0: no x upload)
    ring = 2
    pin = cs.PinnedArrays(2 * ring, bl.shape)
    bhs, xhs = pin.arrays[:ring], pin.arrays[ring:]
    for h in bhs:
        h[...] = bl
    sol.vcycle_stream([bhs[k % ring] for k in range(3)], [xhs[k % ring] for k in range(3)])
    barrier()
    t0 = time.perf_counter()
    sol.vcycle_stream([bhs[k % ring] for k in range(a.steps)], [xhs[k % ring] for k in range(a.steps)])
    ms_e2e = (time.perf_counter() - t0) / a.steps * 1e3
    # (b) one right-hand side at a time: upload b, x = 0, cycle, download x, synchronous
    t_one = 0.0
    for i in range(a.steps + 2):
        barrier()
        t0 = time.perf_counter()
        sol.vcycle_stream([bhs[0]], [xhs[1]])
        if i >= 2:
            t_one += time.perf_counter() - t0
    ms_one = t_one / a.steps * 1e3
    e2e_check = float(np.abs(xhs[(a.steps - 1) % ring] - xhs[1]).max())
    # full solve to tol from x = 0
    barrier()
    sol.upload(rank, "b", bl)
    sol.upload(rank, "x", np.zeros_like(bl))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    hist, _ = sol.solve(maxiter=30, tol=a.tol)
    ms_solve = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop() if rank == 0 else None
    info = {"ndist": sol.ndist, "nlevels": sol.nlevels, "exchanges": sol.exchanges_per_cycle, "ops": sol.ops_per_cycle}
    barrier()
    sol.close()
    del pin
    t = torch.tensor([ms_dev, ms_e2e, ms_solve, ms_one, e2e_check, setup_s], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_dev, ms_e2e, ms_solve, ms_one, e2e_check, setup_s = [float(v) for v in t.tolist()]
    parity = _parity_vs_serial(a, d2, cs, comm, dist, torch, rank, px, py)
    c5 = None
    if not a.no_c5:
        c5 = run_c5(a, d2, cs, comm, dist, torch, world, rank, barrier)
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        own = nxl * nyl * 8
        cyc_bytes = 360 * dofs if (a.pre, a.post) == (1, 1) else None
        cfg = workload(a, world)
        line = {
            "metric": METRIC, "value": dofs / (ms_dev * 1e-3) / 1e6, "unit": "MDOF/s", "n_gpus": world,
            "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_dev, "higher_is_better": True,
            "scaling": "strong" if a.global_n else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": cfg,
            "decomposition": {"blocks": f"{px}x{py}", "halo_lines": a.halo, "distributed_levels": info["ndist"],
                              "overlap": (f"on levels with blocks >= {a.split_min} wide the ghost exchange runs on a second stream "
                                          "beside the interior of the sweep that needs it; the frames follow") if a.split_min else "off",
                              "levels": info["nlevels"], "replicated_below": a.agglomerate,
                              "exchange": "one kernel per exchange: signal, wait, pull out of the neighbours' device memory "
                                          "(CUDA IPC over NVLink); no NCCL on the data path"},
            "e2e": {"value": dofs / (ms_e2e * 1e-3) / 1e6, "unit": "MDOF/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": own * world, "d2h_bytes_per_step": own * world,
                    "call": "cdr2_dist_vcycle_stream: a stream of right-hand sides, every rank's block of b from pinned host "
                            "memory and its block of x back per step (x0 = 0), copies overlapped with the neighbouring "
                            "steps' cycles",
                    "single_call": {"ms_per_step": ms_one, "value": dofs / (ms_one * 1e-3) / 1e6,
                                    "call": "one right-hand side per call (upload b, cycle, download x, synchronous)"},
                    "max_abs_diff_batch_vs_single": e2e_check},
            "gpu_launches": int(launches), "ghost_exchanges_per_step": info["exchanges"],
            "kernels_per_step": launches / a.steps,
            "clocks": clocks,
            "setup": {"ms": setup_s * 1e3, "what": "arena allocation + IPC connect + block upload + distributed interpolation/"
                                                   "Galerkin setup + replicated serial setup of the agglomerated levels "
                                                   "(host-side generation of the operator excluded)",
                      "parts_rank0": setup_parts},
            "solve": {"tol": a.tol, "iterations": len(hist), "ms": ms_solve, "final_rel_residual": float(hist[-1])},
            "parity_vs_serial": parity["result"], "parity_detail": parity,
            "roofline": {"bound": "hbm", "kernel": "whole V-cycle per GPU (k_relax9_fused dominates; single-GPU line has it alone)",
                         "achieved": (cyc_bytes / (ms_dev * 1e-3) / 1e9 / world) if cyc_bytes else None,
                         "peak": hbm_peak, "unit": "GB/s",
                         "frac": (cyc_bytes / (ms_dev * 1e-3) / 1e9 / world / hbm_peak) if cyc_bytes else None,
                         "traffic": None, "note": "algorithmic bytes of the whole cycle (360 B/pt) per GPU"},
        }
        if c5 is not None:
            line["c5"] = c5
        print(json.dumps(line), flush=True)
    barrier()
    killer = threading.Timer(20.0, lambda: os._exit(0))
    killer.daemon = True
    killer.start()
    dist.destroy_process_group()
    killer.cancel()


def run_c5(a, d2, cs, comm, dist, torch, world, rank, barrier):
    """BASELINE configs[4] as specified: 2D 9-point heterogeneous 16385^2, block-decomposed across the N GPUs (strong
    scaling), V(1,1) point relaxation: cycle time, solve to tol"""
    n = a.c5_n
    px, py = d2.grid_decomp((n, n), world)
    sol, bl, setup_s = _dist_problem(a, d2, cs, comm, rank, n, n, px, py, "hetero", 1, 1, a.agglomerate, a.halo)
    sol.upload(rank, "b", bl)
    sol.time_vcycles(3, 1)
    barrier()
    ms = sol.time_vcycles(0, 10)
    barrier()
    sol.upload(rank, "x", np.zeros_like(bl))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    hist, _ = sol.solve(maxiter=300, tol=a.tol)
    ms_solve = (time.perf_counter() - t0) * 1e3
    ex = sol.exchanges_per_cycle
    barrier()
    sol.close()
    t = torch.tensor([ms, ms_solve, setup_s], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_solve, setup_s = [float(v) for v in t.tolist()]
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    return {"workload": f"2D 9-point hetero {n}x{n} (64x64 checkerboard D = 10^U(-3,3)), point relaxation, V(1,1), {px}x{py} blocks",
            "scaling": "strong", "ms_per_cycle": ms, "mdof_per_s": n * n / (ms * 1e-3) / 1e6,
            "vcycle_roofline_frac_per_gpu": 360.0 * n * n / (ms * 1e-3) / 1e9 / world / hbm_peak,
            "ghost_exchanges_per_cycle": ex, "setup_ms": setup_s * 1e3,
            "solve": {"tol": a.tol, "iterations": len(hist), "ms": ms_solve, "final_rel_residual": float(hist[-1]),
                      "max_rel_residual": float(np.max(hist)),
                      "note": "the L2 residual of this operator (jumps of 10^6) first grows, then contracts at ~0.8-0.9 per "
                              "V(1,1) point-relaxation cycle; the serial CPU oracle shows the same history "
                              "(tests/test_gpu_dist2.py::test_c5_operator_history_equals_oracle)"}}


def run_multi_nccl(a, torch, dist, world, rank, local):
    """--transport nccl (round-1 path, kept for comparison and as the fallback when CUDA IPC is not permitted on a box):
    N > 1: ONE block-decomposed problem across the N GPUs (cedar_b200/mpi2.py): blocks of
    (n-1)^2 unknowns per GPU (weak scaling: the global grid is px*(n-1)+1 by py*(n-1)+1),
    ghost exchange over NCCL, coarse levels agglomerated on rank 0."""
    import cedar_b200.mpi2 as m
    nb = a.n - 1
    px, py = m.grid_decomp((nb, nb), world)
    gnx, gny = px * nb + 1, py * nb + 1
    nl = m.compute_num_levels(gnx, gny)
    nd, nn = 0, (gnx, gny)
    while nd < nl - 1 and min(nn) > a.agglomerate:
        nd += 1
        nn = (m.coarse_n(nn[0]), m.coarse_n(nn[1]))
    dec = m.Decomp2(gnx, gny, px, py, nd)
    gen = m.gallery_hetero_block if a.gallery == "hetero" else m.gallery_fe_block
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sol = m.DistSolver2({rank: gen(dec, rank)}, gnx, gny, (px, py), a.pre, a.post, agglomerate=a.agglomerate,
                        transport=m.DistTransport(dec))
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t0
    bh = torch.from_numpy(m.rhs_sin_block(dec, rank)).pin_memory()
    xh = torch.zeros_like(bh).pin_memory()
    b = {rank: bh.cuda()}
    x = {rank: torch.zeros_like(b[rank])}
    dofs = gnx * gny

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    cycle = sol.vcycle if a.no_graph else sol.vcycle_graphed
    for _ in range(max(a.warmup, 3)):
        x[rank].zero_()
        cycle(x, b)
    barrier()
    l0, ex0 = sol.K.launches, sol.exchanges
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps)]
    for i in range(a.steps):
        with torch.cuda.stream(sol.stream):
            x[rank].zero_()
            e0[i].record()
        cycle(x, b)
        with torch.cuda.stream(sol.stream):
            e1[i].record()
    barrier()
    ms_dev = sum(s.elapsed_time(e) for s, e in zip(e0, e1)) / a.steps
    launches = sol.K.launches - l0
    exchanges = sol.exchanges - ex0
    # end to end (a): a stream of right-hand sides, this rank's block of each from pinned host memory and its block
    # of the solution back, every step; copies overlapped with the neighbouring steps' cycles (x0 = 0: no x upload)
    ring = 2
    bhs = [bh] + [bh.clone().pin_memory() for _ in range(ring - 1)]
    xhs = [xh] + [torch.zeros_like(bh).pin_memory() for _ in range(ring - 1)]
    bl = [bhs[k % ring] for k in range(a.steps)]
    xl = [xhs[k % ring] for k in range(a.steps)]
    sol.vcycle_stream(x, b, bl[:3], xl[:3], graphed=not a.no_graph)
    barrier()
    t0 = time.perf_counter()
    sol.vcycle_stream(x, b, bl, xl, graphed=not a.no_graph)
    ms_e2e = (time.perf_counter() - t0) / a.steps * 1e3
    # (b) one right-hand side at a time: H2D b, cycle, D2H x, synchronous
    t_one = 0.0
    for i in range(a.steps + 2):
        barrier()
        t0 = time.perf_counter()
        with torch.cuda.stream(sol.stream):
            b[rank].copy_(bh, non_blocking=True)
            x[rank].zero_()
        cycle(x, b)
        with torch.cuda.stream(sol.stream):
            xh.copy_(x[rank], non_blocking=True)
        sol.stream.synchronize()
        if i >= 2:
            t_one += time.perf_counter() - t0
    ms_one = t_one / a.steps * 1e3
    e2e_check = float((xhs[(a.steps - 1) % ring] - xh).abs().max())
    # full solve to tol
    barrier()
    t0 = time.perf_counter()
    _, hist, _ = sol.solve(b, maxiter=30, tol=a.tol)
    torch.cuda.synchronize()
    ms_solve = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms_dev, ms_e2e, ms_solve, ms_one, e2e_check], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_dev, ms_e2e, ms_solve, ms_one, e2e_check = [float(v) for v in t.tolist()]
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        loc = (nb + 2) * (nb + 2) * 8
        cyc_bytes = 360 * dofs if (a.pre, a.post) == (1, 1) else None
        line = {
            "metric": METRIC, "value": dofs / (ms_dev * 1e-3) / 1e6, "unit": "MDOF/s", "n_gpus": world,
            "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_dev, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload(a, world),
            "decomposition": {"blocks": f"{px}x{py}", "exchange": f"NCCL send/recv, wide-halo fused sweeps, {nd} distributed "
                              f"levels, levels <= {a.agglomerate} agglomerated on rank 0"},
            "e2e": {"value": dofs / (ms_e2e * 1e-3) / 1e6, "unit": "MDOF/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": loc * world, "d2h_bytes_per_step": loc * world,
                    "call": "DistSolver2.vcycle_stream: a stream of right-hand sides, every rank's block of b from pinned "
                            "host memory and its block of x back per step (x0 = 0), copies overlapped with the "
                            "neighbouring steps' cycles",
                    "single_call": {"ms_per_step": ms_one, "value": dofs / (ms_one * 1e-3) / 1e6,
                                    "call": "H2D b block, DistSolver2.vcycle, D2H x block, synchronous"},
                    "max_abs_diff_batch_vs_single": e2e_check},
            "gpu_launches": int(launches), "ghost_exchanges_per_step": exchanges / a.steps,
            "clocks": clocks,
            "setup": {"ms": setup_s * 1e3, "what": "block upload + distributed interpolation/Galerkin setup + agglomerated serial setup"},
            "solve": {"tol": a.tol, "iterations": len(hist), "ms": ms_solve, "final_rel_residual": float(hist[-1])},
            "roofline": {"bound": "hbm", "kernel": "whole V-cycle (per-colour relaxation passes between ghost exchanges)",
                         "achieved": (cyc_bytes / (ms_dev * 1e-3) / 1e9 / world) if cyc_bytes else None,
                         "peak": hbm_peak, "unit": "GB/s",
                         "frac": (cyc_bytes / (ms_dev * 1e-3) / 1e9 / world / hbm_peak) if cyc_bytes else None,
                         "traffic": None, "note": "algorithmic bytes of the whole cycle (360 B/pt) per GPU"},
        }
        print(json.dumps(line), flush=True)
    # captured graphs hold NCCL work: release them first, and never let teardown hang the run
    if hasattr(sol, "_graphs"):
        sol._graphs.clear()
    torch.cuda.synchronize()
    dist.barrier()
    killer = threading.Timer(20.0, lambda: os._exit(0))
    killer.daemon = True
    killer.start()
    dist.destroy_process_group()
    killer.cancel()


def main():
    a = parse()
    if a.impl == "reference":
        return run_reference(a)

    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the BoxMG kernels have no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        if not a.replicas:
            if a.transport == "nccl":
                return run_multi_nccl(a, torch, dist, world, rank, local)
            return run_multi(a, torch, dist, world, rank, local)

    import cedar_b200.dist2 as d2
    import cedar_b200.solver as cs

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    n = a.n
    dofs = n * n
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"

    # ---- problem + setup (timed once; includes operator upload) ------------
    dy = 1e-4 if a.gallery == "diag_diffusion" else 1.0
    op = cs.Operator2(gallery=a.gallery, nx=n, ny=n, dx=1.0, dy=dy)
    b = d2.rhs_sin_block((n, n), (0, 0, n, n))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sol = cs.Solver2(op, relaxation=a.relaxation, nrelax_pre=a.pre, nrelax_post=a.post, maxiter=100, tol=a.tol,
                     cuda_graph=not a.no_graph, coarse_tail=a.coarse_tail)
    torch.cuda.synchronize()
    setup_first_s = time.perf_counter() - t0
    # ---- device-resident V-cycles: value -----------------------------------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    sol.time_vcycles(b, warmup=max(a.warmup, 3), n=1)  # warm-up (graph capture happens here)
    t_pre = time.perf_counter()
    while rank == 0 and sampler.p and not sampler.rows and time.perf_counter() - t_pre < 3.0:
        sol.time_vcycles(b, warmup=0, n=20)             # keep the GPU busy until nvidia-smi delivers its first sample
    barrier()
    l0 = cs.gpu_launches()
    ms_dev = sol.time_vcycles(b, warmup=0, n=a.steps)  # CUDA events around each cycle, mean
    launches = cs.gpu_launches() - l0
    barrier()

    # ---- end to end through the host-buffer API ----------------------------
    # (a) a stream of right-hand sides through cdr2_solver_vcycle_batch: every step uploads its b from pinned host
    #     memory and downloads its x (x0 = 0, so there is no x to upload); the copies of step k+1 / k-1 overlap the
    #     cycle of step k on separate streams.  A ring of 4 pinned b/x pairs stands for the stream.
    ring = 4
    pbs, pxs = sol.pinned_grid_funcs(ring), sol.pinned_grid_funcs(ring)
    for k in range(ring):
        pbs[k][...] = b
    bl = [pbs[k % ring] for k in range(a.steps)]
    xl = [pxs[k % ring] for k in range(a.steps)]
    sol.vcycle_batch(bl[:ring], xl[:ring])          # warm-up (allocates the staging buffers and streams)
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sol.vcycle_batch(bl, xl)                          # returns when the last x has arrived on the host
    ms_e2e = (time.perf_counter() - t0) / a.steps * 1e3
    # (b) one right-hand side at a time (latency of a single synchronous call): H2D b, cycle, D2H x
    hb, hx = sol.buffers()
    hb[...] = b
    for _ in range(2):
        sol.vcycle0_pinned()
    barrier()
    t_one = 0.0
    for _ in range(a.steps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sol.vcycle0_pinned()
        t_one += time.perf_counter() - t0
    ms_one = t_one / a.steps * 1e3
    e2e_check = float(np.max(np.abs(pxs[(a.steps - 1) % ring] - hx)))   # same b, same x0: must be identical
    del pbs, pxs, bl, xl
    bytes_gf = (n + 2) * (n + 2) * 8

    # ---- full solve to tol (device resident) -------------------------------
    ms_solve, iters = sol.time_solves(b, warmup=1, n=2)

    clocks = sampler.stop() if rank == 0 else None   # sampled from warm-up to the end of the timed regions

    # max over ranks
    if world > 1:
        t = torch.tensor([ms_dev, ms_e2e, ms_solve], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, ms_e2e, ms_solve = [float(v) for v in t.tolist()]

    if rank == 0:
        del sol
        torch.cuda.empty_cache()
        per_sweep_ms, alg_bytes, sweep_launches = time_fine_sweep(a, torch)
        ach = alg_bytes / (per_sweep_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": world * dofs / (ms_dev * 1e-3) / 1e6, "unit": "MDOF/s", "n_gpus": world,
            "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_dev, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload(a) if world == 1 else dict(workload(a), parallelism=f"{world} independent replicas (one {n}^2 "
                                                                                    f"problem per GPU)"),
            "e2e": {"value": world * dofs / (ms_e2e * 1e-3) / 1e6, "unit": "MDOF/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": bytes_gf, "d2h_bytes_per_step": bytes_gf,
                    "call": "cdr2_solver_vcycle_batch: a stream of right-hand sides in pinned host memory, x0 = 0; every "
                            "step uploads its b and downloads its x, copies overlapped with the neighbouring steps' cycles",
                    "single_call": {"ms_per_step": ms_one, "value": world * dofs / (ms_one * 1e-3) / 1e6,
                                    "call": "cdr2_solver_vcycle0_pinned (synchronous: H2D b, cycle, D2H x)"},
                    "max_abs_diff_batch_vs_single": e2e_check},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "setup": {"ms": setup_first_s * 1e3, "what": "operator upload + interpolation + Galerkin RAP on all levels "
                      "+ relaxation/coarse-solve setup (first call, includes CUDA context warm-up)"},
            "solve": {"tol": a.tol, "iterations": iters, "ms": ms_solve, "mdof_per_s": dofs / (ms_solve * 1e-3) / 1e6},
            "roofline": {"bound": "hbm", "kernel": "k_relax9_fused / k_relax5_fused: fine-level point-GS sweep, all colours in one launch (bmgx2_relax_GS_oop)",
                         "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                         "peak_source": peak_src, "ms_per_launch": per_sweep_ms, "launches_per_sweep": sweep_launches,
                         "algorithmic_bytes": alg_bytes,
                         "traffic": (ncu_traffic("k_relax9_fused<0>" if a.gallery in ("fe", "hetero") else "k_relax5_fused<0>")
                                     if n == 4097 else None),
                         "traffic_source": f"ncu --set full of this round's build, {NCU_SUMMARY} (bytes per launch, 4097^2)"},
        }
        # whole-cycle roofline: SURVEY 8(d) 360 B/pt (9-pt V(1,1)) / 312 (5-pt fine)
        cyc_bytes = (360 if a.gallery == "fe" else 312) * dofs if (a.pre, a.post) == (1, 1) and a.relaxation == "point" else None
        if cyc_bytes:
            line["roofline"]["vcycle_frac"] = cyc_bytes / (ms_dev * 1e-3) / 1e9 / hbm_peak
        if not a.no_cpu_baseline and world == 1:
            val, ms, cores, setup_s = cpu_vcycles(a, n, n, 5, 1)
            line["cpu_baseline"] = {"value": val, "unit": "MDOF/s", "cores": cores, "kind": "port",
                                    "sample": f"5 full V-cycles of the same workload ({n}^2) after 1 warm-up",
                                    "ms_per_step": ms, "setup_s": setup_s}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
