#!/usr/bin/env python
"""Scratch measurements of the other BASELINE configs (C3 lines, C4 3D) -- V-cycle time,
setup time, solve iterations -- printed as JSON lines.  Not the bench contract (bench.py is)."""
import json
import sys
import time
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import cedar_b200.solver as cs

PEAK = 6548.0


def rhs(shape):
    rng = np.random.default_rng(1)
    b = np.zeros(shape)
    inner = tuple(slice(1, -1) for _ in shape)
    b[inner] = rng.standard_normal(tuple(s - 2 for s in shape)) * 1e-3
    return b


def run2(gallery, n, relax, pre=1, post=1, bpp=None):
    dy = 1e-4 if gallery == "diag_diffusion" else 1.0
    if gallery == "hetero5":       # config C3(b): checkerboard coefficient, jumps up to 1e6 (cedar_b200/gallery.py)
        from cedar_b200 import gallery as gal
        op = cs.Operator2(gal.hetero5(n, n))
    else:
        op = cs.Operator2(gallery=gallery, nx=n, ny=n, dx=1.0, dy=dy)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    sol = cs.Solver2(op, relaxation=relax, nrelax_pre=pre, nrelax_post=post, maxiter=30, tol=1e-10)
    torch.cuda.synchronize(); setup = time.perf_counter() - t0
    b = rhs((n + 2, n + 2))
    ms = sol.time_vcycles(b, warmup=3, n=10)
    mss, it = sol.time_solves(b, warmup=1, n=1)
    out = {"cfg": f"2D {gallery} {n}^2 {relax} V({pre},{post})", "ms_vcycle": ms, "gdof_s": n * n / ms / 1e6,
           "setup_ms": setup * 1e3, "solve_ms": mss, "iters": it}
    if bpp:
        out["frac_cycle_roofline"] = bpp * n * n / (ms * 1e-3) / 1e9 / PEAK
    print(json.dumps(out), flush=True)


def run3(gallery, n, relax, pre=1, post=1, bpp=None):
    op = cs.Operator3(gallery=gallery, nx=n, ny=n, nz=n)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    sol = cs.Solver3(op, relaxation=relax, nrelax_pre=pre, nrelax_post=post, maxiter=30, tol=1e-10)
    torch.cuda.synchronize(); setup = time.perf_counter() - t0
    b = rhs((n + 2, n + 2, n + 2))
    ms = sol.time_vcycles(b, warmup=3, n=5)
    mss, it = sol.time_solves(b, warmup=0, n=1)
    out = {"cfg": f"3D {gallery} {n}^3 {relax} V({pre},{post})", "ms_vcycle": ms, "gdof_s": n ** 3 / ms / 1e6,
           "setup_ms": setup * 1e3, "solve_ms": mss, "iters": it}
    if bpp:
        out["frac_cycle_roofline"] = bpp * n ** 3 / (ms * 1e-3) / 1e9 / PEAK
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    which = sys.argv[1:] or ["c2", "c3", "c4"]
    if "c2" in which:
        run2("fe", 4097, "point", bpp=360)
        run2("poisson", 4097, "point", bpp=312)
    if "c3" in which:
        run2("diag_diffusion", 8193, "line-x")
        run2("diag_diffusion", 8193, "line-y")
        run2("diag_diffusion", 8193, "line-xy")
        run2("fe", 8193, "line-xy")
        run2("hetero5", 8193, "line-xy")
        run2("hetero5", 8193, "point")
    if "c4" in which:
        run3("poisson", 257, "point", bpp=335)
        run3("fe", 257, "point", bpp=575)
        run3("poisson", 513, "point", bpp=335)
        run3("fe", 513, "point", bpp=575)
    if "planes" in which:
        run3("poisson", 129, "plane-xyz")
        run3("fe", 129, "plane-xyz")
        run3("poisson", 257, "plane-xyz")
        run3("poisson", 513, "plane-xyz")
