"""Generates tests/golden/solver2d.json from the CPU oracle (which is itself
pinned to the reference's README history / known-answer tests by
tests/test_oracle_golden.py).  The reference cannot be built on this image
(needs gfortran + MPI + Boost + LAPACK), so these vectors come from the
restatement, not from a reference binary.  Run: python tests/golden/make_golden.py"""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import oracle  # noqa: E402

CASES = [
    dict(gallery="poisson", nx=400, ny=400, relaxation="point", pre=1, post=1, maxiter=10, tol=1e-10, num_levels=7),
    dict(gallery="poisson", nx=400, ny=400, relaxation="point", pre=1, post=1, maxiter=10, tol=1e-10),
    dict(gallery="fe", nx=257, ny=129, relaxation="point", pre=2, post=1, maxiter=10, tol=1e-10),
    dict(gallery="fe", nx=1025, ny=1025, relaxation="point", pre=1, post=1, maxiter=12, tol=1e-10),
    dict(gallery="diag_diffusion", nx=200, ny=120, dx=1.0, dy=1e-4, relaxation="line-x", pre=2, post=1, maxiter=6, tol=1e-10),
    dict(gallery="diag_diffusion", nx=97, ny=130, dx=1e-4, dy=1.0, relaxation="line-y", pre=2, post=1, maxiter=6, tol=1e-10),
    dict(gallery="diag_diffusion", nx=129, ny=129, dx=1.0, dy=1e-2, relaxation="line-xy", pre=2, post=1, maxiter=6, tol=1e-10),
]

out = {"generator": "tests/golden/make_golden.py (oracle/bmg2.c + ml2.c)", "cases": []}
for c in CASES:
    so = oracle.gallery2(c["gallery"], c["nx"], c["ny"], c.get("dx", 1.0), c.get("dy", 1.0))
    b = oracle.rhs_sin2(c["nx"], c["ny"])
    s = oracle.Solver2(so, c["relaxation"], c["pre"], c["post"], num_levels=c.get("num_levels", -1))
    x, h, r0 = s.solve(b, maxiter=c["maxiter"], tol=c["tol"])
    c = dict(c)
    c["nlevels"] = s.nlevels
    c["history"] = [float(v) for v in h]
    c["res0"] = float(r0)
    c["xnorm"] = float((x[1:-1, 1:-1] ** 2).sum() ** 0.5)
    out["cases"].append(c)
with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "solver2d.json"), "w") as f:
    json.dump(out, f, indent=1)
print("wrote", len(out["cases"]), "cases")
