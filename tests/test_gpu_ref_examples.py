"""Run the reference's example mains (built unchanged by examples/Makefile against this
repo's headers and CUDA libraries) on the GPU and check what Cedar's README says they print."""
import os
import re
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE2 = os.path.join(ROOT, "examples", "_ref", "ser-poisson-2d")
EXE3 = os.path.join(ROOT, "examples", "_ref", "ser-poisson-3d")

# reference README.md:52-63
README_HISTORY = [0.388629, 0.0443548, 0.00494131, 0.000513399, 5.44908e-05, 5.60612e-06,
                  5.86933e-07, 6.04942e-08, 6.30975e-09, 6.52713e-10]


def _run(exe, example, tmp_path):
    cfg = open(os.path.join(ROOT, "examples", example, "config.json")).read()
    (tmp_path / "config.json").write_text(cfg)
    p = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout + p.stderr
    hist = [float(m) for m in re.findall(r"Iteration \d+ relative l2 norm: (\S+)", p.stdout)]
    norm = re.search(r"Solution norm: (\S+)", p.stdout)
    return p.stdout, hist, norm.group(1) if norm else None


@pytest.mark.skipif(not os.path.isfile(EXE2), reason="examples/_ref not built (reference tree absent at build time)")
def test_ser_poisson_2d_prints_the_readme(cuda, tmp_path):
    out, hist, norm = _run(EXE2, "basic-2d-ser", tmp_path)
    assert len(hist) == 10, out
    for got, want in zip(hist, README_HISTORY):
        # six printed digits, one unit of the last (9 of 10 print identically; the last is 6.5271358e-10
        # here and in the oracle, printed 6.52714e-10, against the README's 6.52713e-10)
        assert abs(got - want) <= 2e-6 * want, (got, want)
    # grid_func::inf_norm returns the SIGNED entry of largest magnitude; the error field has four lobes of
    # equal magnitude up to the last bits, so which sign wins is below what the README pins
    assert norm.lstrip("-") == "2.04592e-05", out
    assert "Finished Test" in out


@pytest.mark.skipif(not os.path.isfile(EXE3), reason="examples/_ref not built (reference tree absent at build time)")
def test_ser_poisson_3d_converges(cuda, tmp_path):
    out, hist, norm = _run(EXE3, "basic-3d-ser", tmp_path)
    assert hist and hist[-1] < 1e-8, out                       # test/3d/test_poisson.cc: residual < 1e-8
    assert all(b < a for a, b in zip(hist, hist[1:]))
    assert abs(float(norm)) < 1e-3, out                             # second-order error of the 65^3 grid
    assert "Finished test" in out


CAPI = os.path.join(ROOT, "examples", "_ref", "capi-fe2d")


@pytest.mark.skipif(not os.path.isfile(CAPI), reason="examples/_ref not built")
@pytest.mark.parametrize("n", [65, 200])
def test_c_application_against_the_reference_c_interface(cuda, tmp_path, n):
    """examples/capi/fe2d.c: plain C against include/cedar/capi.h (bmg2_topo_create, bmg2_operator_create/set/apply,
    bmg2_solver_create/run by the reference's names and argument orders, include/cedar/2d/interface/c/*.h): the
    vertex-based signed stencil goes through the symmetric-storage conversion, apply equals the stencil, the solve
    recovers x from b = A x"""
    (tmp_path / "config.json").write_text('{"log": ["error"], "solver": {"relaxation": "point", "max-iter": 20, "tol": 1e-12, '
                                          '"cycle": {"nrelax-pre": 2, "nrelax-post": 1}}}')
    p = subprocess.run([CAPI, str(n)], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    m = re.search(r"apply_err (\S+) solve_err (\S+)", p.stdout)
    assert m and float(m.group(1)) < 1e-12 and float(m.group(2)) < 1e-9, p.stdout
    assert (tmp_path / "timer.json").exists()


EXEP = os.path.join(ROOT, "examples", "_ref", "ser-periodic-2d")


@pytest.mark.skipif(not os.path.isfile(EXEP), reason="examples/_ref not built (reference tree absent at build time)")
def test_ser_periodic_2d_history_equals_oracle(cuda, tmp_path):
    """the reference's examples/basic-2d-ser/periodic.cc, built unchanged, with its periodic-config.json (periodic in x,
    300 x 300, point relaxation V(1,1)): the printed residual history equals the CPU oracle's periodic solve of the
    same problem to the six printed digits, and the error norm agrees"""
    import numpy as np
    import oracle
    from tests.util import periodic_problem2
    cfg = open(os.path.join(ROOT, "examples", "basic-2d-ser", "periodic-config.json")).read()
    (tmp_path / "config.json").write_text(cfg)
    p = subprocess.run([EXEP], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout + p.stderr
    hist = [float(m) for m in re.findall(r"Iteration \d+ relative l2 norm: (\S+)", p.stdout)]
    norm = float(re.search(r"Solution norm: (\S+)", p.stdout).group(1))
    so, b, exact = periodic_problem2(300, 300, (True, False))
    xr, hr, _ = oracle.Solver2(so, "point", 1, 1, periodic=(True, False)).solve(b, maxiter=10, tol=1e-10)
    assert len(hist) == len(hr) == 10
    for got, want in zip(hist, hr):
        assert abs(got - want) <= 6e-6 * want, (got, want)          # six printed digits
    err = np.abs((exact - xr)[1:-1, 1:-1]).max()
    assert abs(abs(norm) - err) <= 6e-6 * err, (norm, err)
    assert "Finished Test" in p.stdout


@pytest.mark.skipif(not os.path.isfile(EXEP), reason="examples/_ref not built (reference tree absent at build time)")
@pytest.mark.parametrize("relaxation", ["line-x", "line-y", "line-xy"])
def test_ser_periodic_2d_with_line_relaxation(cuda, tmp_path, relaxation):
    """the same unchanged example main with the config's relaxation switched to zebra lines: lines periodic in x take the
    Sherman-Morrison branch (relax_lines_x.f90:170-306), y-lines the plain branch with one ghost refresh per sweep; the
    printed history follows the CPU oracle (pinned to the reference's Fortran) to the printed digits"""
    import numpy as np
    import oracle
    from tests.util import periodic_problem2
    cfg = open(os.path.join(ROOT, "examples", "basic-2d-ser", "periodic-config.json")).read()
    assert '"relaxation": "point"' in cfg
    (tmp_path / "config.json").write_text(cfg.replace('"relaxation": "point"', f'"relaxation": "{relaxation}"'))
    p = subprocess.run([EXEP], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout + p.stderr
    hist = [float(m) for m in re.findall(r"Iteration \d+ relative l2 norm: (\S+)", p.stdout)]
    so, b, exact = periodic_problem2(300, 300, (True, False))
    xr, hr, _ = oracle.Solver2(so, relaxation, 1, 1, periodic=(True, False)).solve(b, maxiter=10, tol=1e-10)
    assert len(hist) == len(hr) and len(hist) >= 3
    for got, want in zip(hist, hr):
        assert abs(got - want) <= 2e-5 * want + 1e-14, (got, want)
    assert "Finished Test" in p.stdout
