"""The reference's own serial example mains compile and link UNCHANGED against this repo's
include/cedar headers and libraries (solver<>, config and the gallery are
drop-in).  Only where the reference tree is present (this container, not the GPU box)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/examples"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present")
def test_reference_examples_build_unchanged():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "examples"), "-s"])
    for name in ("ser-poisson-2d", "ser-poisson-3d", "ser-periodic-2d"):
        exe = os.path.join(ROOT, "examples", "_ref", name)
        assert os.path.isfile(exe) and os.access(exe, os.X_OK)
        # linked against the host layer and, through it, the CUDA C-ABI library
        out = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
        assert "libcedar_b200.so" in out and "libbmgx.so" in out


@pytest.mark.skipif(not os.path.isfile(os.path.join(ROOT, "examples", "_ref", "ser-poisson-2d")), reason="not built")
def test_example_without_gpu_fails_loudly(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    cfg = open(os.path.join(ROOT, "examples", "basic-2d-ser", "config.json")).read()
    (tmp_path / "config.json").write_text(cfg)
    p = subprocess.run([os.path.join(ROOT, "examples", "_ref", "ser-poisson-2d")], cwd=tmp_path, capture_output=True,
                       text=True, timeout=120)
    assert p.returncode != 0          # no CPU fallback
    assert "Finished Test" not in p.stdout
