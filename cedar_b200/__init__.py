"""cedar_b200 -- B200-native BoxMG (black-box multigrid) kernels and solver.

The product is native code:
  * ``cedar_b200/lib/libbmgx.so``      C ABI over the sm_100a CUDA kernels (``include/bmgx.h``)
  * ``cedar_b200/lib/libcedar_b200.so`` C++ host layer mirroring Cedar's kernel
    registry / multilevel solver (``include/cedar/``) plus a C API for bindings.

This Python package is only the loader used by the tests and ``bench.py``.
There is no CPU fallback: importing :mod:`cedar_b200.abi` without the built
libraries raises.
"""
from . import abi  # noqa: F401

__all__ = ["abi"]
