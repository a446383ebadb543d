"""ctypes loader of oracle/_ref/libbmgref.so (TEST INFRASTRUCTURE): the reference's
own Fortran kernels transliterated at build time by oracle/f2c_ref.py.  The
library is built in the dev container (where /root/reference exists) and
travels to the GPU box as a prebuilt file; `available()` is False elsewhere."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(_HERE, "_ref", "libbmgref.so")
_lib = None


def available():
    return os.path.exists(PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(PATH)
    return _lib


def ip(v):
    """by-reference integer argument (BMG2_SymStd_residual, BMG2_SymStd_SETUP_cg_LU)"""
    return C.byref(C.c_int(v))
