"""ctypes bindings of the C ABI (``include/bmgx.h``) -- the reference-side stub
a maintainer would write to call the B200 kernels from Python.  Device pointers
are plain integers (e.g. ``torch.Tensor.data_ptr()``)."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libbmgx.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(there is no CPU fallback)")

lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)

_p = C.c_void_p
_i = C.c_int
_u = C.c_uint32
_d = C.c_double
_z = C.c_size_t

SIGNATURES = {
    "bmgx_version": (C.c_char_p, []),
    "bmgx_last_error": (C.c_char_p, []),
    "bmgx_launch_count": (C.c_ulonglong, []),
    "bmgx_scratch_trim": (C.c_int, [C.c_size_t]),
    "bmgx_stream_is_capturing": (_i, [_p, C.POINTER(_i)]),
    "bmgx_get_bc": (None, [_i, C.POINTER(_i)]),
    "bmgx_set_batch": (_i, [_u]),
    "bmgx_get_batch": (_u, []),
    "bmgx_fill": (_i, [_p, _z, _d, _p]),
    "bmgx_norm_work_len": (_z, []),
    "bmgx2_sumsq": (_i, [_p, _u, _u, _p, _p, _p]),
    "bmgx3_sumsq": (_i, [_p, _u, _u, _u, _p, _p, _p]),
    "bmgx2_SETUP_recip": (_i, [_p, _p, _u, _u, _i, _i, _p]),
    "bmgx2_relax_GS": (_i, [_i, _p, _p, _p, _p, _u, _u, _i, _i, _i, _i, _i, _i, _i, _p]),
    "bmgx2_relax_GS_oop": (_i, [_i, _p, _p, _p, _p, _p, _u, _u, _i, _i, _i, _i, _i, _i, _i, _p]),
    "bmgx2_relax_GS_oop_rects": (_i, [_i, _p, _p, _p, _p, _p, _u, _u, _i, _i, _i, _i, _i, _i, _i, _p, _i, _p]),
    "bmgx2_SETUP_lines_x": (_i, [_p, _p, _u, _u, _i, _i, _p]),
    "bmgx2_SETUP_lines_y": (_i, [_p, _p, _u, _u, _i, _i, _p]),
    "bmgx2_relax_lines_x": (_i, [_i, _p, _p, _p, _p, _p, _u, _u, _i, _i, _i, _i, _i, _i, _p]),
    "bmgx2_relax_lines_y": (_i, [_i, _p, _p, _p, _p, _p, _u, _u, _i, _i, _i, _i, _i, _i, _p]),
    "bmgx2_relax_lines_x_colour": (_i, [_p, _p, _p, _p, _u, _u, _i, _i, _p]),
    "bmgx2_relax_lines_y_colour": (_i, [_p, _p, _p, _p, _p, _u, _u, _i, _i, _p]),
    "bmgx2_lines_rhs_push": (_i, [_i, _i, _p, _p, _p, _u, _u, _u, _u, _u, _u, _i, _u, _p, _p]),
    "bmgx2_lines_solve_T": (_i, [_p, _p, _u, _u, _i, _p]),
    "bmgx2_lines_scatter_push": (_i, [_i, _p, _u, _u, _i, _u, _p, _p]),
    "bmgx2_relax_lines_y_static": (_i, [_i, _p, _p, _p, _p, _u, _u, _i, _i, _i, _i, _i, _i, _p]),
    "bmgx2_residual": (_i, [_i, _p, _p, _p, _p, _u, _u, _i, _i, _i, _i, _i, _i, _i, _p]),
    "bmgx2_matvec": (_i, [_i, _p, _p, _p, _u, _u, _i, _i, _i, _p]),
    "bmgx2_restrict": (_i, [_p, _p, _p, _u, _u, _u, _u, _i, _p]),
    "bmgx2_interp_add": (_i, [_p, _p, _p, _p, _p, _u, _u, _u, _u, _i, _i, _p]),
    "bmgx2_SETUP_interp_OI": (_i, [_p, _p, _p, _u, _u, _u, _u, _i, _i, _i, _i, _p]),
    "bmgx2_SETUP_ITLI_ex": (_i, [_p, _p, _p, _u, _u, _u, _u, _i, _i, _i, _p]),
    "bmgx2_SETUP_cg_LU": (_i, [_p, _u, _u, _i, _p, _u, _u, _i, _p]),
    "bmgx_cg_LU_info": (_i, [C.POINTER(_i), _p]),
    "bmgx2_SOLVE_cg": (_i, [_p, _p, _u, _u, _p, _p, _u, _u, _i, _p]),
    "bmgx2_relax_GS_pass": (_i, [_i, _p, _p, _p, _p, _u, _u, _i, _i, _i, _i, _i, _i, _i, _i, _p]),
    "bmgx2_SETUP_interp_OI_stage": (_i, [_p, _p, _p, _u, _u, _u, _u, _i, _i, _i, _i, _i, _p]),
    "bmgx2_line_copy": (_i, [_p, _u, _u, _i, _u, _i, _u, _p, _i, _p]),
    "bmgx2_vcycle_tail": (_i, [_p, _i, _p, _p, _u, _i, _i, _p]),
    "bmgx_ipc_export": (_i, [_p, _p]),
    "bmgx_ipc_open": (_i, [_p, C.POINTER(_p)]),
    "bmgx_ipc_close": (_i, [_p]),
    "bmgx_halo_step": (_i, [_p, _i, _i, _p, _p, _i, _i, _i, _i, _d, _p]),
    "bmgx_copy2d": (_i, [_p, _z, _p, _z, _u, _u, _p]),
    "bmgx_memcpy2d_h2d": (_i, [_p, _z, _p, _z, _u, _u, _p]),
    "bmgx_memcpy2d_d2h": (_i, [_p, _z, _p, _z, _u, _u, _p]),
    "bmgx_sumsq_win_work_len": (_z, []),
    "bmgx2_sumsq_win": (_i, [_p, _z, _u, _u, _p, _p, _p]),
    "bmgx_sum_ordered": (_i, [_p, _i, _p, _p]),
    "bmgx2_restrict_v": (_i, [_p, _p, _p, _u, _u, _u, _u, _z, _z, _p]),
    "bmgx2_interp_add_v": (_i, [_p, _p, _p, _p, _p, _u, _u, _u, _u, _z, _z, _p]),
    "bmgx3_plane_copy": (_i, [_p, _p, _u, _u, _u, _i, _u, _i, _p]),
    "bmgx3_plane_coeff": (_i, [_p, _p, _u, _u, _u, _i, _i, _u, _p]),
    "bmgx3_plane_rhs": (_i, [_p, _p, _p, _p, _u, _u, _u, _i, _i, _u, _p]),
    "bmgx3_SETUP_recip": (_i, [_p, _p, _u, _u, _u, _i, _i, _p]),
    "bmgx3_relax_GS": (_i, [_i, _p, _p, _p, _p, _u, _u, _u, _i, _i, _i, _i, _i, _i, _p]),
    "bmgx3_relax_GS_oop": (_i, [_i, _p, _p, _p, _p, _p, _u, _u, _u, _i, _i, _i, _i, _i, _i, _p]),
    "bmgx3_residual": (_i, [_i, _i, _i, _p, _p, _p, _p, _u, _u, _u, _i, _p]),
    "bmgx3_matvec": (_i, [_i, _p, _p, _p, _u, _u, _u, _i, _i, _i, _p]),
    "bmgx3_restrict": (_i, [_p, _p, _p, _u, _u, _u, _u, _u, _u, _i, _p]),
    "bmgx3_interp_add": (_i, [_p, _p, _p, _p, _p, _u, _u, _u, _u, _u, _u, _i, _i, _p]),
    "bmgx3_restrict_v": (_i, [_p, _p, _p, _u, _u, _u, _u, _u, _u, _z, _p]),
    "bmgx3_interp_add_v": (_i, [_p, _p, _p, _p, _p, _u, _u, _u, _u, _u, _u, _z, _p]),
    "bmgx3_SETUP_interp_OI": (_i, [_p, _p, _p, _u, _u, _u, _u, _u, _u, _i, _i, _i, _i, _p, _p]),
    "bmgx3_SETUP_ITLI07_ex": (_i, [_p, _p, _p, _u, _u, _u, _u, _u, _u, _i, _p]),
    "bmgx3_SETUP_ITLI27_ex": (_i, [_p, _p, _p, _u, _u, _u, _u, _u, _u, _i, _p]),
    "bmgx3_SETUP_cg_LU": (_i, [_p, _u, _u, _u, _i, _p, _u, _u, _i, _p]),
    "bmgx3_SOLVE_cg": (_i, [_p, _p, _u, _u, _u, _p, _p, _u, _u, _i, _p]),
}

for _name, (_res, _args) in SIGNATURES.items():
    if hasattr(lib, _name):
        _f = getattr(lib, _name)
        _f.restype = _res
        _f.argtypes = _args


class Seg(C.Structure):
    """bmgx_seg of include/bmgx.h"""
    _fields_ = [("src", _p), ("dst", _p), ("rows", C.c_uint32), ("cols", C.c_uint32), ("src_pitch", C.c_uint64),
                ("dst_pitch", C.c_uint64), ("src_rank", C.c_int32), ("reserved", C.c_int32)]


LINE_GROUP_MAX = 16


class LineGroup(C.Structure):
    """bmgx_line_group of include/bmgx.h"""
    _fields_ = [("nmembers", C.c_int), ("first", C.c_int * (LINE_GROUP_MAX + 1)), ("pitch", C.c_size_t),
                ("lb", _p * LINE_GROUP_MAX)]


class LineOwners(C.Structure):
    """bmgx_line_owners of include/bmgx.h"""
    _fields_ = [("nmembers", C.c_int), ("start", C.c_int * (LINE_GROUP_MAX + 1)), ("w_along", C.c_int * LINE_GROUP_MAX),
                ("w_across", C.c_int), ("pitch", C.c_size_t * LINE_GROUP_MAX), ("q", _p * LINE_GROUP_MAX)]


class Rect(C.Structure):
    """bmgx_rect of include/bmgx.h"""
    _fields_ = [(n, C.c_uint32) for n in ("i0", "j0", "ni", "nj", "wi0", "wi1", "wj0", "wj1")]


class BmgxError(RuntimeError):
    pass


def check(status):
    if status != 0:
        raise BmgxError(f"bmgx status {status}: {lib.bmgx_last_error().decode()}")


def declared_symbols(header=None):
    """Names of every function declared in include/bmgx.h (for the export test)."""
    import re
    header = header or os.path.join(_HERE, "..", "include", "bmgx.h")
    txt = open(header).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(bmgx[0-9a-zA-Z_]*)\s*\(", txt)))
