// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_3D_INTER_PROLONG_OP_H
#define CEDAR_3D_INTER_PROLONG_OP_H
#include <cedar/types.h>
#include <cedar/3d/stencil_op.h>
#include <cedar/3d/grid_func.h>
namespace cedar { namespace cdr3 {
// 26 interpolation weights dimensioned on the COARSE grid; slot order is the Fortran
// lxyl..ltse minus one (reference 3d/prolong_op.h:13-18, BMG_stencils_f90.h:52-64)
enum class inter_dir { XYL = 0, XYR = 1, XYA = 2, XYB = 3, XZA = 4, XZB = 5, XYNE = 6, XYSE = 7, XYSW = 8, XYNW = 9,
                       XZSW = 10, XZNW = 11, XZNE = 12, XZSE = 13, YZSW = 14, YZNW = 15, YZNE = 16, YZSE = 17, BSW = 18,
                       BNW = 19, BNE = 20, BSE = 21, TSW = 22, TNW = 23, TNE = 24, TSE = 25, ndirs };
class prolong_op : public stencil_op<inter_dir>
{
public:
	prolong_op() {}
	prolong_op(len_t nx, len_t ny, len_t nz) : stencil_op<inter_dir>(nx, ny, nz) {}
	// the fine operator this interpolation was built from (set by setup_interp,
	// read by interp_add for the diagonal scaling)
	stencil_op<seven_pt> *fine_op_seven = nullptr;
	stencil_op<xxvii_pt> *fine_op_xxvii = nullptr;
	grid_func *residual = nullptr;
	bool fine_is_seven = false;
};
}}
#endif
