// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_2D_INTER_PROLONG_OP_H
#define CEDAR_2D_INTER_PROLONG_OP_H
#include <iostream>
#include <cedar/types.h>
#include <cedar/2d/stencil_op.h>
#include <cedar/2d/grid_func.h>
namespace cedar { namespace cdr2 {
// interpolation weights, dimensioned on the COARSE grid; slot order is the
// Fortran LL,LR,LA,LB,LSW,LNW,LNE,LSE minus one (reference 2d/prolong_op.h:14-27)
enum class inter_dir { L = 0, R = 1, A = 2, B = 3, SW = 4, NW = 5, NE = 6, SE = 7, ndirs };
class prolong_op : public stencil_op<inter_dir>
{
public:
	prolong_op() {}
	prolong_op(len_t nx, len_t ny) : stencil_op<inter_dir>(nx, ny) {}
	friend std::ostream &operator<<(std::ostream &os, const prolong_op &P);
	// the fine operator this interpolation was built from (set by setup_interp,
	// read by interp_add for the diagonal scaling)
	stencil_op<five_pt> *fine_op_five = nullptr;
	stencil_op<nine_pt> *fine_op_nine = nullptr;
	grid_func *residual = nullptr;
	bool fine_is_five = false;
};
}}
#endif
