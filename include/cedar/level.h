// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/level.h -- what a multilevel solve stores per level (reference: level.h:15-41)
#ifndef CEDAR_LEVEL_H
#define CEDAR_LEVEL_H

#include <array>
#include <functional>

namespace cedar {

template <class sten, class solver_types> struct level
{
	template <class stencil> using stencil_op = typename solver_types::template stencil_op<stencil>;
	using grid_func = typename solver_types::grid_func;
	using prolong_op = typename solver_types::prolong_op;
	using restrict_op = typename solver_types::restrict_op;
	using relax_stencil = typename solver_types::relax_stencil;
	using stypes = solver_types;

	// finest level: wraps the caller's operator
	level(stencil_op<sten> &A) : A(A) {}
	// coarse levels: own everything, dimensioned by the coarse grid
	template <class... Args>
	level(Args &&...args) : Adata(args...), A(Adata), P(args...), x(args...), res(args...), b(args...) {}

	stencil_op<sten> Adata;
	stencil_op<sten> &A;
	prolong_op P;
	restrict_op R;
	grid_func x;
	grid_func res;
	grid_func b;
	std::array<relax_stencil, 2> SOR;

	std::function<void(const stencil_op<sten> &A, grid_func &x, const grid_func &b)> presmoother;
	std::function<void(const stencil_op<sten> &A, grid_func &x, const grid_func &b)> postsmoother;
};

}
#endif
