// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/kernels/interp_add.h -- abstract kernel interface `kernels::interp_add`.
// Signatures follow the reference include/cedar/kernels/interp_add.h (l.17-20); an
// implementation overrides the overloads it supports, anything left out logs
// "routine not provided" instead of throwing (the reference's error behaviour).
#ifndef CEDAR_KERNELS_INTERP_ADD_H
#define CEDAR_KERNELS_INTERP_ADD_H

#include <cedar/kernel.h>
#include <cedar/types.h>

namespace cedar { namespace kernels {

template <class solver_types> class interp_add : public kernel<solver_types>
{
public:
	template <class sten> using stencil_op = typename kernel<solver_types>::template stencil_op<sten>;
	using comp_sten = typename kernel<solver_types>::comp_sten;
	using full_sten = typename kernel<solver_types>::full_sten;
	using grid_func = typename kernel<solver_types>::grid_func;
	using prolong_op = typename kernel<solver_types>::prolong_op;
	using restrict_op = typename kernel<solver_types>::restrict_op;
	using relax_stencil = typename kernel<solver_types>::relax_stencil;
	static const std::string name() { return "interpolation and add"; }
	// fine += P * coarse + residual / diag(A) at F-points; `residual` is scaled in place
	virtual void run(const prolong_op &P, const grid_func &coarse, const grid_func &residual, grid_func &fine) { log::error << name() << ": run routine not provided" << std::endl; }
};

}}
#endif
