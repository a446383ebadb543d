// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/kernels/residual.h -- abstract kernel interface `kernels::residual`.
// Signatures follow the reference include/cedar/kernels/residual.h (l.19-26); an
// implementation overrides the overloads it supports, anything left out logs
// "routine not provided" instead of throwing (the reference's error behaviour).
#ifndef CEDAR_KERNELS_RESIDUAL_H
#define CEDAR_KERNELS_RESIDUAL_H

#include <cedar/kernel.h>
#include <cedar/types.h>

namespace cedar { namespace kernels {

template <class solver_types> class residual : public kernel<solver_types>
{
public:
	template <class sten> using stencil_op = typename kernel<solver_types>::template stencil_op<sten>;
	using comp_sten = typename kernel<solver_types>::comp_sten;
	using full_sten = typename kernel<solver_types>::full_sten;
	using grid_func = typename kernel<solver_types>::grid_func;
	using prolong_op = typename kernel<solver_types>::prolong_op;
	using restrict_op = typename kernel<solver_types>::restrict_op;
	using relax_stencil = typename kernel<solver_types>::relax_stencil;
	static const std::string name() { return "residual"; }
	virtual void run(const stencil_op<comp_sten> &so, const grid_func &x, const grid_func &b, grid_func &r) { log::error << name() << ": run routine not provided" << std::endl; }
	virtual void run(const stencil_op<full_sten> &so, const grid_func &x, const grid_func &b, grid_func &r) { log::error << name() << ": run routine not provided" << std::endl; }
};

}}
#endif
