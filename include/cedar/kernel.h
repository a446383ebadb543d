// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/kernel.h -- base of every kernel implementation (reference: kernel.h:11-38)
#ifndef CEDAR_KERNEL_H
#define CEDAR_KERNEL_H

#include <functional>
#include <memory>
#include <string>

#include <cedar/halo_exchanger_base.h>
#include <cedar/kernel_params.h>
#include <cedar/util/log.h>

namespace cedar {

template <class solver_types> class kernel
{
public:
	template <class sten> using stencil_op = typename solver_types::template stencil_op<sten>;
	using comp_sten = typename solver_types::comp_sten;
	using full_sten = typename solver_types::full_sten;
	using grid_func = typename solver_types::grid_func;
	using prolong_op = typename solver_types::prolong_op;
	using restrict_op = typename solver_types::restrict_op;
	using relax_stencil = typename solver_types::relax_stencil;

	virtual ~kernel() {}
	void add_params(std::shared_ptr<kernel_params> params) { this->params = params; }
	void add_halo(halo_exchanger_base *halof) { this->halof = halof; }
protected:
	std::shared_ptr<kernel_params> params;
	halo_exchanger_base *halof = nullptr;
};

}
#endif
