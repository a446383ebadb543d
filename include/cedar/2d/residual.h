// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/2d/residual.h -- CUDA-backed residual kernel (reference residual_f90,
// include/cedar/2d/residual.h:14-62: k = kf = 0, ifd from the stencil type).
#ifndef CEDAR_2D_RESIDUAL_H
#define CEDAR_2D_RESIDUAL_H
#include <type_traits>
#include <cedar/2d/types.h>
#include <cedar/kernels/residual.h>
#include <cedar/gpu.h>
namespace cedar { namespace cdr2 {
class residual_b200 : public kernels::residual<stypes>
{
public:
	void run(const stencil_op<five_pt> &so, const grid_func &x, const grid_func &b, grid_func &r) override { run_impl(so, x, b, r); }
	void run(const stencil_op<nine_pt> &so, const grid_func &x, const grid_func &b, grid_func &r) override { run_impl(so, x, b, r); }
	template <class sten> void run_impl(const stencil_op<sten> &so, const grid_func &x, const grid_func &b, grid_func &r)
	{
		int ifd = std::is_same<sten, five_pt>::value ? 1 : 0;
		int ibc;
		bmgx_get_bc(params->per_mask(), &ibc);
		gpu::check(bmgx2_residual(0, so.cdev(), b.cdev(), x.cdev(), r.dev(), r.len(0), r.len(1), 0, ifd,
		                          stencil_ndirs<sten>::value, ibc, 0, 0, 0, gpu::stream()), "residual");
	}
};
}}
#endif
