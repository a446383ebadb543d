// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/3d/residual.h -- CUDA-backed 3D residual (reference residual_f90,
// include/cedar/3d/residual.h:16-62: kg = nog = 1, ifd from the stencil type).
#ifndef CEDAR_3D_RESIDUAL_H
#define CEDAR_3D_RESIDUAL_H
#include <type_traits>
#include <cedar/3d/types.h>
#include <cedar/kernels/residual.h>
#include <cedar/gpu.h>
namespace cedar { namespace cdr3 {
class residual_b200 : public kernels::residual<stypes>
{
public:
	void run(const stencil_op<seven_pt> &so, const grid_func &x, const grid_func &b, grid_func &r) override { run_impl(so, x, b, r); }
	void run(const stencil_op<xxvii_pt> &so, const grid_func &x, const grid_func &b, grid_func &r) override { run_impl(so, x, b, r); }
	template <class sten> void run_impl(const stencil_op<sten> &so, const grid_func &x, const grid_func &b, grid_func &r)
	{
		int ifd = std::is_same<sten, seven_pt>::value ? 1 : 0;
		gpu::check(bmgx3_residual(1, 1, ifd, x.cdev(), b.cdev(), so.cdev(), r.dev(), r.len(0), r.len(1), r.len(2),
		                          stencil_ndirs<sten>::value, gpu::stream()), "residual3");
	}
};
}}
#endif
