// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/3d/solve_cg.h -- CUDA-backed coarsest-grid direct solve in 3D (reference
// solve_cg_f90, include/cedar/3d/solve_cg.h, src/3d/solve_cg.cc).
#ifndef CEDAR_3D_SOLVE_CG_H
#define CEDAR_3D_SOLVE_CG_H
#include <cedar/3d/types.h>
#include <cedar/kernels/solve_cg.h>
#include <cedar/gpu.h>
namespace cedar { namespace cdr3 {
class solve_cg_b200 : public kernels::solve_cg<stypes>
{
public:
	void setup(const stencil_op<seven_pt> &so, grid_func &ABD) override { setup_impl(so, ABD); }
	void setup(const stencil_op<xxvii_pt> &so, grid_func &ABD) override { setup_impl(so, ABD); }
	template <class sten> void setup_impl(const stencil_op<sten> &so, grid_func &ABD)
	{
		int ibc;
		bmgx_get_bc(params->per_mask(), &ibc);
		gpu::check(bmgx3_SETUP_cg_LU(so.cdev(), so.len(0), so.len(1), so.len(2), stencil_ndirs<sten>::value, ABD.dev(),
		                             ABD.len(0), ABD.len(1), ibc, gpu::stream()), "SETUP_cg_LU3");
		// DPBTRF's INFO: a half-factored ABD must not be used (the reference logs "Coarse grid Cholesky decomposition
		// failed"); synchronises, setup phase only
		int info = 0;
		gpu::check(bmgx_cg_LU_info(&info, gpu::stream()), "coarse-grid Cholesky");
		bbd_dev.reshape(ABD.len(1));
		bbd_dev.dev();
	}
	void run(grid_func &x, const grid_func &b, const grid_func &ABD, real_t *bbd) override
	{
		(void)bbd;
		int ibc;
		bmgx_get_bc(params->per_mask(), &ibc);
		gpu::check(bmgx3_SOLVE_cg(x.dev(), b.cdev(), x.len(0), x.len(1), x.len(2), ABD.cdev(), bbd_dev.dev(), ABD.len(0),
		                          ABD.len(1), ibc, gpu::stream()), "SOLVE_cg3");
	}
private:
	array<real_t, 1> bbd_dev;
};
}}
#endif
