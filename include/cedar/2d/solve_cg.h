// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/2d/solve_cg.h -- CUDA-backed coarsest-grid direct solve (reference
// solve_cg_f90, include/cedar/2d/solve_cg.h:17-65, src/2d/solve_cg.cc:14-30).
#ifndef CEDAR_2D_SOLVE_CG_H
#define CEDAR_2D_SOLVE_CG_H
#include <cedar/2d/types.h>
#include <cedar/kernels/solve_cg.h>
#include <cedar/gpu.h>
namespace cedar { namespace cdr2 {
class solve_cg_b200 : public kernels::solve_cg<stypes>
{
public:
	void setup(const stencil_op<five_pt> &so, grid_func &ABD) override { setup_impl(so, ABD); }
	void setup(const stencil_op<nine_pt> &so, grid_func &ABD) override { setup_impl(so, ABD); }
	template <class sten> void setup_impl(const stencil_op<sten> &so, grid_func &ABD)
	{
		int ibc;
		bmgx_get_bc(params->per_mask(), &ibc);
		gpu::check(bmgx2_SETUP_cg_LU(so.cdev(), so.len(0), so.len(1), stencil_ndirs<sten>::value, ABD.dev(), ABD.len(0),
		                             ABD.len(1), ibc, gpu::stream()), "SETUP_cg_LU");
		// DPBTRF's INFO: a half-factored ABD must not be used (the reference logs "Coarse grid Cholesky decomposition
		// failed"); synchronises, setup phase only
		int info = 0;
		gpu::check(bmgx_cg_LU_info(&info, gpu::stream()), "coarse-grid Cholesky");
		bbd_dev.reshape(ABD.len(1));
		bbd_dev.dev();
	}
	// the host scratch `bbd` of the reference interface is not used: the
	// right-hand side is packed into device scratch owned by this kernel
	void run(grid_func &x, const grid_func &b, const grid_func &ABD, real_t *bbd) override
	{
		(void)bbd;
		// batch context (plane relaxation): one packed right-hand side per batch entry
		if (bbd_dev.size() < static_cast<std::size_t>(ABD.len(1)) * bmgx_get_batch()) bbd_dev.set_batch(bmgx_get_batch());
		int ibc;
		bmgx_get_bc(params->per_mask(), &ibc);
		gpu::check(bmgx2_SOLVE_cg(x.dev(), b.cdev(), x.len(0), x.len(1), ABD.cdev(), bbd_dev.dev(), ABD.len(0),
		                          ABD.len(1), ibc, gpu::stream()), "SOLVE_cg");
	}
private:
	array<real_t, 1> bbd_dev;
};
}}
#endif
