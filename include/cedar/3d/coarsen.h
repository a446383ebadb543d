// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/3d/coarsen.h -- CUDA-backed 3D Galerkin coarse-operator construction
// (reference class galerkin, include/cedar/3d/coarsen.h:23-62).
#ifndef CEDAR_3D_COARSEN_H
#define CEDAR_3D_COARSEN_H
#include <type_traits>
#include <cedar/3d/types.h>
#include <cedar/kernels/coarsen_op.h>
#include <cedar/gpu.h>
namespace cedar { namespace cdr3 {
class galerkin : public kernels::coarsen_op<stypes>
{
public:
	void run(const prolong_op &P, const stencil_op<seven_pt> &fop, stencil_op<xxvii_pt> &cop) override { run_impl(P, fop, cop); }
	void run(const prolong_op &P, const stencil_op<xxvii_pt> &fop, stencil_op<xxvii_pt> &cop) override { run_impl(P, fop, cop); }
	template <class sten> void run_impl(const prolong_op &P, const stencil_op<sten> &fop, stencil_op<xxvii_pt> &cop)
	{
		int ipn;
		bmgx_get_bc(params->per_mask(), &ipn);
		if (std::is_same<sten, seven_pt>::value)
			gpu::check(bmgx3_SETUP_ITLI07_ex(fop.cdev(), cop.dev(), P.cdev(), fop.len(0), fop.len(1), fop.len(2),
			                                 cop.len(0), cop.len(1), cop.len(2), ipn, gpu::stream()), "SETUP_ITLI07_ex");
		else
			gpu::check(bmgx3_SETUP_ITLI27_ex(fop.cdev(), cop.dev(), P.cdev(), fop.len(0), fop.len(1), fop.len(2),
			                                 cop.len(0), cop.len(1), cop.len(2), ipn, gpu::stream()), "SETUP_ITLI27_ex");
	}
};
}}
#endif
