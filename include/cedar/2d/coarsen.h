// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/2d/coarsen.h -- CUDA-backed Galerkin coarse-operator construction
// (reference class galerkin, include/cedar/2d/coarsen.h:18-61).
#ifndef CEDAR_2D_COARSEN_H
#define CEDAR_2D_COARSEN_H
#include <type_traits>
#include <cedar/2d/types.h>
#include <cedar/kernels/coarsen_op.h>
#include <cedar/gpu.h>
namespace cedar { namespace cdr2 {
class galerkin : public kernels::coarsen_op<stypes>
{
public:
	void run(const prolong_op &P, const stencil_op<five_pt> &fop, stencil_op<nine_pt> &cop) override { run_impl(P, fop, cop); }
	void run(const prolong_op &P, const stencil_op<nine_pt> &fop, stencil_op<nine_pt> &cop) override { run_impl(P, fop, cop); }
	template <class sten> void run_impl(const prolong_op &P, const stencil_op<sten> &fop, stencil_op<nine_pt> &cop)
	{
		int ifd = std::is_same<sten, five_pt>::value ? 1 : 0;
		int ipn;
		bmgx_get_bc(params->per_mask(), &ipn);
		gpu::check(bmgx2_SETUP_ITLI_ex(fop.cdev(), cop.dev(), P.cdev(), fop.len(0), fop.len(1), cop.len(0), cop.len(1),
		                               ifd, stencil_ndirs<sten>::value, ipn, gpu::stream()), "SETUP_ITLI_ex");
	}
};
}}
#endif
