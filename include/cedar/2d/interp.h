// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/2d/interp.h -- CUDA-backed interpolate-and-add and operator-induced
// interpolation setup (reference interp_f90 / setup_interp_f90, src/2d/interp.cc:19-106).
#ifndef CEDAR_2D_INTERP_H
#define CEDAR_2D_INTERP_H
#include <cedar/2d/types.h>
#include <cedar/kernels/interp_add.h>
#include <cedar/kernels/setup_interp.h>
#include <cedar/gpu.h>
namespace cedar { namespace cdr2 {

class interp_b200 : public kernels::interp_add<stypes>
{
public:
	// `residual` is const in the interface but scaled by 1/diag in place, as in
	// the reference (BMG2_SymStd_interp_add.f90:101-105)
	void run(const prolong_op &P, const grid_func &coarse, const grid_func &residual, grid_func &fine) override
	{
		grid_func &res = const_cast<grid_func &>(residual);
		int nstencil;
		const real_t *fop;
		if (P.fine_is_five) { nstencil = 3; fop = P.fine_op_five->cdev(); }
		else { nstencil = 5; fop = P.fine_op_nine->cdev(); }
		int ibc;
		bmgx_get_bc(params->per_mask(), &ibc);
		gpu::check(bmgx2_interp_add(fine.dev(), coarse.cdev(), res.dev(), fop, P.cdev(), coarse.len(0), coarse.len(1),
		                            fine.len(0), fine.len(1), nstencil, ibc, gpu::stream()), "interp_add");
	}
};

class setup_interp_b200 : public kernels::setup_interp<stypes>
{
public:
	void run(const stencil_op<five_pt> &fop, const stencil_op<nine_pt> &cop, prolong_op &P) override
	{
		P.fine_op_five = const_cast<stencil_op<five_pt> *>(&fop);
		P.fine_is_five = true;
		launch(fop.cdev(), fop.len(0), fop.len(1), cop, P, 1, 3);
	}
	void run(const stencil_op<nine_pt> &fop, const stencil_op<nine_pt> &cop, prolong_op &P) override
	{
		P.fine_op_nine = const_cast<stencil_op<nine_pt> *>(&fop);
		P.fine_is_five = false;
		launch(fop.cdev(), fop.len(0), fop.len(1), cop, P, 0, 5);
	}
private:
	void launch(const real_t *fop, len_t iif, len_t jjf, const stencil_op<nine_pt> &cop, prolong_op &P, int ifd, int nstencil)
	{
		int jpn;
		bmgx_get_bc(params->per_mask(), &jpn);
		// the coarse operator is only passed for its extents (it does not exist yet)
		gpu::check(bmgx2_SETUP_interp_OI(fop, nullptr, P.dev(), iif, jjf, cop.len(0), cop.len(1), ifd, nstencil, jpn, 0,
		                                 gpu::stream()), "SETUP_interp_OI");
	}
};

}}
#endif
