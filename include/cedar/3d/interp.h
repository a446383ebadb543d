// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/3d/interp.h -- CUDA-backed 3D interpolate-and-add and operator-induced
// interpolation setup (reference interp_f90 / setup_interp_f90: src/3d/interp.cc:17-45,
// include/cedar/3d/interp.h:20-80).
#ifndef CEDAR_3D_INTERP_H
#define CEDAR_3D_INTERP_H
#include <cedar/3d/types.h>
#include <cedar/kernels/interp_add.h>
#include <cedar/kernels/setup_interp.h>
#include <cedar/gpu.h>
namespace cedar { namespace cdr3 {

class interp_b200 : public kernels::interp_add<stypes>
{
public:
	// `residual` is const in the interface but scaled by 1/diag in place, as in
	// the reference (BMG3_SymStd_interp_add.f90:88-94)
	void run(const prolong_op &P, const grid_func &coarse, const grid_func &residual, grid_func &fine) override
	{
		grid_func &res = const_cast<grid_func &>(residual);
		int nstencil;
		const real_t *fop;
		if (P.fine_is_seven) { nstencil = 4; fop = P.fine_op_seven->cdev(); }
		else { nstencil = 14; fop = P.fine_op_xxvii->cdev(); }
		int ibc;
		bmgx_get_bc(params->per_mask(), &ibc);
		gpu::check(bmgx3_interp_add(fine.dev(), coarse.cdev(), fop, res.dev(), P.cdev(), coarse.len(0), coarse.len(1),
		                            coarse.len(2), fine.len(0), fine.len(1), fine.len(2), nstencil, ibc, gpu::stream()),
		           "interp_add3");
	}
};

class setup_interp_b200 : public kernels::setup_interp<stypes>
{
public:
	void run(const stencil_op<seven_pt> &fop, const stencil_op<xxvii_pt> &cop, prolong_op &P) override
	{
		P.fine_op_seven = const_cast<stencil_op<seven_pt> *>(&fop);
		P.fine_is_seven = true;
		launch(fop.cdev(), fop.len(0), fop.len(1), fop.len(2), cop, P, 1, 4);
	}
	void run(const stencil_op<xxvii_pt> &fop, const stencil_op<xxvii_pt> &cop, prolong_op &P) override
	{
		P.fine_op_xxvii = const_cast<stencil_op<xxvii_pt> *>(&fop);
		P.fine_is_seven = false;
		launch(fop.cdev(), fop.len(0), fop.len(1), fop.len(2), cop, P, 0, 14);
	}
private:
	void launch(const real_t *fop, len_t iif, len_t jjf, len_t kkf, const stencil_op<xxvii_pt> &cop, prolong_op &P,
	            int ifd, int nstencil)
	{
		int jpn;
		bmgx_get_bc(params->per_mask(), &jpn);
		// the coarse operator is only passed for its extents; the reference's `yo`
		// scratch array (interp.h:60-61) is not needed by the device kernels
		gpu::check(bmgx3_SETUP_interp_OI(fop, nullptr, P.dev(), iif, jjf, kkf, cop.len(0), cop.len(1), cop.len(2), ifd,
		                                 nstencil, 0, jpn, nullptr, gpu::stream()), "SETUP_interp_OI3");
	}
};

}}
#endif
