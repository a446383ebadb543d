// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/2d/restrict.h -- CUDA-backed restriction R = P^T (reference restrict_f90, src/2d/restrict.cc:16-32)
#ifndef CEDAR_2D_INTER_RESTRICT_H
#define CEDAR_2D_INTER_RESTRICT_H
#include <cedar/2d/types.h>
#include <cedar/kernels/restrict.h>
#include <cedar/gpu.h>
namespace cedar { namespace cdr2 {
class restrict_b200 : public kernels::restriction<stypes>
{
public:
	void run(const restrict_op &R, const grid_func &fine, grid_func &coarse) override
	{
		const prolong_op &P = R.getP();
		int ibc;
		bmgx_get_bc(params->per_mask(), &ibc);
		// with periodic boundaries the kernel refreshes the ghost layer of the FINE vector first, as the reference does
		// (BMG2_SymStd_restrict.f90:96-111): take the mutable pointer then, so that a host mirror is marked stale
		const bool periodic = params->periodic[0] || params->periodic[1];
		const real_t *fdev = periodic ? const_cast<grid_func &>(fine).dev() : fine.cdev();
		gpu::check(bmgx2_restrict(fdev, coarse.dev(), P.cdev(), fine.len(0), fine.len(1), coarse.len(0),
		                          coarse.len(1), ibc, gpu::stream()), "restrict");
	}
};
}}
#endif
