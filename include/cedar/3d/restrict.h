// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/3d/restrict.h -- CUDA-backed 3D restriction R = P^T (reference restrict_f90, src/3d/restrict.cc:14-32)
#ifndef CEDAR_3D_INTER_RESTRICT_H
#define CEDAR_3D_INTER_RESTRICT_H
#include <cedar/3d/types.h>
#include <cedar/kernels/restrict.h>
#include <cedar/gpu.h>
namespace cedar { namespace cdr3 {
class restrict_b200 : public kernels::restriction<stypes>
{
public:
	void run(const restrict_op &R, const grid_func &fine, grid_func &coarse) override
	{
		const prolong_op &P = R.getP();
		int ibc;
		bmgx_get_bc(params->per_mask(), &ibc);
		gpu::check(bmgx3_restrict(fine.cdev(), coarse.dev(), P.cdev(), fine.len(0), fine.len(1), fine.len(2),
		                          coarse.len(0), coarse.len(1), coarse.len(2), ibc, gpu::stream()), "restrict3");
	}
};
}}
#endif
