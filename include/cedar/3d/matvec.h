// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/3d/matvec.h -- CUDA-backed 3D operator application y = A x (reference matvec_f90 over
// BMG3_SymStd_UTILS_matvec, include/cedar/3d/mpi/matvec.h: kg = nog, ifd from the stencil type).
#ifndef CEDAR_3D_MATVEC_H
#define CEDAR_3D_MATVEC_H
#include <type_traits>
#include <cedar/3d/types.h>
#include <cedar/kernels/matvec.h>
#include <cedar/gpu.h>
namespace cedar { namespace cdr3 {
class matvec_b200 : public kernels::matvec<stypes>
{
public:
	void run(const stencil_op<seven_pt> &so, const grid_func &x, grid_func &y) override { run_impl(so, x, y); }
	void run(const stencil_op<xxvii_pt> &so, const grid_func &x, grid_func &y) override { run_impl(so, x, y); }
	template <class sten> void run_impl(const stencil_op<sten> &so, const grid_func &x, grid_func &y)
	{
		int ifd = std::is_same<sten, seven_pt>::value ? 1 : 0;
		gpu::check(bmgx3_matvec(1, so.cdev(), y.dev(), x.cdev(), y.len(0), y.len(1), y.len(2), 1, ifd,
		                        stencil_ndirs<sten>::value, gpu::stream()), "matvec3");
	}
};
}}
#endif
