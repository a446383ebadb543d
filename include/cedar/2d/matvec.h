// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/2d/matvec.h -- CUDA-backed operator application y = A x (reference matvec_f90 over
// BMG2_SymStd_UTILS_matvec, include/cedar/2d/mpi/matvec.h / include/cedar/2d/matvec.h:20-50: k = kf, ifd from
// the stencil type; the reference offers it on its MPI kernel list, here the serial registry carries it too).
#ifndef CEDAR_2D_MATVEC_H
#define CEDAR_2D_MATVEC_H
#include <type_traits>
#include <cedar/2d/types.h>
#include <cedar/kernels/matvec.h>
#include <cedar/gpu.h>
namespace cedar { namespace cdr2 {
class matvec_b200 : public kernels::matvec<stypes>
{
public:
	void run(const stencil_op<five_pt> &so, const grid_func &x, grid_func &y) override { run_impl(so, x, y); }
	void run(const stencil_op<nine_pt> &so, const grid_func &x, grid_func &y) override { run_impl(so, x, y); }
	template <class sten> void run_impl(const stencil_op<sten> &so, const grid_func &x, grid_func &y)
	{
		int ifd = std::is_same<sten, five_pt>::value ? 1 : 0;
		gpu::check(bmgx2_matvec(1, so.cdev(), y.dev(), x.cdev(), y.len(0), y.len(1), 1, ifd,
		                        stencil_ndirs<sten>::value, gpu::stream()), "matvec");
	}
};
}}
#endif
