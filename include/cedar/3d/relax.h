// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/3d/relax.h -- CUDA-backed 3D point relaxation, registered under the same
// interface as the reference's rbgs (include/cedar/3d/relax.h:22-100): ifd from the
// stencil type, BMG_RELAX_SYM always, boundary code from kernel_params::per_mask().
#ifndef CEDAR_3D_RELAX_H
#define CEDAR_3D_RELAX_H

#include <map>
#include <tuple>
#include <type_traits>

#include <cedar/3d/types.h>
#include <cedar/kernels/point_relax.h>
#include <cedar/gpu.h>

namespace cedar { namespace cdr3 {

class rbgs : public kernels::point_relax<stypes>
{
public:
	void setup(const stencil_op<seven_pt> &so, relax_stencil &sor) override { setup_impl(so, sor); }
	void setup(const stencil_op<xxvii_pt> &so, relax_stencil &sor) override { setup_impl(so, sor); }
	void run(const stencil_op<seven_pt> &so, grid_func &x, const grid_func &b, const relax_stencil &sor,
	         cycle::Dir cdir) override { run_impl(so, x, b, sor, cdir); }
	void run(const stencil_op<xxvii_pt> &so, grid_func &x, const grid_func &b, const relax_stencil &sor,
	         cycle::Dir cdir) override { run_impl(so, x, b, sor, cdir); }

	template <class sten> void setup_impl(const stencil_op<sten> &so, relax_stencil &sor)
	{
		gpu::check(bmgx3_SETUP_recip(so.cdev(), sor.dev(), so.len(0), so.len(1), so.len(2), stencil_ndirs<sten>::value,
		                             2, gpu::stream()), "SETUP_recip3");
	}
	template <class sten>
	void run_impl(const stencil_op<sten> &so, grid_func &x, const grid_func &b, const relax_stencil &sor, cycle::Dir cdir)
	{
		int ifd = std::is_same<sten, seven_pt>::value ? 1 : 0;
		int updown = cdir == cycle::Dir::UP ? BMGX_UP : BMGX_DOWN;
		int ibc;
		bmgx_get_bc(params->per_mask(), &ibc);
		if (ifd == 1) {
			// 7-point: both colours in one pass over memory, out of place into a scratch array of
			// the same shape, then exchange the two device buffers (as cdr2::rbgs does)
			auto &tmp = scratch[std::make_tuple(x.len(0), x.len(1), x.len(2))];
			if (tmp.size() != x.size()) tmp = grid_func(x.shape(0), x.shape(1), x.shape(2));
			const real_t *xin = x.cdev();
			gpu::check(bmgx3_relax_GS_oop(1, so.cdev(), b.cdev(), xin, tmp.dev(), sor.cdev(), so.len(0), so.len(1),
			                              so.len(2), ifd, stencil_ndirs<sten>::value, 2, BMGX_RELAX_SYM, updown, ibc,
			                              gpu::stream()), "relax_GS3_oop");
			x.swap_device(tmp);
			return;
		}
		gpu::check(bmgx3_relax_GS(1, so.cdev(), b.cdev(), x.dev(), sor.cdev(), so.len(0), so.len(1), so.len(2), ifd,
		                          stencil_ndirs<sten>::value, 2, BMGX_RELAX_SYM, updown, ibc, gpu::stream()),
		           "relax_GS3");
	}
private:
	std::map<std::tuple<len_t, len_t, len_t>, grid_func> scratch;
};

}}
#endif
