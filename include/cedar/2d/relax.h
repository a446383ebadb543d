// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/2d/relax.h -- CUDA-backed point and line relaxation kernels, registered
// under the same interfaces as the reference's rbgs / lines<rdir>
// (include/cedar/2d/relax.h:28-190): k = kf = 1, ifd from the stencil type,
// BMG_RELAX_SYM always (the relax_symmetric parameter is not consulted there
// either), boundary code from kernel_params::per_mask().
#ifndef CEDAR_2D_RELAX_H
#define CEDAR_2D_RELAX_H

#include <map>
#include <type_traits>

#include <cedar/2d/types.h>
#include <cedar/kernels/point_relax.h>
#include <cedar/kernels/line_relax.h>
#include <cedar/gpu.h>

namespace cedar { namespace cdr2 {

class rbgs : public kernels::point_relax<stypes>
{
public:
	void setup(const stencil_op<five_pt> &so, relax_stencil &sor) override { setup_impl(so, sor); }
	void setup(const stencil_op<nine_pt> &so, relax_stencil &sor) override { setup_impl(so, sor); }
	void run(const stencil_op<five_pt> &so, grid_func &x, const grid_func &b, const relax_stencil &sor,
	         cycle::Dir cdir) override { run_impl(so, x, b, sor, cdir); }
	void run(const stencil_op<nine_pt> &so, grid_func &x, const grid_func &b, const relax_stencil &sor,
	         cycle::Dir cdir) override { run_impl(so, x, b, sor, cdir); }

	template <class sten> void setup_impl(const stencil_op<sten> &so, relax_stencil &sor)
	{
		gpu::check(bmgx2_SETUP_recip(so.cdev(), sor.dev(), so.len(0), so.len(1), stencil_ndirs<sten>::value, 2,
		                             gpu::stream()), "SETUP_recip");
	}
	template <class sten>
	void run_impl(const stencil_op<sten> &so, grid_func &x, const grid_func &b, const relax_stencil &sor, cycle::Dir cdir)
	{
		int ifd = std::is_same<sten, five_pt>::value ? 1 : 0;
		int updown = cdir == cycle::Dir::UP ? BMGX_UP : BMGX_DOWN;
		int ibc;
		bmgx_get_bc(params->per_mask(), &ibc);
		if (ibc != BMGX_BCs_definite && ibc != BMGX_BCs_indef_nonper) {
			// periodic boundaries: the in-place reference form, which refreshes the ghost columns after every colour
			// and the ghost rows after the sweep (BMG2_SymStd_relax_GS.f90:139-224); the fused single-pass sweep has
			// no periodic form
			gpu::check(bmgx2_relax_GS(1, so.cdev(), b.cdev(), x.dev(), sor.cdev(), so.len(0), so.len(1), 1, ifd,
			                          stencil_ndirs<sten>::value, 2, BMGX_RELAX_SYM, updown, ibc, gpu::stream()), "relax_GS");
			return;
		}
		// one pass over memory for all colours: relax out of place into a scratch
		// array of the same shape, then exchange the two device buffers (x keeps
		// its identity on the host side; its device pointer alternates)
		auto &tmp = scratch[std::make_pair(x.len(0), x.len(1))];
		if (tmp.size() != x.size()) tmp = grid_func(x.shape(0), x.shape(1));
		const real_t *xin = x.cdev();
		gpu::check(bmgx2_relax_GS_oop(1, so.cdev(), b.cdev(), xin, tmp.dev(), sor.cdev(), so.len(0), so.len(1), 1, ifd,
		                              stencil_ndirs<sten>::value, 2, BMGX_RELAX_SYM, updown, ibc, gpu::stream()),
		           "relax_GS_oop");
		x.swap_device(tmp);
	}
private:
	std::map<std::pair<len_t, len_t>, grid_func> scratch;
};

template <relax_dir rdir> class lines : public kernels::line_relax<stypes, rdir>
{
public:
	void setup(const stencil_op<five_pt> &so, relax_stencil &sor) override { setup_impl(so, sor); }
	void setup(const stencil_op<nine_pt> &so, relax_stencil &sor) override { setup_impl(so, sor); }
	void run(const stencil_op<five_pt> &so, grid_func &x, const grid_func &b, const relax_stencil &sor, grid_func &res,
	         cycle::Dir cdir) override { run_impl(so, x, b, sor, res, cdir); }
	void run(const stencil_op<nine_pt> &so, grid_func &x, const grid_func &b, const relax_stencil &sor, grid_func &res,
	         cycle::Dir cdir) override { run_impl(so, x, b, sor, res, cdir); }

	template <class sten> void setup_impl(const stencil_op<sten> &so, relax_stencil &sor)
	{
		int jpn;
		bmgx_get_bc(this->params->per_mask(), &jpn);
		if (rdir == relax_dir::x)
			gpu::check(bmgx2_SETUP_lines_x(so.cdev(), sor.dev(), so.len(0), so.len(1), stencil_ndirs<sten>::value, jpn,
			                               gpu::stream()), "SETUP_lines_x");
		else
			gpu::check(bmgx2_SETUP_lines_y(so.cdev(), sor.dev(), so.len(0), so.len(1), stencil_ndirs<sten>::value, jpn,
			                               gpu::stream()), "SETUP_lines_y");
	}
	template <class sten>
	void run_impl(const stencil_op<sten> &so, grid_func &x, const grid_func &b, const relax_stencil &sor,
	              grid_func &res, cycle::Dir cdir)
	{
		int ifd = std::is_same<sten, five_pt>::value ? 1 : 0;
		int updown = cdir == cycle::Dir::UP ? BMGX_UP : BMGX_DOWN;
		int ibc;
		bmgx_get_bc(this->params->per_mask(), &ibc);
		if (rdir == relax_dir::x)
			gpu::check(bmgx2_relax_lines_x(1, so.cdev(), b.cdev(), x.dev(), sor.cdev(), res.dev(), so.len(0), so.len(1),
			                               1, ifd, stencil_ndirs<sten>::value, BMGX_RELAX_SYM, updown, ibc,
			                               gpu::stream()), "relax_lines_x");
		else {
			// The y sweep needs ii*jj doubles for the transposed right-hand sides.  The reference uses `res` as its 2*jj
			// work vector and clears it again (relax_lines_y.f90:229-231), so the ghost layer of res stays zero -- which
			// the periodic restriction relies on where it does not refresh the fine vector's ghosts.  With periodic
			// boundaries the scratch is therefore an array of the kernel's own; otherwise res serves (nothing reads its
			// ghost layer).
			real_t *work = res.dev();
			if (ibc != BMGX_BCs_definite && ibc != BMGX_BCs_indef_nonper) {
				auto &tmp = scratch[std::make_pair(x.len(0), x.len(1))];
				if (tmp.size() != x.size()) tmp = grid_func(x.shape(0), x.shape(1));
				work = tmp.dev();
			}
			gpu::check(bmgx2_relax_lines_y(1, so.cdev(), b.cdev(), x.dev(), sor.cdev(), work, so.len(0), so.len(1),
			                               1, ifd, stencil_ndirs<sten>::value, BMGX_RELAX_SYM, updown, ibc,
			                               gpu::stream()), "relax_lines_y");
		}
	}
private:
	std::map<std::pair<len_t, len_t>, grid_func> scratch;
};

}}
#endif
