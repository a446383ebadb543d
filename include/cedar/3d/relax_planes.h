// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/3d/relax_planes.h -- plane relaxation: zebra over the planes of one direction,
// each plane "solved" by a 2D BoxMG cycle (reference include/cedar/3d/relax_planes.h:35-261,
// src/3d/relax_planes.cc).
//
// What the reference does, kept here: per plane visit the in-plane iterate is copied out
// with its ghosts (copy32), the right-hand side is b plus the off-plane couplings times x
// (copy_rhs, same term order), the plane's cdr2::solver runs `solve` with
// params->plane_config (default: line-xy relaxation, max-iter 1) and the result is
// copied back (copy23).  Planes are visited odd-first on the way down (ipl = 1,3,..
// then 2,4,..) and even-first on the way up.
//
// Reference quirk preserved: copy_coeff (relax_planes.h:77-161) is
// called without a plane index and loops over ALL planes into the same 2D operator, so
// every plane solver ends up with the coefficients of the LAST interior plane.  All
// plane solvers of one direction and level are therefore identical, and ONE 2D hierarchy
// per direction and level serves every plane here -- the same numbers at 1/N of the
// memory and setup work.
#ifndef CEDAR_3D_RELAX_PLANES_H
#define CEDAR_3D_RELAX_PLANES_H

#include <algorithm>
#include <array>
#include <map>
#include <memory>
#include <type_traits>
#include <vector>

#include <cedar/types.h>
#include <cedar/3d/types.h>
#include <cedar/kernels/plane_relax.h>
#include <cedar/2d/solver.h>

namespace cedar { namespace cdr3 {

template <class sten3>
using sten3_to_sten2 = typename std::conditional<std::is_same<sten3, seven_pt>::value, cdr2::five_pt, cdr2::nine_pt>::type;

template <relax_dir rdir> struct plane_dir { };
template <> struct plane_dir<relax_dir::xy> { static constexpr int code = 0, d1 = 0, d2 = 1, dn = 2; };
template <> struct plane_dir<relax_dir::xz> { static constexpr int code = 1, d1 = 0, d2 = 2, dn = 1; };
template <> struct plane_dir<relax_dir::yz> { static constexpr int code = 2, d1 = 1, d2 = 2, dn = 0; };

// the shared 2D solver of one (direction, level) with its staging buffers
template <class sten2> struct plane_solver
{
	std::unique_ptr<cdr2::stencil_op<sten2>> so2;
	std::unique_ptr<cdr2::solver<sten2>> slv;
	cdr2::grid_func x2, b2;
	len_t nplanes = 0;
	// batched path: all planes of one colour go through the 2D cycle together (one launch per
	// kernel for the whole colour).  The level arrays of `slv` then hold `nbmax` grid functions
	// each, and each colour has its own staging arrays (so that the 2D cycle's captured graphs,
	// keyed by these pointers, keep their own batch size).
	bool batched = false;
	len_t nbmax = 1;
	std::array<cdr2::grid_func, 2> bx2, bb2;
};

// resets the batch context of libbmgx when it goes out of scope
struct batch_guard
{
	explicit batch_guard(unsigned nb) { gpu::check(bmgx_set_batch(nb), "set_batch"); }
	~batch_guard() { bmgx_set_batch(1); }
	batch_guard(const batch_guard &) = delete;
	batch_guard &operator=(const batch_guard &) = delete;
};

template <relax_dir rdir> class planes : public kernels::plane_relax<stypes, rdir>
{
public:
	using pd = plane_dir<rdir>;
	void setup(const stencil_op<seven_pt> &so) override { fine = setup_impl<seven_pt>(so); }
	void setup(const stencil_op<xxvii_pt> &so) override
	{
		level_planes.push_back(setup_impl<xxvii_pt>(so));
		level_map[so.shape(0)] = level_planes.size() - 1;
	}
	void run(const stencil_op<seven_pt> &so, grid_func &x, const grid_func &b, cycle::Dir cdir) override
	{
		relax(so, x, b, cdir, *fine);
	}
	void run(const stencil_op<xxvii_pt> &so, grid_func &x, const grid_func &b, cycle::Dir cdir) override
	{
		relax(so, x, b, cdir, *level_planes[level_map[so.shape(0)]]);
	}

private:
	template <class sten3> std::unique_ptr<plane_solver<sten3_to_sten2<sten3>>> setup_impl(const stencil_op<sten3> &so)
	{
		using sten2 = sten3_to_sten2<sten3>;
		auto ps = std::make_unique<plane_solver<sten2>>();
		len_t n1 = so.shape(pd::d1), n2 = so.shape(pd::d2);
		ps->nplanes = so.shape(pd::dn);
		ps->so2 = std::make_unique<cdr2::stencil_op<sten2>>(n1, n2);
		// the last interior plane's coefficients (see the header comment)
		gpu::check(bmgx3_plane_coeff(so.cdev(), ps->so2->dev(), so.len(0), so.len(1), so.len(2),
		                             stencil_ndirs<sten3>::value, pd::code, ps->nplanes + 1, gpu::stream()), "plane_coeff");
		ps->slv = std::make_unique<cdr2::solver<sten2>>(*ps->so2, this->params->plane_config);
		ps->x2 = cdr2::grid_func(n1, n2);
		ps->b2 = cdr2::grid_func(n1, n2);
		// Batch the planes of a colour when the plane solve is ONE V-cycle with line relaxation
		// (the default plane_config) -- the kernels of that cycle are batch aware (include/bmgx.h);
		// key "plane-batch": false in the plane config keeps the plane-by-plane path
		const auto &st = ps->slv->get_settings();
		using rt = ml_settings::relax_type;
		const bool lines = st.relaxation == rt::line_x || st.relaxation == rt::line_y || st.relaxation == rt::line_xy;
		const bool fits = std::max(n1, n2) + 2 <= 12000;   // a line must fit the shared-memory line solver
		ps->batched = lines && fits && st.maxiter == 1 && st.cycle == ml_settings::cycle_type::v
		              && this->params->plane_config->template get<bool>("plane-batch", true) && ps->nplanes > 1;
		if (ps->batched) {
			ps->nbmax = (ps->nplanes + 1) / 2;
			auto &lv = ps->slv->levels;
			for (std::size_t l = 0; l < lv.size(); l++) {
				if (l == 0) { auto &L = lv.template get<sten2>(0); L.x.set_batch(ps->nbmax); L.b.set_batch(ps->nbmax); L.res.set_batch(ps->nbmax); }
				else { auto &L = lv.get(l); L.x.set_batch(ps->nbmax); L.b.set_batch(ps->nbmax); L.res.set_batch(ps->nbmax); }
			}
			for (int c = 0; c < 2; c++) {
				ps->bx2[c] = cdr2::grid_func(n1, n2); ps->bx2[c].set_batch(ps->nbmax);
				ps->bb2[c] = cdr2::grid_func(n1, n2); ps->bb2[c].set_batch(ps->nbmax);
			}
		}
		return ps;
	}

	template <class sten3, class sten2>
	void relax(const stencil_op<sten3> &so, grid_func &x, const grid_func &b, cycle::Dir cdir, plane_solver<sten2> &ps)
	{
		int lstart, lend, lstride;
		if (cdir == cycle::Dir::DOWN) { lstart = 1; lend = 3; lstride = 1; }
		else { lstart = 2; lend = 0; lstride = -1; }
		const int nst = stencil_ndirs<sten3>::value;
		const bool single_cycle = ps.slv->get_settings().maxiter == 1;
		for (int ipl_beg = lstart; ipl_beg != lend; ipl_beg += lstride) {
			if (ps.batched) {
				if (static_cast<len_t>(ipl_beg) > ps.nplanes) continue;
				const unsigned cnt = (ps.nplanes - ipl_beg) / 2 + 1;   // planes ipl_beg, ipl_beg+2, ...
				auto &x2 = ps.bx2[ipl_beg - 1];
				auto &b2 = ps.bb2[ipl_beg - 1];
				batch_guard bg(cnt);
				gpu::check(bmgx3_plane_copy(x.dev(), x2.dev(), x.len(0), x.len(1), x.len(2), pd::code, ipl_beg + 1, 1,
				                            gpu::stream()), "plane copy32 (batch)");
				gpu::check(bmgx3_plane_rhs(so.cdev(), x.cdev(), b.cdev(), b2.dev(), x.len(0), x.len(1), x.len(2), nst,
				                           pd::code, ipl_beg + 1, gpu::stream()), "plane rhs (batch)");
				ps.slv->vcycle(x2, b2);
				gpu::check(bmgx3_plane_copy(x.dev(), x2.dev(), x.len(0), x.len(1), x.len(2), pd::code, ipl_beg + 1, 0,
				                            gpu::stream()), "plane copy23 (batch)");
				continue;
			}
			for (len_t ipl = ipl_beg; ipl <= ps.nplanes; ipl += 2) {
				// Fortran-style plane index ipl+1 (ipl counts interior planes from 1)
				gpu::check(bmgx3_plane_copy(x.dev(), ps.x2.dev(), x.len(0), x.len(1), x.len(2), pd::code, ipl + 1, 1,
				                            gpu::stream()), "plane copy32");
				gpu::check(bmgx3_plane_rhs(so.cdev(), x.cdev(), b.cdev(), ps.b2.dev(), x.len(0), x.len(1), x.len(2), nst,
				                           pd::code, ipl + 1, gpu::stream()), "plane rhs");
				// max-iter 1 (the default plane_config): solve() = residual/norm bookkeeping
				// around exactly one cycle; the bookkeeping does not touch x, so skip its
				// host round trips.  Any other setting runs the full solve loop.
				if (single_cycle) ps.slv->vcycle(ps.x2, ps.b2);
				else ps.slv->solve(ps.b2, ps.x2);
				gpu::check(bmgx3_plane_copy(x.dev(), ps.x2.dev(), x.len(0), x.len(1), x.len(2), pd::code, ipl + 1, 0,
				                            gpu::stream()), "plane copy23");
			}
		}
	}

	std::vector<std::unique_ptr<plane_solver<cdr2::nine_pt>>> level_planes;
	std::unique_ptr<plane_solver<cdr2::five_pt>> fine;
	std::map<len_t, std::size_t> level_map;
};

}}
#endif
