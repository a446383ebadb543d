// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/2d/solver.h -- cdr2::solver<fsten>: level-count rule, coarse sizes and
// coarse-solve workspace of the reference (include/cedar/2d/solver.h:17-126),
// over device-resident levels.
#ifndef CEDAR_2D_SOLVER_H
#define CEDAR_2D_SOLVER_H

#include <algorithm>
#include <array>
#include <memory>

#include <cedar/multilevel.h>
#include <cedar/level.h>
#include <cedar/2d/level_container.h>
#include <cedar/2d/types.h>
#include <cedar/2d/kernel_manager.h>

namespace cedar { namespace cdr2 {

template <class sten> struct level2 : public level<sten, stypes>
{
	using parent = level<sten, stypes>;
	level2(len_t nx, len_t ny) : parent::level(nx, ny)
	{
		this->SOR = {{relax_stencil(nx, ny), relax_stencil(nx, ny)}};
		this->R.associate(&this->P);
	}
	level2(stencil_op<sten> &A) : parent::level(A)
	{
		this->res = grid_func(A.shape(0), A.shape(1));
		this->SOR = {{relax_stencil(A.shape(0), A.shape(1)), relax_stencil(A.shape(0), A.shape(1))}};
	}
};

template <class fsten>
class solver : public multilevel<exec_mode::serial, level_container<level2, fsten>, fsten, solver<fsten>>
{
public:
	using parent = multilevel<exec_mode::serial, level_container<level2, fsten>, fsten, solver<fsten>>;
	solver(stencil_op<fsten> &fop) : parent::multilevel(fop)
	{
		this->kman = build_kernel_manager(*this->conf);
		parent::setup(fop);
		setup_tail();
	}
	solver(stencil_op<fsten> &fop, std::shared_ptr<config> conf) : parent::multilevel(fop, conf)
	{
		this->kman = build_kernel_manager(*this->conf);
		parent::setup(fop);
		setup_tail();
	}

	// B200 extension: hand the small levels' sub-cycle to one kernel (bmgx2_vcycle_tail, include/bmgx.h).  Point
	// relaxation, non-periodic, V-cycle only; the first tail level is the first coarse level (>= 1, always nine-point)
	// whose larger dimension is at most settings.coarse_tail, and at least two levels must remain.
	void setup_tail()
	{
		auto &st = this->settings;
		auto params = this->kman->get_params();
		const std::size_t nl = this->levels.size();
		if (st.coarse_tail <= 0 || st.relaxation != ml_settings::relax_type::point || st.cycle != ml_settings::cycle_type::v) return;
		if (params->periodic[0] || params->periodic[1] || nl < 3) return;
		std::size_t first = 0;
		for (std::size_t l = 1; l + 1 < nl; l++) {
			auto &A = this->levels.get(l).A;
			if (std::max(A.shape(0), A.shape(1)) <= static_cast<len_t>(st.coarse_tail)
			    && static_cast<std::size_t>(A.len(0)) * A.len(1) <= BMGX_TAIL_MAX_POINTS) { first = l; break; }
		}
		if (first == 0 || nl - first > BMGX_TAIL_MAX) return;
		this->cycle->set_tail(first, [this, first, nl](grid_func &x, const grid_func &b) {
			bmgx_tail_level lv[BMGX_TAIL_MAX];
			for (std::size_t l = first; l < nl; l++) {
				auto &L = this->levels.get(l);
				bmgx_tail_level &d = lv[l - first];
				d.so = L.A.cdev();
				d.sor = l + 1 < nl ? L.SOR[0].cdev() : nullptr;
				d.ci = L.P.cdev();
				// the entry level works on the caller's x and b (the arrays the cycle above restricts into)
				d.q = l == first ? x.dev() : L.x.dev();
				d.qf = l == first ? const_cast<real_t *>(b.cdev()) : L.b.dev();
				d.res = L.res.dev();
				d.ii = L.A.len(0); d.jj = L.A.len(1);
			}
			gpu::check(bmgx2_vcycle_tail(lv, static_cast<int>(nl - first), this->ABD.cdev(), tail_bbd(), this->ABD.len(0),
			                             this->settings.nrelax_pre, this->settings.nrelax_post, gpu::stream()), "vcycle_tail");
		});
	}
	~solver() { delete[] this->bbd; }

	// reference 2d/solver.h:57-73: integer division, float comparison
	std::size_t compute_num_levels(stencil_op<fsten> &fop)
	{
		float nxc, nyc;
		int ng = 0;
		auto min_coarse = this->settings.min_coarse;
		auto nx = fop.shape(0);
		auto ny = fop.shape(1);
		do {
			ng++;
			nxc = (nx - 1) / (1 << ng) + 1;
			nyc = (ny - 1) / (1 << ng) + 1;
		} while (std::min(nxc, nyc) >= min_coarse);
		return ng;
	}

	// reference 2d/solver.h:75-116
	void setup_space(std::size_t nlevels)
	{
		this->levels.init(nlevels);
		auto params = build_kernel_params(*(this->conf));
		for (std::size_t i = 0; i + 1 < nlevels; i++) {
			len_t nx, ny;
			if (i == 0) { auto &fop = this->levels.template get<fsten>(i).A; nx = fop.shape(0); ny = fop.shape(1); }
			else { auto &fop = this->levels.get(i).A; nx = fop.shape(0); ny = fop.shape(1); }
			len_t nxc = (nx - 1) / 2. + 1;
			len_t nyc = (ny - 1) / 2. + 1;
			this->levels.add(nxc, nyc);
			log::debug << "Created coarse grid with dimensions: " << nxc << ", " << nyc << std::endl;
		}
		len_t nxc, nyc;
		if (nlevels == 1) { auto &cop = this->levels.template get<fsten>(0).A; nxc = cop.shape(0); nyc = cop.shape(1); }
		else { auto &cop = this->levels.get(this->levels.size() - 1).A; nxc = cop.shape(0); nyc = cop.shape(1); }
		len_t abd_len_0 = nxc + 2;
		if (params->periodic[0] or params->periodic[1]) abd_len_0 = nxc * nyc;
		this->ABD = grid_func(abd_len_0, nxc * nyc, 0);
		this->bbd = new real_t[this->ABD.len(1)];
	}

	void give_op(std::unique_ptr<stencil_op<fsten>> fop) { fop_ref = std::move(fop); }
protected:
	real_t *tail_bbd()
	{
		if (tail_scratch.size() < static_cast<std::size_t>(this->ABD.len(1))) tail_scratch.reshape(this->ABD.len(1));
		return tail_scratch.dev();
	}
	array<real_t, 1> tail_scratch;
	std::unique_ptr<stencil_op<fsten>> fop_ref;
};

}}
#endif
