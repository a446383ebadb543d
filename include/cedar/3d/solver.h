// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/3d/solver.h -- cdr3::solver<fsten>: level-count rule, coarse sizes and
// coarse-solve workspace of the reference (include/cedar/3d/solver.h:18-130),
// over device-resident levels.
#ifndef CEDAR_3D_SOLVER_H
#define CEDAR_3D_SOLVER_H

#include <algorithm>
#include <array>
#include <memory>

#include <cedar/multilevel.h>
#include <cedar/level.h>
#include <cedar/3d/level_container.h>
#include <cedar/3d/types.h>
#include <cedar/3d/kernel_manager.h>

namespace cedar { namespace cdr3 {

template <class sten> struct level3 : public level<sten, stypes>
{
	using parent = level<sten, stypes>;
	level3(len_t nx, len_t ny, len_t nz) : parent::level(nx, ny, nz)
	{
		this->SOR = {{relax_stencil(nx, ny, nz), relax_stencil(nx, ny, nz)}};
		this->R.associate(&this->P);
	}
	level3(stencil_op<sten> &A) : parent::level(A)
	{
		this->res = grid_func(A.shape(0), A.shape(1), A.shape(2));
		this->SOR = {{relax_stencil(A.shape(0), A.shape(1), A.shape(2)), relax_stencil(A.shape(0), A.shape(1), A.shape(2))}};
	}
};

template <class fsten>
class solver : public multilevel<exec_mode::serial, level_container<level3, fsten>, fsten, solver<fsten>>
{
public:
	using parent = multilevel<exec_mode::serial, level_container<level3, fsten>, fsten, solver<fsten>>;
	solver(stencil_op<fsten> &fop) : parent::multilevel(fop)
	{
		this->kman = build_kernel_manager(*this->conf);
		parent::setup(fop);
	}
	solver(stencil_op<fsten> &fop, std::shared_ptr<config> conf) : parent::multilevel(fop, conf)
	{
		this->kman = build_kernel_manager(*this->conf);
		parent::setup(fop);
	}
	~solver() { delete[] this->bbd; }

	// reference 3d/solver.h:54-72: integer division, float comparison
	std::size_t compute_num_levels(stencil_op<fsten> &fop)
	{
		float nxc, nyc, nzc;
		int ng = 0;
		int min_coarse = this->settings.min_coarse;
		auto nx = fop.shape(0), ny = fop.shape(1), nz = fop.shape(2);
		do {
			ng++;
			nxc = (nx - 1) / (1 << ng) + 1;
			nyc = (ny - 1) / (1 << ng) + 1;
			nzc = (nz - 1) / (1 << ng) + 1;
		} while (std::min({nxc, nyc, nzc}) >= min_coarse);
		return ng;
	}

	// reference 3d/solver.h:75-123
	void setup_space(std::size_t nlevels)
	{
		this->levels.init(nlevels);
		auto params = build_kernel_params(*(this->conf));
		for (std::size_t i = 0; i + 1 < nlevels; i++) {
			len_t nx, ny, nz;
			if (i == 0) { auto &fop = this->levels.template get<fsten>(i).A; nx = fop.shape(0); ny = fop.shape(1); nz = fop.shape(2); }
			else { auto &fop = this->levels.get(i).A; nx = fop.shape(0); ny = fop.shape(1); nz = fop.shape(2); }
			len_t nxc = (nx - 1) / 2. + 1;
			len_t nyc = (ny - 1) / 2. + 1;
			len_t nzc = (nz - 1) / 2. + 1;
			this->levels.add(nxc, nyc, nzc);
			log::debug << "Created coarse grid with dimensions: " << nxc << ", " << nyc << ", " << nzc << std::endl;
		}
		len_t nxc, nyc, nzc;
		if (nlevels == 1) { auto &cop = this->levels.template get<fsten>(0).A; nxc = cop.shape(0); nyc = cop.shape(1); nzc = cop.shape(2); }
		else { auto &cop = this->levels.get(this->levels.size() - 1).A; nxc = cop.shape(0); nyc = cop.shape(1); nzc = cop.shape(2); }
		len_t abd_len_0 = nxc * (nyc + 1) + 2;
		if (params->periodic[0] or params->periodic[1] or params->periodic[2]) abd_len_0 = nxc * nyc * nzc;
		// as in the reference (l.121) this is the 3D constructor (nx, ny, nz = 0): a
		// (abd_len_0+2) x (n+2) x 2 array of which the kernels use the first plane
		this->ABD = grid_func(abd_len_0, nxc * nyc * nzc, 0);
		this->bbd = new real_t[this->ABD.len(1)];
	}

	void give_op(std::unique_ptr<stencil_op<fsten>> fop) { fop_ref = std::move(fop); }
protected:
	std::unique_ptr<stencil_op<fsten>> fop_ref;
};

}}
#endif
