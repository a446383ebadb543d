// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/cycle/types.h -- base class of the cycling strategies (reference: cycle/types.h:9-41)
#ifndef CEDAR_CYCLE_TYPES_H
#define CEDAR_CYCLE_TYPES_H

#include <functional>
#include <memory>

#include <cedar/types.h>
#include <cedar/kernel_registry.h>
#include <cedar/kernel_manager.h>

namespace cedar { namespace cycle {

template <exec_mode emode, class level_container, class fsten> class cycle
{
public:
	template <class sten> using level_t = typename level_container::template level_t<sten>;
	using grid_func = typename level_t<fsten>::grid_func;
	using stypes = typename level_t<fsten>::stypes;
	using manager = kernel_manager<klist<stypes, emode>>;

	cycle(level_container &levels, std::function<void(grid_func &x, const grid_func &b)> &coarse_solver)
	    : levels(levels), coarse_solver(coarse_solver) {}
	virtual ~cycle() {}
	void set_kernels(std::shared_ptr<manager> kregis) { this->kman = kregis; }
	// B200 extension: the levels from `lvl` down (lvl >= 1) may be handed to one kernel that runs their whole sub-cycle
	// (cdr2: bmgx2_vcycle_tail); x arrives zeroed, like the coarse iterate of ncycle_helper
	void set_tail(std::size_t lvl, std::function<void(grid_func &x, const grid_func &b)> f) { tail_level = lvl; tail_solver = std::move(f); }
	virtual void run(grid_func &x, const grid_func &b) = 0;
	void log_residual(std::size_t lvl, const grid_func &res)
	{
		if (log::info.active())
			log::info << "Level " << (levels.size() - lvl - 1) << " residual norm: " << res.template lp_norm<2>()
			          << std::endl;
	}
protected:
	std::shared_ptr<manager> kman;
	level_container &levels;
	std::function<void(grid_func &x, const grid_func &b)> &coarse_solver;
	std::size_t tail_level = 0;   // 0 = none
	std::function<void(grid_func &x, const grid_func &b)> tail_solver;
};

}}
#endif
