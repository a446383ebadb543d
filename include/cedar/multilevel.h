// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/multilevel.h -- generic multilevel driver (reference:
// include/cedar/multilevel.h:26-322).  It only names kernel *types* and talks to
// the registry (kman->setup<K> / kman->run<K>), so the CUDA-backed kernels plug
// in exactly where the reference's Fortran wrappers did.
//
//   setup():  per level  setup_interp -> coarsen_op -> relax setup,  then the
//             coarse-solve setup (multilevel.h:243-265 of the reference)
//   solve():  residual + norm, loop { cycle ; residual ; norm ; tol test }
#ifndef CEDAR_MULTILEVEL_H
#define CEDAR_MULTILEVEL_H

#include <cassert>
#include <functional>
#include <memory>
#include <vector>

#include <cedar/types.h>
#include <cedar/kernel_registry.h>
#include <cedar/kernel_manager.h>
#include <cedar/cycle/types.h>
#include <cedar/cycle/vcycle.h>
#include <cedar/cycle/fcycle.h>
#include <cedar/multilevel_settings.h>
#include <cedar/util/timer.h>

namespace cedar {

template <exec_mode emode, class level_container, class fsten, class child> class multilevel
{
public:
	template <class sten> using level_t = typename level_container::template level_t<sten>;
	template <class sten> using stencil_op = typename level_t<fsten>::template stencil_op<sten>;
	using grid_func = typename level_t<fsten>::grid_func;
	using stypes = typename level_t<fsten>::stypes;
	template <relax_dir rdir> using line_relax = kernels::line_relax<stypes, rdir>;
	template <relax_dir rdir> using plane_relax = kernels::plane_relax<stypes, rdir>;
	using point_relax = kernels::point_relax<stypes>;
	using coarsen_op = kernels::coarsen_op<stypes>;
	using interp_add = kernels::interp_add<stypes>;
	using matvec = kernels::matvec<stypes>;
	using residual = kernels::residual<stypes>;
	using setup_prolong = kernels::setup_interp<stypes>;
	using solve_cg = kernels::solve_cg<stypes>;
	using kern_manager = kernel_manager<klist<stypes, emode>>;
	using conf_ptr = std::shared_ptr<config>;
	using relax_type = ml_settings::relax_type;

	multilevel(stencil_op<fsten> &fop) : levels(fop), conf(std::make_shared<config>("config.json")) { init_settings(); }
	multilevel(stencil_op<fsten> &fop, conf_ptr cfg) : levels(fop), conf(cfg) { init_settings(); }
	virtual ~multilevel() {}

	void setup_cycler()
	{
		if (settings.cycle == ml_settings::cycle_type::v)
			this->cycle = std::make_unique<cycle::vcycle<emode, level_container, fsten>>(levels, coarse_solver,
			                                                                             settings.cuda_graph);
		else
			this->cycle = std::make_unique<cycle::fcycle<emode, level_container, fsten>>(levels, coarse_solver);
	}

	void vcycle(grid_func &x, const grid_func &b)
	{
		assert(settings.cycle == ml_settings::cycle_type::v);
		this->cycle->run(x, b);
	}

	std::shared_ptr<kern_manager> get_kernels() { return kman; }
	std::size_t nlevels() { return levels.size(); }

	virtual void setup_cg_solve()
	{
		if (levels.size() == 1) kman->template setup<solve_cg>(levels.template get<fsten>(0).A, ABD);
		else kman->template setup<solve_cg>(levels.get(levels.size() - 1).A, ABD);
		auto kernels = get_kernels();
		coarse_solver = [&, kernels](grid_func &x, const grid_func &b) {
			kernels->template run<solve_cg>(x, b, ABD, bbd);
		};
	}

	void setup_interp(std::size_t lvl)
	{
		auto &P = levels.get(lvl + 1).P;
		auto &cop = levels.get(lvl + 1).A;
		if (lvl == 0) kman->template run<setup_prolong>(levels.template get<fsten>(lvl).A, cop, P);
		else kman->template run<setup_prolong>(levels.get(lvl).A, cop, P);
	}

	void setup_operator(std::size_t lvl)
	{
		auto &P = levels.get(lvl + 1).P;
		auto &cop = levels.get(lvl + 1).A;
		if (lvl == 0) kman->template run<coarsen_op>(P, levels.template get<fsten>(lvl).A, cop);
		else kman->template run<coarsen_op>(P, levels.get(lvl).A, cop);
	}

	// one relaxation sweep of the configured kind in direction `dir`; line-xy
	// runs x then y on the way down and y then x on the way up, plane-xyz runs
	// xy,yz,xz down and xz,yz,xy up (reference multilevel.h:165-222)
	template <class sten>
	void relax_sweep(level_t<sten> &level, const stencil_op<sten> &A, grid_func &x, const grid_func &b, cycle::Dir dir)
	{
		auto &k = *kman;
		bool down = dir == cycle::Dir::DOWN;
		switch (settings.relaxation) {
		case relax_type::point: k.template run<point_relax>(A, x, b, level.SOR[0], dir); break;
		case relax_type::line_x: k.template run<line_relax<relax_dir::x>>(A, x, b, level.SOR[0], level.res, dir); break;
		case relax_type::line_y: k.template run<line_relax<relax_dir::y>>(A, x, b, level.SOR[0], level.res, dir); break;
		case relax_type::line_xy:
			if (down) {
				k.template run<line_relax<relax_dir::x>>(A, x, b, level.SOR[0], level.res, dir);
				k.template run<line_relax<relax_dir::y>>(A, x, b, level.SOR[1], level.res, dir);
			} else {
				k.template run<line_relax<relax_dir::y>>(A, x, b, level.SOR[1], level.res, dir);
				k.template run<line_relax<relax_dir::x>>(A, x, b, level.SOR[0], level.res, dir);
			}
			break;
		case relax_type::plane_xyz:
			if (down) {
				k.template run<plane_relax<relax_dir::xy>>(A, x, b, dir);
				k.template run<plane_relax<relax_dir::yz>>(A, x, b, dir);
				k.template run<plane_relax<relax_dir::xz>>(A, x, b, dir);
			} else {
				k.template run<plane_relax<relax_dir::xz>>(A, x, b, dir);
				k.template run<plane_relax<relax_dir::yz>>(A, x, b, dir);
				k.template run<plane_relax<relax_dir::xy>>(A, x, b, dir);
			}
			break;
		case relax_type::plane_xy: k.template run<plane_relax<relax_dir::xy>>(A, x, b, dir); break;
		case relax_type::plane_xz: k.template run<plane_relax<relax_dir::xz>>(A, x, b, dir); break;
		case relax_type::plane_yz: k.template run<plane_relax<relax_dir::yz>>(A, x, b, dir); break;
		default: log::error << "Invalid relaxation: " << std::endl;
		}
	}

	template <class sten> void setup_relax_helper(level_t<sten> &level, std::size_t lvl)
	{
		(void)lvl;
		auto &sop = level.A;
		switch (settings.relaxation) {
		case relax_type::point: kman->template setup<point_relax>(sop, level.SOR[0]); break;
		case relax_type::line_x: kman->template setup<line_relax<relax_dir::x>>(sop, level.SOR[0]); break;
		case relax_type::line_y: kman->template setup<line_relax<relax_dir::y>>(sop, level.SOR[0]); break;
		case relax_type::line_xy:
			kman->template setup<line_relax<relax_dir::x>>(sop, level.SOR[0]);
			kman->template setup<line_relax<relax_dir::y>>(sop, level.SOR[1]);
			break;
		case relax_type::plane_xyz:
			kman->template setup<plane_relax<relax_dir::xy>>(sop);
			kman->template setup<plane_relax<relax_dir::xz>>(sop);
			kman->template setup<plane_relax<relax_dir::yz>>(sop);
			break;
		case relax_type::plane_xy: kman->template setup<plane_relax<relax_dir::xy>>(sop); break;
		case relax_type::plane_xz: kman->template setup<plane_relax<relax_dir::xz>>(sop); break;
		case relax_type::plane_yz: kman->template setup<plane_relax<relax_dir::yz>>(sop); break;
		default: log::error << "Invalid relaxation" << std::endl;
		}
		level.presmoother = [this, &level](const stencil_op<sten> &A, grid_func &x, const grid_func &b) {
			for (int i = 0; i < settings.nrelax_pre; i++) relax_sweep(level, A, x, b, cycle::Dir::DOWN);
		};
		level.postsmoother = [this, &level](const stencil_op<sten> &A, grid_func &x, const grid_func &b) {
			for (int i = 0; i < settings.nrelax_post; i++) relax_sweep(level, A, x, b, cycle::Dir::UP);
		};
	}

	void setup_relax(std::size_t lvl)
	{
		if (lvl == 0) setup_relax_helper(levels.template get<fsten>(lvl), lvl);
		else setup_relax_helper(levels.get(lvl), lvl);
	}

	void setup_space(std::size_t nlevels) { static_cast<child *>(this)->setup_space(nlevels); }

	void setup(stencil_op<fsten> &fop)
	{
		this->cycle->set_kernels(this->kman);
		auto num_levels = compute_num_levels(fop);
		auto nlevels_conf = settings.num_levels;
		if (nlevels_conf > 0) {
			if (static_cast<std::size_t>(nlevels_conf) > num_levels) log::error << "too many levels specified" << std::endl;
			else num_levels = nlevels_conf;
		}
		log::debug << "Using a " << num_levels << " level heirarchy" << std::endl;
		setup_space(num_levels);
		timer_begin("setup");
		for (std::size_t i = 0; i + 1 < num_levels; ++i) {
			setup_interp(i);
			setup_operator(i);
			setup_relax(i);
		}
		setup_cg_solve();
		timer_end("setup");
	}

	virtual grid_func solve(const grid_func &b)
	{
		grid_func x = grid_func::zeros_like(b);
		solve(b, x);
		return x;
	}

	virtual void solve(const grid_func &b, grid_func &x)
	{
		auto &level = levels.template get<fsten>(0);
		int maxiter = settings.maxiter;
		real_t tol = settings.tol;
		residual_history.clear();
		kman->template run<residual>(level.A, x, b, level.res);
		real_t res0_l2 = level.res.template lp_norm<2>();
		initial_residual = res0_l2;
		log::info << "Initial residual l2 norm: " << res0_l2 << std::endl;
		timer_begin("solve");
		for (int i = 0; i < maxiter; i++) {
			cycle->run(x, b);
			kman->template run<residual>(level.A, x, b, level.res);
			real_t res_l2 = level.res.template lp_norm<2>();
			real_t rel_l2 = res_l2 / res0_l2;
			residual_history.push_back(rel_l2);
			log::status << "Iteration " << i << " relative l2 norm: " << rel_l2 << std::endl;
			if (rel_l2 < tol) break;
		}
		timer_end("solve");
	}

	std::size_t compute_num_levels(stencil_op<fsten> &fop) { return static_cast<child *>(this)->compute_num_levels(fop); }
	config &get_config() { return *conf; }
	ml_settings &get_settings() { return settings; }
	// relative residual after each cycle of the last solve() (what the
	// reference prints as "Iteration i relative l2 norm"), and ||r_0||
	const std::vector<real_t> &history() const { return residual_history; }
	real_t res0() const { return initial_residual; }

	level_container levels;
protected:
	void init_settings()
	{
		settings.init(*conf);
		log::status << settings << std::endl;
		setup_cycler();
	}
	std::unique_ptr<cycle::cycle<emode, level_container, fsten>> cycle;
	std::function<void(grid_func &x, const grid_func &b)> coarse_solver;
	std::shared_ptr<config> conf;
	std::shared_ptr<kern_manager> kman;
	ml_settings settings;
	grid_func ABD;
	real_t *bbd = nullptr;
	std::vector<real_t> residual_history;
	real_t initial_residual = 0;
};

}
#endif
