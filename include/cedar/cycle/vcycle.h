// cedar/cycle/vcycle.h -- the V-cycle (reference: include/cedar/cycle/vcycle.h:35-115):
//   presmooth, residual, restrict, zero coarse x, (coarse solve | recurse),
//   interpolate-and-add, postsmooth.
// B200-native addition: after one eager pass (which makes every level array
// device resident) the whole cycle for a given (x, b) pair is captured once as a
// CUDA graph and replayed, removing ~10 launches per level per cycle.
#ifndef CEDAR_CYCLE_VCYCLE_H
#define CEDAR_CYCLE_VCYCLE_H

#include <cedar/cycle/types.h>
#include <cedar/gpu.h>

namespace cedar { namespace cycle {

template <exec_mode emode, class level_container, class fsten>
class vcycle : public cycle<emode, level_container, fsten>
{
public:
	template <class sten> using level_t = typename level_container::template level_t<sten>;
	template <class sten> using stencil_op = typename level_t<fsten>::template stencil_op<sten>;
	using grid_func = typename level_t<fsten>::grid_func;
	using stypes = typename level_t<fsten>::stypes;
	using interp_add = kernels::interp_add<stypes>;
	using restriction = kernels::restriction<stypes>;
	using residual = kernels::residual<stypes>;

	using parent = cycle<emode, level_container, fsten>;
	using parent::levels;
	using parent::coarse_solver;
	using parent::log_residual;
	using parent::kman;

	vcycle(level_container &levels, std::function<void(grid_func &, const grid_func &)> &coarse_solver,
	       bool use_graph = true)
	    : parent::cycle(levels, coarse_solver), use_graph(use_graph) {}
	~vcycle() { drop_graph(); }

	void run(grid_func &x, const grid_func &b) override
	{
		if (levels.size() == 1) { coarse_solver(x, b); return; }
		// info-level logging reads norms back inside the cycle: not capturable
		if (!use_graph || log::info.active()) { ncycle(0, x, b); return; }
		const void *kx = x.dev(), *kb = b.cdev();
		if (kx != key_x || kb != key_b) { drop_graph(); key_x = kx; key_b = kb; eager_done = false; }
		if (!eager_done) {
			ncycle(0, x, b);
			eager_done = true;
		} else if (!graph) {
			unsigned long long before = bmgx_launch_count();
			gpu::check(bmgx_graph_begin(gpu::stream()), "vcycle graph capture");
			ncycle(0, x, b);
			gpu::check(bmgx_graph_end(gpu::stream(), &graph), "vcycle graph instantiate");
			graph_nodes = bmgx_launch_count() - before;
			if (x.dev() != kx) {
				// a kernel exchanged device buffers an odd number of times: the
				// captured pointers would not line up on replay -- stay eager
				drop_graph(); use_graph = false; key_x = nullptr;
				log::error << "vcycle: cycle does not preserve buffer identity, CUDA graph disabled" << std::endl;
				ncycle(0, x, b);
				return;
			}
			// the capture already counted this cycle's launches once
			gpu::check(bmgx_graph_launch(graph, gpu::stream()), "vcycle graph launch");
		} else {
			launch_graph();
		}
	}

	void ncycle(std::size_t lvl, grid_func &x, const grid_func &b, int n = 1)
	{
		if (lvl == 0) ncycle_helper(levels.template get<fsten>(lvl), lvl, x, b, n);
		else ncycle_helper(levels.get(lvl), lvl, x, b, n);
	}

	template <class sten> void ncycle_helper(level_t<sten> &level, std::size_t lvl, grid_func &x, const grid_func &b, int n)
	{
		auto &A = level.A;
		level.presmoother(A, x, b);
		grid_func &res = level.res;
		kman->template run<residual>(A, x, b, res);
		log_residual(lvl, res);

		auto &coarse_b = levels.get(lvl + 1).b;
		auto &coarse_x = levels.get(lvl + 1).x;
		kman->template run<restriction>(levels.get(lvl + 1).R, res, coarse_b);
		coarse_x.set(0.0);

		std::size_t coarse_lvl = levels.size() - 1;
		if (lvl + 1 == coarse_lvl) {
			coarse_solver(coarse_x, coarse_b);
			if (log::info.active()) {
				auto &coarse_A = levels.get(levels.size() - 1).A;
				auto &coarse_residual = levels.get(levels.size() - 1).res;
				kman->template run<residual>(coarse_A, coarse_x, coarse_b, coarse_residual);
				log_residual(lvl + 1, coarse_residual);
			}
		} else {
			for (int i = 0; i < n; i++) ncycle(lvl + 1, coarse_x, coarse_b, n);
		}

		kman->template run<interp_add>(levels.get(lvl + 1).P, coarse_x, level.res, x);
		level.postsmoother(A, x, b);

		if (log::info.active()) {
			kman->template run<residual>(A, x, b, res);
			log_residual(lvl, res);
		}
	}

	bool graph_active() const { return graph != nullptr; }
private:
	void launch_graph()
	{
		gpu::check(bmgx_graph_launch(graph, gpu::stream()), "vcycle graph launch");
		gpu::add_launches(graph_nodes);
	}
	void drop_graph()
	{
		if (graph) { bmgx_graph_destroy(graph); graph = nullptr; }
	}
	bool use_graph;
	bool eager_done = false;
	const void *key_x = nullptr, *key_b = nullptr;
	void *graph = nullptr;
	unsigned long long graph_nodes = 0;
};

}}
#endif
