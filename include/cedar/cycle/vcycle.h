// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/cycle/vcycle.h -- the V-cycle (reference: include/cedar/cycle/vcycle.h:35-115):
//   presmooth, residual, restrict, zero coarse x, (coarse solve | recurse),
//   interpolate-and-add, postsmooth.
// B200-native addition: after one eager pass (which makes every level array
// device resident) the whole cycle for a given (x, b) pair is captured once as a
// CUDA graph and replayed, removing ~10 launches per level per cycle.
#ifndef CEDAR_CYCLE_VCYCLE_H
#define CEDAR_CYCLE_VCYCLE_H

#include <map>
#include <vector>

#include <cedar/cycle/types.h>
#include <cedar/array.h>
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
	using parent::tail_level;
	using parent::tail_solver;

	vcycle(level_container &levels, std::function<void(grid_func &, const grid_func &)> &coarse_solver,
	       bool use_graph = true)
	    : parent::cycle(levels, coarse_solver), use_graph(use_graph) {}
	~vcycle() { drop_graph(); }

	void run(grid_func &x, const grid_func &b) override
	{
		if (levels.size() == 1) { coarse_solver(x, b); return; }
		// info-level logging reads norms back inside the cycle: not capturable
		if (!use_graph || log::info.active()) { ncycle(0, x, b); return; }
		// the caller is capturing this stream into its own graph (the block-decomposed
		// solver's whole-cycle graph): record the kernels directly, an executable graph
		// cannot be launched into a capture
		int outer_capture = 0;
		gpu::check(bmgx_stream_is_capturing(gpu::stream(), &outer_capture), "stream capture status");
		if (outer_capture) { ncycle(0, x, b); return; }
		// A captured graph bakes device pointers in.  Kernels may exchange device
		// buffers (the fused relaxation ping-pongs x with a scratch array), so the
		// graph is keyed by the pointer state of (x, b, every level's x): a cycle
		// with an odd number of sweeps simply alternates between two graphs.
		std::vector<const void *> key = state_key(x, b);
		if (!warmed) {
			ncycle(0, x, b);   // first pass: makes every level array (and scratch) device resident
			warmed = true;
			return;
		}
		auto it = graphs.find(key);
		if (it != graphs.end() && !replayable(it->second, x)) { bmgx_graph_destroy(it->second.graph); graphs.erase(it); it = graphs.end(); }
		if (it == graphs.end()) {
			if (graphs.size() >= 8) drop_graph();
			entry e;
			e.xbuf = &x.buffer();
			unsigned long long before = bmgx_launch_count();
			gpu::check(bmgx_graph_begin(gpu::stream()), "vcycle graph capture");
			detail::dual_buffer::set_swap_recorder(&e.swaps);
			try { ncycle(0, x, b); }
			catch (...) { detail::dual_buffer::set_swap_recorder(nullptr); void *g = nullptr; bmgx_graph_end(gpu::stream(), &g); if (g) bmgx_graph_destroy(g); throw; }
			detail::dual_buffer::set_swap_recorder(nullptr);
			gpu::check(bmgx_graph_end(gpu::stream(), &e.graph), "vcycle graph instantiate");
			e.nodes = bmgx_launch_count() - before;
			// capture only recorded the work; the host bookkeeping (buffer exchanges)
			// already reflects the state after the cycle, so one launch makes it true
			gpu::check(bmgx_graph_launch(e.graph, gpu::stream()), "vcycle graph launch");
			graphs.emplace(std::move(key), std::move(e));
			return;
		}
		entry &e = it->second;
		gpu::check(bmgx_graph_launch(e.graph, gpu::stream()), "vcycle graph launch");
		gpu::add_launches(e.nodes);
		for (auto &sw : e.swaps) resolve(e, sw.a, x)->swap_device(*resolve(e, sw.b, x));
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
		} else if (tail_solver && lvl + 1 == tail_level && n == 1 && !log::info.active()) {
			tail_solver(coarse_x, coarse_b);   // the remaining levels' sub-cycle in one launch
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

	bool graph_active() const { return !graphs.empty(); }
private:
	struct entry
	{
		void *graph = nullptr;
		unsigned long long nodes = 0;
		std::vector<detail::dual_buffer::swap_event> swaps;
		detail::dual_buffer *xbuf = nullptr;   // the caller's x at capture time
	};
	std::vector<const void *> state_key(grid_func &x, const grid_func &b)
	{
		std::vector<const void *> key;
		key.push_back(x.dev());
		key.push_back(b.cdev());
		for (std::size_t l = 1; l < levels.size(); l++) key.push_back(levels.get(l).x.cdev());
		return key;
	}
	static detail::dual_buffer *resolve(const entry &e, detail::dual_buffer *p, grid_func &x)
	{
		return p == e.xbuf ? &x.buffer() : p;
	}
	// the recorded exchanges must start from the pointers the graph was captured with
	bool replayable(const entry &e, grid_func &x)
	{
		std::map<detail::dual_buffer *, const void *> cur;
		for (auto &sw : e.swaps) {
			detail::dual_buffer *a = resolve(e, sw.a, x), *b = resolve(e, sw.b, x);
			if (!cur.count(a)) cur[a] = a->raw_dev();
			if (!cur.count(b)) cur[b] = b->raw_dev();
			if (cur[a] != sw.pa || cur[b] != sw.pb) return false;
			std::swap(cur[a], cur[b]);
		}
		return true;
	}
	void drop_graph()
	{
		for (auto &kv : graphs) bmgx_graph_destroy(kv.second.graph);
		graphs.clear();
	}
	bool use_graph;
	bool warmed = false;
	std::map<std::vector<const void *>, entry> graphs;
};

}}
#endif
