// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/multilevel_settings.h -- solver settings read from the JSON config
// (reference: include/cedar/multilevel_settings.h, src/multilevel_settings.cc:15-61).
#ifndef CEDAR_MULTILEVEL_SETTINGS_H
#define CEDAR_MULTILEVEL_SETTINGS_H

#include <iostream>
#include <memory>
#include <cedar/types.h>
#include <cedar/config.h>

namespace cedar {

struct ml_settings
{
	enum class relax_type { point, line_x, line_y, line_xy, plane_xy, plane_xz, plane_yz, plane_xyz };
	enum class cycle_type { v, f };
	enum class cg_type { lu, serial, redist };

	void init(config &conf);
	friend std::ostream &operator<<(std::ostream &os, const ml_settings &obj);

	relax_type relaxation = relax_type::point;
	cycle_type cycle = cycle_type::v;
	cg_type coarse_solver = cg_type::lu;
	std::shared_ptr<config> coarse_config;
	int nrelax_pre = 2;
	int nrelax_post = 1;
	int num_levels = -1;
	int maxiter = 10;
	real_t tol = 1e-8;
	int min_coarse = 3;
	// B200 extension (key "solver.cuda-graph", default true): replay each
	// V-cycle as one CUDA graph instead of ~10 launches per level
	bool cuda_graph = true;
	// B200 extension (key "solver.coarse-tail", default 33; 0 = off): with point relaxation, the levels whose larger
	// dimension is at most this many unknowns run their whole sub-cycle in ONE kernel (bmgx2_vcycle_tail) instead of
	// ~6 launches per level; same arithmetic, bit for bit
	int coarse_tail = 33;
};

}
#endif
