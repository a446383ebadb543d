// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/3d/level_container.h -- level 0 is typed on the fine stencil, every
// coarser level is 27-point (reference 3d/level_container.h).
#ifndef CEDAR_3D_LEVEL_CONTAINER_H
#define CEDAR_3D_LEVEL_CONTAINER_H

#include <deque>
#include <type_traits>

#include <cedar/3d/stencil_op.h>
#include <cedar/util/log.h>

namespace cedar { namespace cdr3 {

template <template <class> class level, class fsten> class level_container
{
public:
	template <class stencil> using stencil_op = typename level<fsten>::template stencil_op<stencil>;
	template <class rsten> using level_t = level<rsten>;

	level_container(stencil_op<fsten> &fine_op) : fine_op(fine_op) {}

	void init(std::size_t nlevels)
	{
		(void)nlevels;  // deques never move their elements: no reserve needed
		if constexpr (std::is_same<fsten, seven_pt>::value) lvls_seven.emplace_back(fine_op);
		else lvls_xxvii.emplace_back(fine_op);
	}
	template <class... Args> void add(Args &&...args) { lvls_xxvii.emplace_back(std::forward<Args>(args)...); }

	template <class rsten = xxvii_pt> level<rsten> &get(std::size_t i)
	{
		if constexpr (std::is_same<rsten, seven_pt>::value) {
			if (i != 0 || lvls_seven.empty()) log::error << "coarse operators are 27 point (not seven)!" << std::endl;
			return lvls_seven[0];
		} else {
			if (i < lvls_seven.size()) log::error << "fine grid operator is seven point (not 27)!" << std::endl;
			return lvls_xxvii[i - lvls_seven.size()];
		}
	}
	std::size_t size() { return lvls_xxvii.size() + lvls_seven.size(); }
protected:
	stencil_op<fsten> &fine_op;
	std::deque<level<xxvii_pt>> lvls_xxvii;
	std::deque<level<seven_pt>> lvls_seven;
};

}}
#endif
