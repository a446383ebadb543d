// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/2d/level_container.h -- level 0 is typed on the fine stencil, every
// coarser level is nine-point (reference 2d/level_container.h:107-141).
#ifndef CEDAR_2D_LEVEL_CONTAINER_H
#define CEDAR_2D_LEVEL_CONTAINER_H

#include <deque>
#include <type_traits>

#include <cedar/2d/stencil_op.h>
#include <cedar/util/log.h>

namespace cedar { namespace cdr2 {

template <template <class> class level, class fsten> class level_container
{
public:
	template <class stencil> using stencil_op = typename level<fsten>::template stencil_op<stencil>;
	template <class rsten> using level_t = level<rsten>;

	level_container(stencil_op<fsten> &fine_op) : fine_op(fine_op) {}

	void init(std::size_t nlevels)
	{
		(void)nlevels;  // deques never move their elements: no reserve needed
		if constexpr (std::is_same<fsten, five_pt>::value) lvls_five.emplace_back(fine_op);
		else lvls_nine.emplace_back(fine_op);
	}
	template <class... Args> void add(Args &&...args) { lvls_nine.emplace_back(std::forward<Args>(args)...); }

	template <class rsten = nine_pt> level<rsten> &get(std::size_t i)
	{
		if constexpr (std::is_same<rsten, five_pt>::value) {
			if (i != 0 || lvls_five.empty()) log::error << "coarse operators are nine point (not five)!" << std::endl;
			return lvls_five[0];
		} else {
			if (i < lvls_five.size()) log::error << "fine grid operator is five point (not nine)!" << std::endl;
			return lvls_nine[i - lvls_five.size()];
		}
	}
	std::size_t size() { return lvls_nine.size() + lvls_five.size(); }
protected:
	stencil_op<fsten> &fine_op;
	std::deque<level<nine_pt>> lvls_nine;
	std::deque<level<five_pt>> lvls_five;
};

}}
#endif
