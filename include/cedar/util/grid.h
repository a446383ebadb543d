// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/util/grid.h -- evenly spaced sample points (reference: include/cedar/util/grid.h,
// src/util/grid.cc).  Only the examples include this header.
#ifndef CEDAR_UTIL_GRID_H
#define CEDAR_UTIL_GRID_H

#include <cedar/types.h>

namespace cedar {

// `num` points from `start` to `end`, both included
inline AlignedVector<real_t> linspace(real_t start, real_t end, len_t num)
{
	AlignedVector<real_t> v(num);
	if (num == 1) { v[0] = start; return v; }
	const real_t h = (end - start) / static_cast<real_t>(num - 1);
	for (len_t i = 0; i < num; i++) v[i] = start + h * static_cast<real_t>(i);
	return v;
}

}
#endif
