// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/2d/grid_func.h -- 2D grid function with one ghost layer (reference:
// include/cedar/2d/grid_func.h:19-53, src/2d/grid_func.cc).  Norms of
// device-resident data are reduced on the GPU (deterministic tree).
#ifndef CEDAR_2D_CORE_GRIDFUNC_H
#define CEDAR_2D_CORE_GRIDFUNC_H

#include <cmath>
#include <fstream>
#include <iostream>

#include <cedar/types.h>
#include <cedar/array.h>
#include <cedar/grid_quantity.h>

namespace cedar { namespace cdr2 {

class grid_func : public array<real_t, 2>, public grid_quantity<len_t, 2>
{
public:
	using array<real_t, 2>::operator();
	grid_func(len_t nx, len_t ny, unsigned int nghosts = 1);
	grid_func() {}
	static grid_func ones(len_t nx, len_t ny);
	static grid_func zeros(len_t nx, len_t ny);
	static grid_func random(len_t nx, len_t ny);
	static grid_func like(const grid_func &likeable);
	static grid_func zeros_like(const grid_func &likeable);
	static grid_func ones_like(const grid_func &likeable);
	// signed entry of largest magnitude over the interior (src/2d/grid_func.cc:118-134)
	virtual real_t inf_norm() const;
	template <int p> real_t lp_norm() const;
	grid_func &operator-=(const grid_func &rhs);
	friend grid_func operator-(grid_func lhs, const grid_func &rhs) { return lhs -= rhs; }
	friend std::ostream &operator<<(std::ostream &os, const grid_func &obj);
private:
	real_t l2_device() const;
};

template <int p> real_t grid_func::lp_norm() const
{
	if (p == 2 && this->on_device() && num_ghosts == 1) return l2_device();
	real_t result = 0;
	const real_t *d = this->cdata();
	for (auto j : this->range(1))
		for (auto i : this->range(0)) result += std::pow(d[static_cast<std::size_t>(j) * len(0) + i], p);
	return std::pow(result, 1. / p);
}

}}
#endif
