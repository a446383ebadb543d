// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/3d/grid_func.h -- 3D grid function with one ghost layer (reference:
// include/cedar/3d/grid_func.h, src/3d/grid_func.cc).  Norms of device-resident
// data are reduced on the GPU (deterministic tree).
#ifndef CEDAR_3D_GRID_FUNC_H
#define CEDAR_3D_GRID_FUNC_H

#include <cmath>
#include <fstream>
#include <iostream>

#include <cedar/types.h>
#include <cedar/array.h>
#include <cedar/grid_quantity.h>

namespace cedar { namespace cdr3 {

class grid_func : public array<real_t, 3>, public grid_quantity<len_t, 3>
{
public:
	using array<real_t, 3>::operator();
	grid_func(len_t nx, len_t ny, len_t nz, unsigned int nghosts = 1);
	grid_func() {}
	static grid_func ones(len_t nx, len_t ny, len_t nz);
	static grid_func zeros(len_t nx, len_t ny, len_t nz);
	static grid_func random(len_t nx, len_t ny, len_t nz);
	static grid_func like(const grid_func &likeable);
	static grid_func zeros_like(const grid_func &likeable);
	static grid_func ones_like(const grid_func &likeable);
	virtual real_t inf_norm() const;
	template <int p> real_t lp_norm() const;
	grid_func &operator-=(const grid_func &rhs);
	friend grid_func operator-(grid_func lhs, const grid_func &rhs) { return lhs -= rhs; }
private:
	real_t l2_device() const;
};

template <int p> real_t grid_func::lp_norm() const
{
	if (p == 2 && this->on_device() && num_ghosts == 1) return l2_device();
	real_t result = 0;
	const real_t *d = this->cdata();
	for (auto k : this->range(2))
		for (auto j : this->range(1))
			for (auto i : this->range(0))
				result += std::pow(d[(static_cast<std::size_t>(k) * len(1) + j) * len(0) + i], p);
	return std::pow(result, 1. / p);
}

}}
#endif
