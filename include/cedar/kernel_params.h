// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/kernel_params.h -- parameters shared by every registered kernel
// (reference: include/cedar/kernel_params.h:9-46, src/kernel_params.cc:61-97).
#ifndef CEDAR_KERNEL_PARAMS_H
#define CEDAR_KERNEL_PARAMS_H

#include <array>
#include <iostream>
#include <memory>
#include <cedar/config.h>

namespace cedar {

struct ml_relax_params
{
	ml_relax_params() : enabled(false), min_gsz(3), factorize(true) {}
	void init(config &conf);
	bool enabled;
	int min_gsz;
	bool factorize;
};

struct kernel_params
{
	enum class halo_lib { tausch, msg };
	friend std::ostream &operator<<(std::ostream &os, const kernel_params &obj);
	std::array<bool, 3> periodic;
	bool relax_symmetric;
	bool definite;
	halo_lib halo;
	ml_relax_params ml_relax;
	std::shared_ptr<config> plane_config;
	int per_mask() const
	{
		int mask = 0;
		for (int i = 0; i < 3; ++i)
			if (periodic[i]) mask |= (1 << i);
		return mask;
	}
};

std::shared_ptr<kernel_params> build_kernel_params(config &conf);

}
#endif
