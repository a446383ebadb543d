// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_3D_RELAX_STENCIL_H
#define CEDAR_3D_RELAX_STENCIL_H
#include <cedar/array.h>
#include <cedar/grid_quantity.h>
namespace cedar { namespace cdr3 {
// relaxation workspace (nx+2, ny+2, nz+2, 2): slot 1 = 1/diag (reference src/3d/relax_stencil.cc:9)
class relax_stencil : public array<real_t, 4>, public grid_quantity<len_t, 3>
{
public:
	relax_stencil() {}
	relax_stencil(len_t nx, len_t ny, len_t nz, unsigned int nghosts = 1)
	    : array<real_t, 4>(nx + 2 * nghosts, ny + 2 * nghosts, nz + 2 * nghosts, 2)
	{
		num_ghosts = nghosts;
		set_ranges();
	}
	using array<real_t, 4>::index;
	using array<real_t, 4>::operator();
};
}}
#endif
