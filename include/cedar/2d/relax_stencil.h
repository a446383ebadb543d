// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_2D_CORE_RELAX_STENCIL_H
#define CEDAR_2D_CORE_RELAX_STENCIL_H
#include <cedar/array.h>
#include <cedar/grid_quantity.h>
namespace cedar { namespace cdr2 {
// relaxation workspace (nx+2, ny+2, 2): point relaxation keeps 1/diag in slot 1;
// line relaxation keeps D and L of L*D*L' per line, y-lines transposed
// (reference src/2d/relax_stencil.cc:5)
class relax_stencil : public array<real_t, 3>, public grid_quantity<len_t, 2>
{
public:
	relax_stencil() {}
	relax_stencil(len_t nx, len_t ny, unsigned int nghosts = 1) : array<real_t, 3>(nx + 2 * nghosts, ny + 2 * nghosts, 2)
	{
		num_ghosts = nghosts;
		set_ranges();
	}
	using array<real_t, 3>::index;
	using array<real_t, 3>::operator();
};
}}
#endif
