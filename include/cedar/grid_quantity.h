// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/grid_quantity.h -- ghost-aware index ranges (reference: grid_quantity.h:15-71)
#ifndef CEDAR_GRID_QUANTITY_H
#define CEDAR_GRID_QUANTITY_H

#include <array>
#include <cedar/types.h>
#include <cedar/array_base.h>

namespace cedar {

template <typename len_type, unsigned short ND>
class grid_quantity : public virtual array_base<len_type>
{
public:
	// number of interior points along dimension i
	virtual len_type shape(int i) const { return this->len(i) - 2 * num_ghosts; }
	// interior indices / all indices (ghosts included) along dimension i
	virtual const range_t<len_type> &range(int i) const { return range_[i]; }
	virtual const range_t<len_type> &grange(int i) const { return grange_[i]; }
protected:
	void set_ranges()
	{
		for (unsigned short i = 0; i < ND; i++) {
			range_[i] = cedar::range(static_cast<len_type>(num_ghosts),
			                         static_cast<len_type>(this->len(i) - num_ghosts));
			grange_[i] = cedar::range(static_cast<len_type>(0), this->len(i));
		}
	}
	int num_ghosts = 1;
	std::array<range_t<len_type>, ND> range_;
	std::array<range_t<len_type>, ND> grange_;
};

}
#endif
