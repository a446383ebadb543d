// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_ARRAY_BASE_H
#define CEDAR_ARRAY_BASE_H
namespace cedar {
// common virtual base of arrays and grid quantities (reference: array_base.h)
template <typename len_type> class array_base
{
public:
	virtual ~array_base() {}
	virtual len_type len(unsigned short i) const = 0;
};
}
#endif
