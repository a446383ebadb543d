// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/stencil_op_nd.h -- structured-grid stencil operator: array of shape
// (n0+2, ..., n_{nd-1}+2, ndirs), i.e. structure-of-arrays by stencil direction
// (reference: stencil_op_nd.h:41-78).  Symmetric storage: a point stores its
// centre and its W/S(/SW/NW/...) couplings, off-diagonals negated.
#ifndef CEDAR_STENCIL_OP_ND_H
#define CEDAR_STENCIL_OP_ND_H

#include <iostream>
#include <utility>

#include <cedar/types.h>
#include <cedar/array.h>
#include <cedar/grid_quantity.h>

namespace cedar {

template <std::size_t nd, class stype>
class stencil_op_nd : public array<real_t, nd + 1>, public grid_quantity<len_t, nd>
{
	using base = array<real_t, nd + 1>;
public:
	stencil_op_nd() {}
	template <typename... N, typename = typename std::enable_if<sizeof...(N) == nd>::type>
	explicit stencil_op_nd(N... n)
	{
		this->num_ghosts = 1;
		base::reshape(static_cast<len_t>(n + 2)..., static_cast<len_t>(stype::ndirs));
		this->set_ranges();
	}
	using base::len;
	using base::data;
	using base::set;
	using base::index;

	template <typename... I> real_t &operator()(I... args) { return at(std::make_index_sequence<sizeof...(I) - 1>(), std::forward_as_tuple(args...)); }
	template <typename... I> real_t operator()(I... args) const { return cat(std::make_index_sequence<sizeof...(I) - 1>(), std::forward_as_tuple(args...)); }
	int ndirs() const { return static_cast<int>(stype::ndirs); }
private:
	template <std::size_t... Is, class Tup> real_t &at(std::index_sequence<Is...>, Tup &&t)
	{
		static_assert(sizeof...(Is) == nd, "stencil_op(i..., dir)");
		stype dir = std::get<nd>(t);  // the last argument must be a direction of this stencil
		return base::operator()(static_cast<len_t>(std::get<Is>(t))..., static_cast<len_t>(dir));
	}
	template <std::size_t... Is, class Tup> real_t cat(std::index_sequence<Is...>, Tup &&t) const
	{
		static_assert(sizeof...(Is) == nd, "stencil_op(i..., dir)");
		stype dir = std::get<nd>(t);
		return base::operator()(static_cast<len_t>(std::get<Is>(t))..., static_cast<len_t>(dir));
	}
};

template <class stype> struct stencil_ndirs { static const int value = static_cast<int>(stype::ndirs); };

}
#endif
