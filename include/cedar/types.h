// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/types.h -- basic types of the B200-native Cedar host layer.
// Mirrors the public names of the reference include/cedar/types.h:23-95
// (len_t, real_t, exec_mode, relax_dir, cycle::Dir, range()) without Boost.
#ifndef CEDAR_UTIL_TYPES_H
#define CEDAR_UTIL_TYPES_H

#include <cassert>
#include <cstddef>
#include <iterator>
#include <memory>
#include <vector>

namespace cedar {

enum class exec_mode { serial, mpi };
enum class relax_dir { x, y, xy, xz, yz };

template <relax_dir rdir> struct relax_dir_name { static constexpr const char *value = ""; };
template <> struct relax_dir_name<relax_dir::xy> { static constexpr const char *value = "xy"; };
template <> struct relax_dir_name<relax_dir::xz> { static constexpr const char *value = "xz"; };
template <> struct relax_dir_name<relax_dir::yz> { static constexpr const char *value = "yz"; };

namespace cycle {
enum class Dir { UP, DOWN };
}

using len_t = unsigned int;
using real_t = double;

// half-open counting range [from, to) usable in range-for, like the reference's
// boost::iterator_range<counting_iterator<T>>.
template <class T> class range_t
{
public:
	class iterator
	{
	public:
		using iterator_category = std::forward_iterator_tag;
		using value_type = T;
		using difference_type = std::ptrdiff_t;
		using pointer = const T *;
		using reference = const T &;
		iterator() : v() {}
		explicit iterator(T v) : v(v) {}
		const T &operator*() const { return v; }
		iterator &operator++() { ++v; return *this; }
		iterator operator++(int) { iterator t(*this); ++v; return t; }
		bool operator==(const iterator &o) const { return v == o.v; }
		bool operator!=(const iterator &o) const { return v != o.v; }
	private:
		T v;
	};
	range_t() : b(), e() {}
	range_t(T from, T to) : b(from), e(to < from ? from : to) {}
	iterator begin() const { return iterator(b); }
	iterator end() const { return iterator(e); }
	T front() const { return b; }
	T back() const { return e - 1; }
	std::size_t size() const { return static_cast<std::size_t>(e - b); }
private:
	T b, e;
};

// the reference's AlignedVector (types.h:46-47, a 16-byte aligned std::vector); operator new already
// aligns doubles to 16 bytes on this platform
template <class T> using AlignedVector = std::vector<T>;

template <class T> range_t<T> range(T to) { return range_t<T>(T(), to); }
template <class T> range_t<T> range(T from, T to) { return range_t<T>(from, to); }

}

#endif
