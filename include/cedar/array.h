// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/array.h -- Fortran-ordered ND array with a host mirror and a
// device-resident buffer.
//
// API of the reference's aarray<len_type,data_type,ND> (include/cedar/array.h:
// 23-184: reshape/init, operator(), index, len, stride, set, scale, data) plus
// the device side the B200 kernels use:
//     dev()        device pointer, read-write  (marks the host mirror stale)
//     cdev()       device pointer, read-only   (host mirror stays valid)
//     swap_device  exchange device buffers with another array of equal size
// The host mirror is allocated lazily (pinned when a GPU is present) and only
// when host code touches the array, so solver-internal levels never pay for
// it; a freshly shaped array is logically all zero on both sides, like the
// reference's zero-initialising vector::resize (array.h:65).
#ifndef CEDAR_ARRAY_H
#define CEDAR_ARRAY_H

#include <array>
#include <cassert>
#include <cstring>
#include <tuple>
#include <vector>

#include <cedar/types.h>
#include <cedar/array_base.h>
#include <cedar/gpu.h>

namespace cedar {

namespace detail {
// untyped storage shared by all array instantiations (src/dual_buffer.cc)
class dual_buffer
{
public:
	dual_buffer() {}
	~dual_buffer() { release(); }
	dual_buffer(const dual_buffer &o);
	dual_buffer &operator=(const dual_buffer &o);
	dual_buffer(dual_buffer &&o) noexcept { steal(o); }
	dual_buffer &operator=(dual_buffer &&o) noexcept { if (this != &o) { release(); steal(o); } return *this; }
	void resize(std::size_t bytes);       // contents become all zero
	std::size_t bytes() const { return nbytes; }
	void *host_rw();                      // valid host copy, device marked stale
	const void *host_ro() const;          // valid host copy
	void *dev_rw();                       // valid device copy, host marked stale
	const void *dev_ro() const;           // valid device copy
	void swap_device(dual_buffer &o);     // both end up device-valid, host stale
	void zero();                          // all zero, wherever the data lives
	void zero_device();                   // all zero on the device (a memset on the stream), host marked stale
	bool on_device() const { return dvalid; }
	const void *raw_dev() const { return d; }
	// While a recorder is installed (CUDA-graph capture of a cycle) every
	// swap_device is logged with the device pointers it exchanged, so that the
	// host bookkeeping can be replayed together with the graph.
	struct swap_event { dual_buffer *a, *b; const void *pa, *pb; };
	static void set_swap_recorder(std::vector<swap_event> *rec);
private:
	void release();
	void steal(dual_buffer &o);
	void ensure_host() const;
	void ensure_dev() const;
	std::size_t nbytes = 0;
	mutable void *h = nullptr;
	mutable void *d = nullptr;
	mutable bool hpinned = false;
	mutable bool hvalid = false, dvalid = false;  // neither valid == logically zero
};
}

template <typename len_type, typename data_type, unsigned short ND>
class aarray : public virtual array_base<len_type>
{
protected:
	detail::dual_buffer buf;
	std::array<len_type, ND> strides;
	std::array<len_type, ND> extents;

	template <typename... T> void set_extents(T... args)
	{
		static_assert(sizeof...(T) == ND, "one extent per dimension");
		len_type tmp[ND] = {static_cast<len_type>(args)...};
		for (unsigned short i = 0; i < ND; i++) extents[i] = tmp[i];
	}
	template <typename... T> std::size_t offset_of(T... args) const
	{
		static_assert(sizeof...(T) == ND, "one index per dimension");
		std::size_t idx[ND] = {static_cast<std::size_t>(args)...};
		std::size_t off = 0, mul = 1;
		for (unsigned short i = 0; i < ND; i++) {
#ifdef BOUNDS_CHECK
			assert(idx[i] < extents[i]);
#endif
			off += idx[i] * mul;
			mul *= extents[i];
		}
		return off;
	}
public:
	aarray() { strides.fill(0); extents.fill(0); }
	template <typename... T> aarray(T... args) { reshape(args...); }
	template <typename... T> void init(T... args) { reshape(args...); }
	template <typename... T> void reshape(T... args)
	{
		set_extents(args...);
		std::size_t n = 1;
		for (unsigned short i = 0; i < ND; i++) {
			strides[i] = static_cast<len_type>(n);
			n *= extents[i];
		}
		buf.resize(n * sizeof(data_type));
	}
	std::size_t size() const { return buf.bytes() / sizeof(data_type); }
	// storage for nb arrays of this shape back to back (the batch context of include/bmgx.h);
	// shape, len() and indexing keep describing ONE array; contents become all zero
	void set_batch(std::size_t nb)
	{
		std::size_t n = 1;
		for (unsigned short i = 0; i < ND; i++) n *= extents[i];
		buf.resize(n * nb * sizeof(data_type));
	}

	template <typename... T> data_type &operator()(T... args)
	{
		return static_cast<data_type *>(buf.host_rw())[offset_of(args...)];
	}
	template <typename... T> const data_type &operator()(T... args) const
	{
		return static_cast<const data_type *>(buf.host_ro())[offset_of(args...)];
	}
	template <typename... T> len_t index(T... args) const { return static_cast<len_t>(offset_of(args...)); }

	virtual len_type len(unsigned short i) const { return extents[i]; }
	virtual len_type &len(unsigned short i) { return extents[i]; }
	len_type stride(unsigned short i) const { return strides[i]; }
	len_type &stride(unsigned short i) { return strides[i]; }

	void set(data_type v);
	void scale(data_type scalar)
	{
		data_type *p = data();
		for (std::size_t i = 0, n = size(); i < n; i++) p[i] *= scalar;
	}
	// host pointers (the reference's data()); cdata() does not dirty the device copy
	data_type *data() { return static_cast<data_type *>(buf.host_rw()); }
	const data_type *cdata() const { return static_cast<const data_type *>(buf.host_ro()); }
	// device pointers
	data_type *dev() { return static_cast<data_type *>(buf.dev_rw()); }
	const data_type *cdev() const { return static_cast<const data_type *>(buf.dev_ro()); }
	void zero_device() { buf.zero_device(); }
	void swap_device(aarray &o)
	{
		assert(size() == o.size());
		buf.swap_device(o.buf);
	}
	bool on_device() const { return buf.on_device(); }
	detail::dual_buffer &buffer() { return buf; }
};

template <typename len_type, typename data_type, unsigned short ND>
void aarray<len_type, data_type, ND>::set(data_type v)
{
	if (v == data_type()) { buf.zero(); return; }
	if (buf.on_device() && sizeof(data_type) == sizeof(double)) {
		double dv = static_cast<double>(v);
		gpu::check(bmgx_fill(reinterpret_cast<double *>(dev()), size(), dv, gpu::stream()), "array::set");
		return;
	}
	data_type *p = data();
	for (std::size_t i = 0, n = size(); i < n; i++) p[i] = v;
}

template <typename data_type, unsigned short ND> using array = aarray<len_t, data_type, ND>;

}
#endif
