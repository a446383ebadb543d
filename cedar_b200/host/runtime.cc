// runtime.cc -- device runtime glue (cedar/gpu.h) and the host/device dual
// buffer behind every cedar array (cedar/array.h).
#include <atomic>
#include <cstdlib>
#include <cstring>

#include <cedar/array.h>
#include <cedar/gpu.h>

namespace cedar { namespace gpu {

// the current stream is per host thread (like the batch context, the last-error text and the swap recorder): two
// solvers driven from two threads never enqueue on each other's stream or into each other's graph capture
static thread_local bmgx_stream_t g_stream = nullptr;
static thread_local bool g_stream_own = false;
static int g_avail = -1;
static std::atomic<unsigned long long> g_extra{0};

bool available()
{
	if (g_avail < 0) g_avail = bmgx_device_count() > 0 ? 1 : 0;
	return g_avail == 1;
}

bmgx_stream_t stream()
{
	if (!g_stream) {
		if (!available()) throw error("no CUDA device: the BoxMG kernels have no CPU fallback");
		check(bmgx_stream_create(&g_stream), "stream create");
		g_stream_own = true;
	}
	return g_stream;
}

void set_stream(bmgx_stream_t s)
{
	if (g_stream && g_stream_own) { bmgx_stream_sync(g_stream); bmgx_stream_destroy(g_stream); }
	g_stream = s;
	g_stream_own = false;
}

stream_guard::stream_guard(bmgx_stream_t s) : prev(g_stream), prev_own(g_stream_own), active(s != nullptr)
{
	if (active) { g_stream = s; g_stream_own = false; }
}
stream_guard::~stream_guard()
{
	if (active) { g_stream = prev; g_stream_own = prev_own; }
}

void sync() { check(bmgx_stream_sync(stream()), "stream sync"); }
unsigned long long launches() { return bmgx_launch_count() + g_extra.load(); }
void add_launches(unsigned long long n) { g_extra.fetch_add(n); }

}}

namespace cedar { namespace detail {

void dual_buffer::release()
{
	if (h) { if (hpinned) bmgx_host_free(h); else std::free(h); }
	if (d) bmgx_free(d);
	h = d = nullptr;
	hvalid = dvalid = false;
	nbytes = 0;
}

void dual_buffer::steal(dual_buffer &o)
{
	nbytes = o.nbytes; h = o.h; d = o.d; hpinned = o.hpinned; hvalid = o.hvalid; dvalid = o.dvalid;
	o.h = o.d = nullptr; o.nbytes = 0; o.hvalid = o.dvalid = false;
}

dual_buffer::dual_buffer(const dual_buffer &o) { *this = o; }

dual_buffer &dual_buffer::operator=(const dual_buffer &o)
{
	if (this == &o) return *this;
	if (nbytes != o.nbytes) { release(); nbytes = o.nbytes; }
	if (o.dvalid) {
		// device-to-device copy keeps solver-internal data off the host
		if (!d) gpu::check(bmgx_malloc(&d, nbytes), "array alloc");
		gpu::check(bmgx_memcpy_d2d(d, o.d, nbytes, gpu::stream()), "array copy");
		dvalid = true; hvalid = false;
	} else if (o.hvalid) {
		ensure_host();
		std::memcpy(h, o.h, nbytes);
		hvalid = true; dvalid = false;
	} else {
		zero();
	}
	return *this;
}

void dual_buffer::resize(std::size_t bytes)
{
	if (bytes != nbytes) { release(); nbytes = bytes; }
	hvalid = dvalid = false;  // logically all zero
}

static void *host_alloc(std::size_t bytes, bool &pinned)
{
	void *p = nullptr;
	// pinned memory makes the host<->device copies of large operators run at
	// full PCIe speed; small arrays and GPU-less hosts use the C heap
	if (bytes >= (1u << 16) && gpu::available() && bmgx_host_alloc(&p, bytes) == BMGX_OK) { pinned = true; return p; }
	pinned = false;
	if (posix_memalign(&p, 64, bytes ? bytes : 64) != 0) throw std::bad_alloc();
	return p;
}

void dual_buffer::ensure_host() const
{
	if (!h) {
		h = host_alloc(nbytes, hpinned);
		if (!dvalid) std::memset(h, 0, nbytes);
	}
	if (!hvalid) {
		if (dvalid) {
			gpu::check(bmgx_memcpy_d2h(h, d, nbytes, gpu::stream()), "array download");
			gpu::sync();
		} else {
			std::memset(h, 0, nbytes);
		}
		hvalid = true;
	}
}

void dual_buffer::ensure_dev() const
{
	if (!d) gpu::check(bmgx_malloc(&d, nbytes), "array alloc");
	if (!dvalid) {
		if (hvalid) {
			gpu::check(bmgx_memcpy_h2d(d, h, nbytes, gpu::stream()), "array upload");
			// pageable uploads return after staging; pinned ones are truly async and
			// the host copy must not be written before the copy ran
			if (hpinned) gpu::sync();
		} else {
			gpu::check(bmgx_memset0(d, nbytes, gpu::stream()), "array zero");
		}
		dvalid = true;
	}
}

void *dual_buffer::host_rw() { ensure_host(); dvalid = false; return h; }
const void *dual_buffer::host_ro() const { ensure_host(); return h; }
void *dual_buffer::dev_rw() { ensure_dev(); hvalid = false; return d; }
const void *dual_buffer::dev_ro() const { ensure_dev(); return d; }

static thread_local std::vector<dual_buffer::swap_event> *g_swap_rec = nullptr;
void dual_buffer::set_swap_recorder(std::vector<swap_event> *rec) { g_swap_rec = rec; }

void dual_buffer::swap_device(dual_buffer &o)
{
	ensure_dev(); o.ensure_dev();
	if (g_swap_rec) g_swap_rec->push_back(swap_event{this, &o, d, o.d});
	std::swap(d, o.d);
	hvalid = o.hvalid = false;
}

void dual_buffer::zero_device()
{
	if (!d) gpu::check(bmgx_malloc(&d, nbytes), "array alloc");
	gpu::check(bmgx_memset0(d, nbytes, gpu::stream()), "array zero");
	dvalid = true; hvalid = false;
}

void dual_buffer::zero()
{
	if (d && (dvalid || !h)) {
		gpu::check(bmgx_memset0(d, nbytes, gpu::stream()), "array zero");
		dvalid = true; hvalid = false;
	} else if (h) {
		std::memset(h, 0, nbytes);
		hvalid = true; dvalid = false;
	} else {
		hvalid = dvalid = false;
	}
}

}}
