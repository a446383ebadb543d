// cedar/gpu.h -- device runtime glue of the host layer (no CUDA headers: every
// call goes through the C ABI of libbmgx.so, include/bmgx.h).
#ifndef CEDAR_GPU_H
#define CEDAR_GPU_H

#include <cstddef>
#include <stdexcept>
#include <string>

#include <bmgx.h>

namespace cedar { namespace gpu {

struct error : std::runtime_error { using std::runtime_error::runtime_error; };

inline void check(int status, const char *what)
{
	if (status != BMGX_OK)
		throw error(std::string(what) + ": bmgx status " + std::to_string(status) + " (" + bmgx_last_error() + ")");
}

// true when a CUDA device is usable.  Host-only code paths (building operators,
// config parsing) work without one; anything that touches dev() throws.
bool available();
// the stream every kernel of this process is enqueued on (created on first use)
bmgx_stream_t stream();
void set_stream(bmgx_stream_t s);
// run on a caller-provided stream for the lifetime of the guard (nullptr = no change);
// the library's own stream, if any, is kept and comes back afterwards
struct stream_guard
{
	explicit stream_guard(bmgx_stream_t s);
	~stream_guard();
	stream_guard(const stream_guard &) = delete;
	stream_guard &operator=(const stream_guard &) = delete;
private:
	bmgx_stream_t prev;
	bool prev_own, active;
};
void sync();
// running count of device launches issued through the host layer, including
// the nodes of replayed CUDA graphs
unsigned long long launches();
void add_launches(unsigned long long n);

}}

#endif
