// batch_stage.h -- a stream of right-hand sides through one solver: uploads, cycles and downloads of
// consecutive right-hand sides overlap (cdr2_/cdr3_solver_vcycle_batch, include/cedar_capi.h).
#ifndef CEDAR_B200_BATCH_STAGE_H
#define CEDAR_B200_BATCH_STAGE_H

#include <cstddef>
#include <bmgx.h>
#include <cedar/gpu.h>

// device staging: right-hand sides arrive in db[], solutions leave from dx[]
struct batch_stage
{
	void *db[2] = {nullptr, nullptr}, *dx[2] = {nullptr, nullptr};
	bmgx_stream_t up = nullptr, down = nullptr;
	void *b_ready[2] = {}, *b_free[2] = {}, *x_ready[2] = {}, *x_free[2] = {};
	bool live = false;
	void init(size_t bytes)
	{
		if (live) return;
		for (int q = 0; q < 2; q++) {
			cedar::gpu::check(bmgx_malloc(&db[q], bytes), "batch staging");
			cedar::gpu::check(bmgx_malloc(&dx[q], bytes), "batch staging");
			cedar::gpu::check(bmgx_event_create(&b_ready[q]), "event"); cedar::gpu::check(bmgx_event_create(&b_free[q]), "event");
			cedar::gpu::check(bmgx_event_create(&x_ready[q]), "event"); cedar::gpu::check(bmgx_event_create(&x_free[q]), "event");
		}
		cedar::gpu::check(bmgx_stream_create(&up), "stream"); cedar::gpu::check(bmgx_stream_create(&down), "stream");
		live = true;
	}
	~batch_stage()
	{
		if (!live) return;
		for (int q = 0; q < 2; q++) {
			bmgx_free(db[q]); bmgx_free(dx[q]);
			bmgx_event_destroy(b_ready[q]); bmgx_event_destroy(b_free[q]);
			bmgx_event_destroy(x_ready[q]); bmgx_event_destroy(x_free[q]);
		}
		bmgx_stream_destroy(up); bmgx_stream_destroy(down);
	}
};


// For k < nrhs: b <- hb[k] (through the upload stream and a staging slot), x = 0, cycle(), hx[k] <- x (through a
// staging slot and the download stream).  `cycle` enqueues one cycle on (x, b) on the library's stream; its CUDA
// graph is bound to the device buffers of x and b, which is why the staging copies are needed.
template <class GF, class Cycle>
void run_batch(batch_stage &st, GF &b, GF &x, int nrhs, const double *const *hb, double *const *hx, Cycle &&cycle)
{
	using cedar::gpu::check;
	const std::size_t bytes = b.size() * sizeof(double);
	st.init(bytes);
	bmgx_stream_t cs = cedar::gpu::stream();
	for (int k = 0; k < nrhs; k++) {
		const int q = k & 1;
		// upload stream: b[k] into its staging slot once the cycle two steps back has taken the slot's contents
		if (k >= 2) check(bmgx_stream_wait_event(st.up, st.b_free[q]), "batch wait");
		check(bmgx_memcpy_h2d(st.db[q], hb[k], bytes, st.up), "batch upload");
		check(bmgx_event_record(st.b_ready[q], st.up), "batch record");
		// compute stream
		check(bmgx_stream_wait_event(cs, st.b_ready[q]), "batch wait");
		check(bmgx_memcpy_d2d(b.dev(), st.db[q], bytes, cs), "batch b");
		check(bmgx_event_record(st.b_free[q], cs), "batch record");
		x.zero_device();
		cycle();
		if (k >= 2) check(bmgx_stream_wait_event(cs, st.x_free[q]), "batch wait");
		check(bmgx_memcpy_d2d(st.dx[q], x.cdev(), bytes, cs), "batch x");
		check(bmgx_event_record(st.x_ready[q], cs), "batch record");
		// download stream
		check(bmgx_stream_wait_event(st.down, st.x_ready[q]), "batch wait");
		check(bmgx_memcpy_d2h(hx[k], st.dx[q], bytes, st.down), "batch download");
		check(bmgx_event_record(st.x_free[q], st.down), "batch record");
	}
	check(bmgx_stream_sync(st.down), "batch sync");
	check(bmgx_stream_sync(cs), "batch sync");
}

#endif
