// bmgx_util.cu -- error state, launch counter, fill, deterministic sum of squares.
#include "bmgx_common.cuh"
#include <cstdlib>
#include <atomic>
#include <cstdio>
#include <cstring>

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

int bmgx_set_cuda_error(cudaError_t e, const char *where)
{
    snprintf(g_err, sizeof(g_err), "%s: %s", where, cudaGetErrorString(e));
    return BMGX_ERR_CUDA;
}
int bmgx_set_error(int code, const char *msg)
{
    snprintf(g_err, sizeof(g_err), "%s", msg);
    return code;
}
void bmgx_count_launch(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

extern "C" const char *bmgx_version(void) { return "bmgx 0.1 (sm_100a)"; }
extern "C" const char *bmgx_last_error(void) { return g_err; }
extern "C" unsigned long long bmgx_launch_count(void) { return g_launches.load(); }

// ---- library-private scratch pool -------------------------------------------------
// Setup kernels that need large transient workspace (3D Galerkin) take it stream-ordered
// from a memory pool owned by this library, one per device, created on first use with
// an unlimited release threshold so that repeated setups reuse the same pages instead
// of paying cudaMalloc/cudaFree of >1 GB per call.  The application's default pool is
// never touched; bmgx_scratch_trim() hands the cached pages back to the driver.
#include <mutex>
static std::mutex g_pool_mu;
static cudaMemPool_t g_pool[64] = {};
static cudaError_t scratch_pool(cudaMemPool_t *out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_pool_mu);
    if (!g_pool[dev]) {
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        e = cudaMemPoolCreate(&g_pool[dev], &props);
        if (e != cudaSuccess) { g_pool[dev] = nullptr; return e; }
        unsigned long long keep = ~0ull;
        e = cudaMemPoolSetAttribute(g_pool[dev], cudaMemPoolAttrReleaseThreshold, &keep);
        if (e != cudaSuccess) return e;
    }
    *out = g_pool[dev];
    return cudaSuccess;
}
cudaError_t bmgx_scratch_alloc(void **p, size_t bytes, cudaStream_t st)
{
    cudaMemPool_t pool;
    cudaError_t e = scratch_pool(&pool);
    if (e != cudaSuccess) return e;
    return cudaMallocFromPoolAsync(p, bytes, pool, st);
}
cudaError_t bmgx_scratch_free(void *p, cudaStream_t st) { return cudaFreeAsync(p, st); }
extern "C" int bmgx_scratch_trim(size_t keep_bytes)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return bmgx_set_cuda_error(e, "scratch_trim");
    std::lock_guard<std::mutex> lk(g_pool_mu);
    if (dev >= 0 && dev < 64 && g_pool[dev]) {
        e = cudaMemPoolTrimTo(g_pool[dev], keep_bytes);
        if (e != cudaSuccess) return bmgx_set_cuda_error(e, "scratch_trim");
    }
    return BMGX_OK;
}

// batch context, per host thread (the host layer is single threaded per solver)
static thread_local unsigned g_batch = 1;
unsigned bmgx_batch() { return g_batch; }
extern "C" int bmgx_set_batch(unsigned nb)
{
    if (nb < 1 || nb > 65535) return bmgx_set_error(BMGX_ERR_ARG, "set_batch: 1 <= nb <= 65535");
    g_batch = nb;
    return BMGX_OK;
}
extern "C" unsigned bmgx_get_batch(void) { return g_batch; }

// src/2d/ftn/BMG_get_bc.f90:13-22
extern "C" void bmgx_get_bc(int per_mask, int *ibc)
{
    static const int bcmap[8] = {0, 2, 1, 3, 5, 6, 7, 8};
    *ibc = bcmap[per_mask & 7];
}

__global__ void k_fill(double *a, size_t n, double v)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) a[i] = v;
}

extern "C" int bmgx_fill(double *a, size_t n, double v, bmgx_stream_t s)
{
    if (n == 0) return BMGX_OK;
    if (v == 0.0) return bmgx_memset0(a, n * sizeof(double), s);
    unsigned nb = cdiv(n, 256);
    if (nb > 148 * 16) nb = 148 * 16;
    k_fill<<<nb, 256, 0, (cudaStream_t)s>>>(a, n, v);
    BMGX_LAUNCH_CHECK("bmgx_fill");
    return BMGX_OK;
}

// ---- deterministic sum of squares over the interior ------------------------
// Stage 1: NB blocks, each sums a fixed, contiguous set of rows with a fixed
// in-block tree; stage 2: one block folds the NB partials in a fixed tree.
// The result does not depend on scheduling (bitwise reproducible run to run).
#define NORM_NB 1184  // 148 SMs x 8
extern "C" size_t bmgx_norm_work_len(void) { return NORM_NB; }

__device__ __forceinline__ double block_sum_256(double v)
{
    __shared__ double sh[8];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) sh[w] = v;
    __syncthreads();
    double r = 0.0;
    if (w == 0) {
        r = l < 8 ? sh[l] : 0.0;
        for (int o = 4; o > 0; o >>= 1) r += __shfl_down_sync(0xffffffffu, r, o);
    }
    __syncthreads();
    return r;  // valid in thread 0
}

__global__ void __launch_bounds__(256) k_sumsq_stage1(const double *__restrict__ q, unsigned II, size_t nrows, size_t rowstride2,
                                                      unsigned JJ, double *__restrict__ part)
{
    // "rows" enumerates interior (j[,k]) lines; line r -> j = r % (JJ-2) + 1 (0-based), k = r / (JJ-2) + 1
    size_t r0 = nrows * blockIdx.x / gridDim.x, r1 = nrows * (blockIdx.x + 1) / gridDim.x;
    double acc = 0.0;
    unsigned nj = JJ - 2;
    for (size_t r = r0; r < r1; r++) {
        size_t j = r % nj + 1, k = r / nj;
        const double *row = q + (rowstride2 ? (k + 1) * rowstride2 : 0) + j * (size_t)II;
        for (unsigned i = 1 + threadIdx.x; i < II - 1; i += 256) {
            double v = row[i];
            acc += v * v;
        }
    }
    double t = block_sum_256(acc);
    if (threadIdx.x == 0) part[blockIdx.x] = t;
}

__global__ void __launch_bounds__(256) k_sumsq_stage2(const double *__restrict__ part, int n, double *out)
{
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) acc += part[i];
    double t = block_sum_256(acc);
    if (threadIdx.x == 0) *out = t;
}

extern "C" int bmgx2_sumsq(const double *q, bmgx_len ii, bmgx_len jj, double *work, double *out, bmgx_stream_t s)
{
    if (ii < 3 || jj < 3) return bmgx_set_error(BMGX_ERR_ARG, "bmgx2_sumsq: empty interior");
    size_t nrows = jj - 2;
    int nb = nrows < NORM_NB ? (int)nrows : NORM_NB;
    k_sumsq_stage1<<<nb, 256, 0, (cudaStream_t)s>>>(q, ii, nrows, 0, jj, work);
    BMGX_LAUNCH_CHECK("bmgx2_sumsq/1");
    k_sumsq_stage2<<<1, 256, 0, (cudaStream_t)s>>>(work, nb, out);
    BMGX_LAUNCH_CHECK("bmgx2_sumsq/2");
    return BMGX_OK;
}

extern "C" int bmgx3_sumsq(const double *q, bmgx_len ii, bmgx_len jj, bmgx_len kk, double *work, double *out, bmgx_stream_t s)
{
    if (ii < 3 || jj < 3 || kk < 3) return bmgx_set_error(BMGX_ERR_ARG, "bmgx3_sumsq: empty interior");
    size_t nrows = (size_t)(jj - 2) * (kk - 2);
    int nb = nrows < NORM_NB ? (int)nrows : NORM_NB;
    k_sumsq_stage1<<<nb, 256, 0, (cudaStream_t)s>>>(q, ii, nrows, (size_t)ii * jj, jj, work);
    BMGX_LAUNCH_CHECK("bmgx3_sumsq/1");
    k_sumsq_stage2<<<1, 256, 0, (cudaStream_t)s>>>(work, nb, out);
    BMGX_LAUNCH_CHECK("bmgx3_sumsq/2");
    return BMGX_OK;
}

// ---- runtime support for the host layer -----------------------------------
#define RT(call, where)                                                \
    do {                                                               \
        cudaError_t e__ = (call);                                      \
        if (e__ != cudaSuccess) return bmgx_set_cuda_error(e__, where); \
        return BMGX_OK;                                                \
    } while (0)

extern "C" int bmgx_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}
extern "C" int bmgx_set_device(int dev) { RT(cudaSetDevice(dev), "bmgx_set_device"); }
extern "C" int bmgx_malloc(void **p, size_t bytes) { RT(cudaMalloc(p, bytes ? bytes : 8), "bmgx_malloc"); }
extern "C" int bmgx_free(void *p) { RT(cudaFree(p), "bmgx_free"); }
extern "C" int bmgx_host_alloc(void **p, size_t bytes) { RT(cudaHostAlloc(p, bytes ? bytes : 8, cudaHostAllocDefault), "bmgx_host_alloc"); }
extern "C" int bmgx_host_free(void *p) { RT(cudaFreeHost(p), "bmgx_host_free"); }
extern "C" int bmgx_memcpy_h2d(void *dst, const void *src, size_t bytes, bmgx_stream_t s)
{
    RT(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)s), "bmgx_memcpy_h2d");
}
extern "C" int bmgx_memcpy_d2h(void *dst, const void *src, size_t bytes, bmgx_stream_t s)
{
    RT(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)s), "bmgx_memcpy_d2h");
}
extern "C" int bmgx_memcpy_d2d(void *dst, const void *src, size_t bytes, bmgx_stream_t s)
{
    RT(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)s), "bmgx_memcpy_d2d");
}
// small clears (the coarse-level iterates of a cycle) as a kernel, so that the chain of programmatic
// dependent launches is not cut by a memset node
__global__ void __launch_bounds__(256) k_zero(double *a, size_t n)
{
    pdl_wait();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) a[i] = 0.0;
}

extern "C" int bmgx_memset0(void *dst, size_t bytes, bmgx_stream_t s)
{
    if (bytes > 0 && bytes <= ((size_t)2 << 20) && bytes % 8 == 0 && ((uintptr_t)dst & 7) == 0) {
        size_t n = bytes / 8;
        launch_pdl(k_zero, dim3(cdiv(n, 1024) > 296 ? 296 : cdiv(n, 1024)), dim3(256), 0, (cudaStream_t)s, (double *)dst, n);
        BMGX_LAUNCH_CHECK("bmgx_memset0");
        return BMGX_OK;
    }
    bmgx_count_launch();
    RT(cudaMemsetAsync(dst, 0, bytes, (cudaStream_t)s), "bmgx_memset0");
}
extern "C" int bmgx_stream_create(bmgx_stream_t *s)
{
    RT(cudaStreamCreateWithFlags((cudaStream_t *)s, cudaStreamNonBlocking), "bmgx_stream_create");
}
extern "C" int bmgx_stream_is_capturing(bmgx_stream_t s, int *capturing)
{
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    cudaError_t e = cudaStreamIsCapturing((cudaStream_t)s, &st);
    if (e != cudaSuccess) return bmgx_set_cuda_error(e, "bmgx_stream_is_capturing");
    *capturing = st == cudaStreamCaptureStatusActive;
    return BMGX_OK;
}
extern "C" int bmgx_stream_destroy(bmgx_stream_t s) { RT(cudaStreamDestroy((cudaStream_t)s), "bmgx_stream_destroy"); }
extern "C" int bmgx_stream_sync(bmgx_stream_t s) { RT(cudaStreamSynchronize((cudaStream_t)s), "bmgx_stream_sync"); }
extern "C" int bmgx_device_sync(void) { RT(cudaDeviceSynchronize(), "bmgx_device_sync"); }
extern "C" int bmgx_event_create(void **ev) { RT(cudaEventCreate((cudaEvent_t *)ev), "bmgx_event_create"); }
extern "C" int bmgx_event_destroy(void *ev) { RT(cudaEventDestroy((cudaEvent_t)ev), "bmgx_event_destroy"); }
extern "C" int bmgx_event_record(void *ev, bmgx_stream_t s) { RT(cudaEventRecord((cudaEvent_t)ev, (cudaStream_t)s), "bmgx_event_record"); }
// periodic ghost refresh of a 2D grid function (used by the periodic branches of relax_GS and restrict)
__global__ void __launch_bounds__(128) k_wrap_x(double *q, unsigned II, unsigned jbeg, unsigned jstep, unsigned nrow)
{
    unsigned r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrow) return;
    double *row = q + (size_t)(jbeg + jstep * r - 1) * II;
    row[0] = row[II - 2];       // Q(1,J)  = Q(I1,J)
    row[II - 1] = row[1];       // Q(II,J) = Q(2,J)
}
__global__ void __launch_bounds__(128) k_wrap_y(double *q, unsigned II, unsigned JJ)
{
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= II) return;
    q[i] = q[(size_t)(JJ - 2) * II + i];                 // Q(I,1)  = Q(I,J1)
    q[(size_t)(JJ - 1) * II + i] = q[(size_t)II + i];    // Q(I,JJ) = Q(I,2)
}

int bmgx_wrap_x(double *q, unsigned ii, unsigned jbeg, unsigned jstep, unsigned nrow, cudaStream_t st)
{
    if (nrow == 0) return BMGX_OK;
    k_wrap_x<<<cdiv(nrow, 128), 128, 0, st>>>(q, ii, jbeg, jstep, nrow);
    BMGX_LAUNCH_CHECK("wrap x");
    return BMGX_OK;
}
int bmgx_wrap_y(double *q, unsigned ii, unsigned jj, cudaStream_t st)
{
    k_wrap_y<<<cdiv(ii, 128), 128, 0, st>>>(q, ii, jj);
    BMGX_LAUNCH_CHECK("wrap y");
    return BMGX_OK;
}

bool bmgx_pdl_enabled()
{
    // measured on the 4097^2 V(1,1) cycle: 1.172 -> 1.168 ms; the coarse-level tail is bound by the kernels'
    // own load round trips, not by launch gaps -- kept opt-in (BMGX_PDL=1)
    static const bool on = [] { const char *e = getenv("BMGX_PDL"); return e && e[0] == '1'; }();
    return on;
}

extern "C" int bmgx_stream_wait_event(bmgx_stream_t s, void *ev) { RT(cudaStreamWaitEvent((cudaStream_t)s, (cudaEvent_t)ev, 0), "bmgx_stream_wait_event"); }
extern "C" int bmgx_event_sync(void *ev) { RT(cudaEventSynchronize((cudaEvent_t)ev), "bmgx_event_sync"); }
extern "C" int bmgx_event_elapsed_ms(void *a, void *b, float *ms) { RT(cudaEventElapsedTime(ms, (cudaEvent_t)a, (cudaEvent_t)b), "bmgx_event_elapsed_ms"); }

extern "C" int bmgx_graph_begin(bmgx_stream_t s)
{
    RT(cudaStreamBeginCapture((cudaStream_t)s, cudaStreamCaptureModeThreadLocal), "bmgx_graph_begin");
}
extern "C" int bmgx_graph_end(bmgx_stream_t s, void **graph_exec)
{
    cudaGraph_t g = nullptr;
    cudaError_t e = cudaStreamEndCapture((cudaStream_t)s, &g);
    if (e != cudaSuccess) return bmgx_set_cuda_error(e, "bmgx_graph_end/capture");
    cudaGraphExec_t ex = nullptr;
    e = cudaGraphInstantiate(&ex, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) return bmgx_set_cuda_error(e, "bmgx_graph_end/instantiate");
    *graph_exec = (void *)ex;
    return BMGX_OK;
}
extern "C" int bmgx_graph_launch(void *graph_exec, bmgx_stream_t s)
{
    RT(cudaGraphLaunch((cudaGraphExec_t)graph_exec, (cudaStream_t)s), "bmgx_graph_launch");
}
extern "C" int bmgx_graph_destroy(void *graph_exec) { RT(cudaGraphExecDestroy((cudaGraphExec_t)graph_exec), "bmgx_graph_destroy"); }

// ---- ghost-line staging for the block-decomposed solver -----------------------
// Copies line `index` (1-based; dir 0: the column i = index, all j; dir 1: the row
// j = index, all i) of every selected plane of A(II,JJ,nplanes) to (to_buf != 0) or
// from a contiguous buffer laid out [selected plane][position along the line].
__global__ void __launch_bounds__(256)
k_line_copy(double *a, unsigned II, unsigned JJ, unsigned nplanes, unsigned mask, int dir, unsigned index,
            double *buf, int to_buf)
{
    const unsigned len = dir == 0 ? JJ : II;
    const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= len) return;
    unsigned slot = 0;
    for (unsigned p = 0; p < nplanes; p++) {
        if (!(mask & (1u << p))) continue;
        const size_t off = (size_t)p * II * JJ + (dir == 0 ? (size_t)t * II + (index - 1) : (size_t)(index - 1) * II + t);
        if (to_buf) buf[(size_t)slot * len + t] = a[off];
        else a[off] = buf[(size_t)slot * len + t];
        slot++;
    }
}

extern "C" int bmgx2_line_copy(double *a, bmgx_len ii, bmgx_len jj, int nplanes, unsigned plane_mask, int dir,
                               bmgx_len index, double *buf, int to_buf, bmgx_stream_t s)
{
    if (dir < 0 || dir > 1 || nplanes < 1 || nplanes > 32 || index < 1 || index > (dir == 0 ? ii : jj))
        return bmgx_set_error(BMGX_ERR_ARG, "line_copy: bad arguments");
    const unsigned len = dir == 0 ? jj : ii;
    k_line_copy<<<cdiv(len, 256), 256, 0, (cudaStream_t)s>>>(a, ii, jj, nplanes, plane_mask, dir, index, buf, to_buf);
    BMGX_LAUNCH_CHECK("line_copy");
    return BMGX_OK;
}
