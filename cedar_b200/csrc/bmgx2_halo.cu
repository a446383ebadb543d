// bmgx2_halo.cu -- ghost-layer exchange over peer memory (NVLink / NVSwitch) for the
// block-decomposed path: the B200 replacement of the reference's MPI halo exchange
// (include/cedar/kernels/halo_exchange.h:22-31, src/2d/mpi/halo.cc: MSG/Tausch
// pack -> MPI_Isend/Irecv -> unpack, called from inside the MPI kernels, e.g.
// src/2d/ftn/mpi/BMG2_SymStd_relax_GS.f90:126,171).
//
// One process per GPU.  Every rank keeps ALL arrays of its block in one device
// allocation (the arena); the ranks open each other's arenas once through CUDA IPC
// (bmgx_ipc_*), so a rank reads the halo lines it needs straight out of the
// neighbour's array through NVLink: no packing, no staging buffers, no library
// send/recv.  One exchange = ONE kernel launch (k_halo_step):
//
//   1. signal: the rank publishes its exchange counter e (all its previous
//      kernels are complete: stream order) into flags[me] of every peer's arena
//      (st.release.sys after a system fence);
//   2. wait:   every CTA spins (ld.acquire.sys) until the rank it reads from has
//      published >= e, i.e. has produced the data of this exchange;
//   3. pull:   the CTAs copy their segment (a rows x cols rectangle) from the
//      peer's array into the local halo with system-scope loads.
//
// A pull needs no "receiver ready" handshake: a rank signals exchange e+1 only after
// its pulls of exchange e are complete, so waiting for the neighbour's signal e+1
// also proves that the neighbour is done reading this rank's data of exchange e (the
// host-side planner never overwrites a pulled region before the next exchange,
// include/cedar/2d/mpi/plan.h).  The counter lives in device memory and is advanced
// by the last CTA of the kernel, so a captured CUDA graph of a whole cycle replays
// with fresh counter values.  A wait that exceeds the time-out records an error code
// in the arena instead of hanging the GPU.
//
// The same launch with phase = 1 (signal only) / 2 (wait + pull) serves several
// blocks living in ONE process (single-GPU tests of the decomposed path).
#include "bmgx_common.cuh"

namespace {

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ double ld_sys(const double *p)
{
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long now_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// sync block of an arena (unsigned long long words):
//   [0, world)   flags: flags[r] = last exchange counter rank r has published to this rank
//   [world]      epoch: number of exchanges this rank has completed
//   [world + 1]  CTAs of the running exchange kernel that have finished
//   [world + 2]  error word (0 = ok; 1 + rank waited for = time-out)
__global__ void __launch_bounds__(256)
k_halo_step(const bmgx_seg *__restrict__ segs, int nseg, unsigned long long *const *__restrict__ peer_sync,
            unsigned long long *my, int me, int world, int phase, int wait_all, unsigned long long timeout_ns)
{
    const unsigned long long e = ld_acquire_sys(my + world) + 1;
    if ((phase & 1) && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x < (unsigned)world) {
        __threadfence_system();
        st_release_sys(peer_sync[threadIdx.x] + me, e);
    }
    if (!(phase & 2)) return;
    const bmgx_seg sg = nseg > 0 ? segs[blockIdx.y] : bmgx_seg{};
    {
        int r = -1;
        if (wait_all) { if (threadIdx.x < (unsigned)world) r = (int)threadIdx.x; }
        else if (threadIdx.x == 0 && nseg > 0) r = sg.src_rank;
        if (r >= 0) {
            const unsigned long long t0 = now_ns();
            while (ld_acquire_sys(my + r) < e) {
                if (now_ns() - t0 > timeout_ns) { atomicExch(my + world + 2, 1ull + (unsigned long long)r); break; }
                __nanosleep(32);
            }
        }
    }
    __syncthreads();
    if (nseg > 0) {
        const size_t n = (size_t)sg.rows * sg.cols;
        const size_t stride = (size_t)gridDim.x * blockDim.x;
        size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        // four independent system-scope loads in flight per thread
        for (; i + 3 * stride < n; i += 4 * stride) {
            double v[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const size_t k = i + u * stride;
                const size_t r = k / sg.cols, c = k - r * sg.cols;
                v[u] = ld_sys(sg.src + r * sg.src_pitch + c);
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const size_t k = i + u * stride;
                const size_t r = k / sg.cols, c = k - r * sg.cols;
                sg.dst[r * sg.dst_pitch + c] = v[u];
            }
        }
        for (; i < n; i += stride) {
            const size_t r = i / sg.cols, c = i - r * sg.cols;
            sg.dst[r * sg.dst_pitch + c] = ld_sys(sg.src + r * sg.src_pitch + c);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned long long total = (unsigned long long)gridDim.x * gridDim.y;
        if (atomicAdd(my + world + 1, 1ull) == total - 1) {
            my[world + 1] = 0;
            __threadfence();
            st_release_sys(my + world, e);
        }
    }
}

__global__ void __launch_bounds__(256)
k_copy2d(double *__restrict__ dst, size_t dpitch, const double *__restrict__ src, size_t spitch, unsigned cols, unsigned rows)
{
    const unsigned c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cols) return;
    for (unsigned r = blockIdx.y; r < rows; r += gridDim.y) dst[(size_t)r * dpitch + c] = src[(size_t)r * spitch + c];
}

// deterministic sum of squares of a rows x cols window (pitch in doubles): fixed rows per CTA, fixed in-CTA tree
#define VN_NB 592
__device__ __forceinline__ double block_sum(double v)
{
    __shared__ double sh[8];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) sh[w] = v;
    __syncthreads();
    double r = 0.0;
    if (w == 0) {
        r = l < 8 ? sh[l] : 0.0;
        for (int o = 4; o > 0; o >>= 1) r += __shfl_down_sync(0xffffffffu, r, o);
    }
    __syncthreads();
    return r;
}
__global__ void __launch_bounds__(256)
k_sumsq_win1(const double *__restrict__ q, size_t pitch, unsigned cols, unsigned rows, double *__restrict__ part)
{
    const size_t r0 = (size_t)rows * blockIdx.x / gridDim.x, r1 = (size_t)rows * (blockIdx.x + 1) / gridDim.x;
    double acc = 0.0;
    for (size_t r = r0; r < r1; r++)
        for (unsigned c = threadIdx.x; c < cols; c += 256) {
            const double v = q[r * pitch + c];
            acc += v * v;
        }
    const double t = block_sum(acc);
    if (threadIdx.x == 0) part[blockIdx.x] = t;
}
__global__ void __launch_bounds__(256) k_sum_fold(const double *__restrict__ part, int n, double *out)
{
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) acc += part[i];
    const double t = block_sum(acc);
    if (threadIdx.x == 0) *out = t;
}
// out = v[0] + v[1] + ... + v[n-1], in this order (the same on every rank)
__global__ void k_sum_ordered(const double *v, int n, double *out)
{
    double acc = 0.0;
    for (int i = 0; i < n; i++) acc += v[i];
    *out = acc;
}

}  // namespace

#define RT(call, where)                                                \
    do {                                                               \
        cudaError_t e__ = (call);                                      \
        if (e__ != cudaSuccess) return bmgx_set_cuda_error(e__, where); \
        return BMGX_OK;                                                \
    } while (0)

extern "C" int bmgx_ipc_export(void *devptr, void *handle64)
{
    static_assert(sizeof(cudaIpcMemHandle_t) == BMGX_IPC_HANDLE_BYTES, "IPC handle size");
    RT(cudaIpcGetMemHandle((cudaIpcMemHandle_t *)handle64, devptr), "bmgx_ipc_export");
}
extern "C" int bmgx_ipc_open(const void *handle64, void **devptr)
{
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    RT(cudaIpcOpenMemHandle(devptr, h, cudaIpcMemLazyEnablePeerAccess), "bmgx_ipc_open");
}
extern "C" int bmgx_ipc_close(void *devptr) { RT(cudaIpcCloseMemHandle(devptr), "bmgx_ipc_close"); }

extern "C" int bmgx_halo_step(const bmgx_seg *segs_dev, int nseg, int max_chunks, void *const *peer_sync_dev, void *my_sync,
                              int me, int world, int phase, int wait_all, double timeout_s, bmgx_stream_t s)
{
    if (world < 1 || world > 256 || me < 0 || me >= world || nseg < 0 || !(phase & 3))
        return bmgx_set_error(BMGX_ERR_ARG, "halo_step: bad arguments");
    if (max_chunks < 1) max_chunks = 1;
    if (max_chunks > 64) max_chunks = 64;
    dim3 grid((phase & 2) ? max_chunks : 1, (phase & 2) && nseg > 0 ? nseg : 1);
    if (nseg == 0) grid.x = 1;
    k_halo_step<<<grid, 256, 0, (cudaStream_t)s>>>(segs_dev, nseg, (unsigned long long *const *)peer_sync_dev,
                                                  (unsigned long long *)my_sync, me, world, phase, wait_all,
                                                  (unsigned long long)(timeout_s * 1e9));
    BMGX_LAUNCH_CHECK("halo_step");
    return BMGX_OK;
}

extern "C" int bmgx_copy2d(double *dst, size_t dpitch, const double *src, size_t spitch, bmgx_len cols, bmgx_len rows,
                           bmgx_stream_t s)
{
    if (cols == 0 || rows == 0) return BMGX_OK;
    dim3 grid(cdiv(cols, 256), rows < 1024 ? rows : 1024);
    k_copy2d<<<grid, 256, 0, (cudaStream_t)s>>>(dst, dpitch, src, spitch, cols, rows);
    BMGX_LAUNCH_CHECK("copy2d");
    return BMGX_OK;
}
extern "C" int bmgx_memcpy2d_h2d(double *dst, size_t dpitch, const double *src, size_t spitch, bmgx_len cols, bmgx_len rows,
                                 bmgx_stream_t s)
{
    RT(cudaMemcpy2DAsync(dst, dpitch * 8, src, spitch * 8, (size_t)cols * 8, rows, cudaMemcpyHostToDevice, (cudaStream_t)s),
       "bmgx_memcpy2d_h2d");
}
extern "C" int bmgx_memcpy2d_d2h(double *dst, size_t dpitch, const double *src, size_t spitch, bmgx_len cols, bmgx_len rows,
                                 bmgx_stream_t s)
{
    RT(cudaMemcpy2DAsync(dst, dpitch * 8, src, spitch * 8, (size_t)cols * 8, rows, cudaMemcpyDeviceToHost, (cudaStream_t)s),
       "bmgx_memcpy2d_d2h");
}

extern "C" size_t bmgx_sumsq_win_work_len(void) { return VN_NB; }
extern "C" int bmgx2_sumsq_win(const double *q, size_t pitch, bmgx_len cols, bmgx_len rows, double *work, double *out,
                               bmgx_stream_t s)
{
    if (cols == 0 || rows == 0) return bmgx_set_error(BMGX_ERR_ARG, "sumsq_win: empty window");
    const int nb = rows < VN_NB ? (int)rows : VN_NB;
    k_sumsq_win1<<<nb, 256, 0, (cudaStream_t)s>>>(q, pitch, cols, rows, work);
    BMGX_LAUNCH_CHECK("sumsq_win/1");
    k_sum_fold<<<1, 256, 0, (cudaStream_t)s>>>(work, nb, out);
    BMGX_LAUNCH_CHECK("sumsq_win/2");
    return BMGX_OK;
}
extern "C" int bmgx_sum_ordered(const double *v, int n, double *out, bmgx_stream_t s)
{
    k_sum_ordered<<<1, 1, 0, (cudaStream_t)s>>>(v, n, out);
    BMGX_LAUNCH_CHECK("sum_ordered");
    return BMGX_OK;
}
