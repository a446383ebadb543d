// bmgx_common.cuh -- shared helpers for the sm_100a BoxMG kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/bmgx.h"

// Fortran-order, 1-based accessors (ghost layer included): the kernels keep the
// reference's index conventions so every formula can be checked against the
// .f90 line it mirrors.
#define IX2(n1, i, j) ((size_t)((j)-1) * (size_t)(n1) + (size_t)((i)-1))
#define IX3(n1, n2, i, j, k) (((size_t)((k)-1) * (size_t)(n2) + (size_t)((j)-1)) * (size_t)(n1) + (size_t)((i)-1))

// stencil slots, src/2d/ftn/BMG_stencils_f90.h:29-37
enum { KO = 1, KW = 2, KS = 3, KSW = 4, KNW = 5 };
enum { LL = 1, LR = 2, LA = 3, LB = 4, LSW = 5, LNW = 6, LNE = 7, LSE = 8 };
enum { MSOR = 2 };

int bmgx_set_cuda_error(cudaError_t e, const char *where);
int bmgx_set_error(int code, const char *msg);
void bmgx_count_launch(unsigned n = 1);
// stream-ordered workspace from the library's own memory pool (bmgx_util.cu)
cudaError_t bmgx_scratch_alloc(void **p, size_t bytes, cudaStream_t st);
cudaError_t bmgx_scratch_free(void *p, cudaStream_t st);
// batch context (bmgx_set_batch): number of grid functions the batch-aware entry points process per call
unsigned bmgx_batch();

#define BMGX_LAUNCH_CHECK(where)                                          \
    do {                                                                  \
        cudaError_t e__ = cudaGetLastError();                             \
        if (e__ != cudaSuccess) return bmgx_set_cuda_error(e__, where);   \
        bmgx_count_launch();                                              \
    } while (0)

// Programmatic dependent launch for the chains of small kernels on the coarse levels of a cycle: a kernel launched
// through launch_pdl may become resident while its predecessor on the stream is still running; it executes
// pdl_wait() before it touches memory (the predecessor has then completed and its writes are visible), so the
// launch latency of kernel N+1 is paid while kernel N runs.  Captured into CUDA graphs as programmatic edges.
// Every kernel launched this way calls pdl_wait() first (completion of N+1 then implies completion of N).
// Opt-in: BMGX_PDL=1 in the environment (default: plain stream order; pdl_wait() is then a no-op).
bool bmgx_pdl_enabled();
template <typename... P, typename... A>
static inline void launch_pdl(void (*kern)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A &&...args)
{
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = bmgx_pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kern, static_cast<A &&>(args)...);
}
__device__ __forceinline__ void pdl_wait()
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// periodic ghost refresh: Q(1,J)=Q(I1,J), Q(II,J)=Q(2,J) on rows jbeg, jbeg+jstep, ... (nrow of them);
// Q(I,1)=Q(I,J1), Q(I,JJ)=Q(I,2) on every column
int bmgx_wrap_x(double *q, unsigned ii, unsigned jbeg, unsigned jstep, unsigned nrow, cudaStream_t st);
int bmgx_wrap_y(double *q, unsigned ii, unsigned jj, cudaStream_t st);

// periodic branches of the interpolation setup and the coarsest-grid solve (bmgx2_periodic.cu)
int bmgx2_periodic_interp_oi(const double *so, double *ci, bmgx_len IIF, bmgx_len JJF, bmgx_len IIC, bmgx_len JJC, int ifd,
                             int jpn, bmgx_stream_t s);
int bmgx2_periodic_cg_lu(const double *so, bmgx_len ii, bmgx_len jj, int nstncl, double *abd, bmgx_len nabd1, bmgx_len nabd2,
                         int ibc, bmgx_stream_t s);
int bmgx2_periodic_solve_cg(double *q, const double *qf, bmgx_len ii, bmgx_len jj, const double *abd, double *bbd, bmgx_len nabd1,
                            int jpn, bmgx_stream_t s);

static inline bool bmgx_nonperiodic(int jpn) { return jpn == BMGX_BCs_definite || jpn == BMGX_BCs_indef_nonper; }
static inline unsigned cdiv(size_t a, size_t b) { return (unsigned)((a + b - 1) / b); }

// streaming (read-once) load: keep it out of L1
__device__ __forceinline__ double ldg_stream(const double *p) { return __ldcs(p); }
