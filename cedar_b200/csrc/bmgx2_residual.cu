// bmgx2_residual.cu -- 2D residual on 5/9-point stencils
// (reference: src/2d/ftn/BMG2_SymStd_residual.f90:87-119).
#include "bmgx_common.cuh"

template <bool NINE>
__global__ void __launch_bounds__(256)
k_residual(const double *__restrict__ so, const double *__restrict__ qf, const double *__restrict__ q,
           double *__restrict__ res, unsigned II, unsigned JJ)
{
    pdl_wait();
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x + 2;
    unsigned j = blockIdx.y * blockDim.y + threadIdx.y + 2;
    if (i > II - 1 || j > JJ - 1) return;
    size_t c = IX2(II, i, j);
    size_t pl = (size_t)II * JJ;
    qf += blockIdx.z * pl; q += blockIdx.z * pl; res += blockIdx.z * pl;   // batch entry
    const double *ko = so, *kw = so + (KW - 1) * pl, *ks = so + (KS - 1) * pl;
    double acc = qf[c] + kw[c] * q[c - 1] + kw[c + 1] * q[c + 1] + ks[c] * q[c - II] + ks[c + II] * q[c + II];
    if (NINE) {
        const double *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
        acc = acc + ksw[c] * q[c - II - 1] + knw[c + 1] * q[c - II + 1] + knw[c + II] * q[c + II - 1]
              + ksw[c + II + 1] * q[c + II + 1];
    }
    res[c] = acc - ko[c] * q[c];
}

extern "C" int bmgx2_residual(int k, const double *so, const double *qf, const double *q, double *res, bmgx_len ii,
                              bmgx_len jj, int kf, int ifd, int nstncl, int ibc, int irelax, int irelax_sym,
                              int updown, bmgx_stream_t s)
{
    (void)nstncl; (void)ibc; (void)irelax; (void)irelax_sym; (void)updown;
    if (ii < 3 || jj < 3) return BMGX_OK;
    dim3 b(128, 2), g(cdiv(ii - 2, 128), cdiv(jj - 2, 2), bmgx_batch());
    if (k < kf || ifd != 1)
        launch_pdl(k_residual<true>, g, b, 0, (cudaStream_t)s, so, qf, q, res, ii, jj);
    else
        launch_pdl(k_residual<false>, g, b, 0, (cudaStream_t)s, so, qf, q, res, ii, jj);
    BMGX_LAUNCH_CHECK("residual");
    return BMGX_OK;
}
