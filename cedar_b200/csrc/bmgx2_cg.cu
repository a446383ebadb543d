// bmgx2_cg.cu -- coarsest-grid direct solve: banded Cholesky on the device.
// Reference: src/2d/ftn/BMG2_SymStd_SETUP_cg_LU.f90:92-118 (9-pt), 236-256 (5-pt)
//            + LAPACK DPBTRF('U') (unblocked DPBTF2 path, KD <= 32 or not),
//            src/2d/ftn/BMG2_SymStd_SOLVE_cg.f90:93-122 + LAPACK DPBTRS('U').
// The factorisation runs in one CTA (the rank-1 trailing update is spread over
// the threads; each entry sees the same operations in the same order as
// DPBTF2, so the factor is bitwise the oracle's).  The solve keeps DTBSV's
// sequential accumulation order in one thread: the coarsest grids of the hot
// path have 9..49 unknowns.
#include "bmgx_common.cuh"
#include <cstdio>

#define ABD(i, j) abd[(size_t)((j)-1) * nabd1 + ((i)-1)]

// INFO of the last banded Cholesky factorisation enqueued by this library (DPBTRF's INFO: 0 = ok, j > 0 = the
// leading minor of order j is not positive definite; the reference logs "Coarse grid Cholesky decomposition
// failed", BMG2_SymStd_SETUP_cg_LU.f90:124-131).  Read back with bmgx_cg_LU_info().
__device__ int g_cg_info = 0;

__global__ void __launch_bounds__(256)
k_cg_assemble(const double *__restrict__ so, unsigned II, unsigned JJ, int nstncl, double *__restrict__ abd,
              unsigned nabd1, int ibc)
{
    unsigned I1 = II - 1, J1 = JJ - 1, I2 = I1 - 1;
    unsigned N = I2 * (J1 - 1);
    for (unsigned kk = blockIdx.x * blockDim.x + threadIdx.x + 1; kk <= N; kk += gridDim.x * blockDim.x) {
        unsigned i = (kk - 1) % I2 + 2, j = (kk - 1) / I2 + 2;
        double d = so[IX3(II, JJ, i, j, KO)];
        if (ibc < 0 && kk == N) d = d + so[IX3(II, JJ, I1, J1, KO)];
        ABD(II, kk) = d;
        ABD(I1, kk) = -so[IX3(II, JJ, i, j, KW)];
        ABD(3, kk) = nstncl == 5 ? -so[IX3(II, JJ, i + 1, j, KNW)] : 0.0;
        ABD(2, kk) = -so[IX3(II, JJ, i, j, KS)];
        ABD(1, kk) = nstncl == 5 ? -so[IX3(II, JJ, i, j, KSW)] : 0.0;
        // rows 4..I1-1 of the band are zero fill (ABD arrives zeroed from the caller)
    }
}

// DPBTF2('U'): AB(kd+1,j) diagonal; one CTA.
__global__ void __launch_bounds__(256) k_dpbtf2_upper(int n, int kd, double *abd, int nabd1, int *info)
{
    int kld = nabd1 - 1 > 1 ? nabd1 - 1 : 1;
    __shared__ double s_r;
    __shared__ int s_bad;
    if (threadIdx.x == 0) s_bad = 0;
    __syncthreads();
    for (int j = 1; j <= n; j++) {
        if (threadIdx.x == 0) {
            double ajj = ABD(kd + 1, j);
            if (ajj <= 0.0) {
                s_bad = j;
            } else {
                ajj = sqrt(ajj);
                ABD(kd + 1, j) = ajj;
                s_r = 1.0 / ajj;
            }
        }
        __syncthreads();
        if (s_bad) break;
        int kn = kd < n - j ? kd : n - j;
        if (kn > 0) {
            double *x = &ABD(kd, j + 1);
            double r = s_r;
            for (int p = threadIdx.x; p < kn; p += blockDim.x) x[(size_t)p * kld] = r * x[(size_t)p * kld];
            __syncthreads();
            double *a = &ABD(kd + 1, j + 1);
            // DSYR upper: a(p,q) += x(p) * (-x(q)), p <= q
            int tot = kn * kn;
            for (int t = threadIdx.x; t < tot; t += blockDim.x) {
                int q = t / kn, p = t % kn;
                if (p <= q) {
                    double xq = x[(size_t)q * kld];
                    if (xq != 0.0) {
                        double temp = -1.0 * xq;
                        a[(size_t)q * kld + p] = a[(size_t)q * kld + p] + x[(size_t)p * kld] * temp;
                    }
                }
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) { g_cg_info = s_bad; if (info) *info = s_bad; }
}

int *bmgx_cg_info_ptr()
{
    int *p = nullptr;
    cudaGetSymbolAddress((void **)&p, g_cg_info);
    return p;
}

extern "C" int bmgx_cg_LU_info(int *info, bmgx_stream_t s)
{
    cudaError_t e = cudaStreamSynchronize((cudaStream_t)s);
    if (e == cudaSuccess) e = cudaMemcpyFromSymbol(info, g_cg_info, sizeof(int));
    if (e != cudaSuccess) return bmgx_set_cuda_error(e, "bmgx_cg_LU_info");
    if (*info != 0) {
        char msg[128];
        snprintf(msg, sizeof(msg), "Coarse grid Cholesky decomposition failed at column %d (operator not positive definite)", *info);
        return bmgx_set_error(BMGX_ERR_ARG, msg);
    }
    return BMGX_OK;
}

extern "C" int bmgx2_SETUP_cg_LU(const double *so, bmgx_len ii, bmgx_len jj, int nstncl, double *abd,
                                 bmgx_len nabd1, bmgx_len nabd2, int ibc, bmgx_stream_t s)
{
    if (!bmgx_nonperiodic(ibc)) return bmgx2_periodic_cg_lu(so, ii, jj, nstncl, abd, nabd1, nabd2, ibc, s);
    if (ii < 3 || jj < 3) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_cg_LU: empty grid");
    unsigned n = (ii - 2) * (jj - 2);
    if (nabd1 < ii || nabd2 < n) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_cg_LU: ABD too small");
    cudaStream_t st = (cudaStream_t)s;
    cudaError_t e = cudaMemsetAsync(abd, 0, sizeof(double) * (size_t)nabd1 * nabd2, st);
    if (e != cudaSuccess) return bmgx_set_cuda_error(e, "SETUP_cg_LU/memset");
    bmgx_count_launch();
    k_cg_assemble<<<cdiv(n, 256) > 1024 ? 1024 : cdiv(n, 256), 256, 0, st>>>(so, ii, jj, nstncl, abd, nabd1, ibc);
    BMGX_LAUNCH_CHECK("SETUP_cg_LU/assemble");
    k_dpbtf2_upper<<<1, 256, 0, st>>>((int)n, (int)(ii - 1), abd, (int)nabd1, nullptr);
    BMGX_LAUNCH_CHECK("SETUP_cg_LU/dpbtf2");
    return BMGX_OK;
}

// copy rhs -> DTBSV(U,T,N) -> DTBSV(U,N,N) -> copy back; single thread keeps
// the reference accumulation order.
__global__ void k_solve_cg(double *q, const double *__restrict__ qf, unsigned II, unsigned JJ,
                           const double *__restrict__ abd, double *bbd, unsigned nabd1)
{
    pdl_wait();
    if (threadIdx.x != 0) return;
    int I1 = II - 1, J1 = JJ - 1;
    q += (size_t)blockIdx.x * II * JJ; qf += (size_t)blockIdx.x * II * JJ;   // batch entry
    bbd += (size_t)blockIdx.x * (I1 - 1) * (J1 - 1);
    int kd = I1, kplus1 = kd + 1;
    int n = 0;
    for (int j = 2; j <= J1; j++)
        for (int i = 2; i <= I1; i++) bbd[n++] = qf[IX2(II, i, j)];
    for (int j = 1; j <= n; j++) {
        double temp = bbd[j - 1];
        int l = kplus1 - j;
        int i0 = j - kd > 1 ? j - kd : 1;
        for (int i = i0; i <= j - 1; i++) temp = temp - ABD(l + i, j) * bbd[i - 1];
        temp = temp / ABD(kplus1, j);
        bbd[j - 1] = temp;
    }
    for (int j = n; j >= 1; j--) {
        if (bbd[j - 1] != 0.0) {
            int l = kplus1 - j;
            bbd[j - 1] = bbd[j - 1] / ABD(kplus1, j);
            double temp = bbd[j - 1];
            int i0 = j - kd > 1 ? j - kd : 1;
            for (int i = j - 1; i >= i0; i--) bbd[i - 1] = bbd[i - 1] - temp * ABD(l + i, j);
        }
    }
    n = 0;
    for (int j = 2; j <= J1; j++)
        for (int i = 2; i <= I1; i++) q[IX2(II, i, j)] = bbd[n++];
}

extern "C" int bmgx2_SOLVE_cg(double *q, const double *qf, bmgx_len ii, bmgx_len jj, const double *abd, double *bbd,
                              bmgx_len nabd1, bmgx_len nabd2, int jpn, bmgx_stream_t s)
{
    (void)nabd2;
    {
        const int ipn = jpn < 0 ? -jpn : jpn;
        if (ipn >= 1 && ipn <= 3) return bmgx2_periodic_solve_cg(q, qf, ii, jj, abd, bbd, nabd1, jpn, s);
    }
    if (jpn != BMGX_BCs_definite) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "SOLVE_cg: indefinite non-periodic not supported");
    launch_pdl(k_solve_cg, dim3(bmgx_batch()), dim3(32), 0, (cudaStream_t)s, q, qf, ii, jj, abd, bbd, nabd1);
    BMGX_LAUNCH_CHECK("SOLVE_cg");
    return BMGX_OK;
}

// ---- 3D ------------------------------------------------------------------------
// BMG3_SymStd_SETUP_cg_LU.f90:92-193: band rows by neighbour offset, bandwidth
// ibw = nx*(ny+1)+1, then the same DPBTF2; BMG3_SymStd_SOLVE_cg.f90:100-135.
enum { KP3 = 1, KPW3, KPS3, KB3, KPSW3, KPNW3, KBW3, KBNW3, KBN3, KBNE3, KBE3, KBSE3, KBS3, KBSW3 };

__global__ void __launch_bounds__(256)
k3_cg_assemble(const double *__restrict__ so, int II, int JJ, int KK, int nstncl, double *__restrict__ abd,
               unsigned nabd1, int ibc)
{
    const int i1 = II - 1, j1 = JJ - 1, k1 = KK - 1, i2 = i1 - 1;
    const int ibw = i2 * j1 + 1;
    const int N = i2 * (j1 - 1) * (k1 - 1);
    const size_t pl = (size_t)II * JJ * KK;
    const bool full = nstncl == 14;
#define SO3(i, j, k, e) so[(size_t)((e)-1) * pl + IX3(II, JJ, i, j, k)]
    for (int kl = blockIdx.x * blockDim.x + threadIdx.x + 1; kl <= N; kl += gridDim.x * blockDim.x) {
        const int i = (kl - 1) % i2 + 2, j = ((kl - 1) / i2) % (j1 - 1) + 2, k = (kl - 1) / (i2 * (j1 - 1)) + 2;
        double d = SO3(i, j, k, KP3);
        if (ibc < 0 && kl == N) d = d + SO3(i1, j1, k1, KP3);
        ABD(ibw + 1, kl) = d;
        ABD(ibw, kl) = -SO3(i, j, k, KPW3);
        ABD(ibw - i1 + 3, kl) = full ? -SO3(i + 1, j, k, KPNW3) : 0.0;
        ABD(ibw - i1 + 2, kl) = -SO3(i, j, k, KPS3);
        ABD(ibw - i1 + 1, kl) = full ? -SO3(i, j, k, KPSW3) : 0.0;
        ABD(ibw - (j1 - 2) * i2 + 2, kl) = full ? -SO3(i + 1, j + 1, k, KBNE3) : 0.0;
        ABD(ibw - (j1 - 2) * i2 + 1, kl) = full ? -SO3(i, j + 1, k, KBN3) : 0.0;
        ABD(ibw - (j1 - 2) * i2, kl) = full ? -SO3(i, j + 1, k, KBNW3) : 0.0;
        ABD(ibw - (j1 - 1) * i2 + 2, kl) = full ? -SO3(i + 1, j, k, KBE3) : 0.0;
        ABD(ibw - (j1 - 1) * i2 + 1, kl) = -SO3(i, j, k, KB3);
        ABD(ibw - (j1 - 1) * i2, kl) = full ? -SO3(i, j, k, KBW3) : 0.0;
        ABD(3, kl) = full ? -SO3(i + 1, j, k, KBSE3) : 0.0;
        ABD(2, kl) = full ? -SO3(i, j, k, KBS3) : 0.0;
        ABD(1, kl) = full ? -SO3(i, j, k, KBSW3) : 0.0;
    }
#undef SO3
}

extern "C" int bmgx3_SETUP_cg_LU(const double *so, bmgx_len ii, bmgx_len jj, bmgx_len kk, int nstncl, double *abd,
                                 bmgx_len nabd1, bmgx_len nabd2, int ibc, bmgx_stream_t s)
{
    if (!bmgx_nonperiodic(ibc)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "SETUP_cg_LU3: periodic BCs not supported");
    if (ii < 3 || jj < 3 || kk < 3) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_cg_LU3: empty grid");
    if (nstncl != 14 && nstncl != 4) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_cg_LU3: NStncl must be 4 or 14");
    const unsigned n = (ii - 2) * (jj - 2) * (kk - 2), ibw = (ii - 2) * (jj - 1) + 1;
    if (nabd1 < ibw + 1 || nabd2 < n) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_cg_LU3: ABD too small");
    cudaStream_t st = (cudaStream_t)s;
    cudaError_t e = cudaMemsetAsync(abd, 0, sizeof(double) * (size_t)nabd1 * nabd2, st);
    if (e != cudaSuccess) return bmgx_set_cuda_error(e, "SETUP_cg_LU3/memset");
    bmgx_count_launch();
    k3_cg_assemble<<<cdiv(n, 256) > 1024 ? 1024 : cdiv(n, 256), 256, 0, st>>>(so, ii, jj, kk, nstncl, abd, nabd1, ibc);
    BMGX_LAUNCH_CHECK("SETUP_cg_LU3/assemble");
    k_dpbtf2_upper<<<1, 256, 0, st>>>((int)n, (int)ibw, abd, (int)nabd1, nullptr);
    BMGX_LAUNCH_CHECK("SETUP_cg_LU3/dpbtf2");
    return BMGX_OK;
}

__global__ void k3_solve_cg(double *q, const double *__restrict__ qf, int II, int JJ, int KK,
                            const double *__restrict__ abd, double *bbd, unsigned nabd1)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int kd = (II - 2) * (JJ - 1) + 1, kplus1 = kd + 1;
    int n = 0;
    for (int k = 2; k <= KK - 1; k++)
        for (int j = 2; j <= JJ - 1; j++)
            for (int i = 2; i <= II - 1; i++) bbd[n++] = qf[IX3(II, JJ, i, j, k)];
    for (int j = 1; j <= n; j++) {
        double temp = bbd[j - 1];
        const int l = kplus1 - j;
        const int i0 = j - kd > 1 ? j - kd : 1;
        for (int i = i0; i <= j - 1; i++) temp = temp - ABD(l + i, j) * bbd[i - 1];
        temp = temp / ABD(kplus1, j);
        bbd[j - 1] = temp;
    }
    for (int j = n; j >= 1; j--) {
        if (bbd[j - 1] != 0.0) {
            const int l = kplus1 - j;
            bbd[j - 1] = bbd[j - 1] / ABD(kplus1, j);
            const double temp = bbd[j - 1];
            const int i0 = j - kd > 1 ? j - kd : 1;
            for (int i = j - 1; i >= i0; i--) bbd[i - 1] = bbd[i - 1] - temp * ABD(l + i, j);
        }
    }
    n = 0;
    for (int k = 2; k <= KK - 1; k++)
        for (int j = 2; j <= JJ - 1; j++)
            for (int i = 2; i <= II - 1; i++) q[IX3(II, JJ, i, j, k)] = bbd[n++];
}

extern "C" int bmgx3_SOLVE_cg(double *q, const double *qf, bmgx_len ii, bmgx_len jj, bmgx_len kk, const double *abd,
                              double *bbd, bmgx_len nabd1, bmgx_len nabd2, int jpn, bmgx_stream_t s)
{
    (void)nabd2;
    if (jpn != BMGX_BCs_definite) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "SOLVE_cg3: only definite, non-periodic");
    k3_solve_cg<<<1, 32, 0, (cudaStream_t)s>>>(q, qf, ii, jj, kk, abd, bbd, nabd1);
    BMGX_LAUNCH_CHECK("SOLVE_cg3");
    return BMGX_OK;
}
