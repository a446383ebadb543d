// bmgx_matvec.cu -- operator application y = A x on the symmetric BoxMG stencils, 2D and 3D
// (reference: src/2d/ftn/mpi/BMG2_SymStd_UTILS_matvec.f90:88-123, src/3d/ftn/mpi/BMG3_SymStd_UTILS_matvec.f90:78-129;
// what the C interface's operator "apply" calls, src/2d/interface/c/operator.cc:88).  Streaming kernels:
// algorithmic traffic = the residual's minus the right-hand side (56 / 40 B per point in 2D, 128 / 48 in 3D).
#include "bmgx_common.cuh"
#include "bmgx3_common.cuh"

template <bool NINE>
__global__ void __launch_bounds__(256)
k_matvec2(const double *__restrict__ so, double *__restrict__ qf, const double *__restrict__ q, unsigned II, unsigned JJ)
{
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x + 2;
    unsigned j = blockIdx.y * blockDim.y + threadIdx.y + 2;
    if (i > II - 1 || j > JJ - 1) return;
    size_t c = IX2(II, i, j);
    size_t pl = (size_t)II * JJ;
    const double *ko = so, *kw = so + (KW - 1) * pl, *ks = so + (KS - 1) * pl;
    // term order of the reference: centre, W, E, S, N (, SW, SE, NW, NE)
    double y = ko[c] * q[c] - kw[c] * q[c - 1] - kw[c + 1] * q[c + 1] - ks[c] * q[c - II] - ks[c + II] * q[c + II];
    if (NINE) {
        const double *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
        y = y - ksw[c] * q[c - II - 1] - knw[c + 1] * q[c - II + 1] - knw[c + II] * q[c + II - 1]
            - ksw[c + II + 1] * q[c + II + 1];
    }
    qf[c] = y;
}

extern "C" int bmgx2_matvec(int k, const double *so, double *qf, const double *q, bmgx_len ii, bmgx_len jj, int kf,
                            int ifd, int nstncl, bmgx_stream_t s)
{
    (void)nstncl;
    if (ii < 3 || jj < 3) return BMGX_OK;
    dim3 b(128, 2), g(cdiv(ii - 2, 128), cdiv(jj - 2, 2));
    if (k < kf || ifd != 1) k_matvec2<true><<<g, b, 0, (cudaStream_t)s>>>(so, qf, q, ii, jj);
    else k_matvec2<false><<<g, b, 0, (cudaStream_t)s>>>(so, qf, q, ii, jj);
    BMGX_LAUNCH_CHECK("matvec");
    return BMGX_OK;
}

namespace {
// {stencil slot, offset of the stencil entry (x,y,z), offset of the x entry (x,y,z)} in the reference's term order
struct Term { int e, sx, sy, sz, dx, dy, dz; };
__device__ constexpr Term MV27[26] = {
    {KPW, 0, 0, 0, -1, 0, 0},   {KPNW, 0, 1, 0, -1, 1, 0},  {KPS, 0, 1, 0, 0, 1, 0},    {KPSW, 1, 1, 0, 1, 1, 0},
    {KPW, 1, 0, 0, 1, 0, 0},    {KPNW, 1, 0, 0, 1, -1, 0},  {KPS, 0, 0, 0, 0, -1, 0},   {KPSW, 0, 0, 0, -1, -1, 0},
    {KB, 0, 0, 0, 0, 0, -1},    {KBW, 0, 0, 0, -1, 0, -1},  {KBNW, 0, 1, 0, -1, 1, -1}, {KBN, 0, 1, 0, 0, 1, -1},
    {KBNE, 1, 1, 0, 1, 1, -1},  {KBE, 1, 0, 0, 1, 0, -1},   {KBSE, 1, 0, 0, 1, -1, -1}, {KBS, 0, 0, 0, 0, -1, -1},
    {KBSW, 0, 0, 0, -1, -1, -1}, {KB, 0, 0, 1, 0, 0, 1},    {KBE, 0, 0, 1, -1, 0, 1},   {KBSE, 0, 1, 1, -1, 1, 1},
    {KBS, 0, 1, 1, 0, 1, 1},    {KBSW, 1, 1, 1, 1, 1, 1},   {KBW, 1, 0, 1, 1, 0, 1},    {KBNW, 1, 0, 1, 1, -1, 1},
    {KBN, 0, 0, 1, 0, -1, 1},   {KBNE, 0, 0, 1, -1, -1, 1}};
__device__ constexpr Term MV7[6] = {
    {KPW, 0, 0, 0, -1, 0, 0}, {KPS, 0, 1, 0, 0, 1, 0}, {KPW, 1, 0, 0, 1, 0, 0},
    {KPS, 0, 0, 0, 0, -1, 0}, {KB, 0, 0, 0, 0, 0, -1}, {KB, 0, 0, 1, 0, 0, 1}};
}  // namespace

template <bool SEVEN>
__global__ void __launch_bounds__(128)
k_matvec3(const double *__restrict__ so, double *__restrict__ qf, const double *__restrict__ q, unsigned II, unsigned JJ,
          unsigned KK)
{
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x + 2;
    const unsigned j = blockIdx.y + 2, k = blockIdx.z + 2;
    if (i > II - 1) return;
    const size_t pl = (size_t)II * JJ * KK;
    const long long sj = II, sk = (long long)II * JJ;
    const long long c = (long long)IX3(II, JJ, i, j, k);
    double y = so[c] * q[c];
    constexpr int NT = SEVEN ? 6 : 26;
#pragma unroll
    for (int t = 0; t < NT; t++) {
        const Term T = SEVEN ? MV7[t] : MV27[t];
        y = y - so[(size_t)(T.e - 1) * pl + c + T.sx + T.sy * sj + T.sz * sk] * q[c + T.dx + T.dy * sj + T.dz * sk];
    }
    qf[c] = y;
}

extern "C" int bmgx3_matvec(int kg, const double *so, double *qf, const double *q, bmgx_len ii, bmgx_len jj,
                            bmgx_len kk, int nog, int ifd, int nstncl, bmgx_stream_t s)
{
    (void)nstncl;
    if (ii < 3 || jj < 3 || kk < 3) return BMGX_OK;
    if (jj - 2 > 65535 || kk - 2 > 65535) return bmgx_set_error(BMGX_ERR_ARG, "matvec: grid too large");
    dim3 g(cdiv(ii - 2, 128), jj - 2, kk - 2);
    if (kg < nog || ifd != 1) k_matvec3<false><<<g, 128, 0, (cudaStream_t)s>>>(so, qf, q, ii, jj, kk);
    else k_matvec3<true><<<g, 128, 0, (cudaStream_t)s>>>(so, qf, q, ii, jj, kk);
    BMGX_LAUNCH_CHECK("matvec3");
    return BMGX_OK;
}
