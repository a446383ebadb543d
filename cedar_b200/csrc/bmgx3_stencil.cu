// bmgx3_stencil.cu -- 3D SETUP_recip, multicolour point Gauss-Seidel, residual.
// Reference: src/3d/ftn/BMG3_SymStd_SETUP_recip.f90, BMG3_SymStd_relax_GS.f90:74-177
// (27-point: 8 colours, 7-point: red-black), BMG3_SymStd_residual.f90:67-120.
// Term order inside every update is the reference's and the file is compiled
// with -fmad=false, so results are bitwise those of the reference arithmetic.
#include "bmgx3_common.cuh"

namespace {

struct G3 {
    const double *so;
    int II, JJ, KK;
    size_t pl;
};

// qf + sum over the neighbours of so*q, in the order of relax_GS.f90:104-131 /
// residual.f90:74-100 (27-point) or 162-168 / 108-114 (7-point)
template <bool SEVEN>
__device__ __forceinline__ double nbr_sum(const G3 &g, const double *q, double qf, size_t c)
{
    const size_t II = g.II, PJ = (size_t)g.II * g.JJ, pl = g.pl;
    const double *so = g.so;
#define S(e, off) __ldg(so + (size_t)((e)-1) * pl + c + (off))
    if (SEVEN) {
        return qf + S(KPW, 0) * q[c - 1] + S(KPS, II) * q[c + II] + S(KPW, 1) * q[c + 1] + S(KPS, 0) * q[c - II]
               + S(KB, 0) * q[c - PJ] + S(KB, PJ) * q[c + PJ];
    }
    double a = qf + S(KPW, 0) * q[c - 1] + S(KPNW, II) * q[c - 1 + II] + S(KPS, II) * q[c + II]
               + S(KPSW, II + 1) * q[c + 1 + II] + S(KPW, 1) * q[c + 1] + S(KPNW, 1) * q[c + 1 - II]
               + S(KPS, 0) * q[c - II] + S(KPSW, 0) * q[c - 1 - II];
    a = a + S(KB, 0) * q[c - PJ] + S(KBW, 0) * q[c - 1 - PJ] + S(KBNW, II) * q[c - 1 + II - PJ]
        + S(KBN, II) * q[c + II - PJ] + S(KBNE, II + 1) * q[c + 1 + II - PJ] + S(KBE, 1) * q[c + 1 - PJ]
        + S(KBSE, 1) * q[c + 1 - II - PJ] + S(KBS, 0) * q[c - II - PJ] + S(KBSW, 0) * q[c - 1 - II - PJ];
    a = a + S(KB, PJ) * q[c + PJ] + S(KBE, PJ) * q[c - 1 + PJ] + S(KBSE, PJ + II) * q[c - 1 + II + PJ]
        + S(KBS, PJ + II) * q[c + II + PJ] + S(KBSW, PJ + II + 1) * q[c + 1 + II + PJ] + S(KBW, PJ + 1) * q[c + 1 + PJ]
        + S(KBNW, PJ + 1) * q[c + 1 - II + PJ] + S(KBN, PJ) * q[c - II + PJ] + S(KBNE, PJ) * q[c - 1 - II + PJ];
    return a;
#undef S
}

__global__ void __launch_bounds__(256)
k3_recip(const double *__restrict__ so, double *__restrict__ sor, int II, int JJ, int KK)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x + 2, j = blockIdx.y + 2, k = blockIdx.z + 2;
    if (i > II - 1) return;
    size_t c = ((size_t)(k - 1) * JJ + (j - 1)) * II + (i - 1);
    sor[(size_t)(MSOR - 1) * II * JJ * KK + c] = 1.0 / so[c];
}

template <bool SEVEN>
__global__ void __launch_bounds__(128)
k3_residual(G3 g, const double *__restrict__ qf, const double *__restrict__ q, double *__restrict__ res)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x + 2, j = blockIdx.y * blockDim.y + threadIdx.y + 2, k = blockIdx.z + 2;
    if (i > g.II - 1 || j > g.JJ - 1) return;
    size_t c = ((size_t)(k - 1) * g.JJ + (j - 1)) * g.II + (i - 1);
    res[c] = nbr_sum<SEVEN>(g, q, qf[c], c) - g.so[c] * q[c];
}

// one colour, in place.  27-point: points (i0+2a, j0+2b, k0+2c); 7-point: all (j,k),
// i from ibeg = mod(j+k+pts,2)+2 in steps of 2
template <bool SEVEN>
__global__ void __launch_bounds__(128)
k3_relax_colour(G3 g, const double *__restrict__ qf, double *q, const double *__restrict__ sor, int i0, int j0, int k0,
                int pts)
{
    int j, k, i;
    if (SEVEN) {
        j = blockIdx.y + 2; k = blockIdx.z + 2;
        i = (j + k + pts) % 2 + 2 + 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    } else {
        j = j0 + 2 * blockIdx.y; k = k0 + 2 * blockIdx.z;
        i = i0 + 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    }
    if (i > g.II - 1 || j > g.JJ - 1 || k > g.KK - 1) return;
    size_t c = ((size_t)(k - 1) * g.JJ + (j - 1)) * g.II + (i - 1);
    q[c] = nbr_sum<SEVEN>(g, q, qf[c], c) * sor[(size_t)(MSOR - 1) * g.pl + c];
}

// 27-point: the two colours of one (j,k) row parity -- they differ only in the parity of
// i and are consecutive in the sweep order -- in ONE launch.  A CTA owns whole rows, so the
// second colour sees the first colour's new values of its own row after a barrier; all other
// neighbours (rows j+-1, planes k+-1) belong to colours that this launch does not touch.
// Both colours of a row pull the same sectors of every coefficient array (the second pass
// from L1/L2), so each launch moves half the DRAM bytes of two single-colour launches.
__global__ void __launch_bounds__(1024)
k3_relax_xpair(G3 g, const double *__restrict__ qf, double *q, const double *__restrict__ sor, int ifirst, int j0,
               int k0)
{
    const int j = j0 + 2 * blockIdx.y, k = k0 + 2 * blockIdx.z;
    const size_t row = ((size_t)(k - 1) * g.JJ + (j - 1)) * g.II;
    const double *rs = sor + (size_t)(MSOR - 1) * g.pl;
    const int i1 = ifirst + 2 * threadIdx.x, i2 = (ifirst == 2 ? 3 : 2) + 2 * threadIdx.x;
    if (i1 <= g.II - 1) {
        const size_t c = row + (i1 - 1);
        q[c] = nbr_sum<false>(g, q, qf[c], c) * rs[c];
    }
    __syncthreads();
    if (i2 <= g.II - 1) {
        const size_t c = row + (i2 - 1);
        q[c] = nbr_sum<false>(g, q, qf[c], c) * rs[c];
    }
}

}  // namespace

extern "C" int bmgx3_SETUP_recip(const double *so, double *sor, bmgx_len nx, bmgx_len ny, bmgx_len nz, int nstncl,
                                 int nsorv, bmgx_stream_t s)
{
    (void)nstncl;
    if (nsorv < 2) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_recip: nsorv < 2");
    if (nx < 3 || ny < 3 || nz < 3) return BMGX_OK;
    dim3 g(cdiv(nx - 2, 256), ny - 2, nz - 2);
    k3_recip<<<g, 256, 0, (cudaStream_t)s>>>(so, sor, nx, ny, nz);
    BMGX_LAUNCH_CHECK("SETUP_recip3");
    return BMGX_OK;
}

extern "C" int bmgx3_residual(int kg, int nog, int ifd, const double *q, const double *qf, const double *so,
                              double *res, bmgx_len ii, bmgx_len jj, bmgx_len kk, int nstncl, bmgx_stream_t s)
{
    (void)nstncl;
    if (ii < 3 || jj < 3 || kk < 3) return BMGX_OK;
    G3 g{so, (int)ii, (int)jj, (int)kk, (size_t)ii * jj * kk};
    dim3 b(64, 2), grid(cdiv(ii - 2, 64), cdiv(jj - 2, 2), kk - 2);
    if (kg < nog || ifd != 1) k3_residual<false><<<grid, b, 0, (cudaStream_t)s>>>(g, qf, q, res);
    else k3_residual<true><<<grid, b, 0, (cudaStream_t)s>>>(g, qf, q, res);
    BMGX_LAUNCH_CHECK("residual3");
    return BMGX_OK;
}

extern "C" int bmgx3_relax_GS(int kg, const double *so, const double *qf, double *q, const double *sor, bmgx_len ii,
                              bmgx_len jj, bmgx_len kk, int ifd, int nstncl, int nsorv, int irelax_sym, int updown,
                              int jpn, bmgx_stream_t s)
{
    (void)kg; (void)nstncl; (void)nsorv;
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_GS3: periodic BCs not supported");
    if (ii < 3 || jj < 3 || kk < 3) return BMGX_OK;
    cudaStream_t st = (cudaStream_t)s;
    G3 g{so, (int)ii, (int)jj, (int)kk, (size_t)ii * jj * kk};
    const bool up = (updown == BMGX_UP) || (irelax_sym == BMGX_RELAX_NONSYM);
    const int I1 = ii - 1, J1 = jj - 1, K1 = kk - 1;
    if (ifd != 1) {
        // 8 colours; UP runs pts = 1..8, DOWN 8..1 (relax_GS.f90:82-96).  Colours 2n-1 and 2n share
        // (j0,k0) and are adjacent in both orders: one launch per pair when a row fits one CTA
        const int npair = (I1 - 2) / 2 + 1;   // points of the larger x colour per row
        if (npair <= 1024) {
            for (int n = 0; n < 4; n++) {
                const int pr = up ? n : 3 - n;                    // pair index: pts = 2 pr + 1, 2 pr + 2
                const int j0 = 2 + pr % 2, k0 = 2 + pr / 2;
                if (j0 > J1 || k0 > K1) continue;
                const int ifirst = up ? 2 : 3;                    // UP: pts odd (i0 = 2) first; DOWN: pts even (i0 = 3)
                dim3 grid(1, (J1 - j0) / 2 + 1, (K1 - k0) / 2 + 1);
                k3_relax_xpair<<<grid, ((npair + 31) / 32) * 32, 0, st>>>(g, qf, q, sor, ifirst, j0, k0);
                BMGX_LAUNCH_CHECK("relax_GS3/27 pair");
            }
            return BMGX_OK;
        }
        for (int n = 0; n < 8; n++) {
            int pts = up ? 1 + n : 8 - n;
            int i0 = 2 + (pts - 1) % 2, j0 = 2 + ((pts - 1) / 2) % 2, k0 = 2 + ((pts - 1) / 4) % 2;
            if (i0 > I1 || j0 > J1 || k0 > K1) continue;
            dim3 grid(cdiv((I1 - i0) / 2 + 1, 128), (J1 - j0) / 2 + 1, (K1 - k0) / 2 + 1);
            k3_relax_colour<false><<<grid, 128, 0, st>>>(g, qf, q, sor, i0, j0, k0, pts);
            BMGX_LAUNCH_CHECK("relax_GS3/27");
        }
    } else {
        // red-black; UP runs pts = 0,1, DOWN 1,0 (relax_GS.f90:144-154)
        for (int n = 0; n < 2; n++) {
            int pts = up ? n : 1 - n;
            dim3 grid(cdiv((I1 - 2) / 2 + 1, 128), J1 - 1, K1 - 1);
            k3_relax_colour<true><<<grid, 128, 0, st>>>(g, qf, q, sor, 0, 0, 0, pts);
            BMGX_LAUNCH_CHECK("relax_GS3/7");
        }
    }
    return BMGX_OK;
}
