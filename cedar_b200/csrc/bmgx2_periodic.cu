// bmgx2_periodic.cu -- periodic boundary conditions: operator-induced interpolation setup and the coarsest-grid
// direct solve (the periodic branches of relaxation, restriction, interpolate-and-add and the Galerkin operator live
// with their non-periodic twins).
// Reference: src/2d/ftn/BMG2_SymStd_SETUP_interp_OI.f90:258-618,
//            src/2d/ftn/BMG2_SymStd_SETUP_cg_LU.f90:121-217 (9-pt), 258-331 (5-pt) + LAPACK DPOTRF('U'),
//            src/2d/ftn/BMG2_SymStd_SOLVE_cg.f90:104-166 + LAPACK DPOTRS('U').
// Boundary codes (BMG_parameters_f90.h:345-352): |code| = 1 periodic in y, 2 in x, 3 in both; negative = indefinite.
//
// Interpolation: the weights are the non-periodic formulas; the periodic branch changes which fine line feeds which
// coarse line.  The reference walks the coarse indices with a running fine index I = MAX(MOD(I+2,IIFC), MIN(I+2,3))
// from IBEG; in closed form the n-th index is t = IBEG + 2n while t < IIFC and MAX(MOD(t,IIFC),3) for the one step
// that wraps (checked against the recurrence for every extent up to 400 and against the oracle).  Same expression
// order as the oracle, -fmad=false => bitwise.
// Coarse solve: the wrap-around couplings leave the band, so the coarsest operator is a dense upper triangle; it has a
// few dozen unknowns, and assembly, factorisation (DPOTF2 order) and solve run in one thread each, literally in the
// reference's statement order.
#include "bmgx_common.cuh"
#include <cfloat>

namespace {

#define SO(i, j, kk) so[IX3(IIF, JJF, i, j, kk)]
#define CI(i, j, kk) ci[IX3(IIC, JJC, i, j, kk)]

struct OiP {
    int IIF, JJF, IIC, JJC;
    int IBEG, IBEGC, IENDC, IBEG_x, IEND_x, IIFC;
    int JBEG, JBEGC, JENDC, JBEG_y, JEND_y, JJFC;
    int perx, pery;
};

__device__ __forceinline__ int oi_map(int beg, int begc, int c, int nfc)
{
    const int t = beg + 2 * (c - begc + 1);
    return t < nfc ? t : max(t % nfc, 3);
}
__device__ __forceinline__ double oi_fix2(double o, double sum, double ep, double base)
{
    const double t = o - (1.0 + ep) * sum;
    return base + (o - sum) * fmax(t, 0.0) / (fabs(t) + DBL_EPSILON);
}

template <bool NINE>
__device__ __forceinline__ void edge_x(const double *__restrict__ so, int IIF, int JJF, int i, int j, double &lr, double &ll)
{
    double a, b;
    if (NINE) {
        a = SO(i, j, KW) + SO(i, j, KNW) + SO(i, j + 1, KSW);
        b = SO(i - 1, j, KW) + SO(i - 1, j, KSW) + SO(i - 1, j + 1, KNW);
    } else {
        a = SO(i, j, KW);
        b = SO(i - 1, j, KW);
    }
    const double o = SO(i - 1, j, KO);
    const double ep = fmin(fabs(a / o), fabs(b / o));
    double sum = a + b + SO(i - 1, j, KS) + SO(i - 1, j + 1, KS);
    sum = oi_fix2(o, sum, ep, a + b);
    sum = 1.0 / sum;
    lr = a * sum;
    ll = b * sum;
}
template <bool NINE>
__device__ __forceinline__ void edge_y(const double *__restrict__ so, int IIF, int JJF, int i, int j, double &la, double &lb)
{
    double a, b;
    if (NINE) {
        a = SO(i, j, KS) + SO(i, j, KNW) + SO(i + 1, j, KSW);
        b = SO(i, j - 1, KS) + SO(i, j - 1, KSW) + SO(i + 1, j - 1, KNW);
    } else {
        a = SO(i, j, KS);
        b = SO(i, j - 1, KS);
    }
    const double o = SO(i, j - 1, KO);
    double sum = a + b + SO(i, j - 1, KW) + SO(i + 1, j - 1, KW);
    const double ep = fmin(fabs(a / o), fabs(b / o));
    sum = oi_fix2(o, sum, ep, a + b);
    sum = 1.0 / sum;
    la = a * sum;
    lb = b * sum;
}

// x-edge weights LR, LL: coarse columns IBEGC..IENDC; coarse rows 2..JJC-1, and rows JJC / 1 from the wrapped fine rows
template <bool NINE>
__global__ void __launch_bounds__(256) kp_edges_x(const double *__restrict__ so, double *ci, OiP p)
{
    const int ic = p.IBEGC + blockIdx.x * blockDim.x + threadIdx.x, jc = 1 + blockIdx.y * blockDim.y + threadIdx.y;
    if (ic > p.IENDC || jc > p.JJC) return;
    int j;
    if (jc >= 2 && jc <= p.JJC - 1) j = 2 * (jc - 1);
    else if (p.pery && jc == p.JJC) j = p.JBEG_y;
    else if (p.pery && jc == 1) j = p.JEND_y;
    else return;
    const int i = oi_map(p.IBEG, p.IBEGC, ic, p.IIFC);
    const int IIC = p.IIC, JJC = p.JJC;
    edge_x<NINE>(so, p.IIF, p.JJF, i, j, CI(ic, jc, LR), CI(ic, jc, LL));
}
template <bool NINE>
__global__ void __launch_bounds__(256) kp_edges_y(const double *__restrict__ so, double *ci, OiP p)
{
    const int ic = 1 + blockIdx.x * blockDim.x + threadIdx.x, jc = p.JBEGC + blockIdx.y * blockDim.y + threadIdx.y;
    if (ic > p.IIC || jc > p.JENDC) return;
    int i;
    if (ic >= 2 && ic <= p.IIC - 1) i = 2 * (ic - 1);
    else if (p.perx && ic == p.IIC) i = p.IBEG_x;
    else if (p.perx && ic == 1) i = p.IEND_x;
    else return;
    const int j = oi_map(p.JBEG, p.JBEGC, jc, p.JJFC);
    const int IIC = p.IIC, JJC = p.JJC;
    edge_y<NINE>(so, p.IIF, p.JJF, i, j, CI(ic, jc, LA), CI(ic, jc, LB));
}
template <bool NINE>
__global__ void __launch_bounds__(256) kp_centres(const double *__restrict__ so, double *ci, OiP p)
{
    const int ic = p.IBEGC + blockIdx.x * blockDim.x + threadIdx.x, jc = p.JBEGC + blockIdx.y * blockDim.y + threadIdx.y;
    if (ic > p.IENDC || jc > p.JENDC) return;
    const int i = oi_map(p.IBEG, p.IBEGC, ic, p.IIFC), j = oi_map(p.JBEG, p.JBEGC, jc, p.JJFC);
    const int IIF = p.IIF, JJF = p.JJF, IIC = p.IIC, JJC = p.JJC;
    const double o = SO(i - 1, j - 1, KO);
    double sum, ep;
    if (NINE) {
        sum = SO(i - 1, j - 1, KW) + SO(i - 1, j, KNW) + SO(i - 1, j, KS) + SO(i, j, KSW) + SO(i, j - 1, KW)
              + SO(i, j - 1, KNW) + SO(i - 1, j - 1, KS) + SO(i - 1, j - 1, KSW);
        ep = fmin(fmin(fabs((SO(i - 1, j - 1, KSW) + SO(i - 1, j - 1, KW) + SO(i - 1, j, KNW)) / o),
                       fabs((SO(i - 1, j, KNW) + SO(i - 1, j, KS) + SO(i, j, KSW)) / o)),
                  fmin(fabs((SO(i, j, KSW) + SO(i, j - 1, KW) + SO(i, j - 1, KNW)) / o),
                       fabs((SO(i, j - 1, KNW) + SO(i - 1, j - 1, KS) + SO(i - 1, j - 1, KSW)) / o)));
    } else {
        sum = SO(i - 1, j - 1, KW) + SO(i - 1, j, KS) + SO(i, j - 1, KW) + SO(i - 1, j - 1, KS);
        ep = fmin(fmin(fabs(SO(i - 1, j - 1, KW) / o), fabs(SO(i - 1, j, KS) / o)),
                  fmin(fabs(SO(i, j - 1, KW) / o), fabs(SO(i - 1, j - 1, KS) / o)));
    }
    sum = oi_fix2(o, sum, ep, sum);
    const double s = 1.0 / sum;
    if (NINE) {
        CI(ic, jc, LSW) = (SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LL) + SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LB)
                           + SO(i - 1, j - 1, KSW)) * s;
        CI(ic, jc, LSE) = (SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR) + SO(i, j - 1, KW) * CI(ic, jc, LB)
                           + SO(i, j - 1, KNW)) * s;
        CI(ic, jc, LNW) = (SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LA) + SO(i - 1, j, KS) * CI(ic, jc, LL)
                           + SO(i - 1, j, KNW)) * s;
        CI(ic, jc, LNE) = (SO(i - 1, j, KS) * CI(ic, jc, LR) + SO(i, j - 1, KW) * CI(ic, jc, LA)
                           + SO(i, j, KSW)) * s;
    } else {
        CI(ic, jc, LSW) = (SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LL) + SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LB)) * s;
        CI(ic, jc, LSE) = (SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR) + SO(i, j - 1, KW) * CI(ic, jc, LB)) * s;
        CI(ic, jc, LNW) = (SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LA) + SO(i - 1, j, KS) * CI(ic, jc, LL)) * s;
        CI(ic, jc, LNE) = (SO(i - 1, j, KS) * CI(ic, jc, LR) + SO(i, j - 1, KW) * CI(ic, jc, LA)) * s;
    }
}
#undef SO
#undef CI

// ---- coarsest grid: dense assembly + DPOTF2('U') (one thread; a few dozen unknowns) ----------------------------
#define SO(i, j, kq) so[IX3(II, JJ, i, j, kq)]
#define ABD(i, j) abd[(size_t)((j)-1) * nabd1 + ((i)-1)]
#define NSO(i, j, kq) (nine ? -SO(i, j, kq) : 0.0)
__global__ void kp_cg_lu(const double *__restrict__ so, int II, int JJ, int nstncl, double *abd, int nabd1, int ibc, int *info)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int ipn = ibc < 0 ? -ibc : ibc;
    const bool perx = ipn == 2 || ipn == 3, pery = ipn == 1 || ipn == 3, perxy = ipn == 3, nine = nstncl == 5;
    const int I1 = II - 1, J1 = JJ - 1, I2 = I1 - 1, N = I2 * (J1 - 1);
    if (ibc == -3 && II == 4 && JJ == 4 && nine) {
        ABD(1, 1) = SO(2, 2, KO);
        ABD(2, 2) = SO(3, 2, KO);
        ABD(1, 2) = -SO(2, 2, KW) - SO(3, 2, KW);
        ABD(3, 3) = SO(2, 3, KO);
        ABD(1, 3) = -SO(2, 2, KS) - SO(2, 3, KS);
        ABD(2, 3) = -SO(3, 3, KNW) - SO(4, 2, KNW) - SO(4, 3, KSW) - SO(3, 2, KSW);
        ABD(4, 4) = SO(3, 3, KO);
        ABD(1, 4) = -SO(3, 3, KSW) - SO(2, 2, KSW) - SO(2, 3, KNW) - SO(3, 2, KNW);
        ABD(2, 4) = -SO(3, 2, KS) - SO(3, 3, KS);
        ABD(3, 4) = -SO(2, 3, KW) - SO(3, 3, KW);
        ABD(4, 4) = ABD(4, 4) + SO(4, 4, KO);
    } else {
        int kk = 1;
        ABD(1, 1) = SO(2, 2, KO);
        for (int i = 3; i <= I1; i++) {
            kk++;
            ABD(kk, kk) = SO(i, 2, KO);
            ABD(kk - 1, kk) = -SO(i, 2, KW);
        }
        if (perx) ABD(kk - I2 + 1, kk) = -SO(II, 2, KW);
        for (int j = 3; j <= J1; j++) {
            if (perx) ABD(kk, kk + 1) = NSO(2, j, KSW);
            for (int i = 2; i <= I1; i++) {
                kk++;
                ABD(kk, kk) = SO(i, j, KO);
                if (i != 2) {
                    ABD(kk - 1, kk) = -SO(i, j, KW);
                    ABD(kk - I2 - 1, kk) = NSO(i, j, KSW);
                }
                ABD(kk - I2 + 1, kk) = NSO(i + 1, j, KNW);
                ABD(kk - I2, kk) = -SO(i, j, KS);
            }
            if (perx) {
                ABD(kk - I2 + 1, kk) = -SO(II, j, KW);
                ABD(kk - 2 * I2 + 1, kk) = NSO(II, j, KNW);
            }
        }
        if (pery) {
            kk = kk - I2;
            const int J2 = (J1 - 2) * I2;
            kk++;
            ABD(kk - J2, kk) = -SO(2, JJ, KS);
            ABD(kk - J2 + 1, kk) = NSO(3, JJ, KSW);
            if (perxy) ABD(I2, kk) = NSO(2, JJ, KNW);
            for (int i = 3; i <= I1; i++) {
                kk++;
                ABD(kk - J2, kk) = -SO(i, JJ, KS);
                ABD(kk - J2 - 1, kk) = NSO(i, JJ, KNW);
                ABD(kk - J2 + 1, kk) = NSO(i + 1, JJ, KSW);
            }
            ABD(kk - J2 + 1, kk) = 0.0;
            if (perxy) ABD(1, kk) = NSO(II, JJ, KSW);
            if (perx) ABD(kk - 2 * I2 + 1, kk) = NSO(II, J1, KNW);
        }
        if (ibc < 0) ABD(kk, kk) = ABD(kk, kk) + SO(I1, J1, KO);
    }
    // DPOTF2('U')
    int bad = 0;
    for (int j = 1; j <= N && !bad; j++) {
        double dot = 0.0;
        for (int k = 1; k <= j - 1; k++) dot = dot + ABD(k, j) * ABD(k, j);
        double ajj = ABD(j, j) - dot;
        if (ajj <= 0.0 || ajj != ajj) { ABD(j, j) = ajj; bad = j; break; }
        ajj = sqrt(ajj);
        ABD(j, j) = ajj;
        if (j < N) {
            for (int c = j + 1; c <= N; c++) {
                double temp = 0.0;
                for (int k = 1; k <= j - 1; k++) temp = temp + ABD(k, c) * ABD(k, j);
                ABD(j, c) = ABD(j, c) + (-1.0) * temp;
            }
            const double r = 1.0 / ajj;
            for (int c = j + 1; c <= N; c++) ABD(j, c) = r * ABD(j, c);
        }
    }
    *info = bad;
}
#undef SO
#undef NSO

// SOLVE_cg, periodic: DPOTRS('U'), mean subtraction, ghost layer of the solution
__global__ void kp_solve_cg(double *q, const double *__restrict__ qf, int II, int JJ, const double *__restrict__ abd, double *bbd,
                            int nabd1, int jpn)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int ipn = jpn < 0 ? -jpn : jpn;
    const int I1 = II - 1, J1 = JJ - 1;
    int n = 0;
    for (int j = 2; j <= J1; j++)
        for (int i = 2; i <= I1; i++) bbd[n++] = qf[IX2(II, i, j)];
    for (int i = 1; i <= n; i++) {
        double temp = bbd[i - 1];
        for (int k = 1; k <= i - 1; k++) temp = temp - ABD(k, i) * bbd[k - 1];
        temp = temp / ABD(i, i);
        bbd[i - 1] = temp;
    }
    for (int k = n; k >= 1; k--) {
        if (bbd[k - 1] != 0.0) {
            bbd[k - 1] = bbd[k - 1] / ABD(k, k);
            for (int i = 1; i <= k - 1; i++) bbd[i - 1] = bbd[i - 1] - bbd[k - 1] * ABD(i, k);
        }
    }
    n = 0;
    for (int j = 2; j <= J1; j++)
        for (int i = 2; i <= I1; i++) q[IX2(II, i, j)] = bbd[n++];
    if (jpn != 0) {
        double cint = 0.0, qint = 0.0;
        for (int j = 2; j <= J1; j++)
            for (int i = 2; i <= I1; i++) { qint = qint + q[IX2(II, i, j)]; cint = cint + 1; }
        const double c = -qint / cint;
        for (int j = 2; j <= J1; j++)
            for (int i = 2; i <= I1; i++) q[IX2(II, i, j)] = q[IX2(II, i, j)] + c;
    }
    if (ipn == 1 || ipn == 3)
        for (int i = 2; i <= I1; i++) { q[IX2(II, i, JJ)] = q[IX2(II, i, 2)]; q[IX2(II, i, 1)] = q[IX2(II, i, J1)]; }
    if (ipn == 2 || ipn == 3)
        for (int j = 2; j <= J1; j++) { q[IX2(II, II, j)] = q[IX2(II, 2, j)]; q[IX2(II, 1, j)] = q[IX2(II, I1, j)]; }
    if (ipn == 3) {
        q[IX2(II, 1, 1)] = q[IX2(II, I1, J1)];
        q[IX2(II, II, 1)] = q[IX2(II, 2, J1)];
        q[IX2(II, 1, JJ)] = q[IX2(II, I1, 2)];
        q[IX2(II, II, JJ)] = q[IX2(II, 2, 2)];
    }
}
#undef ABD

}  // namespace

extern int *bmgx_cg_info_ptr();   // bmgx2_cg.cu: the status word bmgx_cg_LU_info reads

int bmgx2_periodic_interp_oi(const double *so, double *ci, bmgx_len IIF, bmgx_len JJF, bmgx_len IIC, bmgx_len JJC, int ifd,
                             int jpn, bmgx_stream_t s)
{
    const int ipn = jpn < 0 ? -jpn : jpn;
    if (ipn < 1 || ipn > 3) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_interp_OI: unknown boundary code");
    if (IIC < 3 || JJC < 3) return BMGX_OK;
    OiP p{};
    p.IIF = IIF; p.JJF = JJF; p.IIC = IIC; p.JJC = JJC;
    p.perx = ipn == 2 || ipn == 3; p.pery = ipn == 1 || ipn == 3;
    const int IICF = (IIF - 2) / 2 + 3, JJCF = (JJF - 2) / 2 + 3;
    p.IIFC = 2 * (IICF - 2) + 2; p.JJFC = 2 * (JJCF - 2) + 2;
    p.IBEG = 2; p.IBEGC = 3; p.IBEG_x = 2; p.IEND_x = IIF - 2; p.IENDC = IICF - 1; p.JENDC = JJCF - 1;
    if (p.perx) {
        p.IBEG = 0; p.IBEGC = 2;
        if ((int)IIC == IICF) { p.IENDC = IICF; p.IBEG_x = 3; p.IEND_x = IIF - 1; }
    }
    p.JBEG = 2; p.JBEGC = 3; p.JBEG_y = 2; p.JEND_y = JJF - 2;
    if (p.pery) {
        p.JBEG = 0; p.JBEGC = 2;
        if ((int)JJC == JJCF) { p.JENDC = JJCF; p.JBEG_y = 3; p.JEND_y = JJF - 1; }
    }
    cudaStream_t st = (cudaStream_t)s;
    dim3 b(64, 4);
    const int ncx = p.IENDC - p.IBEGC + 1, ncy = p.JENDC - p.JBEGC + 1;
    if (ncx > 0) {
        dim3 g(cdiv(ncx, 64), cdiv(JJC, 4));
        if (ifd != 1) kp_edges_x<true><<<g, b, 0, st>>>(so, ci, p); else kp_edges_x<false><<<g, b, 0, st>>>(so, ci, p);
        BMGX_LAUNCH_CHECK("SETUP_interp_OI/periodic x-edges");
    }
    if (ncy > 0) {
        dim3 g(cdiv(IIC, 64), cdiv(ncy, 4));
        if (ifd != 1) kp_edges_y<true><<<g, b, 0, st>>>(so, ci, p); else kp_edges_y<false><<<g, b, 0, st>>>(so, ci, p);
        BMGX_LAUNCH_CHECK("SETUP_interp_OI/periodic y-edges");
    }
    if (ncx > 0 && ncy > 0) {
        dim3 g(cdiv(ncx, 64), cdiv(ncy, 4));
        if (ifd != 1) kp_centres<true><<<g, b, 0, st>>>(so, ci, p); else kp_centres<false><<<g, b, 0, st>>>(so, ci, p);
        BMGX_LAUNCH_CHECK("SETUP_interp_OI/periodic centres");
    }
    return BMGX_OK;
}

int bmgx2_periodic_cg_lu(const double *so, bmgx_len ii, bmgx_len jj, int nstncl, double *abd, bmgx_len nabd1, bmgx_len nabd2,
                         int ibc, bmgx_stream_t s)
{
    const int ipn = ibc < 0 ? -ibc : ibc;
    if (ipn < 1 || ipn > 3) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_cg_LU: unknown boundary code");
    if (ii < 3 || jj < 3) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_cg_LU: empty grid");
    const unsigned n = (ii - 2) * (jj - 2);
    if (nabd1 < n || nabd2 < n) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_cg_LU: periodic boundaries need a dense ABD(N,N)");
    if (n > 4096) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "SETUP_cg_LU: periodic coarsest grid too large for the dense solve");
    cudaStream_t st = (cudaStream_t)s;
    cudaError_t e = cudaMemsetAsync(abd, 0, sizeof(double) * (size_t)nabd1 * nabd2, st);
    if (e != cudaSuccess) return bmgx_set_cuda_error(e, "SETUP_cg_LU/memset");
    bmgx_count_launch();
    kp_cg_lu<<<1, 32, 0, st>>>(so, (int)ii, (int)jj, nstncl, abd, (int)nabd1, ibc, bmgx_cg_info_ptr());
    BMGX_LAUNCH_CHECK("SETUP_cg_LU/periodic");
    return BMGX_OK;
}

int bmgx2_periodic_solve_cg(double *q, const double *qf, bmgx_len ii, bmgx_len jj, const double *abd, double *bbd, bmgx_len nabd1,
                            int jpn, bmgx_stream_t s)
{
    if (bmgx_batch() != 1) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "SOLVE_cg: periodic BCs not supported in a batch");
    kp_solve_cg<<<1, 32, 0, (cudaStream_t)s>>>(q, qf, (int)ii, (int)jj, abd, bbd, (int)nabd1, jpn);
    BMGX_LAUNCH_CHECK("SOLVE_cg/periodic");
    return BMGX_OK;
}
