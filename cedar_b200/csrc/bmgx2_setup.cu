// bmgx2_setup.cu -- operator-induced interpolation and Galerkin coarse operator.
// Reference: src/2d/ftn/BMG2_SymStd_SETUP_interp_OI.f90:84-256,
//            src/2d/ftn/BMG2_SymStd_SETUP_ITLI_ex.f90:94-213 (9-pt), 252-330 (5-pt).
// Expression order follows the reference term by term (bitwise parity with the
// oracle under -fmad=false).
#include "bmgx_common.cuh"
#include <cfloat>

#define SO(i, j, kk) so[IX3(IIF, JJF, i, j, kk)]
#define CI(i, j, kk) ci[IX3(IIC, JJC, i, j, kk)]
#define SOC(i, j, kk) soc[IX3(IIC, JJC, i, j, kk)]

__device__ __forceinline__ double oi_fix(double o, double sum, double ep, double base)
{
    // base + (o - sum) * max(o - (1+ep)*sum, 0) / (|o - (1+ep)*sum| + eps)
    const double ZEPS = DBL_EPSILON;  // EPSILON(1.D0), SETUP_interp_OI.f90:78
    double t = o - (1.0 + ep) * sum;
    return base + (o - sum) * fmax(t, 0.0) / (fabs(t) + ZEPS);
}

// stages 1 and 2: x-edge weights (LR, LL) and y-edge weights (LA, LB)
template <bool NINE>
__global__ void __launch_bounds__(256)
k_interp_edges(const double *__restrict__ so, double *__restrict__ ci, unsigned IIF, unsigned JJF,
               unsigned IIC, unsigned JJC, unsigned IICF1, unsigned JJCF1)
{
    unsigned ic = blockIdx.x * blockDim.x + threadIdx.x + 2;
    unsigned jc = blockIdx.y * blockDim.y + threadIdx.y + 2;
    unsigned i = 2 * (ic - 1), j = 2 * (jc - 1);
    if (ic >= 3 && ic <= IICF1 && jc <= JJC - 1) {
        double a, b;
        if (NINE) {
            a = SO(i, j, KW) + SO(i, j, KNW) + SO(i, j + 1, KSW);
            b = SO(i - 1, j, KW) + SO(i - 1, j, KSW) + SO(i - 1, j + 1, KNW);
        } else {
            a = SO(i, j, KW);
            b = SO(i - 1, j, KW);
        }
        double o = SO(i - 1, j, KO);
        double ep = fmin(fabs(a / o), fabs(b / o));
        double sum = a + b + SO(i - 1, j, KS) + SO(i - 1, j + 1, KS);
        sum = oi_fix(o, sum, ep, a + b);
        sum = 1.0 / sum;
        CI(ic, jc, LR) = a * sum;
        CI(ic, jc, LL) = b * sum;
    }
    if (jc >= 3 && jc <= JJCF1 && ic <= IIC - 1) {
        double a, b;
        if (NINE) {
            a = SO(i, j, KS) + SO(i, j, KNW) + SO(i + 1, j, KSW);
            b = SO(i, j - 1, KS) + SO(i, j - 1, KSW) + SO(i + 1, j - 1, KNW);
        } else {
            a = SO(i, j, KS);
            b = SO(i, j - 1, KS);
        }
        double o = SO(i, j - 1, KO);
        double ep = fmin(fabs(a / o), fabs(b / o));
        double sum = a + b + SO(i, j - 1, KW) + SO(i + 1, j - 1, KW);
        sum = oi_fix(o, sum, ep, a + b);
        sum = 1.0 / sum;
        CI(ic, jc, LA) = a * sum;
        CI(ic, jc, LB) = b * sum;
    }
}

// stage 3: cell-centre weights, needs the edge weights of the neighbours
template <bool NINE>
__global__ void __launch_bounds__(256)
k_interp_centres(const double *__restrict__ so, double *ci, unsigned IIF, unsigned JJF, unsigned IIC,
                 unsigned JJC, unsigned IICF1, unsigned JJCF1)
{
    unsigned ic = blockIdx.x * blockDim.x + threadIdx.x + 3;
    unsigned jc = blockIdx.y * blockDim.y + threadIdx.y + 3;
    if (ic > IICF1 || jc > JJCF1) return;
    unsigned i = 2 * (ic - 1), j = 2 * (jc - 1);
    double o = SO(i - 1, j - 1, KO);
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
    sum = oi_fix(o, sum, ep, sum);
    double s = 1.0 / sum;
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

// stages: bit 0 = edge F-points (LL,LR,LA,LB), bit 1 = cell centres (LSW..LSE)
static int interp_oi_stages(const double *so, double *ci, bmgx_len IIF, bmgx_len JJF, bmgx_len IIC, bmgx_len JJC, int ifd,
                            int jpn, unsigned stages, bmgx_stream_t s)
{
    if (!bmgx_nonperiodic(jpn)) {
        if (stages != 3u) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "SETUP_interp_OI_stage: periodic BCs not supported on the decomposed path");
        return bmgx2_periodic_interp_oi(so, ci, IIF, JJF, IIC, JJC, ifd, jpn, s);
    }
    if (IIC < 3 || JJC < 3) return BMGX_OK;
    unsigned IICF1 = (IIF - 2) / 2 + 2, JJCF1 = (JJF - 2) / 2 + 2;
    cudaStream_t st = (cudaStream_t)s;
    dim3 b(64, 4);
    if (stages & 1u) {
        // edges: ic in 2..max(IICF1, IIC-1), jc in 2..max(JJCF1, JJC-1)
        unsigned mi = IICF1 > IIC - 1 ? IICF1 : IIC - 1, mj = JJCF1 > JJC - 1 ? JJCF1 : JJC - 1;
        dim3 g1(cdiv(mi - 1, 64), cdiv(mj - 1, 4));
        if (ifd != 1) k_interp_edges<true><<<g1, b, 0, st>>>(so, ci, IIF, JJF, IIC, JJC, IICF1, JJCF1);
        else k_interp_edges<false><<<g1, b, 0, st>>>(so, ci, IIF, JJF, IIC, JJC, IICF1, JJCF1);
        BMGX_LAUNCH_CHECK("SETUP_interp_OI/edges");
    }
    if ((stages & 2u) && IICF1 >= 3 && JJCF1 >= 3) {
        dim3 g2(cdiv(IICF1 - 2, 64), cdiv(JJCF1 - 2, 4));
        if (ifd != 1) k_interp_centres<true><<<g2, b, 0, st>>>(so, ci, IIF, JJF, IIC, JJC, IICF1, JJCF1);
        else k_interp_centres<false><<<g2, b, 0, st>>>(so, ci, IIF, JJF, IIC, JJC, IICF1, JJCF1);
        BMGX_LAUNCH_CHECK("SETUP_interp_OI/centres");
    }
    return BMGX_OK;
}

extern "C" int bmgx2_SETUP_interp_OI(const double *so, const double *soc, double *ci, bmgx_len IIF, bmgx_len JJF,
                                     bmgx_len IIC, bmgx_len JJC, int ifd, int nstncl, int jpn, int irelax,
                                     bmgx_stream_t s)
{
    (void)soc; (void)nstncl; (void)irelax;
    return interp_oi_stages(so, ci, IIF, JJF, IIC, JJC, ifd, jpn, 3u, s);
}

// The two dependent stages separately (stage 1 = edges, 2 = cell centres): the
// block-decomposed solver exchanges the edge weights of the ghost lines in between
// (the reference's MPI kernel does the same between its loops,
// src/2d/ftn/mpi/BMG2_SymStd_SETUP_interp_OI.f90).
extern "C" int bmgx2_SETUP_interp_OI_stage(const double *so, const double *soc, double *ci, bmgx_len IIF, bmgx_len JJF,
                                           bmgx_len IIC, bmgx_len JJC, int ifd, int nstncl, int jpn, int irelax,
                                           int stage, bmgx_stream_t s)
{
    (void)soc; (void)nstncl; (void)irelax;
    if (stage != 1 && stage != 2) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_interp_OI_stage: stage must be 1 or 2");
    return interp_oi_stages(so, ci, IIF, JJF, IIC, JJC, ifd, jpn, 1u << (stage - 1), s);
}

// ---- Galerkin coarse operator ----------------------------------------------
template <bool NINE>
__global__ void __launch_bounds__(128)
k_itli(const double *__restrict__ so, double *__restrict__ soc, const double *__restrict__ ci, unsigned IIF,
       unsigned JJF, unsigned IIC, unsigned JJC)
{
    unsigned ic = blockIdx.x * blockDim.x + threadIdx.x + 2;
    unsigned jc = blockIdx.y * blockDim.y + threadIdx.y + 2;
    if (ic > IIC - 1 || jc > JJC - 1) return;
    unsigned i = 2 * (ic - 1), j = 2 * (jc - 1);
    if (NINE) {
        double co = SO(i, j + 1, KNW) * CI(ic, jc + 1, LSW) + SO(i, j, KW) * CI(ic, jc, LL)
                    + SO(i, j, KSW) * CI(ic, jc, LNW);
        double cs = SO(i, j - 1, KW) * CI(ic, jc, LNW) + SO(i, j, KNW) * CI(ic, jc, LL);
        double csw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LNW) + SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LA)
                     + SO(i - 1, j, KNW) + SO(i - 1, j, KS) * CI(ic, jc, LL);
        double cw = -SO(i - 1, j, KO) * CI(ic, jc, LL) + SO(i - 1, j, KS) * CI(ic, jc, LNW)
                    + SO(i - 1, j, KSW) * CI(ic - 1, jc, LA) + SO(i - 1, j, KW)
                    + SO(i - 1, j + 1, KNW) * CI(ic - 1, jc + 1, LB) + SO(i - 1, j + 1, KS) * CI(ic, jc + 1, LSW);
        double cnw = -SO(i - 1, j + 1, KO) * CI(ic, jc + 1, LSW) + SO(i - 1, j + 1, KS) * CI(ic, jc, LL)
                     + SO(i - 1, j + 1, KSW) + SO(i - 1, j + 1, KW) * CI(ic - 1, jc + 1, LB);
        double cn = SO(i, j + 1, KSW) * CI(ic, jc, LL) + SO(i, j + 1, KW) * CI(ic, jc + 1, LSW);
        SOC(ic, jc, KW) = co + CI(ic, jc, LA) * cs + CI(ic, jc, LNE) * csw + CI(ic, jc, LR) * cw
                          + CI(ic, jc + 1, LSE) * cnw + CI(ic, jc + 1, LB) * cn;
        double cosw = SO(i, j, KSW) * CI(ic, jc, LSW);
        double cssw = SO(i, j - 1, KSW) * CI(ic, jc - 1, LL) + SO(i, j - 1, KW) * CI(ic, jc, LSW);
        double cswsw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LSW) + SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LL)
                       + SO(i - 1, j - 1, KSW) + SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LB);
        double cwsw = SO(i - 1, j, KS) * CI(ic, jc, LSW) + SO(i - 1, j, KSW) * CI(ic - 1, jc, LB);
        SOC(ic, jc, KSW) = cosw + CI(ic, jc, LA) * cssw + CI(ic, jc, LNE) * cswsw + CI(ic, jc, LR) * cwsw;
        double coa = SO(i, j, KSW) * CI(ic, jc, LSE) + SO(i, j, KS) * CI(ic, jc, LB)
                     + SO(i + 1, j, KNW) * CI(ic + 1, jc, LSW);
        double cea = SO(i + 1, j, KS) * CI(ic + 1, jc, LSW) + SO(i + 1, j, KSW) * CI(ic, jc, LB);
        double csea = -SO(i + 1, j - 1, KO) * CI(ic + 1, jc, LSW) + SO(i + 1, j - 1, KS) * CI(ic + 1, jc - 1, LL)
                      + SO(i + 1, j - 1, KSW) + SO(i + 1, j - 1, KW) * CI(ic, jc, LB);
        double csa = -SO(i, j - 1, KO) * CI(ic, jc, LB) + SO(i + 1, j - 1, KW) * CI(ic + 1, jc, LSW)
                     + SO(i + 1, j - 1, KNW) * CI(ic + 1, jc - 1, LL) + SO(i, j - 1, KS)
                     + SO(i, j - 1, KSW) * CI(ic, jc - 1, LR) + SO(i, j - 1, KW) * CI(ic, jc, LSE);
        double cswa = -SO(i - 1, j - 1, KO) * CI(ic, jc, LSE) + SO(i, j - 1, KW) * CI(ic, jc, LB)
                      + SO(i, j - 1, KNW) + SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR);
        double cwa = SO(i, j, KNW) * CI(ic, jc, LB) + SO(i - 1, j, KS) * CI(ic, jc, LSE);
        SOC(ic, jc, KS) = coa + CI(ic + 1, jc, LL) * cea + CI(ic + 1, jc, LNW) * csea + CI(ic, jc, LA) * csa
                          + CI(ic, jc, LNE) * cswa + CI(ic, jc, LR) * cwa;
        double conw = SO(i - 1, j, KNW) * CI(ic, jc, LSE);
        double cenw = SO(i, j, KNW) * CI(ic, jc, LB) + SO(i - 1, j, KS) * CI(ic, jc, LSE);
        double csenw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LSE) + SO(i, j - 1, KW) * CI(ic, jc, LB)
                       + SO(i, j - 1, KNW) + SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR);
        double cfnw = SO(i - 1, j - 1, KW) * CI(ic, jc, LSE) + SO(i - 1, j - 1, KNW) * CI(ic, jc - 1, LR);
        SOC(ic, jc, KNW) = conw + CI(ic, jc, LL) * cenw + CI(ic, jc, LNW) * csenw + CI(ic - 1, jc, LA) * cfnw;

        co = SO(i, j, KW) * CI(ic, jc, LR) + SO(i, j + 1, KNW) * CI(ic, jc + 1, LSE)
             + SO(i, j + 1, KS) * CI(ic, jc + 1, LB) + SO(i + 1, j + 1, KSW) * CI(ic + 1, jc + 1, LSW)
             + SO(i + 1, j, KW) * CI(ic + 1, jc, LL) + SO(i + 1, j, KNW) * CI(ic + 1, jc, LNW)
             + SO(i, j, KS) * CI(ic, jc, LA) + SO(i, j, KSW) * CI(ic, jc, LNE) - SO(i, j, KO);
        cw = -SO(i - 1, j, KO) * CI(ic, jc, LR) + SO(i - 1, j + 1, KS) * CI(ic, jc + 1, LSE)
             + SO(i, j + 1, KSW) * CI(ic, jc + 1, LB) + SO(i, j, KW) + SO(i, j, KNW) * CI(ic, jc, LA)
             + SO(i - 1, j, KS) * CI(ic, jc, LNE);
        cnw = -SO(i - 1, j + 1, KO) * CI(ic, jc + 1, LSE) + SO(i, j + 1, KW) * CI(ic, jc + 1, LB)
              + SO(i, j + 1, KNW) + SO(i - 1, j + 1, KS) * CI(ic, jc, LR);
        cn = -SO(i, j + 1, KO) * CI(ic, jc + 1, LB) + SO(i + 1, j + 1, KNW) * CI(ic + 1, jc, LL)
             + SO(i, j + 1, KS) + SO(i, j + 1, KSW) * CI(ic, jc, LR) + SO(i, j + 1, KW) * CI(ic, jc + 1, LSE)
             + SO(i + 1, j + 1, KW) * CI(ic + 1, jc + 1, LSW);
        double cne = -SO(i + 1, j + 1, KO) * CI(ic + 1, jc + 1, LSW) + SO(i + 1, j + 1, KS) * CI(ic + 1, jc, LL)
                     + SO(i + 1, j + 1, KSW) + SO(i + 1, j + 1, KW) * CI(ic, jc + 1, LB);
        double ce = -SO(i + 1, j, KO) * CI(ic + 1, jc, LL) + SO(i + 1, j, KS) * CI(ic + 1, jc, LNW)
                    + SO(i + 1, j, KSW) * CI(ic, jc, LA) + SO(i + 1, j, KW)
                    + SO(i + 1, j + 1, KNW) * CI(ic, jc + 1, LB) + SO(i + 1, j + 1, KS) * CI(ic + 1, jc + 1, LSW);
        double cse = -SO(i + 1, j - 1, KO) * CI(ic + 1, jc, LNW) + SO(i + 1, j - 1, KW) * CI(ic, jc, LA)
                     + SO(i + 1, j, KNW) + SO(i + 1, j, KS) * CI(ic + 1, jc, LL);
        cs = -SO(i, j - 1, KO) * CI(ic, jc, LA) + SO(i, j - 1, KW) * CI(ic, jc, LNE)
             + SO(i, j, KNW) * CI(ic, jc, LR) + SO(i, j, KS) + SO(i + 1, j, KSW) * CI(ic + 1, jc, LL)
             + SO(i + 1, j - 1, KW) * CI(ic + 1, jc, LNW);
        csw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LNE) + SO(i - 1, j, KS) * CI(ic, jc, LR) + SO(i, j, KSW)
              + SO(i, j - 1, KW) * CI(ic, jc, LA);
        SOC(ic, jc, KO) = -CI(ic, jc + 1, LSE) * cnw - CI(ic, jc + 1, LB) * cn - CI(ic + 1, jc + 1, LSW) * cne
                          - CI(ic, jc, LR) * cw - co - CI(ic + 1, jc, LL) * ce - CI(ic, jc, LNE) * csw
                          - CI(ic, jc, LA) * cs - CI(ic + 1, jc, LNW) * cse;
    } else {
        double co = SO(i, j, KW) * CI(ic, jc, LL);
        double cs = SO(i, j - 1, KW) * CI(ic, jc, LNW);
        double csw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LNW) + SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LA)
                     + SO(i - 1, j, KS) * CI(ic, jc, LL);
        double cw = -SO(i - 1, j, KO) * CI(ic, jc, LL) + SO(i - 1, j, KS) * CI(ic, jc, LNW) + SO(i - 1, j, KW)
                    + SO(i - 1, j + 1, KS) * CI(ic, jc + 1, LSW);
        double cnw = -SO(i - 1, j + 1, KO) * CI(ic, jc + 1, LSW) + SO(i - 1, j + 1, KS) * CI(ic, jc, LL)
                     + SO(i - 1, j + 1, KW) * CI(ic - 1, jc + 1, LB);
        double cn = SO(i, j + 1, KW) * CI(ic, jc + 1, LSW);
        SOC(ic, jc, KW) = co + CI(ic, jc, LA) * cs + CI(ic, jc, LNE) * csw + CI(ic, jc, LR) * cw
                          + CI(ic, jc + 1, LSE) * cnw + CI(ic, jc + 1, LB) * cn;
        double cssw = SO(i, j - 1, KW) * CI(ic, jc, LSW);
        double cswsw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LSW) + SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LL)
                       + SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LB);
        double cwsw = SO(i - 1, j, KS) * CI(ic, jc, LSW);
        SOC(ic, jc, KSW) = CI(ic, jc, LA) * cssw + CI(ic, jc, LNE) * cswsw + CI(ic, jc, LR) * cwsw;
        double coa = SO(i, j, KS) * CI(ic, jc, LB);
        double cea = SO(i + 1, j, KS) * CI(ic + 1, jc, LSW);
        double csea = -SO(i + 1, j - 1, KO) * CI(ic + 1, jc, LSW) + SO(i + 1, j - 1, KS) * CI(ic + 1, jc - 1, LL)
                      + SO(i + 1, j - 1, KW) * CI(ic, jc, LB);
        double csa = -SO(i, j - 1, KO) * CI(ic, jc, LB) + SO(i + 1, j - 1, KW) * CI(ic + 1, jc, LSW)
                     + SO(i, j - 1, KS) + SO(i, j - 1, KW) * CI(ic, jc, LSE);
        double cswa = -SO(i - 1, j - 1, KO) * CI(ic, jc, LSE) + SO(i, j - 1, KW) * CI(ic, jc, LB)
                      + SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR);
        double cwa = SO(i - 1, j, KS) * CI(ic, jc, LSE);
        SOC(ic, jc, KS) = coa + CI(ic + 1, jc, LL) * cea + CI(ic + 1, jc, LNW) * csea + CI(ic, jc, LA) * csa
                          + CI(ic, jc, LNE) * cswa + CI(ic, jc, LR) * cwa;
        double cenw = SO(i - 1, j, KS) * CI(ic, jc, LSE);
        double csenw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LSE) + SO(i, j - 1, KW) * CI(ic, jc, LB)
                       + SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR);
        double csnw = SO(i - 1, j - 1, KW) * CI(ic, jc, LSE);
        SOC(ic, jc, KNW) = CI(ic, jc, LL) * cenw + CI(ic, jc, LNW) * csenw + CI(ic - 1, jc, LA) * csnw;
        co = SO(i, j, KW) * CI(ic, jc, LR) + SO(i, j + 1, KS) * CI(ic, jc + 1, LB)
             + SO(i + 1, j, KW) * CI(ic + 1, jc, LL) + SO(i, j, KS) * CI(ic, jc, LA) - SO(i, j, KO);
        cw = -SO(i - 1, j, KO) * CI(ic, jc, LR) + SO(i - 1, j + 1, KS) * CI(ic, jc + 1, LSE) + SO(i, j, KW)
             + SO(i - 1, j, KS) * CI(ic, jc, LNE);
        cnw = -SO(i - 1, j + 1, KO) * CI(ic, jc + 1, LSE) + SO(i, j + 1, KW) * CI(ic, jc + 1, LB)
              + SO(i - 1, j + 1, KS) * CI(ic, jc, LR);
        cn = -SO(i, j + 1, KO) * CI(ic, jc + 1, LB) + SO(i, j + 1, KS) + SO(i, j + 1, KW) * CI(ic, jc + 1, LSE)
             + SO(i + 1, j + 1, KW) * CI(ic + 1, jc + 1, LSW);
        double cne = -SO(i + 1, j + 1, KO) * CI(ic + 1, jc + 1, LSW) + SO(i + 1, j + 1, KS) * CI(ic + 1, jc, LL)
                     + SO(i + 1, j + 1, KW) * CI(ic, jc + 1, LB);
        double ce = -SO(i + 1, j, KO) * CI(ic + 1, jc, LL) + SO(i + 1, j, KS) * CI(ic + 1, jc, LNW)
                    + SO(i + 1, j, KW) + SO(i + 1, j + 1, KS) * CI(ic + 1, jc + 1, LSW);
        double cse = -SO(i + 1, j - 1, KO) * CI(ic + 1, jc, LNW) + SO(i + 1, j - 1, KW) * CI(ic, jc, LA)
                     + SO(i + 1, j, KS) * CI(ic + 1, jc, LL);
        cs = -SO(i, j - 1, KO) * CI(ic, jc, LA) + SO(i, j - 1, KW) * CI(ic, jc, LNE) + SO(i, j, KS)
             + SO(i + 1, j - 1, KW) * CI(ic + 1, jc, LNW);
        csw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LNE) + SO(i - 1, j, KS) * CI(ic, jc, LR)
              + SO(i, j - 1, KW) * CI(ic, jc, LA);
        SOC(ic, jc, KO) = -CI(ic, jc + 1, LSE) * cnw - CI(ic, jc + 1, LB) * cn - CI(ic + 1, jc + 1, LSW) * cne
                          - CI(ic, jc, LR) * cw - co - CI(ic + 1, jc, LL) * ce - CI(ic, jc, LNE) * csw
                          - CI(ic, jc, LA) * cs - CI(ic + 1, jc, LNW) * cse;
    }
}

extern "C" int bmgx2_SETUP_ITLI_ex(const double *so, double *soc, const double *ci, bmgx_len IIF, bmgx_len JJF,
                                   bmgx_len IIC, bmgx_len JJC, int ifd, int nstncl, int ipn, bmgx_stream_t s)
{
    (void)nstncl;
    const bool periodic = !bmgx_nonperiodic(ipn);
    const int apn = ipn < 0 ? -ipn : ipn;
    if (periodic && (apn < 1 || apn > 3)) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_ITLI_ex: unknown boundary code");
    if (IIC < 3 || JJC < 3) return BMGX_OK;
    dim3 b(32, 4), g(cdiv(IIC - 2, 32), cdiv(JJC - 2, 4));
    if (ifd != 1) k_itli<true><<<g, b, 0, (cudaStream_t)s>>>(so, soc, ci, IIF, JJF, IIC, JJC);
    else k_itli<false><<<g, b, 0, (cudaStream_t)s>>>(so, soc, ci, IIF, JJF, IIC, JJC);
    BMGX_LAUNCH_CHECK("SETUP_ITLI_ex");
    if (periodic) {
        // BMG2_SymStd_SETUP_ITLI_ex.f90:216-247 / 333-364: ghost layer of all five planes of the coarse operator,
        // y rows first, then x columns
        const size_t pc = (size_t)IIC * JJC;
        for (int kk = 0; kk < 5; kk++)
            if ((apn == 1 || apn == 3) && bmgx_wrap_y(soc + kk * pc, IIC, JJC, (cudaStream_t)s)) return BMGX_ERR_CUDA;
        for (int kk = 0; kk < 5; kk++)
            if ((apn == 2 || apn == 3) && bmgx_wrap_x(soc + kk * pc, IIC, 1, 1, JJC, (cudaStream_t)s)) return BMGX_ERR_CUDA;
    }
    return BMGX_OK;
}
