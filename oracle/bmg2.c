/*
 * oracle/bmg2.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, fp64, no FMA contraction: build with
 * -ffp-contract=off) of the serial 2D BoxMG kernels of the reference
 * (OVGULIU/cedar, src/2d/ftn/BMG2_SymStd_*.f90).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library; the product path (cedar_b200/, include/) never does.
 *
 * Every function keeps the loop order and the floating-point expression order
 * of the Fortran routine it cites, so that the CUDA kernels (compiled with
 * -fmad=false) can be compared bit for bit.  Non-periodic branches, plus the
 * periodic branch of point relaxation (the first brick of SURVEY.md section 8(f).2;
 * the other periodic branches are not restated yet).
 *
 * Parity pin: the multilevel driver in ml2.c built on these kernels reproduces
 * the residual history printed in the reference README.md:52-63 (see
 * tests/test_oracle_golden.py); LAPACK restatements are checked against
 * scipy.linalg.lapack in tests/test_oracle_golden.py.
 *
 * Arrays are Fortran ordered, 1-based in the comments, i fastest:
 *   Q(II,JJ), SO(II,JJ,NStncl), SOR(II,JJ,2), CI(IIC,JJC,8).
 */
#include <math.h>
#include <stddef.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "oracle.h"

/* stencil slots, src/2d/ftn/BMG_stencils_f90.h:29-37,70-71 */
enum { KO = 1, KW = 2, KS = 3, KSW = 4, KNW = 5 };
enum { LL = 1, LR = 2, LA = 3, LB = 4, LSW = 5, LNW = 6, LNE = 7, LSE = 8 };
enum { MSOR = 2 };

#define A2(a, n1, i, j) (a)[(size_t)((j)-1) * (n1) + (size_t)((i)-1)]
#define A3(a, n1, n2, i, j, k) \
    (a)[((size_t)((k)-1) * (n2) + (size_t)((j)-1)) * (n1) + (size_t)((i)-1)]

void orc_set_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int orc_get_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------------
 * LAPACK restatements (third-party: system LAPACK, unpinned version; call
 * sites BMG2_SymStd_SETUP_lines_x.f90:86, relax_lines_x.f90:115,145,
 * SETUP_cg_LU.f90:118, SOLVE_cg.f90:104).  Reference LAPACK 3.x algorithms.
 * ---------------------------------------------------------------------- */

/* DPTTRF: L*D*L' of an SPD tridiagonal; d[0..n-1], e[0..n-2] (unblocked form
 * of the 4-way unrolled reference loop; same operations per element). */
int orc_dpttrf(int n, double *d, double *e)
{
    for (int i = 0; i < n - 1; i++) {
        if (d[i] <= 0.0) return i + 1;
        double ei = e[i];
        e[i] = ei / d[i];
        d[i + 1] = d[i + 1] - e[i] * ei;
    }
    if (n > 0 && d[n - 1] <= 0.0) return n;
    return 0;
}

/* DPTTRS/DPTTS2 with NRHS=1: solve L*D*L' x = b in place. */
void orc_dptts2(int n, const double *d, const double *e, double *b)
{
    if (n <= 1) {
        if (n == 1) b[0] = b[0] * (1.0 / d[0]); /* DSCAL(1/d) path of DPTTS2 */
        return;
    }
    for (int i = 1; i < n; i++) b[i] = b[i] - b[i - 1] * e[i - 1];
    b[n - 1] = b[n - 1] / d[n - 1];
    for (int i = n - 2; i >= 0; i--) b[i] = b[i] / d[i] - b[i + 1] * e[i];
}

/* DPBTRF('U') unblocked (DPBTF2; the blocked path needs KD > 32, never the
 * case for the coarsest grids of the hot path).  ab is column major
 * (ldab, n), diagonal in row kd+1 (1-based). */
int orc_dpbtf2_upper(int n, int kd, double *ab, int ldab)
{
#define AB(i, j) ab[(size_t)((j)-1) * ldab + ((i)-1)]
    int kld = ldab - 1 > 1 ? ldab - 1 : 1;
    for (int j = 1; j <= n; j++) {
        double ajj = AB(kd + 1, j);
        if (ajj <= 0.0) return j;
        ajj = sqrt(ajj);
        AB(kd + 1, j) = ajj;
        int kn = kd < n - j ? kd : n - j;
        if (kn > 0) {
            /* x = AB(kd, j+1) with stride kld : row j of U right of diag */
            double *x = &AB(kd, j + 1);
            double r = 1.0 / ajj;
            for (int p = 0; p < kn; p++) x[(size_t)p * kld] = r * x[(size_t)p * kld];
            /* DSYR('Upper', kn, -1, x, kld, AB(kd+1,j+1), kld) */
            double *a = &AB(kd + 1, j + 1);
            for (int q = 0; q < kn; q++) {
                double xq = x[(size_t)q * kld];
                if (xq != 0.0) {
                    double temp = -1.0 * xq;
                    for (int p = 0; p <= q; p++)
                        a[(size_t)q * kld + p] = a[(size_t)q * kld + p] + x[(size_t)p * kld] * temp;
                }
            }
        }
    }
    return 0;
#undef AB
}

/* DPBTRS('U', n, kd, 1, ...) = DTBSV(U,T,N) then DTBSV(U,N,N). */
void orc_dpbtrs_upper(int n, int kd, const double *ab, int ldab, double *x)
{
#define AB(i, j) ab[(size_t)((j)-1) * ldab + ((i)-1)]
    int kplus1 = kd + 1;
    for (int j = 1; j <= n; j++) {
        double temp = x[j - 1];
        int l = kplus1 - j;
        int i0 = j - kd > 1 ? j - kd : 1;
        for (int i = i0; i <= j - 1; i++) temp = temp - AB(l + i, j) * x[i - 1];
        temp = temp / AB(kplus1, j);
        x[j - 1] = temp;
    }
    for (int j = n; j >= 1; j--) {
        if (x[j - 1] != 0.0) {
            int l = kplus1 - j;
            x[j - 1] = x[j - 1] / AB(kplus1, j);
            double temp = x[j - 1];
            int i0 = j - kd > 1 ? j - kd : 1;
            for (int i = j - 1; i >= i0; i--) x[i - 1] = x[i - 1] - temp * AB(l + i, j);
        }
    }
#undef AB
}

/* ------------------------------------------------------------------------
 * BMG2_SymStd_SETUP_recip.f90:63-67
 * ---------------------------------------------------------------------- */
void orc2_setup_recip(const double *so, double *sor, int nx, int ny, int nstncl, int nsorv)
{
    (void)nstncl; (void)nsorv;
#pragma omp parallel for schedule(static)
    for (int j = 2; j <= ny - 1; j++)
        for (int i = 2; i <= nx - 1; i++)
            A3(sor, nx, ny, i, j, MSOR) = 1.0 / A3(so, nx, ny, i, j, KO);
}

/* ------------------------------------------------------------------------
 * BMG2_SymStd_relax_GS.f90:68-135 (non-periodic)
 * ---------------------------------------------------------------------- */
int orc2_relax_gs(int k, const double *so, const double *qf, double *q, const double *sor,
                  int II, int JJ, int kf, int ifd, int nstncl, int nsorv,
                  int irelax_sym, int updown, int jpn)
{
    (void)nstncl; (void)nsorv;
    int J1 = JJ - 1, I1 = II - 1;
    int lstart, lend, lstride;
    if (irelax_sym == 0 || (updown == 0 && irelax_sym == 1)) {
        lstart = 2; lend = 3; lstride = 1;
    } else {
        lstart = 3; lend = 2; lstride = -1;
    }
#define SO(i, j, kk) A3(so, II, JJ, i, j, kk)
#define Q(i, j) A2(q, II, i, j)
#define QF(i, j) A2(qf, II, i, j)
#define SOR(i, j, kk) A3(sor, II, JJ, i, j, kk)
    const int ipn = jpn < 0 ? -jpn : jpn;
    if (ipn != 0 && ipn != 4) {
        /* BMG2_SymStd_relax_GS.f90:139-224, periodic: same updates; the x ghosts of a row are refreshed after
         * each colour of that row (Q(1,J)=Q(I1,J), Q(II,J)=Q(2,J)) when periodic in x (|jpn| = 2, 3), the y ghost
         * rows only ONCE, after the whole sweep, when periodic in y (|jpn| = 1, 3) -- so the first and last
         * interior rows see the previous sweep's wrap.  Literal loop order. */
        const int perx = (ipn == 2 || ipn == 3), pery = (ipn == 1 || ipn == 3);
        if (k < kf || ifd != 1) {
            for (int jbeg = lstart; jbeg != lend + lstride; jbeg += lstride) {
                int jend = 2 * ((J1 - jbeg) / 2) + jbeg;
                for (int j = jbeg; j <= jend; j += 2)
                    for (int ibeg = lstart; ibeg != lend + lstride; ibeg += lstride) {
                        int iend = 2 * ((I1 - ibeg) / 2) + ibeg;
                        for (int i = ibeg; i <= iend; i += 2)
                            Q(i, j) = (QF(i, j) + SO(i, j, KW) * Q(i - 1, j) + SO(i + 1, j, KW) * Q(i + 1, j)
                                       + SO(i, j, KS) * Q(i, j - 1) + SO(i, j + 1, KS) * Q(i, j + 1)
                                       + SO(i, j, KSW) * Q(i - 1, j - 1) + SO(i + 1, j, KNW) * Q(i + 1, j - 1)
                                       + SO(i, j + 1, KNW) * Q(i - 1, j + 1) + SO(i + 1, j + 1, KSW) * Q(i + 1, j + 1))
                                      * SOR(i, j, MSOR);
                        if (perx) { Q(1, j) = Q(I1, j); Q(II, j) = Q(2, j); }
                    }
            }
        } else {
            for (int jo = lstart; jo != lend + lstride; jo += lstride)
                for (int j = 2; j <= J1; j++) {
                    int ibeg = (j + jo) % 2 + 2;
                    int iend = 2 * ((I1 - ibeg) / 2) + ibeg;
                    for (int i = ibeg; i <= iend; i += 2)
                        Q(i, j) = (QF(i, j) + SO(i, j, KW) * Q(i - 1, j) + SO(i + 1, j, KW) * Q(i + 1, j)
                                   + SO(i, j, KS) * Q(i, j - 1) + SO(i, j + 1, KS) * Q(i, j + 1))
                                  * SOR(i, j, MSOR);
                    if (perx) { Q(1, j) = Q(I1, j); Q(II, j) = Q(2, j); }
                }
        }
        if (pery)
            for (int i = 1; i <= II; i++) { Q(i, 1) = Q(i, J1); Q(i, JJ) = Q(i, 2); }
        return 0;
    }
    if (k < kf || ifd != 1) {
        for (int jbeg = lstart; jbeg != lend + lstride; jbeg += lstride) {
            int jend = 2 * ((J1 - jbeg) / 2) + jbeg;
            /* the reference nests IBEG inside J; points of one colour are
             * independent and a row of colour (jbeg,ibeg2) only reads rows of
             * the other j-parity plus its own row's first colour, so running
             * the two i-colours row by row is the literal order. */
#pragma omp parallel for schedule(static)
            for (int j = jbeg; j <= jend; j += 2) {
                for (int ibeg = lstart; ibeg != lend + lstride; ibeg += lstride) {
                    int iend = 2 * ((I1 - ibeg) / 2) + ibeg;
                    for (int i = ibeg; i <= iend; i += 2) {
                        Q(i, j) = (QF(i, j)
                                   + SO(i, j, KW) * Q(i - 1, j)
                                   + SO(i + 1, j, KW) * Q(i + 1, j)
                                   + SO(i, j, KS) * Q(i, j - 1)
                                   + SO(i, j + 1, KS) * Q(i, j + 1)
                                   + SO(i, j, KSW) * Q(i - 1, j - 1)
                                   + SO(i + 1, j, KNW) * Q(i + 1, j - 1)
                                   + SO(i, j + 1, KNW) * Q(i - 1, j + 1)
                                   + SO(i + 1, j + 1, KSW) * Q(i + 1, j + 1))
                                  * SOR(i, j, MSOR);
                    }
                }
            }
        }
    } else {
        for (int jo = lstart; jo != lend + lstride; jo += lstride) {
#pragma omp parallel for schedule(static)
            for (int j = 2; j <= J1; j++) {
                int ibeg = (j + jo) % 2 + 2;
                int iend = 2 * ((I1 - ibeg) / 2) + ibeg;
                for (int i = ibeg; i <= iend; i += 2) {
                    Q(i, j) = (QF(i, j)
                               + SO(i, j, KW) * Q(i - 1, j)
                               + SO(i + 1, j, KW) * Q(i + 1, j)
                               + SO(i, j, KS) * Q(i, j - 1)
                               + SO(i, j + 1, KS) * Q(i, j + 1))
                              * SOR(i, j, MSOR);
                }
            }
        }
    }
    return 0;
#undef SO
#undef Q
#undef QF
#undef SOR
}

/* ------------------------------------------------------------------------
 * BMG2_SymStd_residual.f90:87-119
 * ---------------------------------------------------------------------- */
void orc2_residual(int k, const double *so, const double *qf, const double *q, double *res,
                   int II, int JJ, int kf, int ifd, int nstncl)
{
    (void)nstncl;
    int J1 = JJ - 1, I1 = II - 1;
#define SO(i, j, kk) A3(so, II, JJ, i, j, kk)
#define Q(i, j) A2(q, II, i, j)
#define QF(i, j) A2(qf, II, i, j)
#define RES(i, j) A2(res, II, i, j)
    if (k < kf || ifd != 1) {
#pragma omp parallel for schedule(static)
        for (int j = 2; j <= J1; j++)
            for (int i = 2; i <= I1; i++)
                RES(i, j) = QF(i, j)
                            + SO(i, j, KW) * Q(i - 1, j)
                            + SO(i + 1, j, KW) * Q(i + 1, j)
                            + SO(i, j, KS) * Q(i, j - 1)
                            + SO(i, j + 1, KS) * Q(i, j + 1)
                            + SO(i, j, KSW) * Q(i - 1, j - 1)
                            + SO(i + 1, j, KNW) * Q(i + 1, j - 1)
                            + SO(i, j + 1, KNW) * Q(i - 1, j + 1)
                            + SO(i + 1, j + 1, KSW) * Q(i + 1, j + 1)
                            - SO(i, j, KO) * Q(i, j);
    } else {
#pragma omp parallel for schedule(static)
        for (int j = 2; j <= J1; j++)
            for (int i = 2; i <= I1; i++)
                RES(i, j) = QF(i, j)
                            + SO(i, j, KW) * Q(i - 1, j)
                            + SO(i + 1, j, KW) * Q(i + 1, j)
                            + SO(i, j, KS) * Q(i, j - 1)
                            + SO(i, j + 1, KS) * Q(i, j + 1)
                            - SO(i, j, KO) * Q(i, j);
    }
#undef SO
#undef Q
#undef QF
#undef RES
}

/* ------------------------------------------------------------------------
 * BMG2_SymStd_restrict.f90:73-94 (non-periodic)
 * ---------------------------------------------------------------------- */
int orc2_restrict(const double *q, double *qc, const double *ci,
                  int Nx, int Ny, int Nxc, int Nyc, int jpn)
{
#define CI(i, j, kk) A3(ci, Nxc, Nyc, i, j, kk)
#define Q(i, j) A2(q, Nx, i, j)
#define QC(i, j) A2(qc, Nxc, i, j)
    const int ipn = jpn < 0 ? -jpn : jpn;
    if (ipn != 0 && ipn != 4) {
        /* BMG2_SymStd_restrict.f90:96-111, periodic: the ghost layer of the FINE vector is refreshed first (the
         * argument is mutated, as in the reference) -- y rows, then x columns, each only when that direction is
         * periodic and the coarse grid is the "even" coarsening N/2+1 of the fine one -- then the same stencil. */
        double *qm = (double *)q;
#define QM(i, j) A2(qm, Nx, i, j)
        if ((ipn == 1 || ipn == 3) && Ny / 2 + 1 == Nyc)
            for (int i = 1; i <= Nx; i++) { QM(i, 1) = QM(i, Ny - 1); QM(i, Ny) = QM(i, 2); }
        if ((ipn == 2 || ipn == 3) && Nx / 2 + 1 == Nxc)
            for (int j = 1; j <= Ny; j++) { QM(1, j) = QM(Nx - 1, j); QM(Nx, j) = QM(2, j); }
#undef QM
    }
#pragma omp parallel for schedule(static)
    for (int jc = 2; jc <= Nyc - 1; jc++) {
        int j = 2 * (jc - 1);
        for (int ic = 2; ic <= Nxc - 1; ic++) {
            int i = 2 * (ic - 1);
            QC(ic, jc) = CI(ic, jc, LNE) * Q(i - 1, j - 1)
                         + CI(ic, jc, LA) * Q(i, j - 1)
                         + CI(ic + 1, jc, LNW) * Q(i + 1, j - 1)
                         + CI(ic, jc, LR) * Q(i - 1, j)
                         + Q(i, j)
                         + CI(ic + 1, jc, LL) * Q(i + 1, j)
                         + CI(ic, jc + 1, LSE) * Q(i - 1, j + 1)
                         + CI(ic, jc + 1, LB) * Q(i, j + 1)
                         + CI(ic + 1, jc + 1, LSW) * Q(i + 1, j + 1);
        }
    }
    return 0;
#undef CI
#undef Q
#undef QC
}

/* ------------------------------------------------------------------------
 * BMG2_SymStd_interp_add.f90:88-141 (non-periodic).  RES is scaled in place
 * by 1/SO(.,.,KO) (l.101-105) -- a persistent side effect.
 * ---------------------------------------------------------------------- */
int orc2_interp_add(double *q, const double *qc, double *res, const double *so, const double *ci,
                    int IIC, int JJC, int IIF, int JJF, int nstncl, int jpn)
{
    (void)nstncl;
    int JJF1 = JJF - 1, IIF1 = IIF - 1;
    int IICF = (IIF - 2) / 2 + 3, JJCF = (JJF - 2) / 2 + 3;
    int IICF1 = IICF - 1, JJCF1 = JJCF - 1;
#define CI(i, j, kk) A3(ci, IIC, JJC, i, j, kk)
#define Q(i, j) A2(q, IIF, i, j)
#define QC(i, j) A2(qc, IIC, i, j)
#define RES(i, j) A2(res, IIF, i, j)
#define SO(i, j, kk) A3(so, IIF, JJF, i, j, kk)
#pragma omp parallel for schedule(static)
    for (int j = 2; j <= JJF1; j++)
        for (int i = 2; i <= IIF1; i++)
            RES(i, j) = RES(i, j) / SO(i, j, KO);

    {
        int j = 2, i = 2;
        Q(2, j) = Q(2, j) + QC(2, 2);
        for (int ic = 3; ic <= IICF1; ic++) {
            i = i + 2;
            Q(i, j) = Q(i, j) + QC(ic, 2);
            double a = CI(ic, 2, LR) * QC(ic, 2) + CI(ic, 2, LL) * QC(ic - 1, 2);
            Q(i - 1, j) = Q(i - 1, j) + a + RES(i - 1, j);
        }
    }
#pragma omp parallel for schedule(static)
    for (int jc = 3; jc <= JJCF1; jc++) {
        int j = 2 * (jc - 1);
        int i = 2;
        Q(2, j) = Q(2, j) + QC(2, jc);
        double aq = CI(2, jc, LA) * QC(2, jc) + CI(2, jc, LB) * QC(2, jc - 1);
        Q(2, j - 1) = Q(2, j - 1) + aq + RES(2, j - 1);
        for (int ic = 3; ic <= IICF1; ic++) {
            i = i + 2;
            Q(i, j) = Q(i, j) + QC(ic, jc);
            double a = CI(ic, jc, LR) * QC(ic, jc) + CI(ic, jc, LL) * QC(ic - 1, jc);
            Q(i - 1, j) = Q(i - 1, j) + a + RES(i - 1, j);
            aq = CI(ic, jc, LA) * QC(ic, jc) + CI(ic, jc, LB) * QC(ic, jc - 1);
            Q(i, j - 1) = Q(i, j - 1) + aq + RES(i, j - 1);
            a = CI(ic, jc, LSW) * QC(ic - 1, jc - 1) + CI(ic, jc, LNW) * QC(ic - 1, jc)
                + CI(ic, jc, LNE) * QC(ic, jc) + CI(ic, jc, LSE) * QC(ic, jc - 1);
            Q(i - 1, j - 1) = Q(i - 1, j - 1) + a + RES(i - 1, j - 1);
        }
    }
    {
        /* BMG2_SymStd_interp_add.f90:143-165, periodic: the ghost layer of the corrected fine vector is refreshed
         * at the end, y rows first, then x columns */
        const int ipn = jpn < 0 ? -jpn : jpn;
        if (ipn == 1 || ipn == 3)
            for (int i = 1; i <= IIF; i++) { Q(i, 1) = Q(i, JJF1); Q(i, JJF) = Q(i, 2); }
        if (ipn == 2 || ipn == 3)
            for (int j = 1; j <= JJF; j++) { Q(1, j) = Q(IIF1, j); Q(IIF, j) = Q(2, j); }
    }
    return 0;
#undef CI
#undef Q
#undef QC
#undef RES
#undef SO
}

/* ------------------------------------------------------------------------
 * BMG2_SymStd_SETUP_interp_OI.f90:84-256 (non-periodic)
 * ---------------------------------------------------------------------- */
static inline double oi_min2(double a, double b) { return a < b ? a : b; }
static inline double oi_max2(double a, double b) { return a > b ? a : b; }

/* ------------------------------------------------------------------------
 * BMG2_SymStd_SETUP_interp_OI.f90:258-618, periodic boundary conditions (|jpn| = 1: y, 2: x, 3: both).
 * Same weights as the non-periodic loops; what changes is which fine lines feed which coarse lines: the fine index
 * of coarse index IC is produced by the reference's running update I = MAX(MOD(I+2,IIFC), MIN(I+2,3)) starting from
 * IBEG (l.296-317: periodic directions start one coarse line earlier and may run one further), and the edge weights
 * of the first and last coarse lines (CI(.,1,.), CI(.,JJC,.), CI(1,.,.), CI(IIC,.,.)) are computed from the wrapped
 * fine lines (l.343-373, 394-425).  Loop structure and statement order are the reference's.
 * ---------------------------------------------------------------------- */
#define OI_SO(i, j, kk) A3(so, IIF, JJF, i, j, kk)
static inline void oi_edge_x(const double *so, int IIF, int JJF, int i, int j, int nine, double *lr, double *ll)
{
    const double ZEPS = DBL_EPSILON;
    double a, b;
    if (nine) {
        a = OI_SO(i, j, KW) + OI_SO(i, j, KNW) + OI_SO(i, j + 1, KSW);
        b = OI_SO(i - 1, j, KW) + OI_SO(i - 1, j, KSW) + OI_SO(i - 1, j + 1, KNW);
    } else {
        a = OI_SO(i, j, KW);
        b = OI_SO(i - 1, j, KW);
    }
    double ep = oi_min2(fabs(a / OI_SO(i - 1, j, KO)), fabs(b / OI_SO(i - 1, j, KO)));
    double sum = a + b + OI_SO(i - 1, j, KS) + OI_SO(i - 1, j + 1, KS);
    sum = a + b + (OI_SO(i - 1, j, KO) - sum) * oi_max2(OI_SO(i - 1, j, KO) - (1.0 + ep) * sum, 0.0)
                      / (fabs(OI_SO(i - 1, j, KO) - (1.0 + ep) * sum) + ZEPS);
    sum = 1.0 / sum;
    *lr = a * sum;
    *ll = b * sum;
}
static inline void oi_edge_y(const double *so, int IIF, int JJF, int i, int j, int nine, double *la, double *lb)
{
    const double ZEPS = DBL_EPSILON;
    double a, b;
    if (nine) {
        a = OI_SO(i, j, KS) + OI_SO(i, j, KNW) + OI_SO(i + 1, j, KSW);
        b = OI_SO(i, j - 1, KS) + OI_SO(i, j - 1, KSW) + OI_SO(i + 1, j - 1, KNW);
    } else {
        a = OI_SO(i, j, KS);
        b = OI_SO(i, j - 1, KS);
    }
    /* the nine-point branch forms SUM before EP, the five-point branch EP before SUM (l.383-384 / 533-534): no
     * dependence between the two statements */
    double sum = a + b + OI_SO(i, j - 1, KW) + OI_SO(i + 1, j - 1, KW);
    double ep = oi_min2(fabs(a / OI_SO(i, j - 1, KO)), fabs(b / OI_SO(i, j - 1, KO)));
    sum = a + b + (OI_SO(i, j - 1, KO) - sum) * oi_max2(OI_SO(i, j - 1, KO) - (1.0 + ep) * sum, 0.0)
                      / (fabs(OI_SO(i, j - 1, KO) - (1.0 + ep) * sum) + ZEPS);
    sum = 1.0 / sum;
    *la = a * sum;
    *lb = b * sum;
}
static inline int oi_next(int i, int nfc) { int t = (i + 2) % nfc, m = i + 2 < 3 ? i + 2 : 3; return t > m ? t : m; }

static int oi_periodic(const double *so, double *ci, int IIF, int JJF, int IIC, int JJC, int ifd, int ipn)
{
    const double ZEPS = DBL_EPSILON;
    const int nine = ifd != 1;
    const int IIC1 = IIC - 1, JJC1 = JJC - 1, IIF1 = IIF - 1, IIF2 = IIF - 2, JJF1 = JJF - 1, JJF2 = JJF - 2;
    const int IICF = (IIF - 2) / 2 + 3, JJCF = (JJF - 2) / 2 + 3, IICF1 = IICF - 1, JJCF1 = JJCF - 1;
    const int IIFC = 2 * (IICF - 2) + 2, JJFC = 2 * (JJCF - 2) + 2;
    const int perx = (ipn == 2 || ipn == 3), pery = (ipn == 1 || ipn == 3);
    int IBEG = 2, IBEGC = 3, IBEG_x = 2, IEND_x = IIF2, IENDC = IICF1, JENDC = JJCF1;
    if (perx) {
        IBEG = 0; IBEGC = 2;
        if (IIC == IICF) { IENDC = IICF; IBEG_x = 3; IEND_x = IIF1; }
    }
    int JBEG = 2, JBEGC = 3, JBEG_y = 2, JEND_y = JJF2;
    if (pery) {
        JBEG = 0; JBEGC = 2;
        if (JJC == JJCF) { JENDC = JJCF; JBEG_y = 3; JEND_y = JJF1; }
    }
#define CI(i, j, kk) A3(ci, IIC, JJC, i, j, kk)
    int i, j;
    /* x-edge weights LR, LL (l.322-340 / 475-492) */
    j = 0;
    for (int jc = 2; jc <= JJC1; jc++) {
        j += 2;
        i = IBEG;
        for (int ic = IBEGC; ic <= IENDC; ic++) {
            i = oi_next(i, IIFC);
            oi_edge_x(so, IIF, JJF, i, j, nine, &CI(ic, jc, LR), &CI(ic, jc, LL));
        }
    }
    if (pery) {   /* l.343-373 / 495-524: the last and first coarse lines from the wrapped fine lines */
        i = IBEG; j = JBEG_y;
        for (int ic = IBEGC; ic <= IENDC; ic++) {
            i = oi_next(i, IIFC);
            oi_edge_x(so, IIF, JJF, i, j, nine, &CI(ic, JJC, LR), &CI(ic, JJC, LL));
        }
        i = IBEG; j = JEND_y;
        for (int ic = IBEGC; ic <= IENDC; ic++) {
            i = oi_next(i, IIFC);
            oi_edge_x(so, IIF, JJF, i, j, nine, &CI(ic, 1, LR), &CI(ic, 1, LL));
        }
    }
    /* y-edge weights LA, LB (l.375-392 / 527-544) */
    j = JBEG;
    for (int jc = JBEGC; jc <= JENDC; jc++) {
        j = oi_next(j, JJFC);
        i = 0;
        for (int ic = 2; ic <= IIC1; ic++) {
            i += 2;
            oi_edge_y(so, IIF, JJF, i, j, nine, &CI(ic, jc, LA), &CI(ic, jc, LB));
        }
    }
    if (perx) {   /* l.394-425 / 546-577 */
        i = IBEG_x; j = JBEG;
        for (int jc = JBEGC; jc <= JENDC; jc++) {
            j = oi_next(j, JJFC);
            oi_edge_y(so, IIF, JJF, i, j, nine, &CI(IIC, jc, LA), &CI(IIC, jc, LB));
        }
        i = IEND_x; j = JBEG;
        for (int jc = JBEGC; jc <= JENDC; jc++) {
            j = oi_next(j, JJFC);
            oi_edge_y(so, IIF, JJF, i, j, nine, &CI(1, jc, LA), &CI(1, jc, LB));
        }
    }
    /* cell-centre weights (l.428-464 / 579-606) */
    j = JBEG;
    for (int jc = JBEGC; jc <= JENDC; jc++) {
        j = oi_next(j, JJFC);
        i = IBEG;
        for (int ic = IBEGC; ic <= IENDC; ic++) {
            i = oi_next(i, IIFC);
            double sum, ep, o = OI_SO(i - 1, j - 1, KO);
            if (nine) {
                sum = OI_SO(i - 1, j - 1, KW) + OI_SO(i - 1, j, KNW) + OI_SO(i - 1, j, KS) + OI_SO(i, j, KSW)
                      + OI_SO(i, j - 1, KW) + OI_SO(i, j - 1, KNW) + OI_SO(i - 1, j - 1, KS) + OI_SO(i - 1, j - 1, KSW);
                ep = oi_min2(oi_min2(fabs((OI_SO(i - 1, j - 1, KSW) + OI_SO(i - 1, j - 1, KW) + OI_SO(i - 1, j, KNW)) / o),
                                     fabs((OI_SO(i - 1, j, KNW) + OI_SO(i - 1, j, KS) + OI_SO(i, j, KSW)) / o)),
                             oi_min2(fabs((OI_SO(i, j, KSW) + OI_SO(i, j - 1, KW) + OI_SO(i, j - 1, KNW)) / o),
                                     fabs((OI_SO(i, j - 1, KNW) + OI_SO(i - 1, j - 1, KS) + OI_SO(i - 1, j - 1, KSW)) / o)));
            } else {
                sum = OI_SO(i - 1, j - 1, KW) + OI_SO(i - 1, j, KS) + OI_SO(i, j - 1, KW) + OI_SO(i - 1, j - 1, KS);
                ep = oi_min2(oi_min2(fabs(OI_SO(i - 1, j - 1, KW) / o), fabs(OI_SO(i - 1, j, KS) / o)),
                             oi_min2(fabs(OI_SO(i, j - 1, KW) / o), fabs(OI_SO(i - 1, j - 1, KS) / o)));
            }
            sum = sum + (o - sum) * oi_max2(o - (1.0 + ep) * sum, 0.0) / (fabs(o - (1.0 + ep) * sum) + ZEPS);
            double s = 1.0 / sum;
            if (nine) {
                CI(ic, jc, LSW) = (OI_SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LL) + OI_SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LB)
                                   + OI_SO(i - 1, j - 1, KSW)) * s;
                CI(ic, jc, LSE) = (OI_SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR) + OI_SO(i, j - 1, KW) * CI(ic, jc, LB)
                                   + OI_SO(i, j - 1, KNW)) * s;
                CI(ic, jc, LNW) = (OI_SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LA) + OI_SO(i - 1, j, KS) * CI(ic, jc, LL)
                                   + OI_SO(i - 1, j, KNW)) * s;
                CI(ic, jc, LNE) = (OI_SO(i - 1, j, KS) * CI(ic, jc, LR) + OI_SO(i, j - 1, KW) * CI(ic, jc, LA)
                                   + OI_SO(i, j, KSW)) * s;
            } else {
                CI(ic, jc, LSW) = (OI_SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LL) + OI_SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LB)) * s;
                CI(ic, jc, LSE) = (OI_SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR) + OI_SO(i, j - 1, KW) * CI(ic, jc, LB)) * s;
                CI(ic, jc, LNW) = (OI_SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LA) + OI_SO(i - 1, j, KS) * CI(ic, jc, LL)) * s;
                CI(ic, jc, LNE) = (OI_SO(i - 1, j, KS) * CI(ic, jc, LR) + OI_SO(i, j - 1, KW) * CI(ic, jc, LA)) * s;
            }
        }
    }
    return 0;
#undef CI
}
#undef OI_SO

static int oi_stage_mask = 3;   /* bit 0: edge weights, bit 1: cell-centre weights */
int orc2_setup_interp_oi(const double *so, const double *soc, double *ci,
                         int IIF, int JJF, int IIC, int JJC, int ifd, int nstncl,
                         int jpn, int irelax)
{
    (void)soc; (void)nstncl; (void)irelax;
    if (jpn != 0 && jpn != -4) {
        const int ipn = jpn < 0 ? -jpn : jpn;
        if (ipn < 1 || ipn > 3) return -1;
        return oi_periodic(so, ci, IIF, JJF, IIC, JJC, ifd, ipn);
    }
    const double ZEPS = DBL_EPSILON; /* EPSILON(1.D0), l.78 */
    int IIC1 = IIC - 1, JJC1 = JJC - 1;
    int IICF = (IIF - 2) / 2 + 3, JJCF = (JJF - 2) / 2 + 3;
    int IICF1 = IICF - 1, JJCF1 = JJCF - 1;
#define SO(i, j, kk) A3(so, IIF, JJF, i, j, kk)
#define CI(i, j, kk) A3(ci, IIC, JJC, i, j, kk)
    if (ifd != 1) {
#pragma omp parallel for schedule(static)
        for (int jc = 2; jc <= ((oi_stage_mask & 1) ? JJC1 : 0); jc++) {
            int j = 2 * (jc - 1);
            for (int ic = 3; ic <= IICF1; ic++) {
                int i = 2 * (ic - 1);
                double a = SO(i, j, KW) + SO(i, j, KNW) + SO(i, j + 1, KSW);
                double b = SO(i - 1, j, KW) + SO(i - 1, j, KSW) + SO(i - 1, j + 1, KNW);
                double ep = oi_min2(fabs(a / SO(i - 1, j, KO)), fabs(b / SO(i - 1, j, KO)));
                double sum = a + b + SO(i - 1, j, KS) + SO(i - 1, j + 1, KS);
                sum = a + b + (SO(i - 1, j, KO) - sum)
                                  * oi_max2(SO(i - 1, j, KO) - (1.0 + ep) * sum, 0.0)
                                  / (fabs(SO(i - 1, j, KO) - (1.0 + ep) * sum) + ZEPS);
                sum = 1.0 / sum;
                CI(ic, jc, LR) = a * sum;
                CI(ic, jc, LL) = b * sum;
            }
        }
#pragma omp parallel for schedule(static)
        for (int jc = 3; jc <= ((oi_stage_mask & 1) ? JJCF1 : 0); jc++) {
            int j = 2 * (jc - 1);
            for (int ic = 2; ic <= IIC1; ic++) {
                int i = 2 * (ic - 1);
                double a = SO(i, j, KS) + SO(i, j, KNW) + SO(i + 1, j, KSW);
                double b = SO(i, j - 1, KS) + SO(i, j - 1, KSW) + SO(i + 1, j - 1, KNW);
                double ep = oi_min2(fabs(a / SO(i, j - 1, KO)), fabs(b / SO(i, j - 1, KO)));
                double sum = a + b + SO(i, j - 1, KW) + SO(i + 1, j - 1, KW);
                sum = a + b + (SO(i, j - 1, KO) - sum)
                                  * oi_max2(SO(i, j - 1, KO) - (1.0 + ep) * sum, 0.0)
                                  / (fabs(SO(i, j - 1, KO) - (1.0 + ep) * sum) + ZEPS);
                sum = 1.0 / sum;
                CI(ic, jc, LA) = a * sum;
                CI(ic, jc, LB) = b * sum;
            }
        }
#pragma omp parallel for schedule(static)
        for (int jc = 3; jc <= ((oi_stage_mask & 2) ? JJCF1 : 0); jc++) {
            int j = 2 * (jc - 1);
            for (int ic = 3; ic <= IICF1; ic++) {
                int i = 2 * (ic - 1);
                double sum = SO(i - 1, j - 1, KW) + SO(i - 1, j, KNW) + SO(i - 1, j, KS)
                             + SO(i, j, KSW) + SO(i, j - 1, KW) + SO(i, j - 1, KNW)
                             + SO(i - 1, j - 1, KS) + SO(i - 1, j - 1, KSW);
                double o = SO(i - 1, j - 1, KO);
                double ep = oi_min2(
                    oi_min2(fabs((SO(i - 1, j - 1, KSW) + SO(i - 1, j - 1, KW) + SO(i - 1, j, KNW)) / o),
                            fabs((SO(i - 1, j, KNW) + SO(i - 1, j, KS) + SO(i, j, KSW)) / o)),
                    oi_min2(fabs((SO(i, j, KSW) + SO(i, j - 1, KW) + SO(i, j - 1, KNW)) / o),
                            fabs((SO(i, j - 1, KNW) + SO(i - 1, j - 1, KS) + SO(i - 1, j - 1, KSW)) / o)));
                sum = sum + (o - sum) * oi_max2(o - (1.0 + ep) * sum, 0.0)
                                / (fabs(o - (1.0 + ep) * sum) + ZEPS);
                double s = 1.0 / sum;
                CI(ic, jc, LSW) = (SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LL)
                                   + SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LB)
                                   + SO(i - 1, j - 1, KSW)) * s;
                CI(ic, jc, LSE) = (SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR)
                                   + SO(i, j - 1, KW) * CI(ic, jc, LB)
                                   + SO(i, j - 1, KNW)) * s;
                CI(ic, jc, LNW) = (SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LA)
                                   + SO(i - 1, j, KS) * CI(ic, jc, LL)
                                   + SO(i - 1, j, KNW)) * s;
                CI(ic, jc, LNE) = (SO(i - 1, j, KS) * CI(ic, jc, LR)
                                   + SO(i, j - 1, KW) * CI(ic, jc, LA)
                                   + SO(i, j, KSW)) * s;
            }
        }
    } else {
#pragma omp parallel for schedule(static)
        for (int jc = 2; jc <= ((oi_stage_mask & 1) ? JJC1 : 0); jc++) {
            int j = 2 * (jc - 1);
            for (int ic = 3; ic <= IICF1; ic++) {
                int i = 2 * (ic - 1);
                double a = SO(i, j, KW);
                double b = SO(i - 1, j, KW);
                double ep = oi_min2(fabs(a / SO(i - 1, j, KO)), fabs(b / SO(i - 1, j, KO)));
                double sum = a + b + SO(i - 1, j, KS) + SO(i - 1, j + 1, KS);
                sum = a + b + (SO(i - 1, j, KO) - sum)
                                  * oi_max2(SO(i - 1, j, KO) - (1.0 + ep) * sum, 0.0)
                                  / (fabs(SO(i - 1, j, KO) - (1.0 + ep) * sum) + ZEPS);
                sum = 1.0 / sum;
                CI(ic, jc, LR) = a * sum;
                CI(ic, jc, LL) = b * sum;
            }
        }
#pragma omp parallel for schedule(static)
        for (int jc = 3; jc <= ((oi_stage_mask & 1) ? JJCF1 : 0); jc++) {
            int j = 2 * (jc - 1);
            for (int ic = 2; ic <= IIC1; ic++) {
                int i = 2 * (ic - 1);
                double a = SO(i, j, KS);
                double b = SO(i, j - 1, KS);
                double ep = oi_min2(fabs(a / SO(i, j - 1, KO)), fabs(b / SO(i, j - 1, KO)));
                double sum = a + b + SO(i, j - 1, KW) + SO(i + 1, j - 1, KW);
                sum = a + b + (SO(i, j - 1, KO) - sum)
                                  * oi_max2(SO(i, j - 1, KO) - (1.0 + ep) * sum, 0.0)
                                  / (fabs(SO(i, j - 1, KO) - (1.0 + ep) * sum) + ZEPS);
                sum = 1.0 / sum;
                CI(ic, jc, LA) = a * sum;
                CI(ic, jc, LB) = b * sum;
            }
        }
#pragma omp parallel for schedule(static)
        for (int jc = 3; jc <= ((oi_stage_mask & 2) ? JJCF1 : 0); jc++) {
            int j = 2 * (jc - 1);
            for (int ic = 3; ic <= IICF1; ic++) {
                int i = 2 * (ic - 1);
                double o = SO(i - 1, j - 1, KO);
                double sum = SO(i - 1, j - 1, KW) + SO(i - 1, j, KS) + SO(i, j - 1, KW)
                             + SO(i - 1, j - 1, KS);
                double ep = oi_min2(oi_min2(fabs(SO(i - 1, j - 1, KW) / o), fabs(SO(i - 1, j, KS) / o)),
                                    oi_min2(fabs(SO(i, j - 1, KW) / o), fabs(SO(i - 1, j - 1, KS) / o)));
                sum = sum + (o - sum) * oi_max2(o - (1.0 + ep) * sum, 0.0)
                                / (fabs(o - (1.0 + ep) * sum) + ZEPS);
                double s = 1.0 / sum;
                CI(ic, jc, LSW) = (SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LL)
                                   + SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LB)) * s;
                CI(ic, jc, LSE) = (SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR)
                                   + SO(i, j - 1, KW) * CI(ic, jc, LB)) * s;
                CI(ic, jc, LNW) = (SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LA)
                                   + SO(i - 1, j, KS) * CI(ic, jc, LL)) * s;
                CI(ic, jc, LNE) = (SO(i - 1, j, KS) * CI(ic, jc, LR)
                                   + SO(i, j - 1, KW) * CI(ic, jc, LA)) * s;
            }
        }
    }
    return 0;
#undef SO
#undef CI
}

/* one stage of the above (1 = edges, 2 = cell centres); the block-decomposed solver
 * exchanges the ghost-line edge weights in between */
int orc2_setup_interp_oi_stage(const double *so, const double *soc, double *ci, int IIF, int JJF, int IIC, int JJC,
                               int ifd, int nstncl, int jpn, int irelax, int stage)
{
    oi_stage_mask = 1 << (stage - 1);
    int rc = orc2_setup_interp_oi(so, soc, ci, IIF, JJF, IIC, JJC, ifd, nstncl, jpn, irelax);
    oi_stage_mask = 3;
    return rc;
}

/* ------------------------------------------------------------------------
 * BMG2_SymStd_SETUP_ITLI_ex.f90:94-213 (9-pt fine) and 252-330 (5-pt fine),
 * Galerkin coarse operator, non-periodic.
 * ---------------------------------------------------------------------- */
int orc2_setup_itli_ex(const double *so, double *soc, const double *ci,
                       int IIF, int JJF, int IIC, int JJC, int ifd, int nstncl, int ipn)
{
    (void)nstncl;
    ipn = ipn < 0 ? -ipn : ipn;
    int IIC1 = IIC - 1, JJC1 = JJC - 1;
#define SO(i, j, kk) A3(so, IIF, JJF, i, j, kk)
#define CI(i, j, kk) A3(ci, IIC, JJC, i, j, kk)
#define SOC(i, j, kk) A3(soc, IIC, JJC, i, j, kk)
    if (ifd != 1) {
#pragma omp parallel for schedule(static)
        for (int jc = 2; jc <= JJC1; jc++) {
            int j = 2 * (jc - 1);
            for (int ic = 2; ic <= IIC1; ic++) {
                int i = 2 * (ic - 1);
                double co = SO(i, j + 1, KNW) * CI(ic, jc + 1, LSW)
                            + SO(i, j, KW) * CI(ic, jc, LL)
                            + SO(i, j, KSW) * CI(ic, jc, LNW);
                double cs = SO(i, j - 1, KW) * CI(ic, jc, LNW) + SO(i, j, KNW) * CI(ic, jc, LL);
                double csw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LNW)
                             + SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LA) + SO(i - 1, j, KNW)
                             + SO(i - 1, j, KS) * CI(ic, jc, LL);
                double cw = -SO(i - 1, j, KO) * CI(ic, jc, LL) + SO(i - 1, j, KS) * CI(ic, jc, LNW)
                            + SO(i - 1, j, KSW) * CI(ic - 1, jc, LA)
                            + SO(i - 1, j, KW) + SO(i - 1, j + 1, KNW) * CI(ic - 1, jc + 1, LB)
                            + SO(i - 1, j + 1, KS) * CI(ic, jc + 1, LSW);
                double cnw = -SO(i - 1, j + 1, KO) * CI(ic, jc + 1, LSW)
                             + SO(i - 1, j + 1, KS) * CI(ic, jc, LL) + SO(i - 1, j + 1, KSW)
                             + SO(i - 1, j + 1, KW) * CI(ic - 1, jc + 1, LB);
                double cn = SO(i, j + 1, KSW) * CI(ic, jc, LL) + SO(i, j + 1, KW) * CI(ic, jc + 1, LSW);
                SOC(ic, jc, KW) = co + CI(ic, jc, LA) * cs + CI(ic, jc, LNE) * csw
                                  + CI(ic, jc, LR) * cw + CI(ic, jc + 1, LSE) * cnw
                                  + CI(ic, jc + 1, LB) * cn;
                double cosw = SO(i, j, KSW) * CI(ic, jc, LSW);
                double cssw = SO(i, j - 1, KSW) * CI(ic, jc - 1, LL) + SO(i, j - 1, KW) * CI(ic, jc, LSW);
                double cswsw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LSW)
                               + SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LL) + SO(i - 1, j - 1, KSW)
                               + SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LB);
                double cwsw = SO(i - 1, j, KS) * CI(ic, jc, LSW) + SO(i - 1, j, KSW) * CI(ic - 1, jc, LB);
                SOC(ic, jc, KSW) = cosw + CI(ic, jc, LA) * cssw + CI(ic, jc, LNE) * cswsw
                                   + CI(ic, jc, LR) * cwsw;
                double coa = SO(i, j, KSW) * CI(ic, jc, LSE) + SO(i, j, KS) * CI(ic, jc, LB)
                             + SO(i + 1, j, KNW) * CI(ic + 1, jc, LSW);
                double cea = SO(i + 1, j, KS) * CI(ic + 1, jc, LSW) + SO(i + 1, j, KSW) * CI(ic, jc, LB);
                double csea = -SO(i + 1, j - 1, KO) * CI(ic + 1, jc, LSW)
                              + SO(i + 1, j - 1, KS) * CI(ic + 1, jc - 1, LL) + SO(i + 1, j - 1, KSW)
                              + SO(i + 1, j - 1, KW) * CI(ic, jc, LB);
                double csa = -SO(i, j - 1, KO) * CI(ic, jc, LB) + SO(i + 1, j - 1, KW) * CI(ic + 1, jc, LSW)
                             + SO(i + 1, j - 1, KNW) * CI(ic + 1, jc - 1, LL)
                             + SO(i, j - 1, KS) + SO(i, j - 1, KSW) * CI(ic, jc - 1, LR)
                             + SO(i, j - 1, KW) * CI(ic, jc, LSE);
                double cswa = -SO(i - 1, j - 1, KO) * CI(ic, jc, LSE) + SO(i, j - 1, KW) * CI(ic, jc, LB)
                              + SO(i, j - 1, KNW) + SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR);
                double cwa = SO(i, j, KNW) * CI(ic, jc, LB) + SO(i - 1, j, KS) * CI(ic, jc, LSE);
                SOC(ic, jc, KS) = coa + CI(ic + 1, jc, LL) * cea + CI(ic + 1, jc, LNW) * csea
                                  + CI(ic, jc, LA) * csa + CI(ic, jc, LNE) * cswa
                                  + CI(ic, jc, LR) * cwa;
                double conw = SO(i - 1, j, KNW) * CI(ic, jc, LSE);
                double cenw = SO(i, j, KNW) * CI(ic, jc, LB) + SO(i - 1, j, KS) * CI(ic, jc, LSE);
                double csenw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LSE) + SO(i, j - 1, KW) * CI(ic, jc, LB)
                               + SO(i, j - 1, KNW) + SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR);
                double cfnw = SO(i - 1, j - 1, KW) * CI(ic, jc, LSE)
                              + SO(i - 1, j - 1, KNW) * CI(ic, jc - 1, LR);
                SOC(ic, jc, KNW) = conw + CI(ic, jc, LL) * cenw + CI(ic, jc, LNW) * csenw
                                   + CI(ic - 1, jc, LA) * cfnw;
            }
        }
#pragma omp parallel for schedule(static)
        for (int jc = 2; jc <= JJC1; jc++) {
            int j = 2 * (jc - 1);
            for (int ic = 2; ic <= IIC1; ic++) {
                int i = 2 * (ic - 1);
                double co = SO(i, j, KW) * CI(ic, jc, LR) + SO(i, j + 1, KNW) * CI(ic, jc + 1, LSE)
                            + SO(i, j + 1, KS) * CI(ic, jc + 1, LB)
                            + SO(i + 1, j + 1, KSW) * CI(ic + 1, jc + 1, LSW)
                            + SO(i + 1, j, KW) * CI(ic + 1, jc, LL)
                            + SO(i + 1, j, KNW) * CI(ic + 1, jc, LNW)
                            + SO(i, j, KS) * CI(ic, jc, LA) + SO(i, j, KSW) * CI(ic, jc, LNE)
                            - SO(i, j, KO);
                double cw = -SO(i - 1, j, KO) * CI(ic, jc, LR) + SO(i - 1, j + 1, KS) * CI(ic, jc + 1, LSE)
                            + SO(i, j + 1, KSW) * CI(ic, jc + 1, LB)
                            + SO(i, j, KW) + SO(i, j, KNW) * CI(ic, jc, LA)
                            + SO(i - 1, j, KS) * CI(ic, jc, LNE);
                double cnw = -SO(i - 1, j + 1, KO) * CI(ic, jc + 1, LSE) + SO(i, j + 1, KW) * CI(ic, jc + 1, LB)
                             + SO(i, j + 1, KNW) + SO(i - 1, j + 1, KS) * CI(ic, jc, LR);
                double cn = -SO(i, j + 1, KO) * CI(ic, jc + 1, LB) + SO(i + 1, j + 1, KNW) * CI(ic + 1, jc, LL)
                            + SO(i, j + 1, KS) + SO(i, j + 1, KSW) * CI(ic, jc, LR)
                            + SO(i, j + 1, KW) * CI(ic, jc + 1, LSE)
                            + SO(i + 1, j + 1, KW) * CI(ic + 1, jc + 1, LSW);
                double cne = -SO(i + 1, j + 1, KO) * CI(ic + 1, jc + 1, LSW)
                             + SO(i + 1, j + 1, KS) * CI(ic + 1, jc, LL) + SO(i + 1, j + 1, KSW)
                             + SO(i + 1, j + 1, KW) * CI(ic, jc + 1, LB);
                double ce = -SO(i + 1, j, KO) * CI(ic + 1, jc, LL) + SO(i + 1, j, KS) * CI(ic + 1, jc, LNW)
                            + SO(i + 1, j, KSW) * CI(ic, jc, LA)
                            + SO(i + 1, j, KW) + SO(i + 1, j + 1, KNW) * CI(ic, jc + 1, LB)
                            + SO(i + 1, j + 1, KS) * CI(ic + 1, jc + 1, LSW);
                double cse = -SO(i + 1, j - 1, KO) * CI(ic + 1, jc, LNW) + SO(i + 1, j - 1, KW) * CI(ic, jc, LA)
                             + SO(i + 1, j, KNW) + SO(i + 1, j, KS) * CI(ic + 1, jc, LL);
                double cs = -SO(i, j - 1, KO) * CI(ic, jc, LA) + SO(i, j - 1, KW) * CI(ic, jc, LNE)
                            + SO(i, j, KNW) * CI(ic, jc, LR)
                            + SO(i, j, KS) + SO(i + 1, j, KSW) * CI(ic + 1, jc, LL)
                            + SO(i + 1, j - 1, KW) * CI(ic + 1, jc, LNW);
                double csw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LNE) + SO(i - 1, j, KS) * CI(ic, jc, LR)
                             + SO(i, j, KSW) + SO(i, j - 1, KW) * CI(ic, jc, LA);
                SOC(ic, jc, KO) = -CI(ic, jc + 1, LSE) * cnw - CI(ic, jc + 1, LB) * cn
                                  - CI(ic + 1, jc + 1, LSW) * cne - CI(ic, jc, LR) * cw - co
                                  - CI(ic + 1, jc, LL) * ce - CI(ic, jc, LNE) * csw
                                  - CI(ic, jc, LA) * cs - CI(ic + 1, jc, LNW) * cse;
            }
        }
    } else {
#pragma omp parallel for schedule(static)
        for (int jc = 2; jc <= JJC1; jc++) {
            int j = 2 * (jc - 1);
            for (int ic = 2; ic <= IIC1; ic++) {
                int i = 2 * (ic - 1);
                double co = SO(i, j, KW) * CI(ic, jc, LL);
                double cs = SO(i, j - 1, KW) * CI(ic, jc, LNW);
                double csw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LNW)
                             + SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LA) + SO(i - 1, j, KS) * CI(ic, jc, LL);
                double cw = -SO(i - 1, j, KO) * CI(ic, jc, LL) + SO(i - 1, j, KS) * CI(ic, jc, LNW)
                            + SO(i - 1, j, KW) + SO(i - 1, j + 1, KS) * CI(ic, jc + 1, LSW);
                double cnw = -SO(i - 1, j + 1, KO) * CI(ic, jc + 1, LSW) + SO(i - 1, j + 1, KS) * CI(ic, jc, LL)
                             + SO(i - 1, j + 1, KW) * CI(ic - 1, jc + 1, LB);
                double cn = SO(i, j + 1, KW) * CI(ic, jc + 1, LSW);
                SOC(ic, jc, KW) = co + CI(ic, jc, LA) * cs + CI(ic, jc, LNE) * csw
                                  + CI(ic, jc, LR) * cw + CI(ic, jc + 1, LSE) * cnw
                                  + CI(ic, jc + 1, LB) * cn;
                double cssw = SO(i, j - 1, KW) * CI(ic, jc, LSW);
                double cswsw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LSW)
                               + SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LL)
                               + SO(i - 1, j - 1, KW) * CI(ic - 1, jc, LB);
                double cwsw = SO(i - 1, j, KS) * CI(ic, jc, LSW);
                SOC(ic, jc, KSW) = CI(ic, jc, LA) * cssw + CI(ic, jc, LNE) * cswsw + CI(ic, jc, LR) * cwsw;
                double coa = SO(i, j, KS) * CI(ic, jc, LB);
                double cea = SO(i + 1, j, KS) * CI(ic + 1, jc, LSW);
                double csea = -SO(i + 1, j - 1, KO) * CI(ic + 1, jc, LSW)
                              + SO(i + 1, j - 1, KS) * CI(ic + 1, jc - 1, LL)
                              + SO(i + 1, j - 1, KW) * CI(ic, jc, LB);
                double csa = -SO(i, j - 1, KO) * CI(ic, jc, LB) + SO(i + 1, j - 1, KW) * CI(ic + 1, jc, LSW)
                             + SO(i, j - 1, KS) + SO(i, j - 1, KW) * CI(ic, jc, LSE);
                double cswa = -SO(i - 1, j - 1, KO) * CI(ic, jc, LSE) + SO(i, j - 1, KW) * CI(ic, jc, LB)
                              + SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR);
                double cwa = SO(i - 1, j, KS) * CI(ic, jc, LSE);
                SOC(ic, jc, KS) = coa + CI(ic + 1, jc, LL) * cea + CI(ic + 1, jc, LNW) * csea
                                  + CI(ic, jc, LA) * csa + CI(ic, jc, LNE) * cswa
                                  + CI(ic, jc, LR) * cwa;
                double cenw = SO(i - 1, j, KS) * CI(ic, jc, LSE);
                double csenw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LSE) + SO(i, j - 1, KW) * CI(ic, jc, LB)
                               + SO(i - 1, j - 1, KS) * CI(ic, jc - 1, LR);
                double csnw = SO(i - 1, j - 1, KW) * CI(ic, jc, LSE);
                SOC(ic, jc, KNW) = CI(ic, jc, LL) * cenw + CI(ic, jc, LNW) * csenw
                                   + CI(ic - 1, jc, LA) * csnw;
                co = SO(i, j, KW) * CI(ic, jc, LR) + SO(i, j + 1, KS) * CI(ic, jc + 1, LB)
                     + SO(i + 1, j, KW) * CI(ic + 1, jc, LL)
                     + SO(i, j, KS) * CI(ic, jc, LA) - SO(i, j, KO);
                cw = -SO(i - 1, j, KO) * CI(ic, jc, LR) + SO(i - 1, j + 1, KS) * CI(ic, jc + 1, LSE)
                     + SO(i, j, KW) + SO(i - 1, j, KS) * CI(ic, jc, LNE);
                cnw = -SO(i - 1, j + 1, KO) * CI(ic, jc + 1, LSE) + SO(i, j + 1, KW) * CI(ic, jc + 1, LB)
                      + SO(i - 1, j + 1, KS) * CI(ic, jc, LR);
                cn = -SO(i, j + 1, KO) * CI(ic, jc + 1, LB) + SO(i, j + 1, KS)
                     + SO(i, j + 1, KW) * CI(ic, jc + 1, LSE)
                     + SO(i + 1, j + 1, KW) * CI(ic + 1, jc + 1, LSW);
                double cne = -SO(i + 1, j + 1, KO) * CI(ic + 1, jc + 1, LSW)
                             + SO(i + 1, j + 1, KS) * CI(ic + 1, jc, LL)
                             + SO(i + 1, j + 1, KW) * CI(ic, jc + 1, LB);
                double ce = -SO(i + 1, j, KO) * CI(ic + 1, jc, LL) + SO(i + 1, j, KS) * CI(ic + 1, jc, LNW)
                            + SO(i + 1, j, KW) + SO(i + 1, j + 1, KS) * CI(ic + 1, jc + 1, LSW);
                double cse = -SO(i + 1, j - 1, KO) * CI(ic + 1, jc, LNW) + SO(i + 1, j - 1, KW) * CI(ic, jc, LA)
                             + SO(i + 1, j, KS) * CI(ic + 1, jc, LL);
                cs = -SO(i, j - 1, KO) * CI(ic, jc, LA) + SO(i, j - 1, KW) * CI(ic, jc, LNE)
                     + SO(i, j, KS) + SO(i + 1, j - 1, KW) * CI(ic + 1, jc, LNW);
                csw = -SO(i - 1, j - 1, KO) * CI(ic, jc, LNE) + SO(i - 1, j, KS) * CI(ic, jc, LR)
                      + SO(i, j - 1, KW) * CI(ic, jc, LA);
                SOC(ic, jc, KO) = -CI(ic, jc + 1, LSE) * cnw - CI(ic, jc + 1, LB) * cn
                                  - CI(ic + 1, jc + 1, LSW) * cne - CI(ic, jc, LR) * cw - co
                                  - CI(ic + 1, jc, LL) * ce - CI(ic, jc, LNE) * csw
                                  - CI(ic, jc, LA) * cs - CI(ic + 1, jc, LNW) * cse;
            }
        }
    }
    if (ipn != 0 && ipn != 4) {
        /* BMG2_SymStd_SETUP_ITLI_ex.f90:216-247 and 333-364, periodic: the ghost layer of every plane of the coarse
         * operator is refreshed -- y rows first (row JJC <- row 2, row 1 <- row JJC-1), then x columns */
        if (ipn == 1 || ipn == 3)
            for (int kk = KO; kk <= KNW; kk++)
                for (int ic = 1; ic <= IIC; ic++) { SOC(ic, JJC, kk) = SOC(ic, 2, kk); SOC(ic, 1, kk) = SOC(ic, JJC1, kk); }
        if (ipn == 2 || ipn == 3)
            for (int kk = KO; kk <= KNW; kk++)
                for (int jc = 1; jc <= JJC; jc++) { SOC(IIC, jc, kk) = SOC(2, jc, kk); SOC(1, jc, kk) = SOC(IIC1, jc, kk); }
    }
    return 0;
#undef SO
#undef CI
#undef SOC
}

/* ------------------------------------------------------------------------
 * BMG2_SymStd_SETUP_lines_x.f90:68-87 / SETUP_lines_y.f90:69-88
 * SOR(.,.,1)=D, SOR(.,.,2)=sub-diagonal factor; y-lines stored transposed.
 * ---------------------------------------------------------------------- */
int orc2_setup_lines_x(const double *so, double *sor, int Nx, int Ny, int nstncl, int jpn)
{
    (void)nstncl;
    const int ipn = jpn < 0 ? -jpn : jpn;
    const int per_x = ipn == 2 || ipn == 3;   /* SETUP_lines_x.f90:76-83: periodic in x */
#define SO(i, j, kk) A3(so, Nx, Ny, i, j, kk)
#define SOR(i, j, kk) A3(sor, Nx, Ny, i, j, kk)
#pragma omp parallel for schedule(static)
    for (int j = 2; j <= Ny - 1; j++) {
        for (int i = 2; i <= Nx - 1; i++) {
            SOR(i, j, 2) = -SO(i, j, KW);
            SOR(i, j, 1) = SO(i, j, KO);
        }
        if (per_x) {
            SOR(2, j, 1) = SOR(2, j, 1) + SO(2, j, KW);
            SOR(Nx - 1, j, 1) = SOR(Nx - 1, j, 1) + SO(Nx, j, KW);
        }
        orc_dpttrf(Nx - 2, &SOR(2, j, 1), &SOR(3, j, 2));
    }
    return 0;
#undef SO
#undef SOR
}

int orc2_setup_lines_y(const double *so, double *sor, int Nx, int Ny, int nstncl, int jpn)
{
    (void)nstncl;
    const int ipn = jpn < 0 ? -jpn : jpn;
    const int per_y = ipn == 1 || ipn == 3;   /* SETUP_lines_y.f90:77-84: periodic in y */
#define SO(i, j, kk) A3(so, Nx, Ny, i, j, kk)
#define SOR(j, i, kk) A3(sor, Ny, Nx, j, i, kk)
#pragma omp parallel for schedule(static)
    for (int i = 2; i <= Nx - 1; i++) {
        for (int j = 2; j <= Ny - 1; j++) {
            SOR(j, i, 2) = -SO(i, j, KS);
            SOR(j, i, 1) = SO(i, j, KO);
        }
        if (per_y) {
            SOR(2, i, 1) = SOR(2, i, 1) + SO(i, 2, KS);
            SOR(Ny - 1, i, 1) = SOR(Ny - 1, i, 1) + SO(i, Ny, KS);
        }
        orc_dpttrf(Ny - 2, &SOR(2, i, 1), &SOR(3, i, 2));
    }
    return 0;
#undef SO
#undef SOR
}

/* ghost refresh of a periodic direction (relax_lines_x.f90:279-290 and the like) */
static void lines_wrap_y(double *q, int II, int JJ)
{
    for (int i = 1; i <= II; i++) { A2(q, II, i, 1) = A2(q, II, i, JJ - 1); A2(q, II, i, JJ) = A2(q, II, i, 2); }
}
static void lines_wrap_x(double *q, int II, int JJ)
{
    for (int j = 1; j <= JJ; j++) { A2(q, II, 1, j) = A2(q, II, II - 1, j); A2(q, II, II, j) = A2(q, II, 2, j); }
}

/* ------------------------------------------------------------------------
 * BMG2_SymStd_relax_lines_x.f90:68-310.  Lines not periodic in x (boundary codes definite and periodic in y only,
 * l.73-168): right-hand side, DPTTRS; with y periodic the ghost rows are refreshed ONCE, after both colours
 * (l.162-167).  Otherwise (l.170-306): the factor is that of the tridiagonal part with the two wrap couplings added
 * to the end diagonals (SETUP_lines_x), and every line is corrected by Sherman-Morrison: a second solve with the wrap
 * couplings as right-hand side (l.209-224), ghosts refreshed after every colour.
 * ---------------------------------------------------------------------- */
int orc2_relax_lines_x(int k, const double *so, const double *qf, double *q, const double *sor,
                       double *b, int II, int JJ, int kf, int ifd, int nstncl,
                       int irelax_sym, int updown, int jpn)
{
    (void)nstncl; (void)b;
    const int ipn = jpn < 0 ? -jpn : jpn;
    int J1 = JJ - 1, I1 = II - 1;
    int jb0, jb1, jbs;
    if (irelax_sym == 0 || (updown == 0 && irelax_sym == 1)) {
        jb0 = 3; jb1 = 2; jbs = -1;
    } else {
        jb0 = 2; jb1 = 3; jbs = 1;
    }
    const int nine = (k < kf || ifd != 1);
    /* l.73-75: the first branch is taken for IPN = 0 and PER_y; IPN is compared with BMG_BCs_indef_nonper = -4, which an
     * absolute value never equals, so the code -4 runs the second branch (whose correction vanishes when the ghost
     * couplings are zero) without any ghost refresh */
    const int per_line = !(ipn == 0 || ipn == 1);
    const int wrap_x = ipn == 2 || ipn == 3, wrap_y = ipn == 1 || ipn == 3;
#define SO(i, j, kk) A3(so, II, JJ, i, j, kk)
#define Q(i, j) A2(q, II, i, j)
#define QF(i, j) A2(qf, II, i, j)
#define SOR(i, j, kk) A3(sor, II, JJ, i, j, kk)
    for (int jbeg = jb0; jbeg != jb1 + jbs; jbeg += jbs) {
        int jend = 2 * ((J1 - jbeg) / 2) + jbeg;
#pragma omp parallel
        {
            double *B = per_line ? (double *)__builtin_malloc(sizeof(double) * (size_t)(II + 2)) : 0;
#pragma omp for schedule(static)
            for (int j = jbeg; j <= jend; j += 2) {
                if (nine) {
                    for (int i = 2; i <= I1; i++)
                        Q(i, j) = QF(i, j) + SO(i, j, KS) * Q(i, j - 1) + SO(i, j + 1, KS) * Q(i, j + 1)
                                  + SO(i, j, KSW) * Q(i - 1, j - 1) + SO(i + 1, j, KNW) * Q(i + 1, j - 1)
                                  + SO(i, j + 1, KNW) * Q(i - 1, j + 1)
                                  + SO(i + 1, j + 1, KSW) * Q(i + 1, j + 1);
                } else {
                    for (int i = 2; i <= I1; i++)
                        Q(i, j) = QF(i, j) + SO(i, j, KS) * Q(i, j - 1) + SO(i, j + 1, KS) * Q(i, j + 1);
                }
                orc_dptts2(I1 - 1, &SOR(2, j, 1), &SOR(3, j, 2), &Q(2, j));
                if (per_line) {
                    for (int i = 2; i <= I1; i++) B[i] = 0.0;
                    B[2] = -SO(2, j, KW);
                    B[I1] = -SO(II, j, KW);
                    orc_dptts2(I1 - 1, &SOR(2, j, 1), &SOR(3, j, 2), &B[2]);
                    double alpha = B[2] + B[I1];
                    double beta = Q(2, j) + Q(I1, j);
                    beta = beta / (1.0 + alpha);
                    for (int i = 2; i <= I1; i++) Q(i, j) = Q(i, j) - beta * B[i];
                }
            }
            if (B) __builtin_free(B);
        }
        if (per_line) {
            if (wrap_y) lines_wrap_y(q, II, JJ);
            if (wrap_x) lines_wrap_x(q, II, JJ);
        }
    }
    if (!per_line && wrap_y) lines_wrap_y(q, II, JJ);
    return 0;
#undef SO
#undef Q
#undef QF
#undef SOR
}

/* ------------------------------------------------------------------------
 * BMG2_SymStd_relax_lines_y.f90:68-300; SOR is (JJ,II,2).  Mirror image of the x version: lines not periodic in y
 * (definite, periodic in x only: l.73-176, x ghosts refreshed once at the end), otherwise the Sherman-Morrison
 * correction (l.178-296).
 * ---------------------------------------------------------------------- */
int orc2_relax_lines_y(int k, const double *so, const double *qf, double *q, const double *sor,
                       double *bwork, int II, int JJ, int kf, int ifd, int nstncl,
                       int irelax_sym, int updown, int jpn)
{
    (void)nstncl; (void)bwork;
    const int ipn = jpn < 0 ? -jpn : jpn;
    int J1 = JJ - 1, I1 = II - 1;
    int ib0, ib1, ibs;
    if (irelax_sym == 0 || (updown == 0 && irelax_sym == 1)) {
        ib0 = 3; ib1 = 2; ibs = -1;
    } else {
        ib0 = 2; ib1 = 3; ibs = 1;
    }
    const int nine = (k < kf || ifd != 1);
    const int per_line = !(ipn == 0 || ipn == 2);     /* l.73-75, same quirk for the code -4 as in the x version */
    const int wrap_x = ipn == 2 || ipn == 3, wrap_y = ipn == 1 || ipn == 3;
#define SO(i, j, kk) A3(so, II, JJ, i, j, kk)
#define Q(i, j) A2(q, II, i, j)
#define QF(i, j) A2(qf, II, i, j)
#define SOR(j, i, kk) A3(sor, JJ, II, j, i, kk)
    for (int ibeg = ib0; ibeg != ib1 + ibs; ibeg += ibs) {
        int iend = 2 * ((I1 - ibeg) / 2) + ibeg;
#pragma omp parallel
        {
            double *B = (double *)__builtin_malloc(sizeof(double) * (size_t)(2 * JJ + 2));
#pragma omp for schedule(static)
            for (int i = ibeg; i <= iend; i += 2) {
                if (nine) {
                    for (int j = 2; j <= J1; j++)
                        B[j] = QF(i, j) + SO(i, j, KW) * Q(i - 1, j) + SO(i + 1, j, KW) * Q(i + 1, j)
                               + SO(i, j, KSW) * Q(i - 1, j - 1) + SO(i + 1, j, KNW) * Q(i + 1, j - 1)
                               + SO(i, j + 1, KNW) * Q(i - 1, j + 1)
                               + SO(i + 1, j + 1, KSW) * Q(i + 1, j + 1);
                } else {
                    for (int j = 2; j <= J1; j++)
                        B[j] = QF(i, j) + SO(i, j, KW) * Q(i - 1, j) + SO(i + 1, j, KW) * Q(i + 1, j);
                }
                orc_dptts2(J1 - 1, &SOR(2, i, 1), &SOR(3, i, 2), &B[2]);
                if (per_line) {
                    for (int j = 2; j <= J1; j++) B[JJ + j] = 0.0;
                    B[JJ + 2] = -SO(i, 2, KS);
                    B[JJ + J1] = -SO(i, JJ, KS);
                    orc_dptts2(J1 - 1, &SOR(2, i, 1), &SOR(3, i, 2), &B[JJ + 2]);
                    double alpha = B[JJ + 2] + B[JJ + J1];
                    double beta = B[2] + B[J1];
                    beta = beta / (1.0 + alpha);
                    for (int j = 2; j <= J1; j++) Q(i, j) = B[j] - beta * B[JJ + j];
                } else {
                    for (int j = 2; j <= J1; j++) Q(i, j) = B[j];
                }
            }
            __builtin_free(B);
        }
        if (per_line) {
            if (wrap_y) lines_wrap_y(q, II, JJ);
            if (wrap_x) lines_wrap_x(q, II, JJ);
        }
    }
    if (!per_line && wrap_x) lines_wrap_x(q, II, JJ);
    return 0;
#undef SO
#undef Q
#undef QF
#undef SOR
}

/* LAPACK DPOTRF('U') as its unblocked kernel DPOTF2: for every column j the diagonal minus DDOT of the column above
 * it, square root, then row j of U right of the diagonal by DGEMV('Transpose') and DSCAL.  (LAPACK's blocked path
 * sums in another order for n > NB; the reference pins no LAPACK, see DESIGN.md "Oracle".)  Returns INFO. */
int orc_dpotf2_upper(int n, double *a, int lda)
{
#define AA(i, j) a[(size_t)((j)-1) * lda + ((i)-1)]
    for (int j = 1; j <= n; j++) {
        double dot = 0.0;
        for (int k = 1; k <= j - 1; k++) dot = dot + AA(k, j) * AA(k, j);
        double ajj = AA(j, j) - dot;
        if (ajj <= 0.0 || ajj != ajj) { AA(j, j) = ajj; return j; }
        ajj = sqrt(ajj);
        AA(j, j) = ajj;
        if (j < n) {
            /* DGEMV('T', j-1, n-j, -1, A(1,j+1), lda, A(1,j), 1, 1, A(j,j+1), lda): y(c) = y(c) + (-1) * (column . x) */
            for (int c = j + 1; c <= n; c++) {
                double temp = 0.0;
                for (int k = 1; k <= j - 1; k++) temp = temp + AA(k, c) * AA(k, j);
                AA(j, c) = AA(j, c) + (-1.0) * temp;
            }
            double r = 1.0 / ajj;
            for (int c = j + 1; c <= n; c++) AA(j, c) = r * AA(j, c);
        }
    }
    return 0;
#undef AA
}

/* LAPACK DPOTRS('U', n, 1): DTRSM('L','U','T','N') then DTRSM('L','U','N','N') in the reference BLAS loop order */
void orc_dpotrs_upper(int n, const double *a, int lda, double *b)
{
#define AA(i, j) a[(size_t)((j)-1) * lda + ((i)-1)]
    for (int i = 1; i <= n; i++) {
        double temp = b[i - 1];
        for (int k = 1; k <= i - 1; k++) temp = temp - AA(k, i) * b[k - 1];
        temp = temp / AA(i, i);
        b[i - 1] = temp;
    }
    for (int k = n; k >= 1; k--) {
        if (b[k - 1] != 0.0) {
            b[k - 1] = b[k - 1] / AA(k, k);
            for (int i = 1; i <= k - 1; i++) b[i - 1] = b[i - 1] - b[k - 1] * AA(i, k);
        }
    }
#undef AA
}

/* BMG2_SymStd_SETUP_cg_LU.f90:144-217 (9-pt) / 258-331 (5-pt): periodic in x, y or both -- the operator of the
 * coarsest grid as a DENSE upper triangle ABD(N,N) (the wrap-around couplings leave the band), DPOTRF('U') */
static int cg_lu_periodic(const double *so, int II, int JJ, int nstncl, double *abd, int nabd1, int nabd2, int ibc)
{
    const int ipn = ibc < 0 ? -ibc : ibc;
    const int perx = (ipn == 2 || ipn == 3), pery = (ipn == 1 || ipn == 3), perxy = ipn == 3;
    const int nine = nstncl == 5;
    int I1 = II - 1, J1 = JJ - 1, I2 = I1 - 1, N = I2 * (J1 - 1);
    if (nabd1 < N || nabd2 < N) return -2;
#define SO(i, j, kq) A3(so, II, JJ, i, j, kq)
#define ABD(i, j) abd[(size_t)((j)-1) * nabd1 + ((i)-1)]
#define NSO(i, j, kq) (nine ? -SO(i, j, kq) : 0.0)   /* the five-point branch stores rZERO where the nine-point one has a corner coupling */
    if (ibc == -3 && II == 4 && JJ == 4 && nine) {
        /* l.121-143: periodic in both directions, four computational nodes */
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
        return orc_dpotf2_upper(N, abd, nabd1);
    }
    int kk = 1, j;
    ABD(1, 1) = SO(2, 2, KO);
    for (int i = 3; i <= I1; i++) {
        kk++;
        ABD(kk, kk) = SO(i, 2, KO);
        ABD(kk - 1, kk) = -SO(i, 2, KW);
    }
    if (perx) ABD(kk - I2 + 1, kk) = -SO(II, 2, KW);
    for (j = 3; j <= J1; j++) {
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
    /* the Fortran DO leaves J = J1 + 1 behind; SO(II,J1,KNW) below is written with J1 explicitly (l.209) */
    if (pery) {
        kk = kk - I2;
        int J2 = (J1 - 2) * I2;
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
    return orc_dpotf2_upper(N, abd, nabd1);
#undef SO
#undef ABD
#undef NSO
}

/* ------------------------------------------------------------------------
 * BMG2_SymStd_SETUP_cg_LU.f90:92-118 (9-pt) / 236-256 (5-pt), definite,
 * non-periodic.  ABD(NABD1,NABD2), NABD1 = II.
 * ---------------------------------------------------------------------- */
int orc2_setup_cg_lu(const double *so, int II, int JJ, int nstncl,
                     double *abd, int nabd1, int nabd2, int ibc)
{
    if (ibc != 0 && ibc != -4) return cg_lu_periodic(so, II, JJ, nstncl, abd, nabd1, nabd2, ibc);
    int I1 = II - 1, J1 = JJ - 1, I2 = I1 - 1;
    int N = I2 * (J1 - 1);
    int kk = 0;
    if (nabd1 < II || nabd2 < N) return -2;
#define SO(i, j, kq) A3(so, II, JJ, i, j, kq)
#define ABD(i, j) abd[(size_t)((j)-1) * nabd1 + ((i)-1)]
    for (int j = 2; j <= J1; j++)
        for (int i = 2; i <= I1; i++) {
            kk++;
            ABD(II, kk) = SO(i, j, KO);
            ABD(I1, kk) = -SO(i, j, KW);
            ABD(3, kk) = nstncl == 5 ? -SO(i + 1, j, KNW) : 0.0;
            ABD(2, kk) = -SO(i, j, KS);
            ABD(1, kk) = nstncl == 5 ? -SO(i, j, KSW) : 0.0;
        }
    if (ibc < 0) ABD(II, kk) = ABD(II, kk) + SO(I1, J1, KO);
    return orc_dpbtf2_upper(N, I1, abd, nabd1);
#undef SO
#undef ABD
}

/* BMG2_SymStd_SOLVE_cg.f90:93-122 (definite, non-periodic) */
int orc2_solve_cg(double *q, const double *qf, int II, int JJ,
                  const double *abd, double *bbd, int nabd1, int nabd2, int jpn)
{
    (void)nabd2;
    const int ipn = jpn < 0 ? -jpn : jpn;
    if (ipn > 4) return -1;
    const int periodic = ipn >= 1 && ipn <= 3;
    int I1 = II - 1, J1 = JJ - 1;
    int kk = 0;
    for (int j = 2; j <= J1; j++)
        for (int i = 2; i <= I1; i++) bbd[kk++] = A2(qf, II, i, j);
    if (!periodic) orc_dpbtrs_upper(kk, I1, abd, nabd1, bbd);
    else orc_dpotrs_upper(kk, abd, nabd1, bbd);          /* SOLVE_cg.f90:104-110 */
    kk = 0;
    for (int j = 2; j <= J1; j++)
        for (int i = 2; i <= I1; i++) A2(q, II, i, j) = bbd[kk++];
    if (jpn != 0) {
        /* SOLVE_cg.f90:125-143: subtract the mean (every boundary code but "definite, non-periodic") */
        double cint = 0.0, qint = 0.0;
        for (int j = 2; j <= J1; j++)
            for (int i = 2; i <= I1; i++) { qint = qint + A2(q, II, i, j); cint = cint + 1; }
        double c = -qint / cint;
        for (int j = 2; j <= J1; j++)
            for (int i = 2; i <= I1; i++) A2(q, II, i, j) = A2(q, II, i, j) + c;
    }
    /* SOLVE_cg.f90:145-166: ghost layer of the solution */
    if (ipn == 1 || ipn == 3)
        for (int i = 2; i <= I1; i++) { A2(q, II, i, JJ) = A2(q, II, i, 2); A2(q, II, i, 1) = A2(q, II, i, J1); }
    if (ipn == 2 || ipn == 3)
        for (int j = 2; j <= J1; j++) { A2(q, II, II, j) = A2(q, II, 2, j); A2(q, II, 1, j) = A2(q, II, I1, j); }
    if (ipn == 3) {
        A2(q, II, 1, 1) = A2(q, II, I1, J1);
        A2(q, II, II, 1) = A2(q, II, 2, J1);
        A2(q, II, 1, JJ) = A2(q, II, I1, 2);
        A2(q, II, II, JJ) = A2(q, II, 2, 2);
    }
    return 0;
}

/* grid_func::lp_norm<2>, include/cedar/2d/grid_func.h:42-53: sequential sum,
 * j outer, i inner, over the interior. */
double orc2_l2norm(const double *q, int II, int JJ)
{
    double result = 0;
    for (int j = 2; j <= JJ - 1; j++)
        for (int i = 2; i <= II - 1; i++) {
            double v = A2(q, II, i, j);
            result += v * v; /* std::pow(v,2) == v*v exactly */
        }
    return sqrt(result); /* std::pow(result, 0.5) is exact-sqrt for glibc */
}

/* One colour pass of orc2_relax_gs (pass = position in the sweep order: 9-point
 * 2*a+b for the a-th row parity and b-th column parity, 5-point the a-th colour):
 * the cut points at which the reference's MPI kernel exchanges ghost values
 * (src/2d/ftn/mpi/BMG2_SymStd_relax_GS.f90:101-134).  Used by the CPU tests of the
 * block-decomposed solver's host logic. */
int orc2_relax_gs_pass(int k, const double *so, const double *qf, double *q, const double *sor,
                       int II, int JJ, int kf, int ifd, int nstncl, int nsorv,
                       int irelax_sym, int updown, int jpn, int pass)
{
    (void)nstncl; (void)nsorv;
    if (jpn != 0 && jpn != -4) return -1;
    const int J1 = JJ - 1, I1 = II - 1;
    const int fwd = irelax_sym == 0 || (updown == 0 && irelax_sym == 1);
#define SO(i, j, kk) A3(so, II, JJ, i, j, kk)
#define Q(i, j) A2(q, II, i, j)
    if (k < kf || ifd != 1) {
        const int a = pass / 2, b = pass % 2;
        const int jbeg = fwd ? 2 + a : 3 - a, ibeg = fwd ? 2 + b : 3 - b;
        for (int j = jbeg; j <= J1; j += 2)
            for (int i = ibeg; i <= I1; i += 2)
                Q(i, j) = (A2(qf, II, i, j) + SO(i, j, KW) * Q(i - 1, j) + SO(i + 1, j, KW) * Q(i + 1, j)
                           + SO(i, j, KS) * Q(i, j - 1) + SO(i, j + 1, KS) * Q(i, j + 1)
                           + SO(i, j, KSW) * Q(i - 1, j - 1) + SO(i + 1, j, KNW) * Q(i + 1, j - 1)
                           + SO(i, j + 1, KNW) * Q(i - 1, j + 1) + SO(i + 1, j + 1, KSW) * Q(i + 1, j + 1))
                          * A3(sor, II, JJ, i, j, MSOR);
    } else {
        if (pass > 1) return 0;
        const int jo = fwd ? 2 + pass : 3 - pass;
        for (int j = 2; j <= J1; j++)
            for (int i = (j + jo) % 2 + 2; i <= I1; i += 2)
                Q(i, j) = (A2(qf, II, i, j) + SO(i, j, KW) * Q(i - 1, j) + SO(i + 1, j, KW) * Q(i + 1, j)
                           + SO(i, j, KS) * Q(i, j - 1) + SO(i, j + 1, KS) * Q(i, j + 1))
                          * A3(sor, II, JJ, i, j, MSOR);
    }
    return 0;
#undef SO
#undef Q
}

/* ------------------------------------------------------------------------
 * Operator application y = A x (the C interface's matvec; test infrastructure like the rest of oracle/).
 * src/2d/ftn/mpi/BMG2_SymStd_UTILS_matvec.f90:88-123 (interior loops; the ghost exchange that follows in the MPI
 * kernel has no serial counterpart).  Term order: centre, W, E, S, N, SW, SE, NW, NE.
 * ---------------------------------------------------------------------- */
void orc2_matvec(int k, const double *so, double *qf, const double *q, int II, int JJ, int kf, int ifd, int nstncl)
{
    (void)nstncl;
    const int nine = (k < kf || ifd != 1);
    const size_t pl = (size_t)II * JJ;
    const double *ko = so, *kw = so + (KW - 1) * pl, *ks = so + (KS - 1) * pl;
    const double *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
#pragma omp parallel for schedule(static)
    for (int j = 2; j <= JJ - 1; j++)
        for (int i = 2; i <= II - 1; i++) {
            const size_t c = (size_t)(j - 1) * II + (i - 1);
            double y = ko[c] * q[c] - kw[c] * q[c - 1] - kw[c + 1] * q[c + 1] - ks[c] * q[c - II] - ks[c + II] * q[c + II];
            if (nine)
                y = y - ksw[c] * q[c - II - 1] - knw[c + 1] * q[c - II + 1] - knw[c + II] * q[c + II - 1]
                    - ksw[c + II + 1] * q[c + II + 1];
            qf[c] = y;
        }
}
