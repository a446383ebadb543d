/*
 * oracle/bmg3.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, fp64, -ffp-contract=off) of the serial 3D BoxMG
 * kernels of the reference (OVGULIU/cedar, src/3d/ftn/BMG3_SymStd_*.f90),
 * non-periodic branches.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library.
 *
 * Point relaxation, residual, restriction, interpolate-and-add and the coarse
 * solve keep the loop and expression order of the Fortran they cite (bitwise
 * parity with the reference's arithmetic; checked against the transliterated
 * reference, oracle/_ref, in tests/test_ref_parity3_cpu.py).  The two large
 * setup kernels (operator-induced interpolation, 800 lines of hand-expanded
 * index lists, and the Galerkin product, 3100 lines) are restated in
 * neighbour-offset form: same terms, different summation order, parity with
 * the reference to rounding (tests: 1e-12 relative to the diagonal).
 *
 * Arrays are Fortran ordered, i fastest: Q(II,JJ,KK), SO(II,JJ,KK,NStncl),
 * SOR(II,JJ,KK,2), CI(IIC,JJC,KKC,26).
 */
#include <math.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

#include "oracle.h"

/* src/3d/ftn/BMG_stencils_f90.h:43-64 */
enum { KP = 1, KPW, KPS, KB, KPSW, KPNW, KBW, KBNW, KBN, KBNE, KBE, KBSE, KBS, KBSW };
enum { LXYL = 1, LXYR, LXYA, LXYB, LXZA, LXZB, LXYNE, LXYSE, LXYSW, LXYNW, LXZSW, LXZNW, LXZNE, LXZSE,
       LYZSW, LYZNW, LYZNE, LYZSE, LBSW, LBNW, LBNE, LBSE, LTSW, LTNW, LTNE, LTSE };
enum { MSOR3 = 2 };

#define X3(n1, n2, i, j, k) (((size_t)((k)-1) * (size_t)(n2) + (size_t)((j)-1)) * (size_t)(n1) + (size_t)((i)-1))

/* One row of the 27-point operator in the reference's term order
 * (BMG3_SymStd_relax_GS.f90:104-131 = residual.f90:74-100): the coupling of
 * (i,j,k) to (i+dx,j+dy,k+dz) is SO(i+sx,j+sy,k+sz,e). */
typedef struct { int e, sx, sy, sz, dx, dy, dz; } nbr_t;
static const nbr_t NBR27[26] = {
    {KPW, 0, 0, 0, -1, 0, 0},  {KPNW, 0, 1, 0, -1, 1, 0},  {KPS, 0, 1, 0, 0, 1, 0},   {KPSW, 1, 1, 0, 1, 1, 0},
    {KPW, 1, 0, 0, 1, 0, 0},   {KPNW, 1, 0, 0, 1, -1, 0},  {KPS, 0, 0, 0, 0, -1, 0},  {KPSW, 0, 0, 0, -1, -1, 0},
    {KB, 0, 0, 0, 0, 0, -1},   {KBW, 0, 0, 0, -1, 0, -1},  {KBNW, 0, 1, 0, -1, 1, -1}, {KBN, 0, 1, 0, 0, 1, -1},
    {KBNE, 1, 1, 0, 1, 1, -1}, {KBE, 1, 0, 0, 1, 0, -1},   {KBSE, 1, 0, 0, 1, -1, -1}, {KBS, 0, 0, 0, 0, -1, -1},
    {KBSW, 0, 0, 0, -1, -1, -1},
    {KB, 0, 0, 1, 0, 0, 1},    {KBE, 0, 0, 1, -1, 0, 1},   {KBSE, 0, 1, 1, -1, 1, 1},  {KBS, 0, 1, 1, 0, 1, 1},
    {KBSW, 1, 1, 1, 1, 1, 1},  {KBW, 1, 0, 1, 1, 0, 1},    {KBNW, 1, 0, 1, 1, -1, 1},  {KBN, 0, 0, 1, 0, -1, 1},
    {KBNE, 0, 0, 1, -1, -1, 1}};
/* 7-point order, relax_GS.f90:162-168 = residual.f90:108-114 */
static const nbr_t NBR7[6] = {{KPW, 0, 0, 0, -1, 0, 0}, {KPS, 0, 1, 0, 0, 1, 0}, {KPW, 1, 0, 0, 1, 0, 0},
                              {KPS, 0, 0, 0, 0, -1, 0}, {KB, 0, 0, 0, 0, 0, -1}, {KB, 0, 0, 1, 0, 0, 1}};

static inline double row_sum(const double *so, const double *q, double qf, int II, int JJ, size_t pl, int i, int j,
                             int k, int seven)
{
    const nbr_t *t = seven ? NBR7 : NBR27;
    const int n = seven ? 6 : 26;
    double a = qf;
    for (int m = 0; m < n; m++)
        a = a + so[(size_t)(t[m].e - 1) * pl + X3(II, JJ, i + t[m].sx, j + t[m].sy, k + t[m].sz)]
                    * q[X3(II, JJ, i + t[m].dx, j + t[m].dy, k + t[m].dz)];
    return a;
}

/* BMG3_SymStd_SETUP_recip.f90 */
void orc3_setup_recip(const double *so, double *sor, int II, int JJ, int KK, int nstncl, int nsorv)
{
    (void)nstncl; (void)nsorv;
    const size_t pl = (size_t)II * JJ * KK;
    for (int k = 2; k <= KK - 1; k++)
        for (int j = 2; j <= JJ - 1; j++)
            for (int i = 2; i <= II - 1; i++)
                sor[(size_t)(MSOR3 - 1) * pl + X3(II, JJ, i, j, k)] = 1.0 / so[X3(II, JJ, i, j, k)];
}

/* BMG3_SymStd_relax_GS.f90:74-177 */
int orc3_relax_gs(int kg, const double *so, const double *qf, double *q, const double *sor, int II, int JJ, int KK,
                  int ifd, int nstncl, int nsorv, int irelax_sym, int updown, int jpn)
{
    (void)kg; (void)nstncl; (void)nsorv;
    if (jpn != 0 && abs(jpn) != 4) return 2;
    const int I1 = II - 1, J1 = JJ - 1, K1 = KK - 1;
    const size_t pl = (size_t)II * JJ * KK;
    const double *rs = sor + (size_t)(MSOR3 - 1) * pl;
    const int up = (updown == 1) || (irelax_sym == 0);
    if (ifd != 1) {
        for (int n = 0; n < 8; n++) {
            const int pts = up ? 1 + n : 8 - n;
            const int k0 = 2 + ((pts - 1) / 4) % 2, j0 = 2 + ((pts - 1) / 2) % 2, i0 = 2 + (pts - 1) % 2;
#pragma omp parallel for collapse(2) schedule(static)
            for (int k = k0; k <= K1; k += 2)
                for (int j = j0; j <= J1; j += 2)
                    for (int i = i0; i <= I1; i += 2) {
                        const size_t c = X3(II, JJ, i, j, k);
                        q[c] = row_sum(so, q, qf[c], II, JJ, pl, i, j, k, 0) * rs[c];
                    }
        }
    } else {
        for (int n = 0; n < 2; n++) {
            const int pts = up ? n : 1 - n;
#pragma omp parallel for collapse(2) schedule(static)
            for (int k = 2; k <= K1; k++)
                for (int j = 2; j <= J1; j++)
                    for (int i = (j + k + pts) % 2 + 2; i <= I1; i += 2) {
                        const size_t c = X3(II, JJ, i, j, k);
                        q[c] = row_sum(so, q, qf[c], II, JJ, pl, i, j, k, 1) * rs[c];
                    }
        }
    }
    return 0;
}

/* BMG3_SymStd_residual.f90:67-120 (27-point unless kg == NOG and ifd == 1) */
void orc3_residual(int kg, int nog, int ifd, const double *q, const double *qf, const double *so, double *res,
                   int II, int JJ, int KK, int nstncl)
{
    (void)nstncl;
    const int seven = !(kg < nog || ifd != 1);
    const size_t pl = (size_t)II * JJ * KK;
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 2; k <= KK - 1; k++)
        for (int j = 2; j <= JJ - 1; j++)
            for (int i = 2; i <= II - 1; i++) {
                const size_t c = X3(II, JJ, i, j, k);
                res[c] = row_sum(so, q, qf[c], II, JJ, pl, i, j, k, seven) - so[c] * q[c];
            }
}

/* BMG3_SymStd_restrict.f90:115-160: one 27-term sum, centre weight 1.  Term m:
 * CI(ic+cx, jc+cy, kc+cz, l) * Q(i+dx, j+dy, k+dz), fine (i,j,k) = 2*(c-1). */
typedef struct { int l, cx, cy, cz, dx, dy, dz; } rst_t;
static const rst_t RST[27] = {
    {LXYNE, 0, 0, 0, -1, -1, 0}, {LXYA, 0, 0, 0, 0, -1, 0},  {LXYNW, 1, 0, 0, 1, -1, 0},
    {LXYR, 0, 0, 0, -1, 0, 0},   {0, 0, 0, 0, 0, 0, 0},      {LXYL, 1, 0, 0, 1, 0, 0},
    {LXYSE, 0, 1, 0, -1, 1, 0},  {LXYB, 0, 1, 0, 0, 1, 0},   {LXYSW, 1, 1, 0, 1, 1, 0},
    {LTNE, 0, 0, 0, -1, -1, -1}, {LYZNW, 0, 0, 0, 0, -1, -1}, {LTNW, 1, 0, 0, 1, -1, -1},
    {LXZNE, 0, 0, 0, -1, 0, -1}, {LXZA, 0, 0, 0, 0, 0, -1},  {LXZNW, 1, 0, 0, 1, 0, -1},
    {LTSE, 0, 1, 0, -1, 1, -1},  {LYZNE, 0, 1, 0, 0, 1, -1}, {LTSW, 1, 1, 0, 1, 1, -1},
    {LBNE, 0, 0, 1, -1, -1, 1},  {LYZSW, 0, 0, 1, 0, -1, 1}, {LBNW, 1, 0, 1, 1, -1, 1},
    {LXZSE, 0, 0, 1, -1, 0, 1},  {LXZB, 0, 0, 1, 0, 0, 1},   {LXZSW, 1, 0, 1, 1, 0, 1},
    {LBSE, 0, 1, 1, -1, 1, 1},   {LYZSE, 0, 1, 1, 0, 1, 1},  {LBSW, 1, 1, 1, 1, 1, 1}};

int orc3_restrict(const double *q, double *qc, const double *ci, int Nx, int Ny, int Nz, int Nxc, int Nyc, int Nzc,
                  int jpn)
{
    (void)Nz;
    if (jpn != 0 && abs(jpn) != 4) return 2;
    const size_t pc = (size_t)Nxc * Nyc * Nzc;
#pragma omp parallel for collapse(2) schedule(static)
    for (int kc = 2; kc <= Nzc - 1; kc++)
        for (int jc = 2; jc <= Nyc - 1; jc++)
            for (int ic = 2; ic <= Nxc - 1; ic++) {
                const int i = 2 * (ic - 1), j = 2 * (jc - 1), k = 2 * (kc - 1);
                double r = 0.0;
                for (int m = 0; m < 27; m++) {
                    const rst_t *t = &RST[m];
                    const double qv = q[X3(Nx, Ny, i + t->dx, j + t->dy, k + t->dz)];
                    const double term = t->l ? ci[(size_t)(t->l - 1) * pc + X3(Nxc, Nyc, ic + t->cx, jc + t->cy, kc + t->cz)] * qv
                                             : qv;
                    r = m == 0 ? term : r + term;
                }
                qc[X3(Nxc, Nyc, ic, jc, kc)] = r;
            }
    return 0;
}

/* BMG3_SymStd_interp_add.f90:80-240, loop structure of the Fortran */
int orc3_interp_add(double *q, const double *qc, const double *so, double *res, const double *ci, int IIC, int JJC,
                    int KKC, int IIF, int JJF, int KKF, int nstncl, int jpn)
{
    (void)nstncl;
    if (jpn != 0 && abs(jpn) != 4) return 2;
    const int iic1 = IIC - 1, kkc1 = KKC - 1;
    const int iicf1 = (IIF - 2) / 2 + 2, jjcf1 = (JJF - 2) / 2 + 2, kkcf1 = (KKF - 2) / 2 + 2;
    const size_t pc = (size_t)IIC * JJC * KKC;
#define Q(i, j, k) q[X3(IIF, JJF, i, j, k)]
#define R(i, j, k) res[X3(IIF, JJF, i, j, k)]
#define QC(i, j, k) qc[X3(IIC, JJC, i, j, k)]
#define CI(i, j, k, l) ci[(size_t)((l)-1) * pc + X3(IIC, JJC, i, j, k)]
    for (int k = 2; k <= KKF - 1; k++)
        for (int j = 2; j <= JJF - 1; j++)
            for (int i = 2; i <= IIF - 1; i++) R(i, j, k) = R(i, j, k) / so[X3(IIF, JJF, i, j, k)];
    /* coarse k-planes (f90:100-150) */
    for (int kc = 2; kc <= kkc1; kc++) {
        const int k = 2 * (kc - 1);
        Q(2, 2, k) = Q(2, 2, k) + QC(2, 2, kc);
        for (int ic = 3; ic <= iicf1; ic++) {
            const int i = 2 * (ic - 1);
            Q(i, 2, k) = Q(i, 2, k) + QC(ic, 2, kc);
            double a = CI(ic, 2, kc, LXYR) * QC(ic, 2, kc) + CI(ic, 2, kc, LXYL) * QC(ic - 1, 2, kc);
            Q(i - 1, 2, k) = Q(i - 1, 2, k) + a + R(i - 1, 2, k);
        }
        for (int jc = 3; jc <= jjcf1; jc++) {
            const int j = 2 * (jc - 1);
            Q(2, j, k) = Q(2, j, k) + QC(2, jc, kc);
            double aq = CI(2, jc, kc, LXYA) * QC(2, jc, kc) + CI(2, jc, kc, LXYB) * QC(2, jc - 1, kc);
            Q(2, j - 1, k) = Q(2, j - 1, k) + aq + R(2, j - 1, k);
            for (int ic = 3; ic <= iicf1; ic++) {
                const int i = 2 * (ic - 1);
                Q(i, j, k) = Q(i, j, k) + QC(ic, jc, kc);
                double a = CI(ic, jc, kc, LXYR) * QC(ic, jc, kc) + CI(ic, jc, kc, LXYL) * QC(ic - 1, jc, kc);
                Q(i - 1, j, k) = Q(i - 1, j, k) + a + R(i - 1, j, k);
                aq = CI(ic, jc, kc, LXYA) * QC(ic, jc, kc) + CI(ic, jc, kc, LXYB) * QC(ic, jc - 1, kc);
                Q(i, j - 1, k) = Q(i, j - 1, k) + aq + R(i, j - 1, k);
                a = CI(ic, jc, kc, LXYSW) * QC(ic - 1, jc - 1, kc) + CI(ic, jc, kc, LXYNW) * QC(ic - 1, jc, kc)
                    + CI(ic, jc, kc, LXYNE) * QC(ic, jc, kc) + CI(ic, jc, kc, LXYSE) * QC(ic, jc - 1, kc);
                Q(i - 1, j - 1, k) = Q(i - 1, j - 1, k) + a + R(i - 1, j - 1, k);
            }
        }
    }
    /* planes between coarse planes, coarse x-columns (f90:156-190) */
    for (int kc = 3; kc <= kkcf1; kc++) {
        const int k = 2 * (kc - 1) - 1;
        for (int ic = 2; ic <= iic1; ic++) {
            const int i = 2 * (ic - 1);
            Q(i, 2, k) = Q(i, 2, k) + CI(ic, 2, kc, LXZA) * QC(ic, 2, kc) + CI(ic, 2, kc, LXZB) * QC(ic, 2, kc - 1)
                         + R(i, 2, k);
        }
        for (int jc = 3; jc <= jjcf1; jc++) {
            const int j = 2 * (jc - 1);
            for (int ic = 2; ic <= iic1; ic++) {
                const int i = 2 * (ic - 1);
                Q(i, j, k) = Q(i, j, k) + CI(ic, jc, kc, LXZA) * QC(ic, jc, kc) + CI(ic, jc, kc, LXZB) * QC(ic, jc, kc - 1)
                             + R(i, j, k);
                Q(i, j - 1, k) = Q(i, j - 1, k) + CI(ic, jc, kc, LYZNW) * QC(ic, jc, kc)
                                 + CI(ic, jc, kc, LYZNE) * QC(ic, jc - 1, kc) + CI(ic, jc, kc, LYZSW) * QC(ic, jc, kc - 1)
                                 + CI(ic, jc, kc, LYZSE) * QC(ic, jc - 1, kc - 1) + R(i, j - 1, k);
            }
        }
    }
    /* planes between coarse planes, fine-only x-columns (f90:192-240) */
    for (int kc = 3; kc <= kkcf1; kc++) {
        const int k = 2 * (kc - 1) - 1;
        for (int jc = 2; jc <= jjcf1; jc++) {
            const int j = 2 * (jc - 1);
            for (int ic = 3; ic <= iicf1; ic++) {
                const int i = 2 * (ic - 1) - 1;
                Q(i, j, k) = Q(i, j, k) + CI(ic, jc, kc, LXZNW) * QC(ic - 1, jc, kc) + CI(ic, jc, kc, LXZNE) * QC(ic, jc, kc)
                             + CI(ic, jc, kc, LXZSW) * QC(ic - 1, jc, kc - 1) + CI(ic, jc, kc, LXZSE) * QC(ic, jc, kc - 1)
                             + R(i, j, k);
                if (jc == 2) continue;
                Q(i, j - 1, k) = Q(i, j - 1, k) + CI(ic, jc, kc, LTNW) * QC(ic - 1, jc, kc) + CI(ic, jc, kc, LTNE) * QC(ic, jc, kc)
                                 + CI(ic, jc, kc, LTSW) * QC(ic - 1, jc - 1, kc) + CI(ic, jc, kc, LTSE) * QC(ic, jc - 1, kc)
                                 + CI(ic, jc, kc, LBNW) * QC(ic - 1, jc, kc - 1) + CI(ic, jc, kc, LBNE) * QC(ic, jc, kc - 1)
                                 + CI(ic, jc, kc, LBSW) * QC(ic - 1, jc - 1, kc - 1)
                                 + CI(ic, jc, kc, LBSE) * QC(ic, jc - 1, kc - 1) + R(i, j - 1, k);
            }
        }
    }
#undef Q
#undef R
#undef QC
#undef CI
    return 0;
}

/* ------------------------------------------------------------------------
 * Neighbour-offset helpers for the two setup kernels.
 * ---------------------------------------------------------------------- */
typedef struct { const double *so; int II, JJ, KK, seven; size_t pl; } op3_t;
typedef struct { double *ci; int IIC, JJC, KKC; size_t pl; } ci3_t;

static int NBR_IDX[3][3][3];
static int nbr_idx_ready = 0;
static void nbr_idx_init(void)
{
    if (nbr_idx_ready) return;
    for (int m = 0; m < 26; m++) NBR_IDX[NBR27[m].dx + 1][NBR27[m].dy + 1][NBR27[m].dz + 1] = m;
    NBR_IDX[1][1][1] = -1;
    nbr_idx_ready = 1;
}

/* entry e of the operator at storage point (i,j,k); 0 outside the array */
static inline double op_at(const op3_t *A, int e, int i, int j, int k)
{
    if (i < 1 || i > A->II || j < 1 || j > A->JJ || k < 1 || k > A->KK) return 0.0;
    return A->so[(size_t)(e - 1) * A->pl + X3(A->II, A->JJ, i, j, k)];
}
/* coupling (stored positive) between p = (i,j,k) and p + (dx,dy,dz) */
static inline double cpl3(const op3_t *A, int i, int j, int k, int dx, int dy, int dz)
{
    if (A->seven && (dx != 0) + (dy != 0) + (dz != 0) != 1) return 0.0;
    const nbr_t *t = &NBR27[NBR_IDX[dx + 1][dy + 1][dz + 1]];
    return op_at(A, t->e, i + t->sx, j + t->sy, k + t->sz);
}

/* Name of the weight with which a fine point of parity type (px,py,pz) (1 = odd
 * F-index = between two coarse points along that axis) takes the coarse point
 * U - (a,b,c), U its upper coarse point (read off BMG3_SymStd_interp_add.f90
 * and restrict.f90).  0 = injection. */
static int wname3(int px, int py, int pz, int a, int b, int c)
{
    static const int XY[4] = {LXYNE, LXYNW, LXYSE, LXYSW};          /* s = a + 2b */
    static const int XZ[4] = {LXZNE, LXZNW, LXZSE, LXZSW};          /* s = a + 2c */
    static const int YZ[4] = {LYZNW, LYZNE, LYZSW, LYZSE};          /* s = b + 2c */
    static const int XYZ[8] = {LTNE, LTNW, LTSE, LTSW, LBNE, LBNW, LBSE, LBSW};   /* s = a + 2b + 4c */
    switch (px + 2 * py + 4 * pz) {
    case 0: return 0;
    case 1: return a ? LXYL : LXYR;
    case 2: return b ? LXYB : LXYA;
    case 3: return XY[a + 2 * b];
    case 4: return c ? LXZB : LXZA;
    case 5: return XZ[a + 2 * c];
    case 6: return YZ[b + 2 * c];
    default: return XYZ[a + 2 * b + 4 * c];
    }
}

/* interpolation weight P(f, c): fine point f = (fi,fj,fk) (F-indices) from the
 * coarse point whose fine F-index is (gi,gj,gk) (all even); 0 if c does not
 * interpolate to f */
static double pw3(const ci3_t *W, int fi, int fj, int fk, int gi, int gj, int gk)
{
    const int px = fi & 1, py = fj & 1, pz = fk & 1;
    if (abs(gi - fi) > px || abs(gj - fj) > py || abs(gk - fk) > pz) return 0.0;
    const int l = wname3(px, py, pz, gi < fi, gj < fj, gk < fk);
    if (l == 0) return 1.0;
    const int ui = (fi + 1) / 2 + 1, uj = (fj + 1) / 2 + 1, uk = (fk + 1) / 2 + 1;
    if (ui < 1 || ui > W->IIC || uj < 1 || uj > W->JJC || uk < 1 || uk > W->KKC) return 0.0;
    return W->ci[(size_t)(l - 1) * W->pl + X3(W->IIC, W->JJC, ui, uj, uk)];
}

static inline double oi_fix(double base, double d, double c, double ep)
{
    const double t = d - (1.0 + ep) * c;
    return base + (d - c) * (t > 0.0 ? t : 0.0) / (fabs(t) + 1.0e-13);   /* eMACH, SETUP_interp_OI.f90:78 */
}
static inline double dmin(double a, double b) { return a < b ? a : b; }

/* BMG3_SymStd_SETUP_interp_OI.f90:82-805 in neighbour-offset form.  Every F-point
 * p is handled by its parity type: the operator is collapsed along the axes on
 * which p sits on a coarse line/plane (even index), and the weights of the 2^m
 * coarse corners are operator induced from the couplings and from the weights of
 * the lower-dimensional neighbours (edges before faces before cell centres). */
int orc3_setup_interp_oi(const double *so, const double *soc, double *ci, int IIF, int JJF, int KKF, int IIC,
                         int JJC, int KKC, int ifd, int nstncl, int irelax, int jpn, double *yo)
{
    (void)soc; (void)nstncl; (void)irelax; (void)yo;
    if (jpn != 0 && abs(jpn) != 4) return 2;
    nbr_idx_init();
    const op3_t A = {so, IIF, JJF, KKF, ifd == 1, (size_t)IIF * JJF * KKF};
    const ci3_t W = {ci, IIC, JJC, KKC, (size_t)IIC * JJC * KKC};
    const int cf1[3] = {(IIF - 2) / 2 + 2, (JJF - 2) / 2 + 2, (KKF - 2) / 2 + 2};
    const int c1[3] = {IIC - 1, JJC - 1, KKC - 1};
    for (int stage = 1; stage <= 3; stage++)
        for (int t = 1; t < 8; t++) {
            const int par[3] = {t & 1, (t >> 1) & 1, (t >> 2) & 1};
            if (par[0] + par[1] + par[2] != stage) continue;
            /* upper coarse point U runs 3..cf1 on odd axes, 2..c1 on even ones */
            int lo[3], hi[3];
            for (int a = 0; a < 3; a++) { lo[a] = par[a] ? 3 : 2; hi[a] = par[a] ? cf1[a] : c1[a]; }
            for (int kc = lo[2]; kc <= hi[2]; kc++)
                for (int jc = lo[1]; jc <= hi[1]; jc++)
                    for (int ic = lo[0]; ic <= hi[0]; ic++) {
                        const int p[3] = {2 * (ic - 1) - par[0], 2 * (jc - 1) - par[1], 2 * (kc - 1) - par[2]};
                        const double d = op_at(&A, KP, p[0], p[1], p[2]);
                        /* collapsed couplings D(o), o in {-1,0,1} on odd axes, summed over even axes */
                        double D[3][3][3];
                        memset(D, 0, sizeof D);
                        double offplane = 0.0, tot = 0.0;
                        for (int dz = -1; dz <= 1; dz++)
                            for (int dy = -1; dy <= 1; dy++)
                                for (int dx = -1; dx <= 1; dx++) {
                                    if (!dx && !dy && !dz) continue;
                                    const double c = cpl3(&A, p[0], p[1], p[2], dx, dy, dz);
                                    const int ox = par[0] ? dx : 0, oy = par[1] ? dy : 0, oz = par[2] ? dz : 0;
                                    if (!ox && !oy && !oz) offplane += c; else D[ox + 1][oy + 1][oz + 1] += c;
                                    tot += c;
                                }
                        /* side sums: all collapsed couplings with o_a = -1 / +1, for every odd axis a */
                        double side[3][2] = {{0}};
                        double dp = 0.0;
                        for (int oz = -1; oz <= 1; oz++)
                            for (int oy = -1; oy <= 1; oy++)
                                for (int ox = -1; ox <= 1; ox++) {
                                    const double v = D[ox + 1][oy + 1][oz + 1];
                                    const int o[3] = {ox, oy, oz};
                                    for (int a = 0; a < 3; a++) if (par[a] && o[a]) side[a][o[a] > 0] += v;
                                    dp += v;
                                }
                        double ep = HUGE_VAL;
                        if (stage == 3 && !A.seven) {
                            /* reference quirks (f90:411-446): the north and bottom sums are not
                             * divided by the diagonal; the south sum counts (+,-,-) twice and
                             * misses (-,-,-) */
                            const double fs = side[1][0] - cpl3(&A, p[0], p[1], p[2], -1, -1, -1)
                                              + cpl3(&A, p[0], p[1], p[2], 1, -1, -1);
                            ep = dmin(dmin(fabs(side[0][0]) / d, fabs(side[0][1]) / d), dmin(fabs(side[1][1]), fabs(fs) / d));
                            ep = dmin(ep, dmin(fabs(side[2][0]), fabs(side[2][1]) / d));
                        } else {
                            for (int a = 0; a < 3; a++)
                                if (par[a]) ep = dmin(ep, dmin(fabs(side[a][0] / d), fabs(side[a][1] / d)));
                        }
                        double den;
                        if (stage == 1) den = oi_fix(dp, d, tot, ep);          /* c = a+b+ring; a+b + (d-c)... */
                        else den = oi_fix(dp, d - offplane, dp, ep);            /* sum = d - off-plane couplings */
                        /* weights of the 2^stage corners */
                        for (int cz = (par[2] ? -1 : 0); cz <= (par[2] ? 1 : 0); cz += 2)
                            for (int cy = (par[1] ? -1 : 0); cy <= (par[1] ? 1 : 0); cy += 2)
                                for (int cx = (par[0] ? -1 : 0); cx <= (par[0] ? 1 : 0); cx += 2) {
                                    const int g[3] = {p[0] + cx, p[1] + cy, p[2] + cz};
                                    double acc = 0.0;
                                    /* neighbours of p inside the box spanned by p and the corner */
                                    for (int oz = 0; oz != cz + (cz >= 0 ? 1 : -1); oz += (cz >= 0 ? 1 : -1))
                                        for (int oy = 0; oy != cy + (cy >= 0 ? 1 : -1); oy += (cy >= 0 ? 1 : -1))
                                            for (int ox = 0; ox != cx + (cx >= 0 ? 1 : -1); ox += (cx >= 0 ? 1 : -1)) {
                                                if (!ox && !oy && !oz) continue;
                                                const double w = (ox == cx && oy == cy && oz == cz)
                                                                     ? 1.0
                                                                     : pw3(&W, p[0] + ox, p[1] + oy, p[2] + oz, g[0], g[1], g[2]);
                                                acc += w * D[ox + 1][oy + 1][oz + 1];
                                            }
                                    const int l = wname3(par[0], par[1], par[2], cx < 0, cy < 0, cz < 0);
                                    W.ci[(size_t)(l - 1) * W.pl + X3(IIC, JJC, ic, jc, kc)] = acc / den;
                                }
                    }
        }
    return 0;
}

/* BMG3_SymStd_SETUP_ITLI07_ex.f90 / ITLI27_ex.f90: Galerkin coarse operator
 * A_c = P^T A P, entry by entry.  SOC(S, e) is the coupling between the coarse
 * points c = S - s and c2 = c + d of the table row that names e (stored positive
 * for off-diagonals): sum over f1 in N(c), f2 in N(c2), |f1-f2|_inf <= 1 of
 * P(f1,c) * (-A)(f1,f2) * P(f2,c2). */
int orc3_setup_itli_ex(const double *so, double *soc, const double *ci, int IIF, int JJF, int KKF, int IIC, int JJC,
                       int KKC, int ifd, int nstncl, int ipn)
{
    (void)nstncl;
    if (ipn != 0 && abs(ipn) != 4) return 2;
    nbr_idx_init();
    const op3_t A = {so, IIF, JJF, KKF, ifd == 1, (size_t)IIF * JJF * KKF};
    const ci3_t W = {(double *)ci, IIC, JJC, KKC, (size_t)IIC * JJC * KKC};
    const size_t pc = W.pl;
    int row_of[15];
    for (int m = 25; m >= 0; m--) row_of[NBR27[m].e] = m;   /* first row naming each entry */
#pragma omp parallel for collapse(2) schedule(dynamic)
    for (int kc = 2; kc <= KKC - 1; kc++)
        for (int jc = 2; jc <= JJC - 1; jc++)
            for (int ic = 2; ic <= IIC - 1; ic++)
                for (int e = 1; e <= 14; e++) {
                    int c[3] = {ic, jc, kc}, c2[3] = {ic, jc, kc};
                    if (e != KP) {
                        const nbr_t *t = &NBR27[row_of[e]];
                        c[0] -= t->sx; c[1] -= t->sy; c[2] -= t->sz;
                        c2[0] = c[0] + t->dx; c2[1] = c[1] + t->dy; c2[2] = c[2] + t->dz;
                    }
                    const int g[3] = {2 * (c[0] - 1), 2 * (c[1] - 1), 2 * (c[2] - 1)};
                    const int g2[3] = {2 * (c2[0] - 1), 2 * (c2[1] - 1), 2 * (c2[2] - 1)};
                    double p2[3][3][3];
                    for (int z = -1; z <= 1; z++)
                        for (int y = -1; y <= 1; y++)
                            for (int x = -1; x <= 1; x++)
                                p2[x + 1][y + 1][z + 1] = pw3(&W, g2[0] + x, g2[1] + y, g2[2] + z, g2[0], g2[1], g2[2]);
                    double acc = 0.0;
                    for (int z = -1; z <= 1; z++)
                        for (int y = -1; y <= 1; y++)
                            for (int x = -1; x <= 1; x++) {
                                const int f[3] = {g[0] + x, g[1] + y, g[2] + z};
                                const double w1 = pw3(&W, f[0], f[1], f[2], g[0], g[1], g[2]);
                                if (w1 == 0.0) continue;
                                double row = 0.0;
                                for (int dz = -1; dz <= 1; dz++)
                                    for (int dy = -1; dy <= 1; dy++)
                                        for (int dx = -1; dx <= 1; dx++) {
                                            const int r[3] = {f[0] + dx - g2[0], f[1] + dy - g2[1], f[2] + dz - g2[2]};
                                            if (abs(r[0]) > 1 || abs(r[1]) > 1 || abs(r[2]) > 1) continue;
                                            const double a = (dx || dy || dz) ? cpl3(&A, f[0], f[1], f[2], dx, dy, dz)
                                                                              : -op_at(&A, KP, f[0], f[1], f[2]);
                                            row += a * p2[r[0] + 1][r[1] + 1][r[2] + 1];
                                        }
                                acc += w1 * row;
                            }
                    soc[(size_t)(e - 1) * pc + X3(IIC, JJC, ic, jc, kc)] = e == KP ? -acc : acc;
                }
    return 0;
}

/* BMG3_SymStd_SETUP_cg_LU.f90:92-193 (definite / indefinite non-periodic):
 * upper banded storage, bandwidth ibw = nx*(ny+1)+1, ABD(NABD1,NABD2) with
 * NABD1 = ibw+1 (include/cedar/3d/solver.h:118), then DPBTRF('U'). */
int orc3_setup_cg_lu(const double *so, int II, int JJ, int KK, int nstncl, double *abd, int nabd1, int nabd2, int ibc)
{
    if (ibc != 0 && ibc != -4) return -1;
    if (nstncl != 14 && nstncl != 4) return -3;
    const int i1 = II - 1, j1 = JJ - 1, k1 = KK - 1, i2 = i1 - 1;
    const int ibw = i2 * j1 + 1;
    const size_t pl = (size_t)II * JJ * KK;
    if (nabd1 < ibw + 1 || nabd2 < i2 * (j1 - 1) * (k1 - 1)) return -2;
#define SO(i, j, k, e) so[(size_t)((e)-1) * pl + X3(II, JJ, i, j, k)]
#define ABD(r, c) abd[(size_t)((c)-1) * nabd1 + ((r)-1)]
    const int full = nstncl == 14;
    int kl = 0;
    for (int k = 2; k <= k1; k++)
        for (int j = 2; j <= j1; j++)
            for (int i = 2; i <= i1; i++) {
                kl++;
                ABD(ibw + 1, kl) = SO(i, j, k, KP);
                ABD(ibw, kl) = -SO(i, j, k, KPW);
                ABD(ibw - i1 + 3, kl) = full ? -SO(i + 1, j, k, KPNW) : 0.0;
                ABD(ibw - i1 + 2, kl) = -SO(i, j, k, KPS);
                ABD(ibw - i1 + 1, kl) = full ? -SO(i, j, k, KPSW) : 0.0;
                ABD(ibw - (j1 - 2) * i2 + 2, kl) = full ? -SO(i + 1, j + 1, k, KBNE) : 0.0;
                ABD(ibw - (j1 - 2) * i2 + 1, kl) = full ? -SO(i, j + 1, k, KBN) : 0.0;
                ABD(ibw - (j1 - 2) * i2, kl) = full ? -SO(i, j + 1, k, KBNW) : 0.0;
                ABD(ibw - (j1 - 1) * i2 + 2, kl) = full ? -SO(i + 1, j, k, KBE) : 0.0;
                ABD(ibw - (j1 - 1) * i2 + 1, kl) = -SO(i, j, k, KB);
                ABD(ibw - (j1 - 1) * i2, kl) = full ? -SO(i, j, k, KBW) : 0.0;
                ABD(3, kl) = full ? -SO(i + 1, j, k, KBSE) : 0.0;
                ABD(2, kl) = full ? -SO(i, j, k, KBS) : 0.0;
                ABD(1, kl) = full ? -SO(i, j, k, KBSW) : 0.0;
            }
    if (ibc < 0) ABD(ibw + 1, kl) = ABD(ibw + 1, kl) + SO(i1, j1, k1, KP);
    return orc_dpbtf2_upper(kl, ibw, abd, nabd1);
#undef SO
#undef ABD
}

/* BMG3_SymStd_SOLVE_cg.f90:100-135 (non-periodic) */
int orc3_solve_cg(double *q, const double *qf, int II, int JJ, int KK, const double *abd, double *bbd, int nabd1,
                  int nabd2, int jpn)
{
    (void)nabd2;
    if (jpn != 0) return -1;
    const int ibw = (II - 2) * (JJ - 1) + 1;
    int kt = 0;
    for (int k = 2; k <= KK - 1; k++)
        for (int j = 2; j <= JJ - 1; j++)
            for (int i = 2; i <= II - 1; i++) bbd[kt++] = qf[X3(II, JJ, i, j, k)];
    orc_dpbtrs_upper(kt, ibw, abd, nabd1, bbd);
    kt = 0;
    for (int k = 2; k <= KK - 1; k++)
        for (int j = 2; j <= JJ - 1; j++)
            for (int i = 2; i <= II - 1; i++) q[X3(II, JJ, i, j, k)] = bbd[kt++];
    return 0;
}

/* grid_func::lp_norm<2> (3d): sequential sum over the interior, k outer */
double orc3_l2norm(const double *q, int II, int JJ, int KK)
{
    double result = 0;
    for (int k = 2; k <= KK - 1; k++)
        for (int j = 2; j <= JJ - 1; j++)
            for (int i = 2; i <= II - 1; i++) {
                const double v = q[X3(II, JJ, i, j, k)];
                result += v * v;
            }
    return sqrt(result);
}

/* ----------------------- gallery: src/3d/gallery.cc ----------------------- */
/* fills a zeroed SO(nx+2,ny+2,nz+2,nstncl); F-indices = C++ index + 1 */
#define G3(e, i, j, k) so[(size_t)((e)-1) * pl + X3(II, JJ, i, j, k)]
void orc3_gallery_poisson(double *so, int nx, int ny, int nz)
{
    /* gallery.cc:7-56 */
    const int II = nx + 2, JJ = ny + 2, KK = nz + 2;
    const size_t pl = (size_t)II * JJ * KK;
    const double hx = 1.0 / (II - 1), hy = 1.0 / (JJ - 1), hz = 1.0 / (KK - 1);
    const double xh = hy * hz / hx, yh = hx * hz / hy, zh = hx * hy / hz;
    for (int k = 2; k <= nz + 1; k++)
        for (int j = 2; j <= ny + 1; j++)
            for (int i = 2; i <= nx + 1; i++) {
                if (j >= 3) G3(KPS, i, j, k) = 1.0 * yh;
                if (i >= 3) G3(KPW, i, j, k) = 1.0 * xh;
                if (k >= 3) G3(KB, i, j, k) = 1.0 * zh;
                G3(KP, i, j, k) = 2.0 * xh + 2.0 * yh + 2.0 * zh;
            }
}

void orc3_gallery_fe(double *so, int nx, int ny, int nz)
{
    /* gallery.cc:111-190 */
    const int II = nx + 2, JJ = ny + 2;
    const size_t pl = (size_t)II * JJ * (nz + 2);
    for (int k = 2; k <= nz + 1; k++)
        for (int j = 2; j <= ny + 1; j++)
            for (int i = 2; i <= nx + 1; i++) {
                const int x = i >= 3, y = j >= 3, z = k >= 3;
                if (x) G3(KPW, i, j, k) = 1.0;
                if (y) G3(KPS, i, j, k) = 1.0;
                if (z) G3(KB, i, j, k) = 1.0;
                if (x && y) { G3(KPNW, i, j, k) = 1.0; G3(KPSW, i, j, k) = 1.0; }
                if (x && z) { G3(KBW, i, j, k) = 1.0; G3(KBE, i, j, k) = 1.0; }
                if (y && z) { G3(KBN, i, j, k) = 1.0; G3(KBS, i, j, k) = 1.0; }
                if (x && y && z) {
                    G3(KBNW, i, j, k) = 1.0; G3(KBNE, i, j, k) = 1.0; G3(KBSE, i, j, k) = 1.0; G3(KBSW, i, j, k) = 1.0;
                }
                G3(KP, i, j, k) = 26;
            }
}
#undef G3

/* examples/basic-3d-ser/poisson.cc:16-38 (rhs), 41-65 (exact solution) */
void orc3_rhs_sin(double *b, int nx, int ny, int nz)
{
    const int II = nx + 2, JJ = ny + 2, KK = nz + 2;
    const double pi = M_PI;
    const double hx = 1.0 / (II - 1), hy = 1.0 / (JJ - 1), hz = 1.0 / (KK - 1);
    const double h2 = hx * hy * hz;
    for (int k = 1; k <= nz; k++)
        for (int j = 1; j <= ny; j++)
            for (int i = 1; i <= nx; i++) {
                const double x = i * hx, y = j * hy, z = k * hz;
                b[X3(II, JJ, i + 1, j + 1, k + 1)] = 12 * (pi * pi) * sin(2 * pi * x) * sin(2 * pi * y) * sin(2 * pi * z) * h2;
            }
}
void orc3_exact_sin(double *q, int nx, int ny, int nz)
{
    const int II = nx + 2, JJ = ny + 2, KK = nz + 2;
    const double pi = M_PI;
    const double hx = 1.0 / (II - 1), hy = 1.0 / (JJ - 1), hz = 1.0 / (KK - 1);
    for (int k = 1; k <= nz; k++)
        for (int j = 1; j <= ny; j++)
            for (int i = 1; i <= nx; i++)
                q[X3(II, JJ, i + 1, j + 1, k + 1)] = sin(2 * pi * (i * hx)) * sin(2 * pi * (j * hy)) * sin(2 * pi * (k * hz));
}

/* Operator application y = A x: src/3d/ftn/mpi/BMG3_SymStd_UTILS_matvec.f90:78-129 (interior loops).
 * Each term is {stencil slot, offset of the stencil entry, offset of the x entry}; the list order is the
 * reference's term order (in-plane neighbours, then the plane below, then the plane above). */
static const signed char MV27[26][7] = {
    {KPW, 0, 0, 0, -1, 0, 0},   {KPNW, 0, 1, 0, -1, 1, 0},  {KPS, 0, 1, 0, 0, 1, 0},    {KPSW, 1, 1, 0, 1, 1, 0},
    {KPW, 1, 0, 0, 1, 0, 0},    {KPNW, 1, 0, 0, 1, -1, 0},  {KPS, 0, 0, 0, 0, -1, 0},   {KPSW, 0, 0, 0, -1, -1, 0},
    {KB, 0, 0, 0, 0, 0, -1},    {KBW, 0, 0, 0, -1, 0, -1},  {KBNW, 0, 1, 0, -1, 1, -1}, {KBN, 0, 1, 0, 0, 1, -1},
    {KBNE, 1, 1, 0, 1, 1, -1},  {KBE, 1, 0, 0, 1, 0, -1},   {KBSE, 1, 0, 0, 1, -1, -1}, {KBS, 0, 0, 0, 0, -1, -1},
    {KBSW, 0, 0, 0, -1, -1, -1}, {KB, 0, 0, 1, 0, 0, 1},    {KBE, 0, 0, 1, -1, 0, 1},   {KBSE, 0, 1, 1, -1, 1, 1},
    {KBS, 0, 1, 1, 0, 1, 1},    {KBSW, 1, 1, 1, 1, 1, 1},   {KBW, 1, 0, 1, 1, 0, 1},    {KBNW, 1, 0, 1, 1, -1, 1},
    {KBN, 0, 0, 1, 0, -1, 1},   {KBNE, 0, 0, 1, -1, -1, 1}};
static const signed char MV7[6][7] = {
    {KPW, 0, 0, 0, -1, 0, 0}, {KPS, 0, 1, 0, 0, 1, 0}, {KPW, 1, 0, 0, 1, 0, 0},
    {KPS, 0, 0, 0, 0, -1, 0}, {KB, 0, 0, 0, 0, 0, -1}, {KB, 0, 0, 1, 0, 0, 1}};

void orc3_matvec(int kg, int nog, int ifd, const double *so, double *qf, const double *q, int II, int JJ, int KK,
                 int nstncl)
{
    (void)nstncl;
    const int seven = !(kg < nog || ifd != 1);
    const int nt = seven ? 6 : 26;
    const signed char (*T)[7] = seven ? MV7 : MV27;
    const size_t pl = (size_t)II * JJ * KK;
    const long sj = II, sk = (long)II * JJ;
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 2; k <= KK - 1; k++)
        for (int j = 2; j <= JJ - 1; j++)
            for (int i = 2; i <= II - 1; i++) {
                const long c = ((long)(k - 1) * JJ + (j - 1)) * II + (i - 1);
                double y = so[c] * q[c];
                for (int t = 0; t < nt; t++)
                    y = y - so[(size_t)(T[t][0] - 1) * pl + c + T[t][1] + T[t][2] * sj + T[t][3] * sk]
                                * q[c + T[t][4] + T[t][5] * sj + T[t][6] * sk];
                qf[c] = y;
            }
}
