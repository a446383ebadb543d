/*
 * oracle/ml3.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see bmg3.c header).
 *
 * CPU restatement of the reference's serial 3D multilevel driver:
 *   include/cedar/3d/solver.h:54-123   (level count, coarse sizes, ABD)
 *   include/cedar/multilevel.h:95-298  (setup order, smoothers, solve loop)
 *   include/cedar/cycle/vcycle.h:35-115 (V-cycle)
 * Point relaxation (plane relaxation is driven from Python over the 2D oracle).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "oracle.h"

#define MAXLVL 32

typedef struct {
    int nx, ny, nz;
    int nstncl;      /* 4 (seven_pt, level 0 only) or 14 */
    double *A;
    int ownA;
    double *P;       /* CI on THIS (coarse) level, 26 planes */
    double *x, *res, *b;
    double *SOR;     /* 2 planes */
} olevel3;

struct orc3_solver {
    int nlevels, npre, npost;
    olevel3 lv[MAXLVL];
    double *ABD, *bbd;
    int nabd1, nabd2;
};

static double *zalloc(size_t n) { return (double *)calloc(n ? n : 1, sizeof(double)); }

/* 3d/solver.h:54-72 */
static int compute_num_levels3(int nx, int ny, int nz, int min_coarse)
{
    float nxc, nyc, nzc, m;
    int ng = 0;
    do {
        ng++;
        nxc = (float)((unsigned)(nx - 1) / (1u << ng) + 1);
        nyc = (float)((unsigned)(ny - 1) / (1u << ng) + 1);
        nzc = (float)((unsigned)(nz - 1) / (1u << ng) + 1);
        m = nxc < nyc ? nxc : nyc;
        m = m < nzc ? m : nzc;
    } while (m >= (float)min_coarse);
    return ng;
}

orc3_solver *orc3_solver_create(const double *so, int nx, int ny, int nz, int nstncl, int nrelax_pre, int nrelax_post,
                                int num_levels, int min_coarse)
{
    orc3_solver *s = (orc3_solver *)calloc(1, sizeof(*s));
    s->npre = nrelax_pre;
    s->npost = nrelax_post;
    int nl = compute_num_levels3(nx, ny, nz, min_coarse);
    if (num_levels > 0 && num_levels <= nl) nl = num_levels;
    s->nlevels = nl;
    int cx = nx, cy = ny, cz = nz;
    for (int l = 0; l < nl; l++) {
        olevel3 *L = &s->lv[l];
        L->nx = cx; L->ny = cy; L->nz = cz;
        const size_t n = (size_t)(cx + 2) * (cy + 2) * (cz + 2);
        if (l == 0) {
            L->nstncl = nstncl; L->A = (double *)so; L->ownA = 0;
        } else {
            L->nstncl = 14; L->A = zalloc(n * 14); L->ownA = 1;
            L->P = zalloc(n * 26);
        }
        L->x = zalloc(n); L->res = zalloc(n); L->b = zalloc(n);
        L->SOR = zalloc(n * 2);
        cx = (int)((cx - 1) / 2. + 1);
        cy = (int)((cy - 1) / 2. + 1);
        cz = (int)((cz - 1) / 2. + 1);
    }
    olevel3 *C = &s->lv[nl - 1];
    s->nabd1 = C->nx * (C->ny + 1) + 2;      /* 3d/solver.h:118 */
    s->nabd2 = C->nx * C->ny * C->nz;
    s->ABD = zalloc((size_t)s->nabd1 * s->nabd2);
    s->bbd = zalloc((size_t)s->nabd2);
    /* multilevel::setup, multilevel.h:243-265: per level interp, Galerkin, relaxation setup */
    for (int l = 0; l < nl - 1; l++) {
        olevel3 *F = &s->lv[l], *Cc = &s->lv[l + 1];
        const int ifd = F->nstncl == 4 ? 1 : 0;
        orc3_setup_interp_oi(F->A, Cc->A, Cc->P, F->nx + 2, F->ny + 2, F->nz + 2, Cc->nx + 2, Cc->ny + 2, Cc->nz + 2, ifd,
                             F->nstncl, 0, 0, NULL);
        orc3_setup_itli_ex(F->A, Cc->A, Cc->P, F->nx + 2, F->ny + 2, F->nz + 2, Cc->nx + 2, Cc->ny + 2, Cc->nz + 2, ifd,
                           F->nstncl, 0);
        orc3_setup_recip(F->A, F->SOR, F->nx + 2, F->ny + 2, F->nz + 2, F->nstncl, 2);
    }
    orc3_setup_cg_lu(C->A, C->nx + 2, C->ny + 2, C->nz + 2, C->nstncl, s->ABD, s->nabd1, s->nabd2, 0);
    return s;
}

void orc3_solver_destroy(orc3_solver *s)
{
    if (!s) return;
    for (int l = 0; l < s->nlevels; l++) {
        olevel3 *L = &s->lv[l];
        if (L->ownA) free(L->A);
        free(L->P); free(L->x); free(L->res); free(L->b); free(L->SOR);
    }
    free(s->ABD); free(s->bbd);
    free(s);
}

int orc3_solver_nlevels(const orc3_solver *s) { return s->nlevels; }

void orc3_solver_level_shape(const orc3_solver *s, int lvl, int *nx, int *ny, int *nz)
{
    *nx = s->lv[lvl].nx; *ny = s->lv[lvl].ny; *nz = s->lv[lvl].nz;
}

double *orc3_solver_level_ptr(orc3_solver *s, int lvl, int which)
{
    olevel3 *L = &s->lv[lvl];
    switch (which) {
    case 0: return L->A;
    case 1: return L->P;
    case 2: return L->x;
    case 3: return L->res;
    case 4: return L->b;
    case 5: return L->SOR;
    case 7: return s->ABD;
    }
    return NULL;
}

/* cycle/vcycle.h:57-115; kernel arguments as the wrappers pass them
 * (3d/relax.h:74-99, 3d/residual.h:40-60, src/3d/interp.cc:17-45) */
static void ncycle3(orc3_solver *s, int l, double *x, const double *b)
{
    olevel3 *L = &s->lv[l], *C = &s->lv[l + 1];
    const int II = L->nx + 2, JJ = L->ny + 2, KK = L->nz + 2, IIC = C->nx + 2, JJC = C->ny + 2, KKC = C->nz + 2;
    const int ifd = L->nstncl == 4 ? 1 : 0;
    for (int i = 0; i < s->npre; i++)
        orc3_relax_gs(1, L->A, b, x, L->SOR, II, JJ, KK, ifd, L->nstncl, 2, 1, 0, 0);
    orc3_residual(1, 1, ifd, x, b, L->A, L->res, II, JJ, KK, L->nstncl);
    orc3_restrict(L->res, C->b, C->P, II, JJ, KK, IIC, JJC, KKC, 0);
    memset(C->x, 0, sizeof(double) * (size_t)IIC * JJC * KKC);
    if (l + 1 == s->nlevels - 1)
        orc3_solve_cg(C->x, C->b, IIC, JJC, KKC, s->ABD, s->bbd, s->nabd1, s->nabd2, 0);
    else
        ncycle3(s, l + 1, C->x, C->b);
    orc3_interp_add(x, C->x, L->A, L->res, C->P, IIC, JJC, KKC, II, JJ, KK, L->nstncl, 0);
    for (int i = 0; i < s->npost; i++)
        orc3_relax_gs(1, L->A, b, x, L->SOR, II, JJ, KK, ifd, L->nstncl, 2, 1, 1, 0);
}

void orc3_solver_vcycle(orc3_solver *s, double *x, const double *b)
{
    if (s->nlevels == 1) {
        olevel3 *C = &s->lv[0];
        orc3_solve_cg(x, b, C->nx + 2, C->ny + 2, C->nz + 2, s->ABD, s->bbd, s->nabd1, s->nabd2, 0);
    } else {
        ncycle3(s, 0, x, b);
    }
}

/* multilevel::solve, multilevel.h:278-298 */
int orc3_solver_solve(orc3_solver *s, const double *b, double *x, int maxiter, double tol, double *rel_hist,
                      double *res0)
{
    olevel3 *L = &s->lv[0];
    const int II = L->nx + 2, JJ = L->ny + 2, KK = L->nz + 2;
    const int ifd = L->nstncl == 4 ? 1 : 0;
    orc3_residual(1, 1, ifd, x, b, L->A, L->res, II, JJ, KK, L->nstncl);
    const double r0 = orc3_l2norm(L->res, II, JJ, KK);
    if (res0) *res0 = r0;
    int it = 0;
    for (int i = 0; i < maxiter; i++) {
        orc3_solver_vcycle(s, x, b);
        orc3_residual(1, 1, ifd, x, b, L->A, L->res, II, JJ, KK, L->nstncl);
        const double rel = orc3_l2norm(L->res, II, JJ, KK) / r0;
        if (rel_hist) rel_hist[i] = rel;
        it = i + 1;
        if (rel < tol) break;
    }
    return it;
}
