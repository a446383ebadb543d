/*
 * oracle/ml2.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see bmg2.c header).
 *
 * CPU restatement of the reference's serial 2D multilevel driver:
 *   include/cedar/2d/solver.h:57-116   (level count, coarse sizes, ABD)
 *   include/cedar/multilevel.h:95-298  (setup order, smoothers, solve loop)
 *   include/cedar/cycle/vcycle.h:35-115 (V-cycle)
 *   src/2d/gallery.cc, examples/basic-2d-ser/poisson.cc (problem generators)
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "oracle.h"

#define MAXLVL 32

typedef struct {
    int nx, ny;      /* interior sizes */
    int nstncl;      /* 3 (five_pt, level 0 only) or 5 */
    double *A;       /* (nx+2)(ny+2)nstncl ; level 0 aliases the user's array */
    int ownA;
    double *P;       /* CI on THIS (coarse) level: (nx+2)(ny+2)*8, levels>=1 */
    double *x, *res, *b;
    double *SOR[2];  /* (nx+2)(ny+2)*2 each */
} olevel;

struct orc2_solver {
    int nlevels;
    int relaxation, npre, npost;
    olevel lv[MAXLVL];
    double *ABD, *bbd;
    int nabd1, nabd2;
    double *linebuf;
    int fcycle;   /* 0: V-cycle (default), 1: F-cycle (cycle/fcycle.h) */
    int ibc;      /* BMG_get_bc(per_mask): 0, 2 (periodic x), 1 (periodic y), 3 (both); every kernel wrapper of the
                   * reference passes it on (include/cedar/2d/relax.h:84-86, src/2d/restrict.cc:27, interp.cc:28,46, ...) */
};

static double *zalloc(size_t n) { return (double *)calloc(n ? n : 1, sizeof(double)); }

/* 2d/solver.h:57-73 */
static int compute_num_levels(int nx, int ny, int min_coarse)
{
    float nxc, nyc;
    int ng = 0;
    do {
        ng++;
        nxc = (float)((unsigned)(nx - 1) / (1u << ng) + 1);
        nyc = (float)((unsigned)(ny - 1) / (1u << ng) + 1);
    } while ((nxc < nyc ? nxc : nyc) >= (float)min_coarse);
    return ng;
}

static void relax_level(orc2_solver *s, int l, double *x, const double *b, int updown)
{
    olevel *L = &s->lv[l];
    int II = L->nx + 2, JJ = L->ny + 2;
    int ifd = L->nstncl == 3 ? 1 : 0;
    /* rbgs::run_impl / lines::run_impl, include/cedar/2d/relax.h:72-102,157-190:
     * k = kf = 1, BMG_RELAX_SYM always. */
    switch (s->relaxation) {
    case ORC_RELAX_POINT:
        orc2_relax_gs(1, L->A, b, x, L->SOR[0], II, JJ, 1, ifd, L->nstncl, 2, 1, updown, s->ibc);
        break;
    case ORC_RELAX_LINE_X:
        orc2_relax_lines_x(1, L->A, b, x, L->SOR[0], s->linebuf, II, JJ, 1, ifd, L->nstncl, 1, updown, s->ibc);
        break;
    case ORC_RELAX_LINE_Y:
        orc2_relax_lines_y(1, L->A, b, x, L->SOR[0], s->linebuf, II, JJ, 1, ifd, L->nstncl, 1, updown, s->ibc);
        break;
    case ORC_RELAX_LINE_XY:
        /* multilevel.h:175-178 (DOWN: x then y), 204-207 (UP: y then x) */
        if (updown == 0) {
            orc2_relax_lines_x(1, L->A, b, x, L->SOR[0], s->linebuf, II, JJ, 1, ifd, L->nstncl, 1, updown, s->ibc);
            orc2_relax_lines_y(1, L->A, b, x, L->SOR[1], s->linebuf, II, JJ, 1, ifd, L->nstncl, 1, updown, s->ibc);
        } else {
            orc2_relax_lines_y(1, L->A, b, x, L->SOR[1], s->linebuf, II, JJ, 1, ifd, L->nstncl, 1, updown, s->ibc);
            orc2_relax_lines_x(1, L->A, b, x, L->SOR[0], s->linebuf, II, JJ, 1, ifd, L->nstncl, 1, updown, s->ibc);
        }
        break;
    }
}

orc2_solver *orc2_solver_create(const double *so, int nx, int ny, int nstncl,
                                int relaxation, int nrelax_pre, int nrelax_post,
                                int num_levels, int min_coarse)
{
    return orc2_solver_create_per(so, nx, ny, nstncl, relaxation, nrelax_pre, nrelax_post, num_levels, min_coarse, 0);
}

/* per_mask: bit 0 = periodic in x, bit 1 = periodic in y (kernel_params::per_mask, include/cedar/kernel_params.h:33-41) */
orc2_solver *orc2_solver_create_per(const double *so, int nx, int ny, int nstncl,
                                    int relaxation, int nrelax_pre, int nrelax_post,
                                    int num_levels, int min_coarse, int per_mask)
{
    static const int bcmap[4] = {0, 2, 1, 3};   /* BMG_get_bc.f90:13-22 */
    orc2_solver *s = (orc2_solver *)calloc(1, sizeof(*s));
    s->ibc = bcmap[per_mask & 3];
    s->relaxation = relaxation;
    s->npre = nrelax_pre;
    s->npost = nrelax_post;
    int nl = compute_num_levels(nx, ny, min_coarse);
    if (num_levels > 0 && num_levels <= nl) nl = num_levels; /* multilevel.h:246-253 */
    s->nlevels = nl;
    /* setup_space, 2d/solver.h:75-116 */
    int cx = nx, cy = ny;
    int maxlen = 0;
    for (int l = 0; l < nl; l++) {
        olevel *L = &s->lv[l];
        L->nx = cx; L->ny = cy;
        size_t n = (size_t)(cx + 2) * (cy + 2);
        if (cx + 2 > maxlen) maxlen = cx + 2;
        if (cy + 2 > maxlen) maxlen = cy + 2;
        if (l == 0) {
            L->nstncl = nstncl; L->A = (double *)so; L->ownA = 0;
        } else {
            L->nstncl = 5; L->A = zalloc(n * 5); L->ownA = 1;
            L->P = zalloc(n * 8);
        }
        L->x = zalloc(n); L->res = zalloc(n); L->b = zalloc(n);
        L->SOR[0] = zalloc(n * 2); L->SOR[1] = zalloc(n * 2);
        cx = (int)((cx - 1) / 2. + 1);
        cy = (int)((cy - 1) / 2. + 1);
    }
    s->linebuf = zalloc((size_t)maxlen + 2);
    olevel *C = &s->lv[nl - 1];
    s->nabd1 = s->ibc ? C->nx * C->ny : C->nx + 2;   /* 2d/solver.h:111-114: dense for periodic */
    s->nabd2 = C->nx * C->ny;
    s->ABD = zalloc((size_t)s->nabd1 * s->nabd2);
    s->bbd = zalloc((size_t)s->nabd2);

    /* multilevel::setup, multilevel.h:243-265 */
    for (int l = 0; l < nl - 1; l++) {
        olevel *F = &s->lv[l], *Cc = &s->lv[l + 1];
        int ifd = F->nstncl == 3 ? 1 : 0;
        orc2_setup_interp_oi(F->A, Cc->A, Cc->P, F->nx + 2, F->ny + 2, Cc->nx + 2, Cc->ny + 2,
                             ifd, F->nstncl, s->ibc, 0);
        orc2_setup_itli_ex(F->A, Cc->A, Cc->P, F->nx + 2, F->ny + 2, Cc->nx + 2, Cc->ny + 2,
                           ifd, F->nstncl, s->ibc);
        switch (relaxation) {
        case ORC_RELAX_POINT:
            orc2_setup_recip(F->A, F->SOR[0], F->nx + 2, F->ny + 2, F->nstncl, 2);
            break;
        case ORC_RELAX_LINE_X:
            orc2_setup_lines_x(F->A, F->SOR[0], F->nx + 2, F->ny + 2, F->nstncl, s->ibc);
            break;
        case ORC_RELAX_LINE_Y:
            orc2_setup_lines_y(F->A, F->SOR[0], F->nx + 2, F->ny + 2, F->nstncl, s->ibc);
            break;
        case ORC_RELAX_LINE_XY:
            orc2_setup_lines_x(F->A, F->SOR[0], F->nx + 2, F->ny + 2, F->nstncl, s->ibc);
            orc2_setup_lines_y(F->A, F->SOR[1], F->nx + 2, F->ny + 2, F->nstncl, s->ibc);
            break;
        }
    }
    orc2_setup_cg_lu(C->A, C->nx + 2, C->ny + 2, C->nstncl, s->ABD, s->nabd1, s->nabd2, s->ibc);
    return s;
}

void orc2_solver_destroy(orc2_solver *s)
{
    if (!s) return;
    for (int l = 0; l < s->nlevels; l++) {
        olevel *L = &s->lv[l];
        if (L->ownA) free(L->A);
        free(L->P); free(L->x); free(L->res); free(L->b);
        free(L->SOR[0]); free(L->SOR[1]);
    }
    free(s->ABD); free(s->bbd); free(s->linebuf);
    free(s);
}

int orc2_solver_nlevels(const orc2_solver *s) { return s->nlevels; }

void orc2_solver_level_shape(const orc2_solver *s, int lvl, int *nx, int *ny)
{
    *nx = s->lv[lvl].nx; *ny = s->lv[lvl].ny;
}

double *orc2_solver_level_ptr(orc2_solver *s, int lvl, int which)
{
    olevel *L = &s->lv[lvl];
    switch (which) {
    case 0: return L->A;
    case 1: return L->P;
    case 2: return L->x;
    case 3: return L->res;
    case 4: return L->b;
    case 5: return L->SOR[0];
    case 6: return L->SOR[1];
    case 7: return s->ABD;
    }
    return NULL;
}

/* cycle/vcycle.h:57-115 */
static void ncycle(orc2_solver *s, int l, double *x, const double *b)
{
    olevel *L = &s->lv[l], *C = &s->lv[l + 1];
    int II = L->nx + 2, JJ = L->ny + 2, IIC = C->nx + 2, JJC = C->ny + 2;
    int ifd = L->nstncl == 3 ? 1 : 0;
    for (int i = 0; i < s->npre; i++) relax_level(s, l, x, b, 0);
    orc2_residual(0, L->A, b, x, L->res, II, JJ, 0, ifd, L->nstncl);
    orc2_restrict(L->res, C->b, C->P, II, JJ, IIC, JJC, s->ibc);
    memset(C->x, 0, sizeof(double) * (size_t)IIC * JJC);
    if (l + 1 == s->nlevels - 1)
        orc2_solve_cg(C->x, C->b, IIC, JJC, s->ABD, s->bbd, s->nabd1, s->nabd2, s->ibc);
    else
        ncycle(s, l + 1, C->x, C->b);
    orc2_interp_add(x, C->x, L->res, L->A, C->P, IIC, JJC, II, JJ, L->nstncl, s->ibc);
    for (int i = 0; i < s->npost; i++) relax_level(s, l, x, b, 1);
}

/* cycle/fcycle.h:49-83: fmg_cycle + fmg_cycle_helper.  The right-hand side itself is restricted down to the coarsest
 * level, which is solved directly; on the way up every level starts from x = 0, takes the interpolated coarse
 * solution (interp_add with a zero residual) and runs one V-cycle from that level. */
static void fmg_cycle(orc2_solver *s, int l, double *x, const double *b)
{
    olevel *L = &s->lv[l];
    int II = L->nx + 2, JJ = L->ny + 2;
    if (l == s->nlevels - 1) {
        orc2_solve_cg(x, b, II, JJ, s->ABD, s->bbd, s->nabd1, s->nabd2, 0);
        return;
    }
    olevel *C = &s->lv[l + 1];
    int IIC = C->nx + 2, JJC = C->ny + 2;
    orc2_restrict(b, C->b, C->P, II, JJ, IIC, JJC, 0);
    fmg_cycle(s, l + 1, C->x, C->b);
    memset(x, 0, sizeof(double) * (size_t)II * JJ);
    memset(L->res, 0, sizeof(double) * (size_t)II * JJ);
    orc2_interp_add(x, C->x, L->res, L->A, C->P, IIC, JJC, II, JJ, L->nstncl, 0);
    ncycle(s, l, x, b);
}

void orc2_solver_set_cycle(orc2_solver *s, int fcycle) { s->fcycle = fcycle; }

void orc2_solver_vcycle(orc2_solver *s, double *x, const double *b)
{
    if (s->fcycle && s->nlevels > 1) { fmg_cycle(s, 0, x, b); return; }
    if (s->nlevels == 1) {
        olevel *C = &s->lv[0];
        orc2_solve_cg(x, b, C->nx + 2, C->ny + 2, s->ABD, s->bbd, s->nabd1, s->nabd2, s->ibc);
    } else {
        ncycle(s, 0, x, b);
    }
}

/* multilevel::solve, multilevel.h:278-298 */
int orc2_solver_solve(orc2_solver *s, const double *b, double *x, int maxiter, double tol,
                      double *rel_hist, double *res0)
{
    olevel *L = &s->lv[0];
    int II = L->nx + 2, JJ = L->ny + 2;
    int ifd = L->nstncl == 3 ? 1 : 0;
    orc2_residual(0, L->A, b, x, L->res, II, JJ, 0, ifd, L->nstncl);
    double res0_l2 = orc2_l2norm(L->res, II, JJ);
    if (res0) *res0 = res0_l2;
    int it = 0;
    for (int i = 0; i < maxiter; i++) {
        orc2_solver_vcycle(s, x, b);
        orc2_residual(0, L->A, b, x, L->res, II, JJ, 0, ifd, L->nstncl);
        double rel = orc2_l2norm(L->res, II, JJ) / res0_l2;
        if (rel_hist) rel_hist[i] = rel;
        it = i + 1;
        if (rel < tol) break;
    }
    return it;
}

/* ----------------------- gallery: src/2d/gallery.cc ----------------------- */
#define S3(so, II, JJ, i, j, k) (so)[((size_t)(k) * (JJ) + (size_t)(j)) * (II) + (size_t)(i)]
/* 0-based C++ indices as in gallery.cc; dirs c=0,w=1,s=2,sw=3,nw=4 */

void orc2_gallery_diag_diffusion(double *so, int nx, int ny, double dx, double dy)
{
    int II = nx + 2, JJ = ny + 2;
    double hx = 1.0 / (II - 1), hy = 1.0 / (JJ - 1);
    double xh = hy / hx, yh = hx / hy;
    for (int j = 2; j < ny + 1; j++)
        for (int i = 1; i < nx + 1; i++) S3(so, II, JJ, i, j, 2) = dy * yh;
    for (int j = 1; j < ny + 1; j++)
        for (int i = 2; i < nx + 1; i++) S3(so, II, JJ, i, j, 1) = dx * xh;
    for (int j = 1; j <= ny; j++)
        for (int i = 1; i <= nx; i++) S3(so, II, JJ, i, j, 0) = 2 * dx * xh + 2 * dy * yh;
}

void orc2_gallery_poisson(double *so, int nx, int ny)
{
    /* gallery.cc:7-39: identical to diag_diffusion(1,1) term by term
     * (1.0*yh, 1.0*xh, 2*xh+2*yh). */
    int II = nx + 2, JJ = ny + 2;
    double hx = 1.0 / (II - 1), hy = 1.0 / (JJ - 1);
    double xh = hy / hx, yh = hx / hy;
    for (int j = 2; j < ny + 1; j++)
        for (int i = 1; i < nx + 1; i++) S3(so, II, JJ, i, j, 2) = 1.0 * yh;
    for (int j = 1; j < ny + 1; j++)
        for (int i = 2; i < nx + 1; i++) S3(so, II, JJ, i, j, 1) = 1.0 * xh;
    for (int j = 1; j <= ny; j++)
        for (int i = 1; i <= nx; i++) S3(so, II, JJ, i, j, 0) = 2 * xh + 2 * yh;
}

void orc2_gallery_fe(double *so, int nx, int ny)
{
    int II = nx + 2, JJ = ny + 2;
    for (int j = 2; j < ny + 1; j++)
        for (int i = 1; i < nx + 1; i++) S3(so, II, JJ, i, j, 2) = 1.0;
    for (int j = 1; j < ny + 1; j++)
        for (int i = 2; i < nx + 1; i++) S3(so, II, JJ, i, j, 1) = 1.0;
    for (int j = 2; j < ny + 1; j++)
        for (int i = 2; i < nx + 1; i++) {
            S3(so, II, JJ, i, j, 3) = 1.0;
            S3(so, II, JJ, i, j, 4) = 1.0;
        }
    for (int j = 1; j <= ny; j++)
        for (int i = 1; i <= nx; i++) S3(so, II, JJ, i, j, 0) = 8.0;
}

void orc2_rhs_sin(double *b, int nx, int ny)
{
    const double pi = M_PI;
    int II = nx + 2, JJ = ny + 2;
    double hx = 1.0 / (II - 1), hy = 1.0 / (JJ - 1);
    double h2 = hx * hy;
    memset(b, 0, sizeof(double) * (size_t)II * JJ);
    for (int j = 1; j <= ny; j++)
        for (int i = 1; i <= nx; i++) {
            double x = i * hx, y = j * hy;
            b[(size_t)j * II + i] = 8 * (pi * pi) * sin(2 * pi * x) * sin(2 * pi * y) * h2;
        }
}

void orc2_exact_sin(double *q, int nx, int ny)
{
    const double pi = M_PI;
    int II = nx + 2, JJ = ny + 2;
    double hx = 1.0 / (II - 1), hy = 1.0 / (JJ - 1);
    for (int j = 0; j < JJ; j++)
        for (int i = 0; i < II; i++) {
            double x = i * hx, y = j * hy;
            q[(size_t)j * II + i] = sin(2 * pi * x) * sin(2 * pi * y);
        }
}
