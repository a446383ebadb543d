/*
 * oracle/oracle.h -- TEST INFRASTRUCTURE (see bmg2.c header).  Prototypes of
 * the CPU restatement of the reference's serial BoxMG path.
 */
#ifndef BMG_ORACLE_H
#define BMG_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

void orc_set_threads(int n);
int orc_get_max_threads(void);

int orc_dpttrf(int n, double *d, double *e);
void orc_dptts2(int n, const double *d, const double *e, double *b);
/* LAPACK DPOTRF('U') as unblocked DPOTF2 and DPOTRS('U', nrhs = 1) in reference-BLAS loop order (dense Cholesky of the
 * periodic coarse-grid operator) */
int orc_dpotf2_upper(int n, double *a, int lda);
void orc_dpotrs_upper(int n, const double *a, int lda, double *b);
int orc_dpbtf2_upper(int n, int kd, double *ab, int ldab);
void orc_dpbtrs_upper(int n, int kd, const double *ab, int ldab, double *x);

/* 2D kernels: argument lists follow the BIND(C) interfaces of
 * src/2d/ftn/BMG2_SymStd_*.f90. */
void orc2_setup_recip(const double *so, double *sor, int nx, int ny, int nstncl, int nsorv);
int orc2_relax_gs(int k, const double *so, const double *qf, double *q, const double *sor,
                  int II, int JJ, int kf, int ifd, int nstncl, int nsorv,
                  int irelax_sym, int updown, int jpn);
int orc2_relax_gs_pass(int k, const double *so, const double *qf, double *q, const double *sor,
                       int II, int JJ, int kf, int ifd, int nstncl, int nsorv,
                       int irelax_sym, int updown, int jpn, int pass);
void orc2_matvec(int k, const double *so, double *qf, const double *q, int II, int JJ, int kf, int ifd, int nstncl);
void orc2_residual(int k, const double *so, const double *qf, const double *q, double *res,
                   int II, int JJ, int kf, int ifd, int nstncl);
int orc2_restrict(const double *q, double *qc, const double *ci,
                  int Nx, int Ny, int Nxc, int Nyc, int jpn);
int orc2_interp_add(double *q, const double *qc, double *res, const double *so, const double *ci,
                    int IIC, int JJC, int IIF, int JJF, int nstncl, int jpn);
int orc2_setup_interp_oi(const double *so, const double *soc, double *ci,
                         int IIF, int JJF, int IIC, int JJC, int ifd, int nstncl,
                         int jpn, int irelax);
int orc2_setup_interp_oi_stage(const double *so, const double *soc, double *ci, int IIF, int JJF, int IIC, int JJC,
                               int ifd, int nstncl, int jpn, int irelax, int stage);
int orc2_setup_itli_ex(const double *so, double *soc, const double *ci,
                       int IIF, int JJF, int IIC, int JJC, int ifd, int nstncl, int ipn);
int orc2_setup_lines_x(const double *so, double *sor, int Nx, int Ny, int nstncl, int jpn);
int orc2_setup_lines_y(const double *so, double *sor, int Nx, int Ny, int nstncl, int jpn);
int orc2_relax_lines_x(int k, const double *so, const double *qf, double *q, const double *sor,
                       double *b, int II, int JJ, int kf, int ifd, int nstncl,
                       int irelax_sym, int updown, int jpn);
int orc2_relax_lines_y(int k, const double *so, const double *qf, double *q, const double *sor,
                       double *b, int II, int JJ, int kf, int ifd, int nstncl,
                       int irelax_sym, int updown, int jpn);
int orc2_setup_cg_lu(const double *so, int II, int JJ, int nstncl,
                     double *abd, int nabd1, int nabd2, int ibc);
int orc2_solve_cg(double *q, const double *qf, int II, int JJ,
                  const double *abd, double *bbd, int nabd1, int nabd2, int jpn);
double orc2_l2norm(const double *q, int II, int JJ);

/* gallery (src/2d/gallery.cc) -- fills a zeroed SO(nx+2,ny+2,nstncl) */
void orc2_gallery_poisson(double *so, int nx, int ny);
void orc2_gallery_diag_diffusion(double *so, int nx, int ny, double dx, double dy);
void orc2_gallery_fe(double *so, int nx, int ny);
void orc2_rhs_sin(double *b, int nx, int ny);      /* examples/basic-2d-ser/poisson.cc:15-37 */
void orc2_exact_sin(double *q, int nx, int ny);    /* ...poisson.cc:40-58 */

/* multilevel driver (include/cedar/multilevel.h, cycle/vcycle.h, 2d/solver.h) */
enum { ORC_RELAX_POINT = 0, ORC_RELAX_LINE_X = 1, ORC_RELAX_LINE_Y = 2, ORC_RELAX_LINE_XY = 3 };
typedef struct orc2_solver orc2_solver;
orc2_solver *orc2_solver_create(const double *so, int nx, int ny, int nstncl,
                                int relaxation, int nrelax_pre, int nrelax_post,
                                int num_levels, int min_coarse);
/* per_mask: bit 0 periodic in x, bit 1 periodic in y; point relaxation only (NULL otherwise) */
orc2_solver *orc2_solver_create_per(const double *so, int nx, int ny, int nstncl, int relaxation, int nrelax_pre,
                                    int nrelax_post, int num_levels, int min_coarse, int per_mask);
void orc2_solver_destroy(orc2_solver *s);
int orc2_solver_nlevels(const orc2_solver *s);
void orc2_solver_level_shape(const orc2_solver *s, int lvl, int *nx, int *ny);
/* which: 0=A 1=P(ci) 2=x 3=res 4=b 5=SOR0 6=SOR1 7=ABD */
double *orc2_solver_level_ptr(orc2_solver *s, int lvl, int which);
void orc2_solver_vcycle(orc2_solver *s, double *x, const double *b);
/* 1: orc2_solver_vcycle / _solve run the F-cycle of include/cedar/cycle/fcycle.h:49-83 instead of the V-cycle */
void orc2_solver_set_cycle(orc2_solver *s, int fcycle);
/* returns #iterations done; rel_hist[i] = ||r_i||/||r_0|| after cycle i */
int orc2_solver_solve(orc2_solver *s, const double *b, double *x, int maxiter, double tol,
                      double *rel_hist, double *res0);

/* 3D kernels: argument lists follow the BIND(C) interfaces of
 * src/3d/ftn/BMG3_SymStd_*.f90 (include/cedar/3d/relax.h:11-16, residual.h:9-12,
 * src/3d/restrict.cc:7-10, src/3d/interp.cc:7-11, 3d/interp.h:11-15, 3d/coarsen.h:12-16). */
void orc3_setup_recip(const double *so, double *sor, int II, int JJ, int KK, int nstncl, int nsorv);
int orc3_relax_gs(int kg, const double *so, const double *qf, double *q, const double *sor, int II, int JJ, int KK,
                  int ifd, int nstncl, int nsorv, int irelax_sym, int updown, int jpn);
void orc3_matvec(int kg, int nog, int ifd, const double *so, double *qf, const double *q, int II, int JJ, int KK,
                 int nstncl);
void orc3_residual(int kg, int nog, int ifd, const double *q, const double *qf, const double *so, double *res,
                   int II, int JJ, int KK, int nstncl);
int orc3_restrict(const double *q, double *qc, const double *ci, int Nx, int Ny, int Nz, int Nxc, int Nyc, int Nzc,
                  int jpn);
int orc3_interp_add(double *q, const double *qc, const double *so, double *res, const double *ci, int IIC, int JJC,
                    int KKC, int IIF, int JJF, int KKF, int nstncl, int jpn);
int orc3_setup_interp_oi(const double *so, const double *soc, double *ci, int IIF, int JJF, int KKF, int IIC,
                         int JJC, int KKC, int ifd, int nstncl, int irelax, int jpn, double *yo);
/* ITLI07_ex (ifd == 1) and ITLI27_ex in one entry point */
int orc3_setup_itli_ex(const double *so, double *soc, const double *ci, int IIF, int JJF, int KKF, int IIC, int JJC,
                       int KKC, int ifd, int nstncl, int ipn);
int orc3_setup_cg_lu(const double *so, int II, int JJ, int KK, int nstncl, double *abd, int nabd1, int nabd2, int ibc);
int orc3_solve_cg(double *q, const double *qf, int II, int JJ, int KK, const double *abd, double *bbd, int nabd1,
                  int nabd2, int jpn);
double orc3_l2norm(const double *q, int II, int JJ, int KK);
void orc3_gallery_poisson(double *so, int nx, int ny, int nz);
void orc3_gallery_fe(double *so, int nx, int ny, int nz);
void orc3_rhs_sin(double *b, int nx, int ny, int nz);
void orc3_exact_sin(double *q, int nx, int ny, int nz);

typedef struct orc3_solver orc3_solver;
orc3_solver *orc3_solver_create(const double *so, int nx, int ny, int nz, int nstncl, int nrelax_pre, int nrelax_post,
                                int num_levels, int min_coarse);
void orc3_solver_destroy(orc3_solver *s);
int orc3_solver_nlevels(const orc3_solver *s);
void orc3_solver_level_shape(const orc3_solver *s, int lvl, int *nx, int *ny, int *nz);
double *orc3_solver_level_ptr(orc3_solver *s, int lvl, int which);
void orc3_solver_vcycle(orc3_solver *s, double *x, const double *b);
int orc3_solver_solve(orc3_solver *s, const double *b, double *x, int maxiter, double tol, double *rel_hist,
                      double *res0);

#ifdef __cplusplus
}
#endif
#endif
