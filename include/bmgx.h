/*
 * bmgx.h -- C ABI of the B200-native BoxMG kernels (libbmgx.so).
 *
 * This is the drop-in boundary: one entry point per `BIND(C)` Fortran routine
 * that the reference's kernel wrapper classes call (cedar include/cedar/2d headers,
 * src/2d sources).  Argument order and meaning follow the
 * reference interface cited at each prototype, with three changes:
 *   - every array pointer is a DEVICE pointer (fp64, Fortran order, i fastest,
 *     1 ghost layer: Q(II,JJ), SO(II,JJ,NStncl), SOR(II,JJ,2), CI(IIC,JJC,8));
 *   - every function takes a trailing `bmgx_stream_t` (a cudaStream_t; 0 = the
 *     legacy default stream) and only enqueues work: no host synchronisation,
 *     so all of them are CUDA-graph capturable;
 *   - every function returns an int status (0 = ok) instead of calling back
 *     into a logger; nothing throws across this boundary.
 * Scalars the Fortran takes by reference (BMG2_SymStd_residual) are by value.
 *
 * There is no CPU fallback: if no CUDA device is usable the calls return
 * BMGX_ERR_CUDA.
 */
#ifndef BMGX_H
#define BMGX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef double   bmgx_real;   /* cedar real_t, include/cedar/types.h:44 */
typedef uint32_t bmgx_len;    /* cedar len_t,  include/cedar/types.h:43 */
typedef void    *bmgx_stream_t;

enum {
    BMGX_OK = 0,
    BMGX_ERR_CUDA = 1,        /* CUDA runtime error, see bmgx_last_error() */
    BMGX_ERR_UNSUPPORTED = 2, /* e.g. periodic boundary conditions in 3D (not built) */
    BMGX_ERR_ARG = 3
};

/* src/2d/ftn/BMG_parameters_f90.h:345-413 */
enum { BMGX_BCs_definite = 0, BMGX_BCs_indef_nonper = -4 };
enum { BMGX_RELAX_NONSYM = 0, BMGX_RELAX_SYM = 1 };
enum { BMGX_DOWN = 0, BMGX_UP = 1 };

const char *bmgx_version(void);
const char *bmgx_last_error(void);
/* number of kernels this library has launched so far (process wide) */
unsigned long long bmgx_launch_count(void);
/* Transient setup workspace (3D Galerkin) comes stream-ordered from a memory pool this library owns (one per
 * device, cached between calls; the application's default pool is never reconfigured).  Return what is cached beyond
 * keep_bytes to the driver. */
int bmgx_scratch_trim(size_t keep_bytes);

/* BMG_get_bc, src/2d/ftn/BMG_get_bc.f90:1-24 */
void bmgx_get_bc(int per_mask, int *ibc);

/* ---- runtime support for the host layer ---------------------------------- */
/* Thin, C-callable wrappers so the C++ host layer (include/cedar) needs no
 * CUDA headers.  All return BMGX_OK or BMGX_ERR_CUDA. */
int bmgx_device_count(void);
int bmgx_set_device(int dev);
int bmgx_malloc(void **p, size_t bytes);              /* cudaMalloc */
int bmgx_free(void *p);
int bmgx_host_alloc(void **p, size_t bytes);          /* pinned host memory */
int bmgx_host_free(void *p);
int bmgx_memcpy_h2d(void *dst, const void *src, size_t bytes, bmgx_stream_t s);  /* async */
int bmgx_memcpy_d2h(void *dst, const void *src, size_t bytes, bmgx_stream_t s);  /* async */
int bmgx_memcpy_d2d(void *dst, const void *src, size_t bytes, bmgx_stream_t s);  /* async */
int bmgx_memset0(void *dst, size_t bytes, bmgx_stream_t s);
int bmgx_stream_create(bmgx_stream_t *s);
int bmgx_stream_destroy(bmgx_stream_t s);
int bmgx_stream_sync(bmgx_stream_t s);
/* 1 if the stream is being captured into a CUDA graph by the caller */
int bmgx_stream_is_capturing(bmgx_stream_t s, int *capturing);
int bmgx_device_sync(void);
/* CUDA events for device-side timing on the launching stream */
int bmgx_event_create(void **ev);
int bmgx_event_destroy(void *ev);
int bmgx_event_record(void *ev, bmgx_stream_t s);
int bmgx_event_sync(void *ev);
int bmgx_stream_wait_event(bmgx_stream_t s, void *ev);   /* work enqueued on s after this call waits for ev */
int bmgx_event_elapsed_ms(void *start, void *stop, float *ms);
/* CUDA-graph capture of a sequence of bmgx calls (one V-cycle) */
int bmgx_graph_begin(bmgx_stream_t s);
int bmgx_graph_end(bmgx_stream_t s, void **graph_exec);
int bmgx_graph_launch(void *graph_exec, bmgx_stream_t s);
int bmgx_graph_destroy(void *graph_exec);

/* ---- batch context ---------------------------------------------------------
 * Plane relaxation solves many independent 2D problems with ONE operator (the planes of
 * a colour).  bmgx_set_batch(nb) makes the batch-aware entry points below treat every
 * grid-function argument (q, qf, res, the line scratch b, bbd) as nb arrays stored back to
 * back (stride = the array's own size, e.g. ii*jj doubles) and process all of them in one
 * launch; operator-like arguments (so, sor, ci, abd) are shared.  Batch-aware:
 * bmgx2_relax_lines_x/_y, bmgx2_residual, bmgx2_restrict, bmgx2_interp_add, bmgx2_SOLVE_cg,
 * and bmgx3_plane_copy / bmgx3_plane_rhs, which then stage planes ipl, ipl+2, ... (one per
 * batch entry).  Every other entry point ignores the context.  Per host thread; read when a
 * call is enqueued (so it is baked into a captured CUDA graph); reset with bmgx_set_batch(1). */
int bmgx_set_batch(unsigned nb);
unsigned bmgx_get_batch(void);

/* ---- utilities used by the host layer (grid_func::set / lp_norm) ---------- */
/* array<>::set(v) over n doubles */
int bmgx_fill(bmgx_real *a, size_t n, bmgx_real v, bmgx_stream_t s);
/* sum of squares of the interior of Q(II,JJ[,KK]) -> *out (device double),
 * grid_func::lp_norm<2>, include/cedar/2d/grid_func.h:42-53 (the caller takes
 * the sqrt).  Deterministic (fixed reduction tree).  work: >= bmgx_norm_work_len()
 * doubles of device scratch. */
size_t bmgx_norm_work_len(void);
int bmgx2_sumsq(const bmgx_real *q, bmgx_len ii, bmgx_len jj, bmgx_real *work,
                bmgx_real *out, bmgx_stream_t s);
int bmgx3_sumsq(const bmgx_real *q, bmgx_len ii, bmgx_len jj, bmgx_len kk, bmgx_real *work,
                bmgx_real *out, bmgx_stream_t s);

/* ---- 2D ------------------------------------------------------------------ */

/* BMG2_SymStd_SETUP_recip(so, sor, nx, ny, nstncl, nsor_v)
 * include/cedar/2d/relax.h:11-12, src/2d/ftn/BMG2_SymStd_SETUP_recip.f90 */
int bmgx2_SETUP_recip(const bmgx_real *so, bmgx_real *sor, bmgx_len nx, bmgx_len ny,
                      int nstncl, int nsor_v, bmgx_stream_t s);

/* BMG2_SymStd_relax_GS(k, so, qf, q, sor, ii, jj, kf, ifd, nstncl, nsorv,
 *                      irelax_sym, updown, jpn)
 * include/cedar/2d/relax.h:13-15, src/2d/ftn/BMG2_SymStd_relax_GS.f90 */
int bmgx2_relax_GS(int k, const bmgx_real *so, const bmgx_real *qf, bmgx_real *q,
                   const bmgx_real *sor, bmgx_len ii, bmgx_len jj, int kf, int ifd,
                   int nstncl, int nsorv, int irelax_sym, int updown, int jpn,
                   bmgx_stream_t s);

/* Out-of-place form of the same sweep (B200-native fast path): reads q_in,
 * writes the relaxed iterate (interior + copied ghost ring) to q_out; all
 * colours in ONE pass over memory.  q_in and q_out must not alias. */
int bmgx2_relax_GS_oop(int k, const bmgx_real *so, const bmgx_real *qf, const bmgx_real *q_in,
                       bmgx_real *q_out, const bmgx_real *sor, bmgx_len ii, bmgx_len jj,
                       int kf, int ifd, int nstncl, int nsorv, int irelax_sym, int updown,
                       int jpn, bmgx_stream_t s);

/* The same out-of-place sweep restricted to rectangles of the array (at most 5 per launch): each rectangle is a
 * sub-array [i0, i0+ni) x [j0, j0+nj) (0-based, i0 and j0 even so that the colouring by index parity is kept) that is
 * swept as if it were a grid of its own -- its outer lines play the ghost ring and come from q_in -- and only the
 * values inside the window [wi0, wi1] x [wj0, wj1] (1-based, inclusive, sub-array coordinates) are stored to q_out.
 * A window whose points lie at least ncolours lines inside the sub-array (or on a side where the sub-array ends
 * with the array) receives exactly the values of the whole-array sweep: this is how the block-decomposed path
 * relaxes the interior of a block while its halos are still travelling, and the frames afterwards. */
typedef struct {
    bmgx_len i0, j0, ni, nj;
    bmgx_len wi0, wi1, wj0, wj1;
} bmgx_rect;
int bmgx2_relax_GS_oop_rects(int k, const bmgx_real *so, const bmgx_real *qf, const bmgx_real *q_in, bmgx_real *q_out,
                             const bmgx_real *sor, bmgx_len ii, bmgx_len jj, int kf, int ifd, int nstncl, int nsorv,
                             int irelax_sym, int updown, int jpn, const bmgx_rect *rects, int nrects, bmgx_stream_t s);

/* BMG2_SymStd_SETUP_lines_x / _y (so, sor, nx, ny, nstncl, jpn)
 * include/cedar/2d/relax.h:16-19; LAPACK DPTTRF is replaced on device. */
int bmgx2_SETUP_lines_x(const bmgx_real *so, bmgx_real *sor, bmgx_len nx, bmgx_len ny,
                        int nstncl, int jpn, bmgx_stream_t s);
int bmgx2_SETUP_lines_y(const bmgx_real *so, bmgx_real *sor, bmgx_len nx, bmgx_len ny,
                        int nstncl, int jpn, bmgx_stream_t s);

/* BMG2_SymStd_relax_lines_x / _y (k, so, qf, q, sor, b, ii, jj, kf, ifd,
 *                                 nstncl, irelax_sym, updown, jpn)
 * include/cedar/2d/relax.h:20-25; `b` is scratch of >= ii*jj doubles.  All boundary codes: lines periodic in their own
 * direction get the Sherman-Morrison second solve (relax_lines_x.f90:209-224) inside the line solver, the ghost layer
 * is refreshed as the reference does (per colour, or once after both colours for the other direction); periodic lines
 * of more than ~8500 unknowns are refused (BMGX_ERR_UNSUPPORTED). */
int bmgx2_relax_lines_x(int k, const bmgx_real *so, const bmgx_real *qf, bmgx_real *q,
                        const bmgx_real *sor, bmgx_real *b, bmgx_len ii, bmgx_len jj,
                        int kf, int ifd, int nstncl, int irelax_sym, int updown, int jpn,
                        bmgx_stream_t s);
int bmgx2_relax_lines_y(int k, const bmgx_real *so, const bmgx_real *qf, bmgx_real *q,
                        const bmgx_real *sor, bmgx_real *b, bmgx_len ii, bmgx_len jj,
                        int kf, int ifd, int nstncl, int irelax_sym, int updown, int jpn,
                        bmgx_stream_t s);

/* bmgx2_relax_lines_y with NO allocation anywhere: the ii*jj doubles of scratch live in a device array built into the
 * library.  No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid it).  Limit: ii*jj (times the batch count)
 * <= BMGX_LINES_Y_STATIC_MAX doubles (8 MiB, held by every GPU that loads the library until exit); larger grids are
 * refused with BMGX_ERR_ARG.  All calls share the one array: calls that overlap on the GPU (other streams, other host
 * threads) race.  Same kernels, same result bit for bit as bmgx2_relax_lines_y. */
#define BMGX_LINES_Y_STATIC_MAX (1u << 20)
int bmgx2_relax_lines_y_static(int k, const bmgx_real *so, const bmgx_real *qf, bmgx_real *q, const bmgx_real *sor,
                               bmgx_len ii, bmgx_len jj, int kf, int ifd, int nstncl, int irelax_sym, int updown,
                               int jpn, bmgx_stream_t s);

/* one colour of the sweeps above: the lines with F-index first, first + 2, ... (first = 2 or 3) */
int bmgx2_relax_lines_x_colour(const bmgx_real *so, const bmgx_real *qf, bmgx_real *q, const bmgx_real *sor, bmgx_len ii,
                               bmgx_len jj, int nine_pt, int first, bmgx_stream_t s);
int bmgx2_relax_lines_y_colour(const bmgx_real *so, const bmgx_real *qf, bmgx_real *q, const bmgx_real *sor, bmgx_real *b,
                               bmgx_len ii, bmgx_len jj, int nine_pt, int first, bmgx_stream_t s);

/* Lines that cross block cuts (the decomposed solver).  Counterpart of the reference's interface-system line solve
 * BMG2_SymStd_relax_lines_x/_y of src/2d/ftn/mpi/ (LineSolve_A/B/C, :156-215) -- here the blocks that share lines
 * transpose them over peer memory and the serial solve runs on whole lines (cedar_b200/csrc/bmgx2_lines.cu).
 * A line buffer has rows 1 .. (lines its block solves), row pitch = global line length + 2, the unknown with global
 * F-index p + 1 in column p; the factor `sor` (2 planes of the same shape) comes from bmgx2_SETUP_lines_x/_y on a
 * strip of gathered diagonals. */
#define BMGX_LINE_GROUP_MAX 16
typedef struct {
    int nmembers;                         /* blocks that share the lines */
    int first[BMGX_LINE_GROUP_MAX + 1];   /* member m solves the group's lines first[m] .. first[m+1]-1 (even bounds) */
    size_t pitch;                         /* global line length + 2 */
    bmgx_real *lb[BMGX_LINE_GROUP_MAX];   /* line buffers of the members as mapped in this process */
} bmgx_line_group;
typedef struct {
    int nmembers;
    int start[BMGX_LINE_GROUP_MAX + 1];   /* member m owns the positions start[m] .. start[m+1]-1 (0-based) of a line */
    int w_along[BMGX_LINE_GROUP_MAX];     /* array index of member m's first own point along the line */
    int w_across;                         /* array index of the group's first line (the same for every member) */
    size_t pitch[BMGX_LINE_GROUP_MAX];    /* row pitch of member m's array */
    bmgx_real *q[BMGX_LINE_GROUP_MAX];    /* the members' iterates as mapped in this process */
} bmgx_line_owners;
/* right-hand sides of the own points (nx x ny, first at array entry [wb][wl]) of the lines of one colour -- dirn 0:
 * x-lines, own rows t with t = first_findex (mod 2); dirn 1: y-lines, own columns likewise -- stored at column
 * col0 + (own index along the line) of the line's row in its solver's buffer.  The expression is the serial
 * kernels' (BMG2_SymStd_relax_lines_x.f90:100-128, _y.f90:99-127). */
int bmgx2_lines_rhs_push(int dirn, int nine, const bmgx_real *so, const bmgx_real *qf, const bmgx_real *q, bmgx_len ii,
                         bmgx_len jj, bmgx_len wl, bmgx_len wb, bmgx_len nx, bmgx_len ny, int first_findex,
                         bmgx_len col0, const bmgx_line_group *g, bmgx_stream_t s);
/* DPTTRS on the rows ibeg-1, ibeg+1, ... of a line buffer of nrows2 = lines + 2 rows and pitch len2 = length + 2, in
 * place: the second step of bmgx2_relax_lines_y */
int bmgx2_lines_solve_T(bmgx_real *lb, const bmgx_real *sor, bmgx_len nrows2, bmgx_len len2, int ibeg, bmgx_stream_t s);
/* the solved rows of that colour -> the owners' iterates; tfirst = the group line held by row 1 */
int bmgx2_lines_scatter_push(int dirn, const bmgx_real *lb, bmgx_len nrows2, bmgx_len len2, int ibeg, bmgx_len tfirst,
                             const bmgx_line_owners *o, bmgx_stream_t s);

/* BMG2_SymStd_residual(k, so, qf, q, res, ii, jj, kf, ifd, nstncl, ibc,
 *                      irelax, irelax_sym, updown)   [by reference there]
 * include/cedar/2d/residual.h:9-10 */
int bmgx2_residual(int k, const bmgx_real *so, const bmgx_real *qf, const bmgx_real *q,
                   bmgx_real *res, bmgx_len ii, bmgx_len jj, int kf, int ifd, int nstncl,
                   int ibc, int irelax, int irelax_sym, int updown, bmgx_stream_t s);

/* BMG2_SymStd_UTILS_matvec(k, so, qf, q, ii, jj, kf, ifd, nstncl): qf = A q on the interior
 * (src/2d/ftn/mpi/BMG2_SymStd_UTILS_matvec.f90:88-123; the operator "apply" of the C interface,
 * src/2d/interface/c/operator.cc:88, include/cedar/2d/matvec.h) */
int bmgx2_matvec(int k, const bmgx_real *so, bmgx_real *qf, const bmgx_real *q, bmgx_len ii, bmgx_len jj,
                 int kf, int ifd, int nstncl, bmgx_stream_t s);

/* BMG2_SymStd_restrict(q, qc, ci, nx, ny, nxc, nyc, jpn), src/2d/restrict.cc:6-9 */
int bmgx2_restrict(const bmgx_real *q, bmgx_real *qc, const bmgx_real *ci, bmgx_len nx,
                   bmgx_len ny, bmgx_len nxc, bmgx_len nyc, int jpn, bmgx_stream_t s);

/* BMG2_SymStd_interp_add(q, qc, res, so, ci, iic, jjc, iif, jjf, nstncl, jpn)
 * src/2d/interp.cc:8-10.  `res` is scaled in place by 1/diag as in the
 * reference (BMG2_SymStd_interp_add.f90:101-105). */
int bmgx2_interp_add(bmgx_real *q, const bmgx_real *qc, bmgx_real *res, const bmgx_real *so,
                     const bmgx_real *ci, bmgx_len iic, bmgx_len jjc, bmgx_len iif,
                     bmgx_len jjf, int nstncl, int jpn, bmgx_stream_t s);

/* BMG2_SymStd_SETUP_interp_OI(so, soc, ci, iif, jjf, iic, jjc, ifd, nstncl,
 *                             jpn, irelax), src/2d/interp.cc:11-13 */
int bmgx2_SETUP_interp_OI(const bmgx_real *so, const bmgx_real *soc, bmgx_real *ci,
                          bmgx_len iif, bmgx_len jjf, bmgx_len iic, bmgx_len jjc, int ifd,
                          int nstncl, int jpn, int irelax, bmgx_stream_t s);

/* BMG2_SymStd_SETUP_ITLI_ex(so, soc, ci, iif, jjf, iic, jjc, ifd, nstncl, ipn)
 * include/cedar/2d/coarsen.h:10-12 (Galerkin coarse operator) */
int bmgx2_SETUP_ITLI_ex(const bmgx_real *so, bmgx_real *soc, const bmgx_real *ci,
                        bmgx_len iif, bmgx_len jjf, bmgx_len iic, bmgx_len jjc, int ifd,
                        int nstncl, int ipn, bmgx_stream_t s);

/* BMG2_SymStd_SETUP_cg_LU(so, ii, jj, nstncl, abd, nabd1, nabd2, ibc)
 * include/cedar/2d/solve_cg.h:10 ; LAPACK DPBTRF replaced on device. */
int bmgx2_SETUP_cg_LU(const bmgx_real *so, bmgx_len ii, bmgx_len jj, int nstncl,
                      bmgx_real *abd, bmgx_len nabd1, bmgx_len nabd2, int ibc,
                      bmgx_stream_t s);

/* INFO of the last factorisation enqueued by bmgx2_/bmgx3_SETUP_cg_LU (DPBTRF's INFO; the reference logs "Coarse grid
 * Cholesky decomposition failed", BMG2_SymStd_SETUP_cg_LU.f90:124-131).  Synchronises the stream.  Returns
 * BMGX_ERR_ARG with the failing column in bmgx_last_error() when *info != 0: the factor is unusable. */
int bmgx_cg_LU_info(int *info, bmgx_stream_t s);

/* BMG2_SymStd_SOLVE_cg(q, qf, ii, jj, abd, bbd, nabd1, nabd2, jpn)
 * src/2d/solve_cg.cc:6 ; bbd is DEVICE scratch of nabd2 doubles. */
int bmgx2_SOLVE_cg(bmgx_real *q, const bmgx_real *qf, bmgx_len ii, bmgx_len jj,
                   const bmgx_real *abd, bmgx_real *bbd, bmgx_len nabd1, bmgx_len nabd2,
                   int jpn, bmgx_stream_t s);

/* ---- the coarse tail of a V-cycle in one launch (cedar_b200/csrc/bmgx2_tail.cu) --------------------------------
 * Runs vcycle::ncycle (include/cedar/cycle/vcycle.h; reference cycle/vcycle.h:57-115) for `nlevels` successive
 * nine-point levels, finest first, with point relaxation and the direct solve (bmgx2_SOLVE_cg's ABD, bbd) on the last:
 * presmooth, residual, restrict, zero coarse x, ..., interpolate-and-add, postsmooth -- one CTA, no launches in
 * between.  Same arithmetic as the per-level entry points (bitwise).  Entry l: so/sor (5 and 2 planes), q (iterate,
 * in place; the caller provides q of entry 0, the others are cleared), qf (right-hand side; entries > 0 are
 * overwritten), res (scratch), ci = interpolation between level l-1 and l (8 planes on level l's grid, unused for
 * l = 0); ii, jj with ghost layer.  Levels are limited to BMGX_TAIL_MAX_POINTS entries. */
#define BMGX_TAIL_MAX 8
#define BMGX_TAIL_MAX_POINTS (160 * 160)
typedef struct {
    const bmgx_real *so, *sor, *ci;
    bmgx_real *q, *qf, *res;
    bmgx_len ii, jj;
} bmgx_tail_level;
int bmgx2_vcycle_tail(const bmgx_tail_level *levels, int nlevels, const bmgx_real *abd, bmgx_real *bbd, bmgx_len nabd1,
                      int nrelax_pre, int nrelax_post, bmgx_stream_t s);

/* ---- entry points of the block-decomposed (multi-GPU) 2D path ----------------
 * The reference's MPI kernels (src/2d/ftn/mpi/) exchange ghost values between the
 * colour passes of a sweep and between the stages of the interpolation setup; these
 * entry points expose the same cut points of the serial kernels. */
/* one colour pass (0..3 nine-point, 0..1 five-point, in sweep order) of bmgx2_relax_GS */
int bmgx2_relax_GS_pass(int k, const bmgx_real *so, const bmgx_real *qf, bmgx_real *q,
                        const bmgx_real *sor, bmgx_len ii, bmgx_len jj, int kf, int ifd,
                        int nstncl, int nsorv, int irelax_sym, int updown, int jpn, int pass,
                        bmgx_stream_t s);
/* stage 1 (edge weights LL,LR,LA,LB) or 2 (cell-centre weights) of bmgx2_SETUP_interp_OI */
int bmgx2_SETUP_interp_OI_stage(const bmgx_real *so, const bmgx_real *soc, bmgx_real *ci,
                                bmgx_len iif, bmgx_len jjf, bmgx_len iic, bmgx_len jjc, int ifd,
                                int nstncl, int jpn, int irelax, int stage, bmgx_stream_t s);
/* ghost-line staging: copy column (dir 0) / row (dir 1) `index` (1-based) of the planes
 * selected by plane_mask of A(II,JJ,nplanes) to (to_buf != 0) or from a contiguous
 * buffer [selected plane][position]; what halo-2d's pack/unpack did on the host */
int bmgx2_line_copy(bmgx_real *a, bmgx_len ii, bmgx_len jj, int nplanes, unsigned plane_mask,
                    int dir, bmgx_len index, bmgx_real *buf, int to_buf, bmgx_stream_t s);

/* ---- peer-memory ghost exchange of the block-decomposed path (cedar_b200/csrc/bmgx2_halo.cu) ----
 * Replaces kernels::halo_exchange (include/cedar/kernels/halo_exchange.h:22-31; MSG/Tausch + MPI in
 * src/2d/mpi/halo.cc): every rank keeps its arrays in one device allocation, the ranks open each other's
 * allocations once through CUDA IPC, and one exchange is ONE kernel that signals, waits for the ranks it
 * reads from and pulls the halo rectangles out of their arrays over NVLink. */
#define BMGX_IPC_HANDLE_BYTES 64
int bmgx_ipc_export(void *devptr, void *handle64);            /* cudaIpcGetMemHandle of a bmgx_malloc'd block */
int bmgx_ipc_open(const void *handle64, void **devptr);       /* maps a peer's block (enables peer access) */
int bmgx_ipc_close(void *devptr);
/* one rectangle of an exchange: rows x cols doubles from src (in rank src_rank's allocation, as mapped in
 * this process) to dst (local); pitches in doubles */
typedef struct {
    const bmgx_real *src;
    bmgx_real *dst;
    uint32_t rows, cols;
    uint64_t src_pitch, dst_pitch;
    int32_t src_rank, reserved;
} bmgx_seg;
/* One exchange step.  my_sync: this rank's sync block (world + 3 zero-initialised 64-bit words at a fixed place of
 * its allocation: flags[world], exchange counter, CTA counter, error word); peer_sync_dev: DEVICE array of
 * `world` pointers to the sync blocks of all ranks (own included) as mapped here.  phase: 1 = signal, 2 = wait
 * + pull, 3 = both (one rank per process).  wait_all != 0: wait for every rank, not only the ranks read from.
 * nseg may be 0 (a pure synchronisation step).  A wait longer than timeout_s stores 1 + rank in the error word
 * and carries on instead of hanging.  max_chunks: CTAs per rectangle (1..64). */
int bmgx_halo_step(const bmgx_seg *segs_dev, int nseg, int max_chunks, void *const *peer_sync_dev, void *my_sync,
                   int me, int world, int phase, int wait_all, double timeout_s, bmgx_stream_t s);
/* rectangle copies (pitches in doubles): device to device (a kernel), host <-> device (cudaMemcpy2DAsync) */
int bmgx_copy2d(bmgx_real *dst, size_t dpitch, const bmgx_real *src, size_t spitch, bmgx_len cols, bmgx_len rows,
                bmgx_stream_t s);
int bmgx_memcpy2d_h2d(bmgx_real *dst, size_t dpitch, const bmgx_real *src, size_t spitch, bmgx_len cols,
                      bmgx_len rows, bmgx_stream_t s);
int bmgx_memcpy2d_d2h(bmgx_real *dst, size_t dpitch, const bmgx_real *src, size_t spitch, bmgx_len cols,
                      bmgx_len rows, bmgx_stream_t s);
/* deterministic sum of squares of a rows x cols window (a block's own points inside its halo-extended array);
 * work: >= bmgx_sumsq_win_work_len() doubles */
size_t bmgx_sumsq_win_work_len(void);
int bmgx2_sumsq_win(const bmgx_real *q, size_t pitch, bmgx_len cols, bmgx_len rows, bmgx_real *work, bmgx_real *out,
                    bmgx_stream_t s);
/* *out = v[0] + ... + v[n-1] in index order (the rank-ordered fold of the partial norms: same bits on every rank) */
int bmgx_sum_ordered(const bmgx_real *v, int n, bmgx_real *out, bmgx_stream_t s);
/* bmgx2_restrict / bmgx2_interp_add (non-periodic) with the COARSE arrays (qc, ci) given as a window of a larger
 * array: pointers to the window's first entry, row pitch cpitch and plane stride cplane in doubles.  The fine
 * arrays are plain.  Used by the block-decomposed path, whose coarse arrays carry wider halos than the fine
 * array's coarsening reaches. */
int bmgx2_restrict_v(const bmgx_real *q, bmgx_real *qc, const bmgx_real *ci, bmgx_len nx, bmgx_len ny, bmgx_len nxc,
                     bmgx_len nyc, size_t cpitch, size_t cplane, bmgx_stream_t s);
int bmgx2_interp_add_v(bmgx_real *q, const bmgx_real *qc, bmgx_real *res, const bmgx_real *so, const bmgx_real *ci,
                       bmgx_len iic, bmgx_len jjc, bmgx_len iif, bmgx_len jjf, size_t cpitch, size_t cplane,
                       bmgx_stream_t s);

/* ---- 3D ------------------------------------------------------------------
 * Arrays: Q(II,JJ,KK), SO(II,JJ,KK,NStncl) with NStncl = 4 (seven_pt: p,pw,ps,b)
 * or 14 (xxvii_pt), SOR(II,JJ,KK,2), CI(IIC,JJC,KKC,26); slot order of
 * src/3d/ftn/BMG_stencils_f90.h:43-64. */

/* BMG3_SymStd_SETUP_recip(so, sor, nx, ny, nz, nstncl, nsorv), include/cedar/3d/relax.h:11-13 */
int bmgx3_SETUP_recip(const bmgx_real *so, bmgx_real *sor, bmgx_len nx, bmgx_len ny, bmgx_len nz,
                      int nstncl, int nsorv, bmgx_stream_t s);

/* BMG3_SymStd_relax_GS(kg, so, qf, q, sor, ii, jj, kk, ifd, nstncl, nsorv, irelax_sym,
 *                      updown, jpn), include/cedar/3d/relax.h:14-16 */
int bmgx3_relax_GS(int kg, const bmgx_real *so, const bmgx_real *qf, bmgx_real *q,
                   const bmgx_real *sor, bmgx_len ii, bmgx_len jj, bmgx_len kk, int ifd, int nstncl,
                   int nsorv, int irelax_sym, int updown, int jpn, bmgx_stream_t s);

/* Out-of-place form of the same sweep for the 7-point operator (ifd == 1; B200-native fast
 * path): reads q_in, writes the relaxed iterate (interior + copied ghost layer) to q_out; both
 * colours in ONE pass over memory.  q_in and q_out must not alias.  27-point operators
 * (ifd != 1) return BMGX_ERR_UNSUPPORTED: use bmgx3_relax_GS. */
int bmgx3_relax_GS_oop(int kg, const bmgx_real *so, const bmgx_real *qf, const bmgx_real *q_in,
                       bmgx_real *q_out, const bmgx_real *sor, bmgx_len ii, bmgx_len jj, bmgx_len kk,
                       int ifd, int nstncl, int nsorv, int irelax_sym, int updown, int jpn,
                       bmgx_stream_t s);

/* BMG3_SymStd_residual(kg, nog, ifd, q, qf, so, res, ii, jj, kk, nstncl),
 * include/cedar/3d/residual.h:9-12 */
int bmgx3_residual(int kg, int nog, int ifd, const bmgx_real *q, const bmgx_real *qf,
                   const bmgx_real *so, bmgx_real *res, bmgx_len ii, bmgx_len jj, bmgx_len kk,
                   int nstncl, bmgx_stream_t s);

/* BMG3_SymStd_UTILS_matvec(kg, so, qf, q, ii, jj, kk, nog, ifd, nstncl): qf = A q on the interior
 * (src/3d/ftn/mpi/BMG3_SymStd_UTILS_matvec.f90:78-129, include/cedar/3d/mpi/matvec.h) */
int bmgx3_matvec(int kg, const bmgx_real *so, bmgx_real *qf, const bmgx_real *q, bmgx_len ii, bmgx_len jj,
                 bmgx_len kk, int nog, int ifd, int nstncl, bmgx_stream_t s);

/* BMG3_SymStd_restrict(q, qc, ci, nx, ny, nz, nxc, nyc, nzc, jpn), src/3d/restrict.cc:7-10 */
int bmgx3_restrict(const bmgx_real *q, bmgx_real *qc, const bmgx_real *ci, bmgx_len nx, bmgx_len ny,
                   bmgx_len nz, bmgx_len nxc, bmgx_len nyc, bmgx_len nzc, int jpn, bmgx_stream_t s);

/* BMG3_SymStd_interp_add(q, qc, so, res, ci, iic, jjc, kkc, iif, jjf, kkf, nstncl, jpn),
 * src/3d/interp.cc:7-11; `res` is scaled in place by 1/diag as in the reference */
int bmgx3_interp_add(bmgx_real *q, const bmgx_real *qc, const bmgx_real *so, bmgx_real *res,
                     const bmgx_real *ci, bmgx_len iic, bmgx_len jjc, bmgx_len kkc, bmgx_len iif,
                     bmgx_len jjf, bmgx_len kkf, int nstncl, int jpn, bmgx_stream_t s);

/* bmgx3_restrict / bmgx3_interp_add (non-periodic) with the coarse arrays given as a window of PLANES of a thicker array
 * (the slab-decomposed multi-GPU path, whose coarse arrays carry more halo planes than the fine array's coarsening
 * reaches): qc and ci point at the window's first plane, kkc is the window's thickness, cstride the stride in doubles
 * between the 26 weight arrays of ci (iic*jjc*KKC of the thick array). */
int bmgx3_restrict_v(const bmgx_real *q, bmgx_real *qc, const bmgx_real *ci, bmgx_len nx, bmgx_len ny, bmgx_len nz,
                     bmgx_len nxc, bmgx_len nyc, bmgx_len nzc, size_t cstride, bmgx_stream_t s);
int bmgx3_interp_add_v(bmgx_real *q, const bmgx_real *qc, const bmgx_real *so, bmgx_real *res, const bmgx_real *ci,
                       bmgx_len iic, bmgx_len jjc, bmgx_len kkc, bmgx_len iif, bmgx_len jjf, bmgx_len kkf, size_t cstride,
                       bmgx_stream_t s);

/* BMG3_SymStd_SETUP_interp_OI(so, soc, ci, iif, jjf, kkf, iic, jjc, kkc, ifd, nstncl, irelax,
 *                             jpn, yo), include/cedar/3d/interp.h:11-15; the `yo` scratch of
 * the reference is not needed (may be NULL) */
int bmgx3_SETUP_interp_OI(const bmgx_real *so, const bmgx_real *soc, bmgx_real *ci, bmgx_len iif,
                          bmgx_len jjf, bmgx_len kkf, bmgx_len iic, bmgx_len jjc, bmgx_len kkc,
                          int ifd, int nstncl, int irelax, int jpn, bmgx_real *yo, bmgx_stream_t s);

/* BMG3_SymStd_SETUP_ITLI07_ex / ITLI27_ex(so, soc, ci, iif, jjf, kkf, iic, jjc, kkc, ipn),
 * include/cedar/3d/coarsen.h:12-16 (Galerkin coarse operator, 7- / 27-point fine operator) */
int bmgx3_SETUP_ITLI07_ex(const bmgx_real *so, bmgx_real *soc, const bmgx_real *ci, bmgx_len iif,
                          bmgx_len jjf, bmgx_len kkf, bmgx_len iic, bmgx_len jjc, bmgx_len kkc,
                          int ipn, bmgx_stream_t s);
int bmgx3_SETUP_ITLI27_ex(const bmgx_real *so, bmgx_real *soc, const bmgx_real *ci, bmgx_len iif,
                          bmgx_len jjf, bmgx_len kkf, bmgx_len iic, bmgx_len jjc, bmgx_len kkc,
                          int ipn, bmgx_stream_t s);

/* BMG3_SymStd_SETUP_cg_LU(so, ii, jj, kk, nstncl, abd, nabd1, nabd2, ibc) and
 * BMG3_SymStd_SOLVE_cg(q, qf, ii, jj, kk, abd, bbd, nabd1, nabd2, ibc),
 * include/cedar/3d/solve_cg.h; bbd is DEVICE scratch of nabd2 doubles */
int bmgx3_SETUP_cg_LU(const bmgx_real *so, bmgx_len ii, bmgx_len jj, bmgx_len kk, int nstncl,
                      bmgx_real *abd, bmgx_len nabd1, bmgx_len nabd2, int ibc, bmgx_stream_t s);
int bmgx3_SOLVE_cg(bmgx_real *q, const bmgx_real *qf, bmgx_len ii, bmgx_len jj, bmgx_len kk,
                   const bmgx_real *abd, bmgx_real *bbd, bmgx_len nabd1, bmgx_len nabd2, int jpn,
                   bmgx_stream_t s);

/* ---- plane relaxation staging (3D <-> 2D) ---------------------------------
 * dir: 0 = xy planes (plane index k), 1 = xz (index j), 2 = yz (index i); ipl is the
 * 1-based plane index; the 2D arrays are (II,JJ), (II,KK), (JJ,KK) with ghosts. */
/* copy32 (to2d != 0) / copy23, src/3d/relax_planes.cc:181-238 */
int bmgx3_plane_copy(bmgx_real *x3, bmgx_real *x2, bmgx_len ii, bmgx_len jj, bmgx_len kk, int dir,
                     bmgx_len ipl, int to2d, bmgx_stream_t s);
/* copy_coeff for one plane, include/cedar/3d/relax_planes.h:77-161: so2 (3 or 5 planes) */
int bmgx3_plane_coeff(const bmgx_real *so3, bmgx_real *so2, bmgx_len ii, bmgx_len jj, bmgx_len kk,
                      int nstncl, int dir, bmgx_len ipl, bmgx_stream_t s);
/* copy_rhs, src/3d/relax_planes.cc:26-178: b2 = b + (off-plane couplings) * x */
int bmgx3_plane_rhs(const bmgx_real *so, const bmgx_real *x, const bmgx_real *b, bmgx_real *b2,
                    bmgx_len ii, bmgx_len jj, bmgx_len kk, int nstncl, int dir, bmgx_len ipl,
                    bmgx_stream_t s);

#ifdef __cplusplus
}
#endif
#endif /* BMGX_H */
