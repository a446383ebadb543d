/*
 * cedar_capi.h -- C API of the C++ host layer (libcedar_b200.so) for bindings
 * (Python ctypes in tests/ and bench.py).  It plays the role of the reference's
 * include/cedar/capi.h (bmg2_operator_* / bmg2_solver_*): create an operator,
 * fill it on the host, build a solver (= setup phase), run solves.
 * Arrays handed in or out are HOST arrays in Cedar layout ((nx+2)(ny+2)[ndirs],
 * i fastest); the library moves them to and from the GPU.
 * All functions return 0 on success; the message of the last failure is
 * available from cdr_last_error().
 */
#ifndef CEDAR_B200_CAPI_H
#define CEDAR_B200_CAPI_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cdr2_operator cdr2_operator;
typedef struct cdr2_solver cdr2_solver;

const char *cdr_last_error(void);
unsigned long long cdr_gpu_launches(void);
/* log levels like the "log" key of config.json, comma separated ("status,error"; "" = silent) */
void cdr_set_log(const char *levels);

/* nstencil = 3 (five_pt) or 5 (nine_pt).  The returned host pointer views the
 * operator storage (nstencil planes of (nx+2)(ny+2)); fill it before building
 * the solver. */
int cdr2_operator_create(int nstencil, unsigned nx, unsigned ny, cdr2_operator **op);
double *cdr2_operator_data(cdr2_operator *op);
/* gallery: "poisson", "diag_diffusion" (uses dx, dy), "fe" */
int cdr2_operator_gallery(const char *name, unsigned nx, unsigned ny, double dx, double dy, cdr2_operator **op);
void cdr2_operator_destroy(cdr2_operator *op);

/* setup phase.  config_json: text of a Cedar config.json ("" or NULL = defaults) */
int cdr2_solver_create(cdr2_operator *op, const char *config_json, cdr2_solver **solver);
void cdr2_solver_destroy(cdr2_solver *solver);
int cdr2_solver_nlevels(cdr2_solver *solver);
int cdr2_solver_level_shape(cdr2_solver *solver, int lvl, unsigned *nx, unsigned *ny);
/* copies a level array to host memory: which = "A","P","x","res","b","SOR0","SOR1","ABD";
 * n = capacity of out in doubles; *written = doubles copied */
int cdr2_solver_level_array(cdr2_solver *solver, int lvl, const char *which, double *out, size_t n, size_t *written);

/* host-buffer solve: uploads b (and x as initial guess), runs multilevel::solve,
 * downloads x.  hist receives the relative residual after each cycle. */
int cdr2_solver_solve(cdr2_solver *solver, const double *b, double *x, double *hist, int hist_cap, int *niter,
                      double *res0);
/* one V-cycle on host buffers (upload x,b; cycle; download x) */
int cdr2_solver_vcycle(cdr2_solver *solver, const double *b, double *x);

/* y = A x with the fine-level operator (host buffers, ghost ring included; the operator "apply" of the
 * reference's C interface, src/2d/interface/c/operator.cc:76-90, through the registry's matvec kernel) */
int cdr2_solver_matvec(cdr2_solver *solver, const double *x, double *y);

/* one V-cycle on DEVICE buffers ((nx+2)(ny+2) doubles each), enqueued on `stream`
 * (a cudaStream_t; NULL = the library's own stream): no host traffic, no synchronisation */
int cdr2_solver_vcycle_dev(cdr2_solver *solver, const double *b_dev, double *x_dev, void *stream);

/* The solver's own device-resident b and x ((nx+2)(ny+2) doubles each) and one V-cycle on them, enqueued on `stream`:
 * for device-side callers that fill b and x themselves (the block-decomposed solver's replicated coarse levels).
 * The fused relaxation exchanges x's device buffer with a scratch array, so ask for the pointers again after a
 * cycle: *x_dev is where the iterate is NOW. */
int cdr2_solver_resident_ptrs(cdr2_solver *solver, double **b_dev, double **x_dev);
int cdr2_solver_vcycle_resident(cdr2_solver *solver, void *stream);

/* Pinned host staging buffers owned by the solver ((nx+2)(ny+2) doubles each):
 * fill *b (and *x as initial guess) in place, then call the _pinned variants,
 * which upload b and x, run on the device and download x into the same buffer
 * -- no intermediate host copy. */
int cdr2_solver_buffers(cdr2_solver *solver, double **b, double **x);
int cdr2_solver_vcycle_pinned(cdr2_solver *solver);
int cdr2_solver_solve_pinned(cdr2_solver *solver, double *hist, int hist_cap, int *niter, double *res0);
/* zero initial guess (what multilevel::solve(b) of the reference starts from, include/cedar/multilevel.h:278-286):
 * uploads b only, clears x on the device, one cycle, downloads x into the pinned buffer */
int cdr2_solver_vcycle0_pinned(cdr2_solver *solver);

/* A stream of right-hand sides.  For k < nrhs: one cycle from a zero initial guess on b[k], result in x[k]
 * (host pointers, (nx+2)(ny+2) doubles each).  The upload of b[k+1], the cycle on b[k] and the download of x[k-1]
 * run concurrently (two copy streams, two device staging buffers each way), provided the host arrays are pinned
 * (cdr_pinned_alloc); pageable arrays work but serialise.  Returns when x[nrhs-1] has arrived.  Every x[k] is
 * identical to what cdr2_solver_vcycle0_pinned gives for b[k]. */
int cdr2_solver_vcycle_batch(cdr2_solver *solver, int nrhs, const double *const *b, double *const *x);
int cdr_pinned_alloc(void **p, size_t bytes);
int cdr_pinned_free(void *p);

/* device-resident timing: keeps b and x on the GPU, runs `warmup` + `n` cycles
 * (x reset to 0 before each), returns the mean device time per cycle in ms
 * measured with CUDA events on the launching stream. */
int cdr2_solver_time_vcycles(cdr2_solver *solver, const double *b, int warmup, int n, float *ms_per_cycle);
/* device-resident full solve (x0 = 0) repeated `n` times after `warmup`; mean ms per solve */
int cdr2_solver_time_solves(cdr2_solver *solver, const double *b, int warmup, int n, float *ms_per_solve, int *niter);

/* one-shot end-to-end call: host operator + rhs in, solution out (operator
 * upload, setup, solve to tolerance, download) */
int cdr2_solve_host(int nstencil, unsigned nx, unsigned ny, const double *so, const double *b, double *x,
                    const char *config_json, double *hist, int hist_cap, int *niter);


/* ---- block-decomposed 2D solver across GPUs (cdr2::mpi::solver of the reference, include/cedar/2d/mpi/solver.h) ----
 * One process per GPU (or several blocks in one process on one GPU).  The global nx x ny grid is cut into px x py
 * blocks (block id = cy*px + cx, include/cedar/decomp.h); each process creates the solver with the ids it owns,
 * uploads the own entries of its blocks' operators, hands the others its IPC handles (the caller moves these bytes
 * between the processes: one all-gather), connects, and runs setup.  Ghost exchanges then go through peer memory
 * over NVLink (include/bmgx.h, bmgx_halo_step); collective calls (setup, vcycle, solve, time_vcycles,
 * vcycle_stream) must be made by all processes in the same order.  relaxation: "point", "line-x", "line-y" or
 * "line-xy" (lines that cross block cuts are transposed between the blocks that share them and solved whole,
 * include/cedar/2d/mpi/plan.h); results are bit for bit those of the serial cdr2::solver on the global problem.
 * Levels of global size <= agglomerate are solved replicated on every rank.  halo_width: odd, lines of the neighbours every array carries (15 = one exchange per
 * level per V(1,1) cycle). */
typedef struct cdr2_dist cdr2_dist;
/* split_min: on the levels where every block has at least this many own points per dimension, a sweep that directly
 * follows the exchange of its operands is split: the exchange runs on a second stream while the interior is relaxed,
 * the frames follow when the halos have arrived (0 = never; CDR_DIST_OVERLAP=0 in the environment runs the same
 * operation list without the second stream). */
int cdr2_dist_create(unsigned nx, unsigned ny, int px, int py, int nstencil, int nowned, const int *owned_blocks,
                     int nrelax_pre, int nrelax_post, const char *relaxation, int agglomerate, int halo_width, int split_min,
                     int cuda_graph, cdr2_dist **out);
void cdr2_dist_destroy(cdr2_dist *d);
int cdr2_dist_info(cdr2_dist *d, int *ndist, int *nlevels, int *halo_width, int *exchanges_per_cycle, int *ops_per_cycle);
/* first own point (0-based global interior index) and own counts of a block on a distributed level */
int cdr2_dist_block_extent(cdr2_dist *d, int lvl, int block, int *x0, int *y0, int *nx, int *ny);
/* handles: nblocks * 64 bytes indexed by block id; export fills the entries of the owned blocks */
int cdr2_dist_export_handles(cdr2_dist *d, void *handles);
int cdr2_dist_connect(cdr2_dist *d, const void *handles);
/* own entries of a block's fine operator: host array (nstencil, nyl+2, nxl+2) in Cedar layout, ghost ring ignored */
int cdr2_dist_set_operator(cdr2_dist *d, int block, const double *so);
int cdr2_dist_setup(cdr2_dist *d);
/* grid functions of a block, host arrays (nyl+2, nxl+2), own points only; which = "x" | "b" (download also "res") */
int cdr2_dist_upload(cdr2_dist *d, int block, const char *which, const double *host);
int cdr2_dist_download(cdr2_dist *d, int block, const char *which, double *host);
/* own points of a level array -> host (planes, nyl, nxl); which = "A" | "P" | "SOR" | "x" | "b" | "res" */
int cdr2_dist_level_array(cdr2_dist *d, int block, int lvl, const char *which, double *out);
/* own points of an array of any distributed level <- host: grid functions (nyl, nxl), which = "x" | "b" | "res";
 * the operator (planes, nyl, nxl), which = "A" */
int cdr2_dist_upload_level(cdr2_dist *d, int block, int lvl, const char *which, const double *host);
/* the whole halo-extended array of a level -> host (planes, jj, ii), extents as cdr2_dist_geometry reports them */
int cdr2_dist_extended_array(cdr2_dist *d, int block, int lvl, const char *which, double *out);
/* The reference's ghost-exchange seam (include/cedar/halo_exchanger_base.h; reference
 * include/cedar/kernels/halo_exchange.h:22-31: exchange_func(k, gf), exchange_sten(k, so), called from inside its MPI
 * kernels) over the peer-memory exchange.  k: the reference's level number (distributed levels + 1 = finest, 1 = the
 * agglomerated level); the pointer is the device array of an owned block (cdr2_dist_level_ptr); collective.
 * cdr2_dist_halo_exchanger returns the object (a cedar::halo_exchanger_base *, for kernel_manager::add_halo),
 * cdr2_dist_exchange makes one call through it. */
double *cdr2_dist_level_ptr(cdr2_dist *d, int block, int lvl, const char *which);
void *cdr2_dist_halo_exchanger(cdr2_dist *d);
int cdr2_dist_exchange(cdr2_dist *d, int k, int stencil, double *ptr);
/* one V-cycle on the resident x, b (enqueued; cdr2_dist_sync waits and reports exchange time-outs) */
int cdr2_dist_vcycle(cdr2_dist *d);
int cdr2_dist_sync(cdr2_dist *d);
int cdr2_dist_solve(cdr2_dist *d, int maxiter, double tol, double *hist, int hist_cap, int *niter, double *res0);
int cdr2_dist_time_vcycles(cdr2_dist *d, int warmup, int n, float *ms_per_cycle);
int cdr2_dist_vcycle_stream(cdr2_dist *d, int nrhs, const double *const *b, double *const *x);
/* diagnostic: device time of every operation of the cycle list, eager with a host synchronisation per operation */
int cdr2_dist_profile_ops(cdr2_dist *d, int n, float *ms_per_op, int cap, int *nops);
/* host-only views of the schedule (no device needed): operation list as JSON, block geometry
 * (x0,y0,nx,ny, wl,wr,wb,wt, coarse window ox,oy,iic,jjc) */
int cdr2_dist_plan_json(unsigned nx, unsigned ny, int px, int py, int nstencil, int nrelax_pre, int nrelax_post,
                        const char *relaxation, int agglomerate, int halo_width, int split_min, const char *what, char *out,
                        size_t cap, int *ndist);
/* rectangles of a split sweep (part 1 = interior, 2 = frames) of a block: 8 ints each (i0, j0, ni, nj, wi0, wi1, wj0, wj1),
 * the bmgx_rect of include/bmgx.h */
int cdr2_dist_sweep_rects(unsigned nx, unsigned ny, int px, int py, int ndist, int halo_width, int lvl, int block,
                          int ncolours, int part, int *out8, int cap, int *nrects);
int cdr2_dist_geometry(unsigned nx, unsigned ny, int px, int py, int ndist, int halo_width, int lvl, int block, int *out12);

/* ---- 3D (cdr3::solver): same protocol; nstencil = 4 (seven_pt) or 14 (xxvii_pt),
 * arrays (nx+2)(ny+2)(nz+2)[ndirs] ---- */
typedef struct cdr3_operator cdr3_operator;
typedef struct cdr3_solver cdr3_solver;
int cdr3_operator_create(int nstencil, unsigned nx, unsigned ny, unsigned nz, cdr3_operator **op);
double *cdr3_operator_data(cdr3_operator *op);
/* gallery: "poisson" (7-point), "fe" (27-point) */
int cdr3_operator_gallery(const char *name, unsigned nx, unsigned ny, unsigned nz, cdr3_operator **op);
void cdr3_operator_destroy(cdr3_operator *op);
int cdr3_solver_create(cdr3_operator *op, const char *config_json, cdr3_solver **solver);
void cdr3_solver_destroy(cdr3_solver *solver);
int cdr3_solver_nlevels(cdr3_solver *solver);
int cdr3_solver_level_shape(cdr3_solver *solver, int lvl, unsigned *nx, unsigned *ny, unsigned *nz);
int cdr3_solver_level_array(cdr3_solver *solver, int lvl, const char *which, double *out, size_t n, size_t *written);
int cdr3_solver_solve(cdr3_solver *solver, const double *b, double *x, double *hist, int hist_cap, int *niter,
                      double *res0);
int cdr3_solver_vcycle(cdr3_solver *solver, const double *b, double *x);
int cdr3_solver_buffers(cdr3_solver *solver, double **b, double **x);
int cdr3_solver_vcycle_pinned(cdr3_solver *solver);
int cdr3_solver_matvec(cdr3_solver *solver, const double *x, double *y);
int cdr3_solver_vcycle0_pinned(cdr3_solver *solver);   /* x0 = 0: uploads b only */
int cdr3_solver_vcycle_batch(cdr3_solver *solver, int nrhs, const double *const *b, double *const *x);
int cdr3_solver_time_vcycles(cdr3_solver *solver, const double *b, int warmup, int n, float *ms_per_cycle);
int cdr3_solver_time_solves(cdr3_solver *solver, const double *b, int warmup, int n, float *ms_per_solve, int *niter);
/* one plane-relaxation sweep (dir = "xy" | "xz" | "yz"; updown 0 = down, 1 = up) through the kernel registry, the call
 * sequence of the reference's test/3d/test_planes.cc:13-52 (build_kernel_manager, setup<plane_relax<dir>>, run);
 * config_json carries "plane-config" as in test/3d/test-planes-ser.json.  Host arrays; x is updated in place. */
int cdr3_plane_relax(cdr3_operator *op, const char *config_json, const char *dir, int updown, const double *b, double *x);

/* the solver's own device-resident b and x and one V-cycle on them (see cdr2_solver_resident_ptrs) */
int cdr3_solver_resident_ptrs(cdr3_solver *solver, double **b_dev, double **x_dev);
int cdr3_solver_vcycle_resident(cdr3_solver *solver, void *stream);

/* ---- slab-decomposed 3D solver across GPUs (cdr3::mpi::solver of the reference, include/cedar/3d/mpi/solver.h) ----
 * The nx x ny x nz grid is cut into `nslabs` slabs of xy planes, one per GPU (include/cedar/3d/mpi/plan.h); protocol
 * and semantics as cdr2_dist_*: create with the owned slab ids, set the own planes of the operator, move the IPC
 * handles, connect, setup; ghost planes are pulled out of the two neighbours' device memory over NVLink; levels with
 * at most `agglomerate` planes are solved replicated.  Point relaxation (7-point red-black fused, 27-point 8 colours).
 * Host arrays of a slab: (entries,) nzl+2, ny+2, nx+2 with the own planes in 1..nzl. */
typedef struct cdr3_dist cdr3_dist;
int cdr3_dist_create(unsigned nx, unsigned ny, unsigned nz, int nslabs, int nstencil, int nowned, const int *owned_slabs,
                     int nrelax_pre, int nrelax_post, int agglomerate, int halo_width, int cuda_graph, cdr3_dist **out);
void cdr3_dist_destroy(cdr3_dist *d);
int cdr3_dist_info(cdr3_dist *d, int *ndist, int *nlevels, int *halo_width, int *exchanges_per_cycle, int *ops_per_cycle);
int cdr3_dist_slab_extent(cdr3_dist *d, int lvl, int slab, int *nx, int *ny, int *z0, int *nz);
int cdr3_dist_export_handles(cdr3_dist *d, void *handles);
int cdr3_dist_connect(cdr3_dist *d, const void *handles);
int cdr3_dist_set_operator(cdr3_dist *d, int slab, const double *so);
int cdr3_dist_setup(cdr3_dist *d);
int cdr3_dist_upload(cdr3_dist *d, int slab, const char *which, const double *host);
int cdr3_dist_download(cdr3_dist *d, int slab, const char *which, double *host);
/* own planes of a level array -> host (entries, nzl, ny+2, nx+2); which = "A" | "P" */
int cdr3_dist_level_array(cdr3_dist *d, int slab, int lvl, const char *which, double *out);
/* as in 2D: own planes of any level <- host (entries, nzl, ny+2, nx+2); the whole halo-extended array -> host (entries, kk,
 * ny+2, nx+2); the reference's halo_exchanger_base seam over the slab exchange (k = distributed levels + 1 = finest) */
int cdr3_dist_upload_level(cdr3_dist *d, int slab, int lvl, const char *which, const double *host);
int cdr3_dist_extended_array(cdr3_dist *d, int slab, int lvl, const char *which, double *out);
double *cdr3_dist_level_ptr(cdr3_dist *d, int slab, int lvl, const char *which);
void *cdr3_dist_halo_exchanger(cdr3_dist *d);
int cdr3_dist_exchange(cdr3_dist *d, int k, int stencil, double *ptr);
int cdr3_dist_vcycle(cdr3_dist *d);
int cdr3_dist_sync(cdr3_dist *d);
int cdr3_dist_solve(cdr3_dist *d, int maxiter, double tol, double *hist, int hist_cap, int *niter, double *res0);
int cdr3_dist_time_vcycles(cdr3_dist *d, int warmup, int n, float *ms_per_cycle);
int cdr3_dist_plan_json(unsigned nx, unsigned ny, unsigned nz, int nslabs, int nstencil, int nrelax_pre, int nrelax_post,
                        int agglomerate, int halo_width, const char *what, char *out, size_t cap, int *ndist);
/* nx, ny, z0, nz, wb, wt, coarse window first plane and thickness */
int cdr3_dist_geometry(unsigned nx, unsigned ny, unsigned nz, int nslabs, int ndist, int halo_width, int lvl, int slab, int *out8);

#ifdef __cplusplus
}
#endif
#endif
