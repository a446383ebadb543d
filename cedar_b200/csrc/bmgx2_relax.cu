// bmgx2_relax.cu -- 2D multicolour point Gauss-Seidel (reference:
// src/2d/ftn/BMG2_SymStd_relax_GS.f90:68-135) and SETUP_recip.
//
// Colour order (F-index parity, first interior index is 2):
//   9-pt: (j even,i even) (j even,i odd) (j odd,i even) (j odd,i odd) on DOWN,
//         exact reverse on UP (relax_GS.f90:78-114);
//   5-pt: red-black, pass JO=2 then 3 on DOWN, IBEG = mod(J+JO,2)+2
//         (relax_GS.f90:120-135).
// Term order inside the update is the reference's, and the file is compiled
// with -fmad=false, so results are bitwise those of the oracle.
#include "bmgx_common.cuh"

__global__ void k_setup_recip(const double *__restrict__ so, double *__restrict__ sor, unsigned nx, unsigned ny)
{
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x + 2;
    unsigned j = blockIdx.y + 2;
    if (i <= nx - 1 && j <= ny - 1) sor[IX3(nx, ny, i, j, MSOR)] = 1.0 / so[IX3(nx, ny, i, j, KO)];
}

extern "C" int bmgx2_SETUP_recip(const double *so, double *sor, bmgx_len nx, bmgx_len ny, int nstncl, int nsor_v,
                                 bmgx_stream_t s)
{
    (void)nstncl;
    if (nsor_v < 2) return bmgx_set_error(BMGX_ERR_ARG, "SETUP_recip: nsor_v < 2");
    if (nx < 3 || ny < 3) return BMGX_OK;
    if (ny - 2 > 65535) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "SETUP_recip: more than 65535 interior rows (one row per CTA row of the launch grid)");
    dim3 g(cdiv(nx - 2, 256), ny - 2);
    k_setup_recip<<<g, 256, 0, (cudaStream_t)s>>>(so, sor, nx, ny);
    BMGX_LAUNCH_CHECK("SETUP_recip");
    return BMGX_OK;
}

// ---- in-place, one launch per colour ---------------------------------------
template <bool NINE>
__global__ void __launch_bounds__(128)
k_relax_colour(const double *__restrict__ so, const double *__restrict__ qf, double *q,
               const double *__restrict__ sor, unsigned II, unsigned JJ,
               unsigned ibeg9, unsigned jbeg, unsigned jstep, unsigned jo5)
{
    unsigned j = jbeg + jstep * blockIdx.y;
    unsigned ibeg = NINE ? ibeg9 : ((j + jo5) % 2 + 2);
    unsigned i = ibeg + 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    if (i > II - 1 || j > JJ - 1) return;
    size_t c = IX2(II, i, j);
    size_t pl = (size_t)II * JJ;
    const double *kw = so + (KW - 1) * pl, *ks = so + (KS - 1) * pl;
    double acc = qf[c] + kw[c] * q[c - 1] + kw[c + 1] * q[c + 1] + ks[c] * q[c - II] + ks[c + II] * q[c + II];
    if (NINE) {
        const double *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
        acc = acc + ksw[c] * q[c - II - 1] + knw[c + 1] * q[c - II + 1] + knw[c + II] * q[c + II - 1]
              + ksw[c + II + 1] * q[c + II + 1];
    }
    q[c] = acc * sor[(MSOR - 1) * pl + c];
}

// 9-point: the two colours of one row class (they differ in the parity of i and are consecutive in the sweep order) in
// ONE launch.  A CTA owns a whole row: the second colour sees the first colour's new values of its own row after a
// barrier, every other neighbour (rows j-1, j+1) belongs to the other row class, which this launch does not touch.
// Both colours pull the same sectors of every array (the second pass from L1/L2), so a sweep moves half the DRAM bytes of
// four single-colour launches.  PERX: the row's ghost columns are refreshed after each colour, as the reference does
// (relax_GS.f90:163-166).  Rows are mapped to blockIdx.x (no 65535 limit).
template <bool PERX>
__global__ void __launch_bounds__(256)
k_relax_rowpair9(const double *__restrict__ so, const double *__restrict__ qf, double *q, const double *__restrict__ sor,
                 unsigned II, unsigned JJ, unsigned ifirst, unsigned jbeg)
{
    const unsigned j = jbeg + 2 * blockIdx.x;
    if (j > JJ - 1) return;
    const size_t pl = (size_t)II * JJ, row = (size_t)(j - 1) * II;
    const double *kw = so + (KW - 1) * pl, *ks = so + (KS - 1) * pl, *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
    const double *rs = sor + (MSOR - 1) * pl;
#pragma unroll 1
    for (int pass = 0; pass < 2; pass++) {
        const unsigned ibeg = pass == 0 ? ifirst : 5 - ifirst;   // 2 <-> 3
        for (unsigned i = ibeg + 2 * threadIdx.x; i <= II - 1; i += 2 * blockDim.x) {
            const size_t c = row + (i - 1);
            double acc = qf[c] + kw[c] * q[c - 1] + kw[c + 1] * q[c + 1] + ks[c] * q[c - II] + ks[c + II] * q[c + II];
            acc = acc + ksw[c] * q[c - II - 1] + knw[c + 1] * q[c - II + 1] + knw[c + II] * q[c + II - 1]
                  + ksw[c + II + 1] * q[c + II + 1];
            q[c] = acc * rs[c];
        }
        __syncthreads();
        if (PERX) {
            if (threadIdx.x == 0) { q[row] = q[row + II - 2]; q[row + II - 1] = q[row + 1]; }
            __syncthreads();
        }
    }
}

// Periodic boundaries (BMG2_SymStd_relax_GS.f90:139-224): the reference refreshes the x ghosts of a row after each
// colour of that row and the y ghost rows once, after the whole sweep.  One colour never reads its own rows' x
// ghosts written in the same pass, so refreshing them for all rows of the pass right after the pass is the same.
// passes: bit n set = run the n-th colour pass of the sweep (9-point: 4 passes, 5-point: 2)
static int relax_gs_passes(int k, const double *so, const double *qf, double *q, const double *sor, bmgx_len ii,
                           bmgx_len jj, int kf, int ifd, int irelax_sym, int updown, int jpn, unsigned passes,
                           bmgx_stream_t s)
{
    const int ipn = jpn < 0 ? -jpn : jpn;
    const bool periodic = !bmgx_nonperiodic(jpn);
    if (periodic && (ipn < 1 || ipn > 3)) return bmgx_set_error(BMGX_ERR_ARG, "relax_GS: unknown boundary code");
    if (periodic && passes != 0xfu)
        return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_GS_pass: periodic BCs not supported on the decomposed path");
    const bool perx = periodic && (ipn == 2 || ipn == 3), pery = periodic && (ipn == 1 || ipn == 3);
    if (ii < 3 || jj < 3) return BMGX_OK;
    const bool pair9 = (k < kf || ifd != 1) && passes == 0xfu && ii - 1 >= 3;
    if (!pair9 && jj - 2 > 65535)
        return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_GS (in place, 5-point or single passes): more than 65535 interior rows; use relax_GS_oop, which walks the rows in bands");
    cudaStream_t st = (cudaStream_t)s;
    bool fwd = (irelax_sym == BMGX_RELAX_NONSYM) || (updown == BMGX_DOWN && irelax_sym == BMGX_RELAX_SYM);
    unsigned I1 = ii - 1, J1 = jj - 1;
    if (pair9) {
        // whole sweep: one launch per row class, both x colours in it
        for (int a = 0; a < 2; a++) {
            unsigned jbeg = fwd ? 2 + a : 3 - a;
            if (jbeg > J1) continue;
            unsigned nrow = (J1 - jbeg) / 2 + 1;
            if (perx) k_relax_rowpair9<true><<<nrow, 256, 0, st>>>(so, qf, q, sor, ii, jj, fwd ? 2 : 3, jbeg);
            else k_relax_rowpair9<false><<<nrow, 256, 0, st>>>(so, qf, q, sor, ii, jj, fwd ? 2 : 3, jbeg);
            BMGX_LAUNCH_CHECK("relax_GS/9 pair");
        }
    } else if (k < kf || ifd != 1) {
        for (int a = 0; a < 2; a++) {
            unsigned jbeg = fwd ? 2 + a : 3 - a;
            if (jbeg > J1) continue;
            unsigned nrow = (J1 - jbeg) / 2 + 1;
            for (int b = 0; b < 2; b++) {
                if (!(passes & (1u << (2 * a + b)))) continue;
                unsigned ibeg = fwd ? 2 + b : 3 - b;
                if (ibeg > I1) continue;
                unsigned ncol = (I1 - ibeg) / 2 + 1;
                dim3 g(cdiv(ncol, 128), nrow);
                k_relax_colour<true><<<g, 128, 0, st>>>(so, qf, q, sor, ii, jj, ibeg, jbeg, 2, 0);
                BMGX_LAUNCH_CHECK("relax_GS/9");
                if (perx) {
                    if (bmgx_wrap_x(q, ii, jbeg, 2, nrow, st)) return BMGX_ERR_CUDA;
                }
            }
        }
    } else {
        for (int a = 0; a < 2; a++) {
            if (!(passes & (1u << a))) continue;
            unsigned jo = fwd ? 2 + a : 3 - a;
            unsigned ncol = (I1 - 2) / 2 + 1;  // upper bound on points per row of one colour
            dim3 g(cdiv(ncol, 128), J1 - 1);
            k_relax_colour<false><<<g, 128, 0, st>>>(so, qf, q, sor, ii, jj, 0, 2, 1, jo);
            BMGX_LAUNCH_CHECK("relax_GS/5");
            if (perx) {
                if (bmgx_wrap_x(q, ii, 2, 1, J1 - 1, st)) return BMGX_ERR_CUDA;
            }
        }
    }
    if (pery) {
        if (bmgx_wrap_y(q, ii, jj, st)) return BMGX_ERR_CUDA;
    }
    return BMGX_OK;
}

extern "C" int bmgx2_relax_GS(int k, const double *so, const double *qf, double *q, const double *sor, bmgx_len ii,
                              bmgx_len jj, int kf, int ifd, int nstncl, int nsorv, int irelax_sym, int updown,
                              int jpn, bmgx_stream_t s)
{
    (void)nstncl; (void)nsorv;
    return relax_gs_passes(k, so, qf, q, sor, ii, jj, kf, ifd, irelax_sym, updown, jpn, 0xfu, s);
}

// One colour pass of the sweep (pass = 0..3 for 9-point, 0..1 for 5-point, in sweep order):
// the block-decomposed solver exchanges ghost values between the passes, like the
// reference's MPI kernel (src/2d/ftn/mpi/BMG2_SymStd_relax_GS.f90:101-134).
extern "C" int bmgx2_relax_GS_pass(int k, const double *so, const double *qf, double *q, const double *sor, bmgx_len ii,
                                   bmgx_len jj, int kf, int ifd, int nstncl, int nsorv, int irelax_sym, int updown,
                                   int jpn, int pass, bmgx_stream_t s)
{
    (void)nstncl; (void)nsorv;
    if (pass < 0 || pass > 3) return bmgx_set_error(BMGX_ERR_ARG, "relax_GS_pass: pass out of range");
    return relax_gs_passes(k, so, qf, q, sor, ii, jj, kf, ifd, irelax_sym, updown, jpn, 1u << pass, s);
}
