// bmgx2_lines.cu -- zebra line relaxation with batched tridiagonal solves.
// Reference: src/2d/ftn/BMG2_SymStd_SETUP_lines_x.f90:68-87, _y.f90:69-88 (+DPTTRF),
//            src/2d/ftn/BMG2_SymStd_relax_lines_x.f90:76-170, _y.f90:75-166 (+DPTTRS).
//
// SOR layout is the reference's: x-lines SOR(i,j,1)=D, SOR(i,j,2)=L sub-diagonal
// factor; y-lines stored transposed SOR(j,i,.) so every line is contiguous.
//
// Setup: the LDL' recurrence d(i+1) = d(i+1) - e(i)^2/d(i) is sequential per
// line; one lane per line, 32x32 tiles staged through shared memory so global
// traffic is coalesced.  Bitwise DPTTRF.
//
// Solve: with the factor known, DPTTS2 is two first-order linear recurrences,
//     y(i) = b(i) - e(i-1) y(i-1),      x(i) = y(i)/d(i) - e(i) x(i+1),
// evaluated here as warp-wide affine scans (parallel cyclic reduction of a
// bidiagonal system): one warp per line, 32 unknowns per step, 5 shuffle
// rounds, carry to the next 32.  The operation order differs from the
// sequential LAPACK loop, so parity with the oracle is to rounding (the tests
// use 1e-12 relative), not bitwise.
#include "bmgx_common.cuh"

// ---------------------------------------------------------------- setup ----
// line-major view: element (p, l) of line l at position p lives at
//   src: so_d[p*sp + l*sl] (diag), so_o[...] (off-diag, negated on load)
//   dst: sor[(size_t)l*dl + p] for D and sor[plane + l*dl + p] for L
__global__ void __launch_bounds__(128)
k_setup_lines(const double *__restrict__ so_d, const double *__restrict__ so_o, size_t sp, size_t sl,
              double *__restrict__ sor, size_t plane, size_t dl, unsigned n, unsigned nlines)
{
    // lines are 1..nlines-2 (0-based interior 1..), positions 1..n-2 interior
    unsigned l = blockIdx.x * blockDim.x + threadIdx.x + 1;
    if (l > nlines - 2) return;
    const double *pd = so_d + l * sl, *po = so_o + l * sl;
    double *D = sor + (size_t)l * dl, *L = sor + plane + (size_t)l * dl;
    // DPTTRF(n-2, D(1..), L(2..)) on d(p)=diag(p), e(p)=-off(p+1)
    double d = pd[(size_t)1 * sp];
    for (unsigned p = 1; p + 1 <= n - 2; p++) {
        double ei = -po[(size_t)(p + 1) * sp];
        double dn = pd[(size_t)(p + 1) * sp];
        D[p] = d;
        double e = ei / d;
        L[p + 1] = e;
        d = dn - e * ei;
    }
    D[n - 2] = d;
    L[1] = -po[(size_t)1 * sp];  // SOR(2,j,2) = -SO(2,j,KW): stored, never used by the solve
}

extern "C" int bmgx2_SETUP_lines_x(const double *so, double *sor, bmgx_len nx, bmgx_len ny, int nstncl, int jpn,
                                   bmgx_stream_t s)
{
    (void)nstncl;
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "SETUP_lines_x: periodic BCs not supported");
    if (nx < 3 || ny < 3) return BMGX_OK;
    size_t pl = (size_t)nx * ny;
    // line l = j (0-based), position p = i (0-based): src stride sp=1, sl=nx; dst line stride nx
    k_setup_lines<<<cdiv(ny - 2, 128), 128, 0, (cudaStream_t)s>>>(so + (KO - 1) * pl, so + (KW - 1) * pl, 1, nx, sor,
                                                                  pl, nx, nx, ny);
    BMGX_LAUNCH_CHECK("SETUP_lines_x");
    return BMGX_OK;
}

extern "C" int bmgx2_SETUP_lines_y(const double *so, double *sor, bmgx_len nx, bmgx_len ny, int nstncl, int jpn,
                                   bmgx_stream_t s)
{
    (void)nstncl;
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "SETUP_lines_y: periodic BCs not supported");
    if (nx < 3 || ny < 3) return BMGX_OK;
    size_t pl = (size_t)nx * ny;
    // line l = i, position p = j: src stride sp=nx, sl=1; dst SOR(j,i,.) line stride ny
    k_setup_lines<<<cdiv(nx - 2, 128), 128, 0, (cudaStream_t)s>>>(so + (KO - 1) * pl, so + (KS - 1) * pl, nx, 1, sor,
                                                                  pl, ny, ny, nx);
    BMGX_LAUNCH_CHECK("SETUP_lines_y");
    return BMGX_OK;
}

// ---------------------------------------------------------------- solve ----
// affine map f(t) = a*t + b ; compose(g after f)
struct Aff { double a, b; };

__device__ __forceinline__ Aff scan_up(Aff v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        double pa = __shfl_up_sync(0xffffffffu, v.a, o);
        double pb = __shfl_up_sync(0xffffffffu, v.b, o);
        if (lane >= o) { v.b = v.a * pb + v.b; v.a = v.a * pa; }
    }
    return v;
}
__device__ __forceinline__ Aff scan_down(Aff v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        double pa = __shfl_down_sync(0xffffffffu, v.a, o);
        double pb = __shfl_down_sync(0xffffffffu, v.b, o);
        if (lane + o < 32) { v.b = v.a * pb + v.b; v.a = v.a * pa; }
    }
    return v;
}

// One warp per line.  v holds the right-hand side on entry (positions 1..n-2,
// 0-based, contiguous) and the solution on exit.  D, L as stored by the setup.
__device__ __forceinline__ void warp_ldl_solve(double *v, const double *__restrict__ D,
                                               const double *__restrict__ L, unsigned n, int lane)
{
    unsigned m = n - 2;  // unknowns at positions 1..m
    double carry = 0.0;
    for (unsigned base = 1; base <= m; base += 32) {
        unsigned p = base + lane;
        Aff f;
        if (p <= m) { f.a = (p > 1) ? -L[p] : 0.0; f.b = v[p]; } else { f.a = 1.0; f.b = 0.0; }
        f = scan_up(f, lane);
        double y = f.a * carry + f.b;
        if (p <= m) v[p] = y;
        carry = __shfl_sync(0xffffffffu, y, 31);
    }
    __syncwarp();
    carry = 0.0;
    unsigned nseg = (m + 31) / 32;
    for (unsigned sgi = nseg; sgi-- > 0;) {
        unsigned p = 1 + sgi * 32 + lane;
        Aff f;
        if (p <= m) { f.a = (p < m) ? -L[p + 1] : 0.0; f.b = v[p] / D[p]; } else { f.a = 1.0; f.b = 0.0; }
        f = scan_down(f, lane);
        double x = f.a * carry + f.b;
        if (p <= m) v[p] = x;
        carry = __shfl_sync(0xffffffffu, x, 0);
    }
}

// x-lines: warp per line j of one colour; builds the right-hand side in place
// (rows j-1, j+1 belong to the other colour and are not written in this pass).
template <bool NINE>
__global__ void __launch_bounds__(128)
k_relax_lines_x(const double *__restrict__ so, const double *__restrict__ qf, double *q,
                const double *__restrict__ sor, unsigned II, unsigned JJ, unsigned jbeg)
{
    int lane = threadIdx.x & 31;
    unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    unsigned j = jbeg + 2 * w;  // F-index
    if (j > JJ - 1) return;
    size_t pl = (size_t)II * JJ;
    const double *ks = so + (KS - 1) * pl, *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
    size_t row = (size_t)(j - 1) * II;
    for (unsigned i0 = 1; i0 <= II - 2; i0 += 32) {  // 0-based i
        unsigned i = i0 + lane;
        if (i <= II - 2) {
            size_t c = row + i;
            double r = qf[c] + ks[c] * q[c - II] + ks[c + II] * q[c + II];
            if (NINE)
                r = r + ksw[c] * q[c - II - 1] + knw[c + 1] * q[c - II + 1] + knw[c + II] * q[c + II - 1]
                    + ksw[c + II + 1] * q[c + II + 1];
            q[c] = r;
        }
    }
    __syncwarp();
    warp_ldl_solve(q + row, sor + row, sor + pl + row, II, lane);
}

// y-lines, step 1: right-hand side, written transposed into B(j,i)
template <bool NINE>
__global__ void __launch_bounds__(256)
k_lines_y_rhs(const double *__restrict__ so, const double *__restrict__ qf, const double *__restrict__ q,
              double *__restrict__ B, unsigned II, unsigned JJ, unsigned ibeg)
{
    __shared__ double tile[32][33];
    // tile: 32 lines (i = ibeg + 2*(bx*32 + tx)) x 32 positions j
    unsigned tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    size_t pl = (size_t)II * JJ;
    const double *kw = so + (KW - 1) * pl, *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
    unsigned i = ibeg + 2 * (blockIdx.x * 32 + tx);  // F-index
    for (unsigned r = ty; r < 32; r += 8) {
        unsigned j = 2 + blockIdx.y * 32 + r;  // F-index
        double v = 0.0;
        if (i <= II - 1 && j <= JJ - 1) {
            size_t c = IX2(II, i, j);
            v = qf[c] + kw[c] * q[c - 1] + kw[c + 1] * q[c + 1];
            if (NINE)
                v = v + ksw[c] * q[c - II - 1] + knw[c + 1] * q[c - II + 1] + knw[c + II] * q[c + II - 1]
                    + ksw[c + II + 1] * q[c + II + 1];
        }
        tile[r][tx] = v;
    }
    __syncthreads();
    for (unsigned r = ty; r < 32; r += 8) {
        unsigned il = ibeg + 2 * (blockIdx.x * 32 + r);
        unsigned j = 2 + blockIdx.y * 32 + tx;
        if (il <= II - 1 && j <= JJ - 1) B[(size_t)(il - 1) * JJ + (j - 1)] = tile[tx][r];
    }
}

// step 2: warp per line on the transposed data
__global__ void __launch_bounds__(128)
k_lines_y_solve(double *B, const double *__restrict__ sor, unsigned II, unsigned JJ, unsigned ibeg)
{
    int lane = threadIdx.x & 31;
    unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    unsigned i = ibeg + 2 * w;
    if (i > II - 1) return;
    size_t row = (size_t)(i - 1) * JJ;
    warp_ldl_solve(B + row, sor + row, sor + (size_t)II * JJ + row, JJ, lane);
}

// step 3: scatter back Q(i,j) = B(j,i)
__global__ void __launch_bounds__(256)
k_lines_y_scatter(const double *__restrict__ B, double *__restrict__ q, unsigned II, unsigned JJ, unsigned ibeg)
{
    __shared__ double tile[32][33];
    unsigned tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (unsigned r = ty; r < 32; r += 8) {
        unsigned il = ibeg + 2 * (blockIdx.x * 32 + r);
        unsigned j = 2 + blockIdx.y * 32 + tx;
        tile[r][tx] = (il <= II - 1 && j <= JJ - 1) ? B[(size_t)(il - 1) * JJ + (j - 1)] : 0.0;
    }
    __syncthreads();
    unsigned i = ibeg + 2 * (blockIdx.x * 32 + tx);
    for (unsigned r = ty; r < 32; r += 8) {
        unsigned j = 2 + blockIdx.y * 32 + r;
        if (i <= II - 1 && j <= JJ - 1) q[IX2(II, i, j)] = tile[tx][r];
    }
}

extern "C" int bmgx2_relax_lines_x(int k, const double *so, const double *qf, double *q, const double *sor,
                                   double *b, bmgx_len ii, bmgx_len jj, int kf, int ifd, int nstncl,
                                   int irelax_sym, int updown, int jpn, bmgx_stream_t s)
{
    (void)nstncl; (void)b;
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_lines_x: periodic BCs not supported");
    if (ii < 3 || jj < 3) return BMGX_OK;
    bool fwd = (irelax_sym == BMGX_RELAX_NONSYM) || (updown == BMGX_DOWN && irelax_sym == BMGX_RELAX_SYM);
    for (int a = 0; a < 2; a++) {
        unsigned jbeg = fwd ? 3 - a : 2 + a;  // DOWN: odd lines (3,5,..) first (relax_lines_x.f90:83-99)
        if (jbeg > jj - 1) continue;
        unsigned nl = (jj - 1 - jbeg) / 2 + 1;
        if (k < kf || ifd != 1)
            k_relax_lines_x<true><<<cdiv((size_t)nl * 32, 128), 128, 0, (cudaStream_t)s>>>(so, qf, q, sor, ii, jj, jbeg);
        else
            k_relax_lines_x<false><<<cdiv((size_t)nl * 32, 128), 128, 0, (cudaStream_t)s>>>(so, qf, q, sor, ii, jj, jbeg);
        BMGX_LAUNCH_CHECK("relax_lines_x");
    }
    return BMGX_OK;
}

extern "C" int bmgx2_relax_lines_y(int k, const double *so, const double *qf, double *q, const double *sor,
                                   double *b, bmgx_len ii, bmgx_len jj, int kf, int ifd, int nstncl,
                                   int irelax_sym, int updown, int jpn, bmgx_stream_t s)
{
    (void)nstncl;
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_lines_y: periodic BCs not supported");
    if (!b) return bmgx_set_error(BMGX_ERR_ARG, "relax_lines_y: scratch b (ii*jj doubles) required");
    if (ii < 3 || jj < 3) return BMGX_OK;
    cudaStream_t st = (cudaStream_t)s;
    bool fwd = (irelax_sym == BMGX_RELAX_NONSYM) || (updown == BMGX_DOWN && irelax_sym == BMGX_RELAX_SYM);
    for (int a = 0; a < 2; a++) {
        unsigned ibeg = fwd ? 3 - a : 2 + a;
        if (ibeg > ii - 1) continue;
        unsigned nl = (ii - 1 - ibeg) / 2 + 1;
        dim3 g(cdiv(nl, 32), cdiv(jj - 2, 32));
        if (k < kf || ifd != 1) k_lines_y_rhs<true><<<g, 256, 0, st>>>(so, qf, q, b, ii, jj, ibeg);
        else k_lines_y_rhs<false><<<g, 256, 0, st>>>(so, qf, q, b, ii, jj, ibeg);
        BMGX_LAUNCH_CHECK("relax_lines_y/rhs");
        k_lines_y_solve<<<cdiv((size_t)nl * 32, 128), 128, 0, st>>>(b, sor, ii, jj, ibeg);
        BMGX_LAUNCH_CHECK("relax_lines_y/solve");
        k_lines_y_scatter<<<g, 256, 0, st>>>(b, q, ii, jj, ibeg);
        BMGX_LAUNCH_CHECK("relax_lines_y/scatter");
    }
    return BMGX_OK;
}
