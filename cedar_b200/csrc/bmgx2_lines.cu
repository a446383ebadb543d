// bmgx2_lines.cu -- zebra line relaxation with batched tridiagonal solves.
// Reference: src/2d/ftn/BMG2_SymStd_SETUP_lines_x.f90:68-87, _y.f90:69-88 (+DPTTRF),
//            src/2d/ftn/BMG2_SymStd_relax_lines_x.f90:76-170, _y.f90:75-166 (+DPTTRS).
//
// SOR layout is the reference's: x-lines SOR(i,j,1)=D, SOR(i,j,2)=L sub-diagonal
// factor; y-lines stored transposed SOR(j,i,.) so every line is contiguous.
//
// Setup: the LDL' recurrence d(i+1) = d(i+1) - e(i)^2/d(i) is sequential per
// line; one lane per line, operands fetched 16 positions ahead of the chain.  The kernel is bound by that chain, not
// by memory: n dependent fp64 divisions per line (~200 cycles each: 4097 x 200 cycles = 0.4 ms at 1.9 GHz) however
// the operands arrive; 0.80 ms at 4097^2 = 10 % of the copy roofline.  A variant that turned 32 x 32 tiles in shared
// memory so that loads AND stores were coalesced (round 2) kept DPTTRF's bits but ran in 1.26 ms -- the transposes sit
// on the chain's critical path -- and was dropped.  Going faster means evaluating the continued fraction as a scan of
// 2 x 2 matrices, which gives up bitwise DPTTRF.  Setup only.
//
// Solve: with the factor known, DPTTS2 is two first-order linear recurrences,
//     y(i) = b(i) - e(i-1) y(i-1),      x(i) = y(i)/d(i) - e(i) x(i+1),
// i.e. a chain of affine maps.  One CTA per line, the line staged in shared
// memory (right-hand side and sub-diagonal factor, chunk-padded so both the
// coalesced "interleaved" phases and the per-thread "chunk" phases are free of
// bank conflicts: chunks have odd length):
//   1. interleaved: build the right-hand side (reference term order), stage L;
//   2. chunk: every thread composes the affine maps of its C consecutive
//      unknowns, a block-wide scan of the chunk maps (warp shuffles + one
//      cross-warp step) yields the value entering each chunk, and the thread
//      re-runs the plain sequential recurrence over its chunk from that value;
//   3. interleaved: divide by D (coalesced read);  4. chunk: the same, backwards;
//   5. interleaved: store the line.
// Inside a chunk the arithmetic is DPTTS2's; only the chunk-entry values are
// rounded differently, so parity with the oracle is to rounding (tests: 1e-12
// relative), not bitwise.  Lines too long for shared memory (> ~13000 unknowns)
// fall back to the warp-per-line scan kernels further down.
#include "bmgx_common.cuh"

// ---------------------------------------------------------------- setup ----
// line-major view: element (p, l) of line l at position p lives at
//   src: so_d[p*sp + l*sl] (diag), so_o[...] (off-diag, negated on load)
//   dst: sor[(size_t)l*dl + p] for D and sor[plane + l*dl + p] for L
__global__ void __launch_bounds__(128)
k_setup_lines(const double *__restrict__ so_d, const double *__restrict__ so_o, size_t sp, size_t sl,
              double *__restrict__ sor, size_t plane, size_t dl, unsigned n, unsigned nlines, int per)
{
    // lines are 1..nlines-2 (0-based interior 1..), positions 1..n-2 interior
    unsigned l = blockIdx.x * blockDim.x + threadIdx.x + 1;
    if (l > nlines - 2) return;
    const double *pd = so_d + l * sl, *po = so_o + l * sl;
    double *D = sor + (size_t)l * dl, *L = sor + plane + (size_t)l * dl;
    // DPTTRF(n-2, D(1..), L(2..)) on d(p)=diag(p), e(p)=-off(p+1).  The recurrence is sequential, its operands
    // are not: they are fetched 16 positions ahead so that the chain never waits for memory
    // lines periodic in their own direction (SETUP_lines_x.f90:76-83): the wrap couplings SO(2,.,KW) and SO(Nx,.,KW) join
    // the first and the last diagonal entry before the factorisation
    const unsigned last = n - 2;
    double d = pd[(size_t)1 * sp];
    if (per) {
        d = d + po[(size_t)1 * sp];
        if (last == 1) d = d + po[(size_t)(n - 1) * sp];
    }
    const double wrap_r = per ? po[(size_t)(n - 1) * sp] : 0.0;
    constexpr unsigned U = 16;
    for (unsigned p0 = 1; p0 + 1 <= n - 2; p0 += U) {
        double eo[U], dn[U];
#pragma unroll
        for (unsigned u = 0; u < U; u++) {
            const unsigned p = p0 + u;
            const bool ok = p + 1 <= n - 2;
            eo[u] = ok ? __ldg(po + (size_t)(p + 1) * sp) : 0.0;
            dn[u] = ok ? __ldg(pd + (size_t)(p + 1) * sp) : 0.0;
        }
#pragma unroll
        for (unsigned u = 0; u < U; u++) {
            const unsigned p = p0 + u;
            if (p + 1 <= n - 2) {
                const double ei = -eo[u];
                D[p] = d;
                const double e = ei / d;
                L[p + 1] = e;
                d = ((per && p + 1 == last) ? dn[u] + wrap_r : dn[u]) - e * ei;
            }
        }
    }
    D[n - 2] = d;
    L[1] = -po[(size_t)1 * sp];  // SOR(2,j,2) = -SO(2,j,KW): stored, never used by the solve
}

extern "C" int bmgx2_SETUP_lines_x(const double *so, double *sor, bmgx_len nx, bmgx_len ny, int nstncl, int jpn,
                                   bmgx_stream_t s)
{
    (void)nstncl;
    if (nx < 3 || ny < 3) return BMGX_OK;
    const int ipn = jpn < 0 ? -jpn : jpn, per = (ipn == 2 || ipn == 3) ? 1 : 0;
    size_t pl = (size_t)nx * ny;
    // line l = j (0-based), position p = i (0-based): src stride sp=1, sl=nx; dst line stride nx
    k_setup_lines<<<cdiv(ny - 2, 128), 128, 0, (cudaStream_t)s>>>(so + (KO - 1) * pl, so + (KW - 1) * pl, 1, nx, sor,
                                                                  pl, nx, nx, ny, per);
    BMGX_LAUNCH_CHECK("SETUP_lines_x");
    return BMGX_OK;
}

extern "C" int bmgx2_SETUP_lines_y(const double *so, double *sor, bmgx_len nx, bmgx_len ny, int nstncl, int jpn,
                                   bmgx_stream_t s)
{
    (void)nstncl;
    if (nx < 3 || ny < 3) return BMGX_OK;
    const int ipn = jpn < 0 ? -jpn : jpn, per = (ipn == 1 || ipn == 3) ? 1 : 0;
    size_t pl = (size_t)nx * ny;
    // line l = i, position p = j: src stride sp=nx, sl=1; dst SOR(j,i,.) line stride ny
    k_setup_lines<<<cdiv(nx - 2, 128), 128, 0, (cudaStream_t)s>>>(so + (KO - 1) * pl, so + (KS - 1) * pl, nx, 1, sor,
                                                                  pl, ny, ny, nx, per);
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

// ---- block-per-line solver ----------------------------------------------------
// The line lives in shared memory at slot p-1 (positions p = 1..m).  Thread t owns
// the chunk of C consecutive positions starting at 1 + t*C; C is odd, so the
// per-thread chunk walks (stride C doubles between threads) are bank-conflict free
// without padding, and the interleaved phases are trivially so.

// exclusive block scan of affine maps in thread order (map of thread 0 applied
// first): returns the composition of the maps of all threads before this one
__device__ __forceinline__ Aff block_scan_excl(Aff v, Aff *wsum, Aff *wpre)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    Aff inc = scan_up(v, lane);
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        Aff w;
        if (lane < nw) w = wsum[lane]; else { w.a = 1.0; w.b = 0.0; }
        Aff wi = scan_up(w, lane);
        Aff ex;
        ex.a = __shfl_up_sync(0xffffffffu, wi.a, 1);
        ex.b = __shfl_up_sync(0xffffffffu, wi.b, 1);
        if (lane == 0) { ex.a = 1.0; ex.b = 0.0; }
        wpre[lane] = ex;
    }
    __syncthreads();
    const Aff pre = wpre[warp];
    Aff e;
    e.a = __shfl_up_sync(0xffffffffu, inc.a, 1);
    e.b = __shfl_up_sync(0xffffffffu, inc.b, 1);
    if (lane == 0) { e.a = 1.0; e.b = 0.0; }
    Aff r;
    r.a = e.a * pre.a;
    r.b = e.a * pre.b + e.b;
    return r;
}

// one bidiagonal sweep over the staged line: forward  v(p) = v(p) - L(p)   v(p-1)  (p = 2..m)
//                                            backward v(p) = v(p) - L(p+1) v(p+1)  (p = m-1..1)
// sv[p-1] = v(p), se[p-1] = L(p)
template <bool BWD>
__device__ __forceinline__ void block_bidiag(double *sv, const double *se, unsigned m, unsigned C, Aff *wsum, Aff *wpre)
{
    const unsigned t = threadIdx.x;
    const unsigned ch = BWD ? blockDim.x - 1 - t : t;   // backward: thread 0 owns the last chunk
    const unsigned p0 = 1 + ch * C;
    const bool act = p0 <= m;
    const unsigned p1 = act ? min(m, p0 + C - 1) : 0;
    Aff f;
    f.a = 1.0; f.b = 0.0;
    if (act) {
        if (!BWD) {
            for (unsigned p = p0; p <= p1; p++) {
                const double l = p > 1 ? se[p - 1] : 0.0;
                f.a = -(l * f.a);
                f.b = sv[p - 1] - l * f.b;
            }
        } else {
            for (unsigned p = p1; p >= p0; p--) {
                const double l = p < m ? se[p] : 0.0;
                f.a = -(l * f.a);
                f.b = sv[p - 1] - l * f.b;
            }
        }
    }
    const Aff pre = block_scan_excl(f, wsum, wpre);
    if (act) {
        double y = pre.b;   // the value entering the line is 0
        if (!BWD) {
            for (unsigned p = p0; p <= p1; p++) {
                const double l = p > 1 ? se[p - 1] : 0.0;
                y = sv[p - 1] - l * y;
                sv[p - 1] = y;
            }
        } else {
            for (unsigned p = p1; p >= p0; p--) {
                const double l = p < m ? se[p] : 0.0;
                y = sv[p - 1] - l * y;
                sv[p - 1] = y;
            }
        }
    }
}

// the staged solve: sv holds the right-hand side, se the L factor; D read coalesced
__device__ __forceinline__ void block_ldl_solve(double *sv, const double *se, const double *__restrict__ D,
                                                unsigned m, unsigned C, Aff *wsum, Aff *wpre)
{
    __syncthreads();
    block_bidiag<false>(sv, se, m, C, wsum, wpre);
    __syncthreads();
    for (unsigned p0 = 1 + threadIdx.x; p0 <= m; p0 += 4 * blockDim.x) {
        double d[4];
#pragma unroll
        for (int u = 0; u < 4; u++) { const unsigned p = p0 + u * blockDim.x; d[u] = __ldg(D + (p <= m ? p : 1)); }
#pragma unroll
        for (int u = 0; u < 4; u++) { const unsigned p = p0 + u * blockDim.x; if (p <= m) sv[p - 1] = sv[p - 1] / d[u]; }
    }
    __syncthreads();
    block_bidiag<true>(sv, se, m, C, wsum, wpre);
    __syncthreads();
}

// Lines periodic in their own direction (relax_lines_x.f90:209-224): the factor is that of the tridiagonal part with the
// wrap couplings added to its end diagonals; Sherman-Morrison puts the wrap back: a second solve with the couplings as
// right-hand side (sz), then x = y - (y_first + y_last) / (1 + z_first + z_last) * z.  sv holds y on entry, x on exit.
__device__ __forceinline__ void block_periodic_fix(double *sv, double *sz, const double *se, const double *__restrict__ D,
                                                   unsigned m, unsigned C, double wl, double wr, Aff *wsum, Aff *wpre)
{
    for (unsigned p = 1 + threadIdx.x; p <= m; p += blockDim.x) sz[p - 1] = 0.0;
    __syncthreads();
    if (threadIdx.x == 0) { sz[0] = -wl; sz[m - 1] = -wr; }
    block_ldl_solve(sz, se, D, m, C, wsum, wpre);
    const double alpha = sz[0] + sz[m - 1];
    double beta = sv[0] + sv[m - 1];
    beta = beta / (1.0 + alpha);
    __syncthreads();
    for (unsigned p = 1 + threadIdx.x; p <= m; p += blockDim.x) sv[p - 1] = sv[p - 1] - beta * sz[p - 1];
    __syncthreads();
}

// pull n doubles starting at base into L2 (one request per 32-byte sector, no destination
// register, never waited on): the demand loads that follow are scheduled by ptxas in small
// groups, each group paying one full memory latency -- after the prefetch that is an L2 hit
__device__ __forceinline__ void l2_prefetch_row(const double *base, unsigned n)
{
    for (unsigned o = 4 * threadIdx.x; o < n; o += 4 * blockDim.x) asm volatile("prefetch.global.L2 [%0];" ::"l"(base + o));
}

// x-lines: one CTA per line j of one colour (rows j-1, j+1 belong to the other colour
// and are not written in this pass)
template <bool NINE>
__global__ void __launch_bounds__(1024)
k_linex_block(const double *__restrict__ so, const double *__restrict__ qf, double *q,
              const double *__restrict__ sor, unsigned II, unsigned JJ, unsigned jbeg, unsigned C, int per)
{
    extern __shared__ double lsm[];
    __shared__ Aff wsum[32], wpre[32];
    const unsigned m = II - 2;
    double *sv = lsm, *se = lsm + m + 1;
    const unsigned j = jbeg + 2 * blockIdx.x;  // F-index
    const size_t pl = (size_t)II * JJ;
    q += blockIdx.y * pl; qf += blockIdx.y * pl;   // batch entry
    const double *ks = so + (KS - 1) * pl, *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
    const size_t row = (size_t)(j - 1) * II;
    const double *Lf = sor + pl + row;
    l2_prefetch_row(qf + row, II); l2_prefetch_row(q + row - II, II); l2_prefetch_row(q + row + II, II);
    l2_prefetch_row(ks + row, 2 * II);
    if (NINE) { l2_prefetch_row(ksw + row, 2 * II); l2_prefetch_row(knw + row, 2 * II); }
    l2_prefetch_row(Lf, II); l2_prefetch_row(sor + row, II);
    // rows j-1 and j+1 are not written by this launch: every operand is a read-only global load,
    // issued for two positions at once before any of them is consumed (the loop is latency bound)
    const double *qs = q + row - II, *qn = q + row + II;
    for (unsigned i0 = 1 + threadIdx.x; i0 <= m; i0 += 2 * blockDim.x) {  // 0-based i = position
        double r[2], l[2];
        double v[2][NINE ? 13 : 5];
#pragma unroll
        for (int u = 0; u < 2; u++) {
            const unsigned i = i0 + u * blockDim.x;
            const bool ok = i <= m;
            const size_t c = row + (ok ? i : 1);
            const unsigned ii = ok ? i : 1;
            v[u][0] = __ldg(qf + c); v[u][1] = __ldg(ks + c); v[u][2] = __ldg(qs + ii);
            v[u][3] = __ldg(ks + c + II); v[u][4] = __ldg(qn + ii);
            if (NINE) {
                v[u][5] = __ldg(ksw + c); v[u][6] = __ldg(qs + ii - 1); v[u][7] = __ldg(knw + c + 1);
                v[u][8] = __ldg(qs + ii + 1); v[u][9] = __ldg(knw + c + II); v[u][10] = __ldg(qn + ii - 1);
                v[u][11] = __ldg(ksw + c + II + 1); v[u][12] = __ldg(qn + ii + 1);
            }
            l[u] = __ldg(Lf + ii);
        }
#pragma unroll
        for (int u = 0; u < 2; u++) {
            r[u] = v[u][0] + v[u][1] * v[u][2] + v[u][3] * v[u][4];
            if (NINE) r[u] = r[u] + v[u][5] * v[u][6] + v[u][7] * v[u][8] + v[u][9] * v[u][10] + v[u][11] * v[u][12];
            const unsigned i = i0 + u * blockDim.x;
            if (i <= m) { sv[i - 1] = r[u]; se[i - 1] = l[u]; }
        }
    }
    block_ldl_solve(sv, se, sor + row, m, C, wsum, wpre);
    if (per) {
        const double *kw = so + (KW - 1) * pl;
        block_periodic_fix(sv, lsm + 2 * (m + 1), se, sor + row, m, C, kw[row + 1], kw[row + II - 1], wsum, wpre);
    }
    for (unsigned i = 1 + threadIdx.x; i <= m; i += blockDim.x) q[row + i] = sv[i - 1];
}

// y-lines, step 2: one CTA per line i on the transposed right-hand side B(j,i), in place
__global__ void __launch_bounds__(1024)
k_liney_block(double *__restrict__ B, const double *__restrict__ sor, unsigned II, unsigned JJ, unsigned ibeg,
              unsigned C, const double *__restrict__ ks_per)
{
    extern __shared__ double lsm[];
    __shared__ Aff wsum[32], wpre[32];
    const unsigned m = JJ - 2;
    double *sv = lsm, *se = lsm + m + 1;
    const unsigned i = ibeg + 2 * blockIdx.x;  // F-index
    const size_t row = (size_t)(i - 1) * JJ;
    const double *Lf = sor + (size_t)II * JJ + row;
    B += (size_t)blockIdx.y * II * JJ;   // batch entry
    l2_prefetch_row(B + row, JJ); l2_prefetch_row(Lf, JJ); l2_prefetch_row(sor + row, JJ);
    for (unsigned p0 = 1 + threadIdx.x; p0 <= m; p0 += 4 * blockDim.x) {
        double bv[4], lv[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const unsigned p = p0 + u * blockDim.x, pp = p <= m ? p : 1;
            bv[u] = __ldcg(B + row + pp);
            lv[u] = __ldg(Lf + pp);
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const unsigned p = p0 + u * blockDim.x;
            if (p <= m) { sv[p - 1] = bv[u]; se[p - 1] = lv[u]; }
        }
    }
    block_ldl_solve(sv, se, sor + row, m, C, wsum, wpre);
    if (ks_per)   // the KS plane of the operator: SO(i,2,KS), SO(i,JJ,KS) are the wrap couplings of line i (relax_lines_y.f90:215-219)
        block_periodic_fix(sv, lsm + 2 * (m + 1), se, sor + row, m, C, ks_per[(size_t)II + (i - 1)],
                           ks_per[(size_t)(JJ - 1) * II + (i - 1)], wsum, wpre);
    for (unsigned p = 1 + threadIdx.x; p <= m; p += blockDim.x) B[row + p] = sv[p - 1];
}

// odd chunk length C and CTA width for a line of m unknowns; false if it does not fit.
// Shared memory (16 B per unknown) bounds the lines resident per SM; the CTA width is
// chosen so that those CTAs together keep ~1280 threads (<= 48 registers each) on the SM.
static bool line_plan(unsigned m, unsigned *C, unsigned *threads, size_t *smem, unsigned nvec = 2)
{
    *smem = nvec * (size_t)(m + 1) * sizeof(double);
    if (*smem > 200 * 1024) return false;
    unsigned resident = (unsigned)((220 * 1024) / (*smem + 1024));
    if (resident < 1) resident = 1;
    if (resident > 16) resident = 16;
    unsigned want = 1280 / resident;
    if (want > 1024) want = 1024;
    if (want < 64) want = 64;
    unsigned c = (m + want - 1) / want;
    if (c < 3 && m > 32) c = 3;
    c |= 1u;
    unsigned chunks = (m + c - 1) / c;
    *C = c;
    *threads = ((chunks + 31) / 32) * 32;
    return *threads <= 1024;
}

template <bool NINE>
static cudaError_t launch_linex(unsigned C, unsigned nl, unsigned threads, size_t smem, cudaStream_t st, const double *so,
                                const double *qf, double *q, const double *sor, unsigned II, unsigned JJ, unsigned jbeg,
                                int per = 0)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_linex_block<NINE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    k_linex_block<NINE><<<dim3(nl, bmgx_batch()), threads, smem, st>>>(so, qf, q, sor, II, JJ, jbeg, C, per);
    return cudaSuccess;
}

static cudaError_t launch_liney(unsigned C, unsigned nl, unsigned threads, size_t smem, cudaStream_t st, double *B,
                                const double *sor, unsigned II, unsigned JJ, unsigned ibeg, const double *ks_per = nullptr)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_liney_block, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    k_liney_block<<<dim3(nl, bmgx_batch()), threads, smem, st>>>(B, sor, II, JJ, ibeg, C, ks_per);
    return cudaSuccess;
}

// ---- warp-per-line fallback (lines too long for shared memory) ---------------
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
    qf += blockIdx.z * pl; q += blockIdx.z * pl; B += blockIdx.z * pl;   // batch entry
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
    B += (size_t)blockIdx.z * II * JJ; q += (size_t)blockIdx.z * II * JJ;   // batch entry
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

// ---- y-lines, both colours without touching q in between ----------------------------------------------------
// After the first colour's solve the scratch B holds the new values of those lines, line-contiguous; the second colour's
// right-hand sides need exactly them (columns i-1, i+1).  Instead of scattering them into q (half-written sectors:
// 303 MB moved for 134 MB of payload, ncu) and reading q back, the second colour takes them straight from B through a
// shared-memory tile, and ONE transpose of the whole of B (every line is solved by then) writes q in full sectors.
// Same operands, same expression: bit for bit the result of the three-kernel form.
template <bool NINE>
__global__ void __launch_bounds__(256)
k_lines_y_rhs_fromB(const double *__restrict__ so, const double *__restrict__ qf, const double *__restrict__ q,
                    double *B, unsigned II, unsigned JJ, unsigned ibeg)
{
    __shared__ double nb[33][35];   // neighbour lines ibeg-1+2k (k = 0..32) x positions j0-1 .. j0+32
    __shared__ double tile[32][33];
    const unsigned tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    const size_t pl = (size_t)II * JJ;
    const double *kw = so + (KW - 1) * pl, *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
    const unsigned i0 = ibeg + 2 * blockIdx.x * 32;   // first line of the tile (F-index)
    const unsigned j0 = 2 + blockIdx.y * 32;          // first position of the tile (F-index)
    // new values of the other colour's lines out of B; ghost columns / rows (never in B) out of q
    for (unsigned e = threadIdx.x; e < 33 * 34; e += 256) {
        const unsigned k = e / 34, c = e - k * 34;
        const unsigned il = i0 - 1 + 2 * k, j = j0 - 1 + c;   // F-indices
        double v = 0.0;
        if (il >= 1 && il <= II && j >= 1 && j <= JJ) {
            if (il >= 2 && il <= II - 1 && j >= 2 && j <= JJ - 1) v = __ldcg(B + (size_t)(il - 1) * JJ + (j - 1));
            else v = q[(size_t)(j - 1) * II + (il - 1)];
        }
        nb[k][c] = v;
    }
    __syncthreads();
    const unsigned i = i0 + 2 * tx;
    for (unsigned r = ty; r < 32; r += 8) {
        const unsigned j = j0 + r;
        double v = 0.0;
        if (i <= II - 1 && j <= JJ - 1) {
            const size_t c = IX2(II, i, j);
            // q(i-1, .) = nb[tx][.], q(i+1, .) = nb[tx + 1][.]; position j sits in column r + 1
            v = qf[c] + kw[c] * nb[tx][r + 1] + kw[c + 1] * nb[tx + 1][r + 1];
            if (NINE)
                v = v + ksw[c] * nb[tx][r] + knw[c + 1] * nb[tx + 1][r] + knw[c + II] * nb[tx][r + 2]
                    + ksw[c + II + 1] * nb[tx + 1][r + 2];
        }
        tile[r][tx] = v;
    }
    __syncthreads();
    for (unsigned r = ty; r < 32; r += 8) {
        const unsigned il = i0 + 2 * r;
        const unsigned j = j0 + tx;
        if (il <= II - 1 && j <= JJ - 1) B[(size_t)(il - 1) * JJ + (j - 1)] = tile[tx][r];
    }
}

// q(i,j) = B(j,i) for all interior points
__global__ void __launch_bounds__(256)
k_lines_y_transpose_back(const double *__restrict__ B, double *__restrict__ q, unsigned II, unsigned JJ)
{
    __shared__ double tile[32][33];
    const unsigned tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (unsigned r = ty; r < 32; r += 8) {
        const unsigned il = 2 + blockIdx.x * 32 + r, j = 2 + blockIdx.y * 32 + tx;
        tile[r][tx] = (il <= II - 1 && j <= JJ - 1) ? __ldcs(B + (size_t)(il - 1) * JJ + (j - 1)) : 0.0;
    }
    __syncthreads();
    const unsigned i = 2 + blockIdx.x * 32 + tx;
    for (unsigned r = ty; r < 32; r += 8) {
        const unsigned j = 2 + blockIdx.y * 32 + r;
        if (i <= II - 1 && j <= JJ - 1) q[IX2(II, i, j)] = tile[tx][r];
    }
}

// one non-periodic y sweep, both colours, q written once at the end; false: not applicable (use the per-colour form)
static int liney_sweep_merged(const double *so, const double *qf, double *q, const double *sor, double *b, bmgx_len ii,
                              bmgx_len jj, int nine_pt, bool fwd, bmgx_stream_t s, bool *done)
{
    *done = false;
    unsigned lg, threads; size_t smem;
    if (bmgx_batch() != 1 || ii < 5 || !line_plan(jj - 2, &lg, &threads, &smem)) return BMGX_OK;
    cudaStream_t st = (cudaStream_t)s;
    for (int a = 0; a < 2; a++) {
        const unsigned ibeg = fwd ? 3 - a : 2 + a;
        const unsigned nl = (ii - 1 - ibeg) / 2 + 1;
        dim3 g(cdiv(nl, 32), cdiv(jj - 2, 32));
        if (a == 0) {
            if (nine_pt) k_lines_y_rhs<true><<<g, 256, 0, st>>>(so, qf, q, b, ii, jj, ibeg);
            else k_lines_y_rhs<false><<<g, 256, 0, st>>>(so, qf, q, b, ii, jj, ibeg);
        } else {
            if (nine_pt) k_lines_y_rhs_fromB<true><<<g, 256, 0, st>>>(so, qf, q, b, ii, jj, ibeg);
            else k_lines_y_rhs_fromB<false><<<g, 256, 0, st>>>(so, qf, q, b, ii, jj, ibeg);
        }
        BMGX_LAUNCH_CHECK("relax_lines_y/rhs");
        cudaError_t e = launch_liney(lg, nl, threads, smem, st, b, sor, ii, jj, ibeg);
        if (e != cudaSuccess) return bmgx_set_cuda_error(e, "relax_lines_y/attr");
        BMGX_LAUNCH_CHECK("relax_lines_y/solve");
    }
    k_lines_y_transpose_back<<<dim3(cdiv(ii - 2, 32), cdiv(jj - 2, 32)), 256, 0, st>>>(b, q, ii, jj);
    BMGX_LAUNCH_CHECK("relax_lines_y/transpose");
    *done = true;
    return BMGX_OK;
}

// the lines jbeg, jbeg + 2, ... (one colour) of an x-line sweep; per: the lines are periodic (Sherman-Morrison)
static int linex_colour(const double *so, const double *qf, double *q, const double *sor, bmgx_len ii, bmgx_len jj,
                        int nine_pt, int jbeg_, int per, bmgx_stream_t s)
{
    if (ii < 3 || jj < 3) return BMGX_OK;
    {
        const unsigned jbeg = (unsigned)jbeg_;
        const bool nine = nine_pt != 0;
        if (jbeg < 2 || jbeg > jj - 1) return BMGX_OK;
        unsigned nl = (jj - 1 - jbeg) / 2 + 1;
        unsigned lg, threads; size_t smem;
        if (line_plan(ii - 2, &lg, &threads, &smem, per ? 3 : 2)) {
            cudaError_t e = nine ? launch_linex<true>(lg, nl, threads, smem, (cudaStream_t)s, so, qf, q, sor, ii, jj, jbeg, per)
                                 : launch_linex<false>(lg, nl, threads, smem, (cudaStream_t)s, so, qf, q, sor, ii, jj, jbeg, per);
            if (e != cudaSuccess) return bmgx_set_cuda_error(e, "relax_lines_x/attr");
        } else if (per)
            return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_lines_x: periodic lines too long for shared memory (> 8500 unknowns)");
        else if (bmgx_batch() > 1)
            return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_lines_x: lines too long for the batched path");
        else if (nine)
            k_relax_lines_x<true><<<cdiv((size_t)nl * 32, 128), 128, 0, (cudaStream_t)s>>>(so, qf, q, sor, ii, jj, jbeg);
        else
            k_relax_lines_x<false><<<cdiv((size_t)nl * 32, 128), 128, 0, (cudaStream_t)s>>>(so, qf, q, sor, ii, jj, jbeg);
        BMGX_LAUNCH_CHECK("relax_lines_x");
    }
    return BMGX_OK;
}

extern "C" int bmgx2_relax_lines_x_colour(const double *so, const double *qf, double *q, const double *sor, bmgx_len ii,
                                          bmgx_len jj, int nine_pt, int jbeg_, bmgx_stream_t s)
{
    return linex_colour(so, qf, q, sor, ii, jj, nine_pt, jbeg_, 0, s);
}

extern "C" int bmgx2_relax_lines_x(int k, const double *so, const double *qf, double *q, const double *sor,
                                   double *b, bmgx_len ii, bmgx_len jj, int kf, int ifd, int nstncl,
                                   int irelax_sym, int updown, int jpn, bmgx_stream_t s)
{
    (void)nstncl; (void)b;
    if (ii < 3 || jj < 3) return BMGX_OK;
    // relax_lines_x.f90:73-75, 170: x periodic => Sherman-Morrison per line and a ghost refresh after every colour;
    // y periodic only => the plain solve and ONE refresh of the ghost rows after both colours (l.162-167).  (The code
    // -4 also enters the reference's second branch, with zero wrap couplings: a correction of exactly zero.)
    const int ipn = jpn < 0 ? -jpn : jpn;
    const bool wrap_x = ipn == 2 || ipn == 3, wrap_y = ipn == 1 || ipn == 3;
    bool fwd = (irelax_sym == BMGX_RELAX_NONSYM) || (updown == BMGX_DOWN && irelax_sym == BMGX_RELAX_SYM);
    for (int a = 0; a < 2; a++) {
        // DOWN: odd lines (3,5,..) first (relax_lines_x.f90:83-99)
        int st = linex_colour(so, qf, q, sor, ii, jj, (k < kf || ifd != 1) ? 1 : 0, fwd ? 3 - a : 2 + a, wrap_x ? 1 : 0, s);
        if (st != BMGX_OK) return st;
        if (wrap_x) {
            for (unsigned nb = 0; nb < bmgx_batch(); nb++) {
                double *qb = q + (size_t)nb * ii * jj;
                if (wrap_y && bmgx_wrap_y(qb, ii, jj, (cudaStream_t)s)) return BMGX_ERR_CUDA;
                if (bmgx_wrap_x(qb, ii, 1, 1, jj, (cudaStream_t)s)) return BMGX_ERR_CUDA;
            }
        }
    }
    if (!wrap_x && wrap_y)
        for (unsigned nb = 0; nb < bmgx_batch(); nb++)
            if (bmgx_wrap_y(q + (size_t)nb * ii * jj, ii, jj, (cudaStream_t)s)) return BMGX_ERR_CUDA;
    return BMGX_OK;
}

// the lines ibeg, ibeg + 2, ... (one colour) of a y-line sweep; per: the lines are periodic (Sherman-Morrison)
static int liney_colour(const double *so, const double *qf, double *q, const double *sor, double *b, bmgx_len ii, bmgx_len jj,
                        int nine_pt, int ibeg_, int per, bmgx_stream_t s)
{
    if (!b) return bmgx_set_error(BMGX_ERR_ARG, "relax_lines_y: scratch b (ii*jj doubles) required");
    if (ii < 3 || jj < 3) return BMGX_OK;
    cudaStream_t st = (cudaStream_t)s;
    {
        const unsigned ibeg = (unsigned)ibeg_;
        if (ibeg < 2 || ibeg > ii - 1) return BMGX_OK;
        unsigned nl = (ii - 1 - ibeg) / 2 + 1;
        dim3 g(cdiv(nl, 32), cdiv(jj - 2, 32), bmgx_batch());
        if (nine_pt) k_lines_y_rhs<true><<<g, 256, 0, st>>>(so, qf, q, b, ii, jj, ibeg);
        else k_lines_y_rhs<false><<<g, 256, 0, st>>>(so, qf, q, b, ii, jj, ibeg);
        BMGX_LAUNCH_CHECK("relax_lines_y/rhs");
        unsigned lg, threads; size_t smem;
        if (line_plan(jj - 2, &lg, &threads, &smem, per ? 3 : 2)) {
            cudaError_t e = launch_liney(lg, nl, threads, smem, st, b, sor, ii, jj, ibeg,
                                         per ? so + (size_t)(KS - 1) * ii * jj : nullptr);
            if (e != cudaSuccess) return bmgx_set_cuda_error(e, "relax_lines_y/attr");
        } else if (per)
            return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_lines_y: periodic lines too long for shared memory (> 8500 unknowns)");
        else if (bmgx_batch() > 1)
            return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_lines_y: lines too long for the batched path");
        else
            k_lines_y_solve<<<cdiv((size_t)nl * 32, 128), 128, 0, st>>>(b, sor, ii, jj, ibeg);
        BMGX_LAUNCH_CHECK("relax_lines_y/solve");
        k_lines_y_scatter<<<g, 256, 0, st>>>(b, q, ii, jj, ibeg);
        BMGX_LAUNCH_CHECK("relax_lines_y/scatter");
    }
    return BMGX_OK;
}

extern "C" int bmgx2_relax_lines_y_colour(const double *so, const double *qf, double *q, const double *sor, double *b,
                                          bmgx_len ii, bmgx_len jj, int nine_pt, int ibeg_, bmgx_stream_t s)
{
    return liney_colour(so, qf, q, sor, b, ii, jj, nine_pt, ibeg_, 0, s);
}

extern "C" int bmgx2_relax_lines_y(int k, const double *so, const double *qf, double *q, const double *sor,
                                   double *b, bmgx_len ii, bmgx_len jj, int kf, int ifd, int nstncl,
                                   int irelax_sym, int updown, int jpn, bmgx_stream_t s)
{
    (void)nstncl;
    if (!b) return bmgx_set_error(BMGX_ERR_ARG, "relax_lines_y: scratch b (ii*jj doubles) required");
    if (ii < 3 || jj < 3) return BMGX_OK;
    // relax_lines_y.f90:73-75, 178: the mirror image of the x version
    const int ipn = jpn < 0 ? -jpn : jpn;
    const bool wrap_x = ipn == 2 || ipn == 3, wrap_y = ipn == 1 || ipn == 3;
    if (wrap_y && bmgx_batch() > 1) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_lines_y: periodic lines are not batched");
    bool fwd = (irelax_sym == BMGX_RELAX_NONSYM) || (updown == BMGX_DOWN && irelax_sym == BMGX_RELAX_SYM);
    if (!wrap_x && !wrap_y) {
        bool done = false;
        int st = liney_sweep_merged(so, qf, q, sor, b, ii, jj, (k < kf || ifd != 1) ? 1 : 0, fwd, s, &done);
        if (st != BMGX_OK || done) return st;
    }
    for (int a = 0; a < 2; a++) {
        int st = liney_colour(so, qf, q, sor, b, ii, jj, (k < kf || ifd != 1) ? 1 : 0, fwd ? 3 - a : 2 + a, wrap_y ? 1 : 0, s);
        if (st != BMGX_OK) return st;
        if (wrap_y) {
            if (bmgx_wrap_y(q, ii, jj, (cudaStream_t)s)) return BMGX_ERR_CUDA;
            if (wrap_x && bmgx_wrap_x(q, ii, 1, 1, jj, (cudaStream_t)s)) return BMGX_ERR_CUDA;
        }
    }
    if (!wrap_y && wrap_x)
        for (unsigned nb = 0; nb < bmgx_batch(); nb++)
            if (bmgx_wrap_x(q + (size_t)nb * ii * jj, ii, 1, 1, jj, (cudaStream_t)s)) return BMGX_ERR_CUDA;
    return BMGX_OK;
}

// ---- lines across block cuts (the decomposed solver, cedar_b200/host/dist2.cc) ------------------------------
// The blocks of a process row (x-lines) or column (y-lines) transpose their lines: every block computes the
// right-hand side of its OWN points -- the serial kernels' expression, term for term -- and stores it straight into
// the line buffer of the block that solves the line (peer memory over NVLink); the solver runs the serial solve
// (k_liney_block / k_lines_y_solve on the line-major buffer, the very kernels of relax_lines_y's second step) on
// whole lines and stores the solution into the owners' iterate.  Line buffer of a solver: rows 1 .. its lines,
// row pitch len + 2, position p of a line (global F-index p + 1) at column p -- the layout of the serial y-lines'
// transposed scratch, so the factor (SETUP_lines on a strip of gathered diagonals) and the solve are the serial ones.
namespace {

// destination of line t (0-based within the group) at line-buffer column c
__device__ __forceinline__ double *line_slot(const bmgx_line_group &g, unsigned t, unsigned c)
{
    int m = 0;
#pragma unroll 1
    while (m + 1 < g.nmembers && t >= (unsigned)g.first[m + 1]) m++;
    return g.lb[m] + (size_t)(1 + t - g.first[m]) * g.pitch + c;
}

// x-lines: thread per own point of a line of the colour (own lines t = t0, t0 + 2, ...)
template <bool NINE>
__global__ void __launch_bounds__(256)
k_lines_rhs_push_x(const double *__restrict__ so, const double *__restrict__ qf, const double *__restrict__ q, unsigned II,
                   unsigned JJ, unsigned wl, unsigned wb, unsigned nx, unsigned t0, unsigned col0, bmgx_line_group g)
{
    const unsigned il = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned t = t0 + 2 * blockIdx.y;
    if (il >= nx) return;
    const size_t pl = (size_t)II * JJ;
    const double *ks = so + (KS - 1) * pl, *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
    const size_t c = (size_t)(wb + t) * II + wl + il;
    double r = qf[c] + ks[c] * q[c - II] + ks[c + II] * q[c + II];
    if (NINE)
        r = r + ksw[c] * q[c - II - 1] + knw[c + 1] * q[c - II + 1] + knw[c + II] * q[c + II - 1]
            + ksw[c + II + 1] * q[c + II + 1];
    *line_slot(g, t, col0 + il) = r;
}

// y-lines: 32 lines (own columns of the colour) x 32 positions (own rows) per CTA, turned in shared memory so that
// the stores into the line buffers run along the lines
template <bool NINE>
__global__ void __launch_bounds__(256)
k_lines_rhs_push_y(const double *__restrict__ so, const double *__restrict__ qf, const double *__restrict__ q, unsigned II,
                   unsigned JJ, unsigned wl, unsigned wb, unsigned nx, unsigned ny, unsigned t0, unsigned col0,
                   bmgx_line_group g)
{
    __shared__ double tile[32][33];
    const unsigned tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const size_t pl = (size_t)II * JJ;
    const double *kw = so + (KW - 1) * pl, *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
    const unsigned t = t0 + 2 * (blockIdx.x * 32 + tx);   // own column
    for (unsigned r = ty; r < 32; r += 8) {
        const unsigned jl = blockIdx.y * 32 + r;          // own row
        double v = 0.0;
        if (t < nx && jl < ny) {
            const size_t c = (size_t)(wb + jl) * II + wl + t;
            v = qf[c] + kw[c] * q[c - 1] + kw[c + 1] * q[c + 1];
            if (NINE)
                v = v + ksw[c] * q[c - II - 1] + knw[c + 1] * q[c - II + 1] + knw[c + II] * q[c + II - 1]
                    + ksw[c + II + 1] * q[c + II + 1];
        }
        tile[r][tx] = v;
    }
    __syncthreads();
    for (unsigned r = ty; r < 32; r += 8) {
        const unsigned tl = t0 + 2 * (blockIdx.x * 32 + r);
        const unsigned jl = blockIdx.y * 32 + tx;
        if (tl < nx && jl < ny) *line_slot(g, tl, col0 + jl) = tile[tx][r];
    }
}

// solved lines -> the owners' iterate.  x-lines: CTA row = one of this solver's lines of the colour
__global__ void __launch_bounds__(256)
k_lines_scatter_push_x(const double *__restrict__ lb, size_t pitch, unsigned r0, unsigned tfirst, unsigned len,
                       bmgx_line_owners o)
{
    const unsigned p = blockIdx.x * blockDim.x + threadIdx.x;   // 0-based global position along the line
    const unsigned r = r0 + 2 * blockIdx.y;                     // line-buffer row
    if (p >= len) return;
    int m = 0;
#pragma unroll 1
    while (m + 1 < o.nmembers && p >= (unsigned)o.start[m + 1]) m++;
    const unsigned t = tfirst + r - 1;                          // line within the group
    o.q[m][(size_t)(o.w_across + t) * o.pitch[m] + o.w_along[m] + (p - o.start[m])] = __ldcg(lb + (size_t)r * pitch + 1 + p);
}

// y-lines: 32 of this solver's lines x 32 positions per CTA, turned so that the stores run along the owners' rows
__global__ void __launch_bounds__(256)
k_lines_scatter_push_y(const double *__restrict__ lb, size_t pitch, unsigned r0, unsigned nrows, unsigned tfirst, unsigned len,
                       bmgx_line_owners o)
{
    __shared__ double tile[32][33];
    const unsigned tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (unsigned k = ty; k < 32; k += 8) {
        const unsigned r = r0 + 2 * (blockIdx.x * 32 + k), p = blockIdx.y * 32 + tx;
        tile[k][tx] = (r <= nrows && p < len) ? __ldcg(lb + (size_t)r * pitch + 1 + p) : 0.0;
    }
    __syncthreads();
    const unsigned r = r0 + 2 * (blockIdx.x * 32 + tx);
    for (unsigned k = ty; k < 32; k += 8) {
        const unsigned p = blockIdx.y * 32 + k;
        if (r > nrows || p >= len) continue;
        int m = 0;
#pragma unroll 1
        while (m + 1 < o.nmembers && p >= (unsigned)o.start[m + 1]) m++;
        const unsigned t = tfirst + r - 1;
        o.q[m][(size_t)(o.w_along[m] + (p - o.start[m])) * o.pitch[m] + o.w_across + t] = tile[tx][k];
    }
}

}  // namespace

extern "C" int bmgx2_lines_rhs_push(int dirn, int nine, const double *so, const double *qf, const double *q, bmgx_len ii,
                                    bmgx_len jj, bmgx_len wl, bmgx_len wb, bmgx_len nx, bmgx_len ny, int first_findex,
                                    bmgx_len col0, const bmgx_line_group *g, bmgx_stream_t s)
{
    if (!g || g->nmembers < 1 || g->nmembers > BMGX_LINE_GROUP_MAX)
        return bmgx_set_error(BMGX_ERR_ARG, "lines_rhs_push: 1 .. BMGX_LINE_GROUP_MAX blocks share a line");
    cudaStream_t st = (cudaStream_t)s;
    // own lines of the colour: the group's first line has an even global index (F-index even), so line t has the
    // parity of t
    const unsigned t0 = (unsigned)(first_findex & 1);
    const unsigned nl_all = dirn == 0 ? ny : nx;
    if (t0 >= nl_all) return BMGX_OK;
    const unsigned nl = (nl_all - t0 + 1) / 2;
    if (dirn == 0) {
        dim3 grid(cdiv(nx, 256), nl);
        if (nine) k_lines_rhs_push_x<true><<<grid, 256, 0, st>>>(so, qf, q, ii, jj, wl, wb, nx, t0, col0, *g);
        else k_lines_rhs_push_x<false><<<grid, 256, 0, st>>>(so, qf, q, ii, jj, wl, wb, nx, t0, col0, *g);
    } else {
        dim3 grid(cdiv(nl, 32), cdiv(ny, 32));
        if (nine) k_lines_rhs_push_y<true><<<grid, 256, 0, st>>>(so, qf, q, ii, jj, wl, wb, nx, ny, t0, col0, *g);
        else k_lines_rhs_push_y<false><<<grid, 256, 0, st>>>(so, qf, q, ii, jj, wl, wb, nx, ny, t0, col0, *g);
    }
    BMGX_LAUNCH_CHECK("lines_rhs_push");
    return BMGX_OK;
}

extern "C" int bmgx2_lines_solve_T(double *lb, const double *sor, bmgx_len nrows2, bmgx_len len2, int ibeg, bmgx_stream_t s)
{
    // second step of relax_lines_y on a line-major buffer: "ii" = rows of the buffer, "jj" = its pitch
    if (nrows2 < 3 || len2 < 3 || (unsigned)ibeg > nrows2 - 1) return BMGX_OK;
    cudaStream_t st = (cudaStream_t)s;
    const unsigned nl = (nrows2 - 1 - ibeg) / 2 + 1;
    unsigned lg, threads; size_t smem;
    if (line_plan(len2 - 2, &lg, &threads, &smem)) {
        const unsigned nb = bmgx_batch();
        if (nb != 1) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "lines_solve_T: not batched");
        cudaError_t e = launch_liney(lg, nl, threads, smem, st, lb, sor, nrows2, len2, ibeg);
        if (e != cudaSuccess) return bmgx_set_cuda_error(e, "lines_solve_T/attr");
    } else
        k_lines_y_solve<<<cdiv((size_t)nl * 32, 128), 128, 0, st>>>(lb, sor, nrows2, len2, ibeg);
    BMGX_LAUNCH_CHECK("lines_solve_T");
    return BMGX_OK;
}

extern "C" int bmgx2_lines_scatter_push(int dirn, const double *lb, bmgx_len nrows2, bmgx_len len2, int ibeg, bmgx_len tfirst,
                                        const bmgx_line_owners *o, bmgx_stream_t s)
{
    if (!o || o->nmembers < 1 || o->nmembers > BMGX_LINE_GROUP_MAX)
        return bmgx_set_error(BMGX_ERR_ARG, "lines_scatter_push: 1 .. BMGX_LINE_GROUP_MAX blocks share a line");
    if (nrows2 < 3 || len2 < 3 || (unsigned)ibeg > nrows2 - 1) return BMGX_OK;
    cudaStream_t st = (cudaStream_t)s;
    const unsigned r0 = ibeg - 1, nrows = nrows2 - 2, nl = (nrows - r0) / 2 + 1, len = len2 - 2;
    if (dirn == 0) k_lines_scatter_push_x<<<dim3(cdiv(len, 256), nl), 256, 0, st>>>(lb, len2, r0, tfirst, len, *o);
    else k_lines_scatter_push_y<<<dim3(cdiv(nl, 32), cdiv(len, 32)), 256, 0, st>>>(lb, len2, r0, nrows, tfirst, len, *o);
    BMGX_LAUNCH_CHECK("lines_scatter_push");
    return BMGX_OK;
}

// ---- y-line relaxation without any allocation on the caller's side -------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  bmgx2_relax_lines_y needs ii*jj
// doubles of scratch for the transposed right-hand sides; here that scratch is a file-scope device array of
// BMGX_LINES_Y_STATIC_MAX doubles (8 MiB), held by every GPU that loads the library until the process exits.
// The array's name is a host-side stand-in in host code: its device address comes from cudaGetSymbolAddress, once per
// device, with that device current.  Every call shares the array and finds the previous call's contents in it; nothing
// in it has to start at zero (every entry the solve reads is written by the right-hand-side kernel of the same colour
// first), so nothing is cleared.  Calls that overlap on the GPU (different streams, or several host threads) RACE on it.
static __device__ __align__(256) double g_lines_y_scratch[BMGX_LINES_Y_STATIC_MAX];

extern "C" int bmgx2_relax_lines_y_static(int k, const double *so, const double *qf, double *q, const double *sor,
                                          bmgx_len ii, bmgx_len jj, int kf, int ifd, int nstncl, int irelax_sym,
                                          int updown, int jpn, bmgx_stream_t s)
{
    static double *scratch_of[64] = {};   // per device
    if ((size_t)ii * jj * bmgx_batch() > (size_t)BMGX_LINES_Y_STATIC_MAX)
        return bmgx_set_error(BMGX_ERR_ARG, "relax_lines_y_static: ii*jj (times the batch count) exceeds "
                                            "BMGX_LINES_Y_STATIC_MAX = 1048576 doubles of built-in scratch; use "
                                            "bmgx2_relax_lines_y with caller scratch");
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return bmgx_set_cuda_error(e, "relax_lines_y_static/device");
    if (dev < 0 || dev >= 64) return bmgx_set_error(BMGX_ERR_ARG, "relax_lines_y_static: device ordinal out of range");
    if (!scratch_of[dev]) {
        double *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_lines_y_scratch);
        if (e != cudaSuccess || !p) return bmgx_set_cuda_error(e, "relax_lines_y_static/cudaGetSymbolAddress");
        scratch_of[dev] = p;
    }
    return bmgx2_relax_lines_y(k, so, qf, q, sor, scratch_of[dev], ii, jj, kf, ifd, nstncl, irelax_sym, updown, jpn, s);
}
