// bmgx2_relax_fused.cu -- multicolour point Gauss-Seidel, ALL colours in ONE
// pass over memory (out of place: q_in -> q_out).
//
// Same arithmetic as bmgx2_relax.cu (reference: src/2d/ftn/BMG2_SymStd_relax_GS.f90
// :93-135), same term order, -fmad=false => bitwise the oracle's result.  What
// changes is the schedule.  Points of one colour are independent, so any order
// inside a colour is legal; the kernel streams down the grid so that every
// array is read from HBM exactly once per sweep:
//
//   * a CTA owns a strip of RF_TW columns and a band of rows; thread t owns the
//     column pair (ie, io) = (x0+2t, x0+2t+1) and walks down the rows;
//   * the iterate lives in a shared-memory ring of 8 rows; stencil coefficients
//     stay in registers (each coefficient row is loaded once and carried from
//     its use as the "north" row to its use as the "own" row);
//   * 9-point, DOWN (colours (je,ie) (je,io) (jo,ie) (jo,io), F-index parity):
//     per step the even row j gets colour 1 then 2, then the odd row j-1
//     (whose north and south rows are now final) gets colour 3 then 4;
//     UP is the mirror image (odd rows lead, odd columns first);
//   * 5-point red-black: per step red on rows j, j+1, then black on rows j-1, j;
//   * strips overlap by RF_MARGIN columns on each side (values there are
//     recomputed, never stored) and bands by one leading row, so CTAs are
//     independent: no inter-CTA ordering, no atomics.
//
// Algorithmic traffic: 64 B/point (9-pt), 48 B/point (5-pt); redundant reads
// are 8/256 of the columns and ~3 rows per band.
#include "bmgx_common.cuh"

#define RF_NT 128
#define RF_W (2 * RF_NT)
#define RF_MARGIN 4
#define RF_TW (RF_W - 2 * RF_MARGIN)
#define RF_RING 8

namespace {

struct RowC {
    double kw0, kw1, kw2, ks0, ks1, ksw0, ksw1, ksw2, knw0, knw1, knw2, sor0, sor1, qf0, qf1;
};

// One (sub-)array: II x JJ entries (ghost ring included) with row pitch P; every pointer addresses its entry (1,1).
// Finished values are stored only inside the window [wi0, wi1] x [wj0, wj1] (1-based, inclusive): the whole array for
// a plain sweep, a part of it when a sweep is split into rectangles that are relaxed with recomputed margins
// (bmgx2_relax_GS_oop_rects).
struct Grid {
    const double *kw, *ks, *ksw, *knw, *sor, *qf;
    int II, JJ;
    long long P;
    int wi0, wi1, wj0, wj1;
};

template <bool NINE>
__device__ __forceinline__ void load_row(RowC &R, const Grid &g, int r, int ie)
{
    const bool rok = r >= 1 && r <= g.JJ;
    const bool c0 = rok && ie >= 1 && ie <= g.II;
    const bool c1 = rok && ie + 1 >= 1 && ie + 1 <= g.II;
    const bool c2 = rok && ie + 2 >= 1 && ie + 2 <= g.II;
    const long long b = (long long)(r - 1) * g.P + (ie - 1);
    R.kw0 = c0 ? __ldg(g.kw + b) : 0.0;
    R.kw1 = c1 ? __ldg(g.kw + b + 1) : 0.0;
    R.kw2 = c2 ? __ldg(g.kw + b + 2) : 0.0;
    R.ks0 = c0 ? __ldg(g.ks + b) : 0.0;
    R.ks1 = c1 ? __ldg(g.ks + b + 1) : 0.0;
    if (NINE) {
        R.ksw0 = c0 ? __ldg(g.ksw + b) : 0.0;
        R.ksw1 = c1 ? __ldg(g.ksw + b + 1) : 0.0;
        R.ksw2 = c2 ? __ldg(g.ksw + b + 2) : 0.0;
        R.knw0 = c0 ? __ldg(g.knw + b) : 0.0;
        R.knw1 = c1 ? __ldg(g.knw + b + 1) : 0.0;
        R.knw2 = c2 ? __ldg(g.knw + b + 2) : 0.0;
    }
    R.sor0 = c0 ? __ldg(g.sor + b) : 0.0;
    R.sor1 = c1 ? __ldg(g.sor + b + 1) : 0.0;
    R.qf0 = c0 ? __ldg(g.qf + b) : 0.0;
    R.qf1 = c1 ? __ldg(g.qf + b + 1) : 0.0;
}

// the reference update, term by term (relax_GS.f90:99-109 / 126-131);
// rm / r0 / rp: ring rows r-1, r, r+1; s: ring column of the point
template <bool NINE>
__device__ __forceinline__ double gs_update(const double *rm, const double *r0, const double *rp, int s, double qf,
                                            double kw_c, double kw_e, double ks_c, double ks_n, double ksw_c,
                                            double knw_e, double knw_n, double ksw_ne, double sor)
{
    double acc = qf + kw_c * r0[s - 1] + kw_e * r0[s + 1] + ks_c * rm[s] + ks_n * rp[s];
    if (NINE) acc = acc + ksw_c * rm[s - 1] + knw_e * rm[s + 1] + knw_n * rp[s - 1] + ksw_ne * rp[s + 1];
    return acc * sor;
}

// point at the even (ODD=false) or odd (ODD=true) column of the thread's pair, row
// with own coefficients R and north-row coefficients N
template <bool NINE, bool ODD>
__device__ __forceinline__ double gs_point(const double *rm, const double *r0, const double *rp, int se,
                                           const RowC &R, const RowC &N)
{
    if (!ODD)
        return gs_update<NINE>(rm, r0, rp, se, R.qf0, R.kw0, R.kw1, R.ks0, N.ks0, R.ksw0, R.knw1, N.knw0, N.ksw1, R.sor0);
    return gs_update<NINE>(rm, r0, rp, se + 1, R.qf1, R.kw1, R.kw2, R.ks1, N.ks1, R.ksw1, R.knw2, N.knw1, N.ksw2, R.sor1);
}

#define RING(r) ring[((r) + 8) & (RF_RING - 1)]

__device__ __forceinline__ double ld_q(const double *__restrict__ q, int r, int c, const Grid &g)
{
    return (r >= 1 && r <= g.JJ && c >= 1 && c <= g.II) ? __ldg(q + (long long)(r - 1) * g.P + (c - 1)) : 0.0;
}

// store one finished row from the ring, coalesced (thread t: columns x0+t and x0+NT+t)
__device__ __forceinline__ void st_row(double *__restrict__ qout, double (*ring)[RF_W + 2], int r, int x0, int t,
                                       const Grid &g, int jb, int je)
{
    if (r < jb || r > je || r < g.wj0 || r > g.wj1) return;
    const double *row = RING(r);
#pragma unroll
    for (int h = 0; h < 2; h++) {
        int cl = t + h * RF_NT;          // 0..W-1
        int c = x0 + cl;
        if (cl >= RF_MARGIN && cl < RF_W - RF_MARGIN && c >= g.wi0 && c <= g.wi1)
            qout[(long long)(r - 1) * g.P + (c - 1)] = row[cl + 1];
    }
}

// ---- 9-point, four colours ---------------------------------------------------
// ODDFIRST=false: DOWN (even rows lead, even column first); true: UP.
template <bool ODDFIRST>
__device__ __forceinline__ void relax9_body(const Grid &g, const double *__restrict__ qin, double *__restrict__ qout,
                                            int band_h, int bx, int by, double (*ring)[RF_W + 2])
{
    const int t = threadIdx.x;
    const int II = g.II, JJ = g.JJ, I1 = II - 1, J1 = JJ - 1;
    const int x0 = -RF_MARGIN + bx * RF_TW;
    const int ie = x0 + 2 * t, se = 2 * t + 1;
    const int jb = 1 + by * band_h;
    const int je = min(JJ, jb + band_h - 1);
    constexpr int pA = ODDFIRST ? 1 : 0;
    if (t < RF_RING) { ring[t][0] = 0.0; ring[t][RF_W + 1] = 0.0; }
    int j0 = max(jb - 1, 1);
    if ((j0 & 1) != pA) j0++;
    for (int r = j0 - 1; r <= j0 + 1; r++) {
        RING(r)[t + 1] = ld_q(qin, r, x0 + t, g);
        RING(r)[t + RF_NT + 1] = ld_q(qin, r, x0 + t + RF_NT, g);
    }
    RowC Rm;
    load_row<true>(Rm, g, j0 - 1, ie);
    const bool eok = ie >= 2 && ie <= I1, ook = ie + 1 >= 2 && ie + 1 <= I1;
    const bool fok = ODDFIRST ? ook : eok, sok = ODDFIRST ? eok : ook;
    __syncthreads();
    for (int j = j0; j <= je + 1; j += 2) {
        RowC R0, Rp;
        load_row<true>(R0, g, j, ie);
        load_row<true>(Rp, g, j + 1, ie);
        double n0a = ld_q(qin, j + 2, x0 + t, g), n0b = ld_q(qin, j + 2, x0 + t + RF_NT, g);
        double n1a = ld_q(qin, j + 3, x0 + t, g), n1b = ld_q(qin, j + 3, x0 + t + RF_NT, g);
        st_row(qout, ring, j - 3, x0, t, g, jb, je);   // the B row finished by the previous step
        const bool updA = j >= 2 && j <= J1;
        const bool updB = j - 1 >= max(2, jb) && j - 1 <= min(J1, je);
        double *rA = RING(j), *rB = RING(j - 1);
        const double *rAp = RING(j + 1), *rBm = RING(j - 2);
        if (updA && fok) rA[se + (ODDFIRST ? 1 : 0)] = gs_point<true, ODDFIRST>(rB, rA, rAp, se, R0, Rp);
        __syncthreads();
        if (updA && sok) rA[se + (ODDFIRST ? 0 : 1)] = gs_point<true, !ODDFIRST>(rB, rA, rAp, se, R0, Rp);
        __syncthreads();
        st_row(qout, ring, j, x0, t, g, jb, je);
        if (updB && fok) rB[se + (ODDFIRST ? 1 : 0)] = gs_point<true, ODDFIRST>(rBm, rB, rA, se, Rm, R0);
        __syncthreads();
        if (updB && sok) rB[se + (ODDFIRST ? 0 : 1)] = gs_point<true, !ODDFIRST>(rBm, rB, rA, se, Rm, R0);
        RING(j + 2)[t + 1] = n0a; RING(j + 2)[t + RF_NT + 1] = n0b;
        RING(j + 3)[t + 1] = n1a; RING(j + 3)[t + RF_NT + 1] = n1b;
        Rm = Rp;
        __syncthreads();
    }
    // B row finished by the last step, and ghost/trailing rows of the band that no step stored
    int jl = j0 + 2 * ((je + 1 - j0) / 2);   // last j executed
    st_row(qout, ring, jl - 1, x0, t, g, jb, je);
}

// ---- 5-point, red-black ------------------------------------------------------
// BLACKFIRST=false: DOWN (red = (i+j) even first); true: UP.
template <bool BLACKFIRST>
__device__ __forceinline__ void relax5_body(const Grid &g, const double *__restrict__ qin, double *__restrict__ qout,
                                            int band_h, int bx, int by, double (*ring)[RF_W + 2])
{
    const int t = threadIdx.x;
    const int II = g.II, JJ = g.JJ, I1 = II - 1, J1 = JJ - 1;
    const int x0 = -RF_MARGIN + bx * RF_TW;
    const int ie = x0 + 2 * t, se = 2 * t + 1;
    const int jb = 1 + by * band_h;
    const int je = min(JJ, jb + band_h - 1);
    if (t < RF_RING) { ring[t][0] = 0.0; ring[t][RF_W + 1] = 0.0; }
    int j0 = ((jb - 1) & 1) ? jb - 2 : jb - 1;   // even, covers row jb-1 with its first colour
    for (int r = j0 - 1; r <= j0 + 2; r++) {
        RING(r)[t + 1] = ld_q(qin, r, x0 + t, g);
        RING(r)[t + RF_NT + 1] = ld_q(qin, r, x0 + t + RF_NT, g);
    }
    RowC Rm, R0;   // rows j-1, j (carried)
    load_row<false>(Rm, g, j0 - 1, ie);
    load_row<false>(R0, g, j0, ie);
    const bool eok = ie >= 2 && ie <= I1, ook = ie + 1 >= 2 && ie + 1 <= I1;
    __syncthreads();
    for (int j = j0; j <= je + 1; j += 2) {
        RowC Rp, Rpp;
        load_row<false>(Rp, g, j + 1, ie);
        load_row<false>(Rpp, g, j + 2, ie);
        double n0a = ld_q(qin, j + 3, x0 + t, g), n0b = ld_q(qin, j + 3, x0 + t + RF_NT, g);
        double n1a = ld_q(qin, j + 4, x0 + t, g), n1b = ld_q(qin, j + 4, x0 + t + RF_NT, g);
        st_row(qout, ring, j - 3, x0, t, g, jb, je);
        st_row(qout, ring, j - 2, x0, t, g, jb, je);
        double *rm = RING(j - 1), *r0 = RING(j), *rp = RING(j + 1);
        const double *rmm = RING(j - 2), *rpp = RING(j + 2);
        // first colour on rows j (even) and j+1 (odd)
        const bool x0ok = j >= max(2, jb - 1) && j <= min(J1, je + 1);
        const bool x1ok = j + 1 >= max(2, jb - 1) && j + 1 <= min(J1, je + 1);
        if (!BLACKFIRST) {
            if (x0ok && eok) r0[se] = gs_point<false, false>(rm, r0, rp, se, R0, Rp);
            if (x1ok && ook) rp[se + 1] = gs_point<false, true>(r0, rp, rpp, se, Rp, Rpp);
        } else {
            if (x0ok && ook) r0[se + 1] = gs_point<false, true>(rm, r0, rp, se, R0, Rp);
            if (x1ok && eok) rp[se] = gs_point<false, false>(r0, rp, rpp, se, Rp, Rpp);
        }
        __syncthreads();
        // second colour on rows j-1 (odd) and j (even)
        const bool y0ok = j - 1 >= max(2, jb) && j - 1 <= min(J1, je);
        const bool y1ok = j >= max(2, jb) && j <= min(J1, je);
        if (!BLACKFIRST) {
            if (y0ok && eok) rm[se] = gs_point<false, false>(rmm, rm, r0, se, Rm, R0);
            if (y1ok && ook) r0[se + 1] = gs_point<false, true>(rm, r0, rp, se, R0, Rp);
        } else {
            if (y0ok && ook) rm[se + 1] = gs_point<false, true>(rmm, rm, r0, se, Rm, R0);
            if (y1ok && eok) r0[se] = gs_point<false, false>(rm, r0, rp, se, R0, Rp);
        }
        RING(j + 3)[t + 1] = n0a; RING(j + 3)[t + RF_NT + 1] = n0b;
        RING(j + 4)[t + 1] = n1a; RING(j + 4)[t + RF_NT + 1] = n1b;
        Rm = Rp; R0 = Rpp;
        __syncthreads();
    }
    int jl = j0 + 2 * ((je + 1 - j0) / 2);
    st_row(qout, ring, jl - 1, x0, t, g, jb, je);
    st_row(qout, ring, jl, x0, t, g, jb, je);
}

// one (sub-)array per launch
template <bool ODDFIRST>
__global__ void __launch_bounds__(RF_NT, 4)
k_relax9_fused(const __grid_constant__ Grid g, const double *__restrict__ qin, double *__restrict__ qout, int band_h)
{
    pdl_wait();
    __shared__ double ring[RF_RING][RF_W + 2];
    relax9_body<ODDFIRST>(g, qin, qout, band_h, (int)blockIdx.x, (int)blockIdx.y, ring);
}
template <bool BLACKFIRST>
__global__ void __launch_bounds__(RF_NT, 4)
k_relax5_fused(const __grid_constant__ Grid g, const double *__restrict__ qin, double *__restrict__ qout, int band_h)
{
    pdl_wait();
    __shared__ double ring[RF_RING][RF_W + 2];
    relax5_body<BLACKFIRST>(g, qin, qout, band_h, (int)blockIdx.x, (int)blockIdx.y, ring);
}

// several rectangles of one array in one launch (blockIdx.z = rectangle): the frames of a sweep that was split so
// that its interior could run while the halos were still travelling (bmgx2_relax_GS_oop_rects)
#define RF_MAXRECT 5
struct RectSet {
    Grid g[RF_MAXRECT];
    const double *qin[RF_MAXRECT];
    double *qout[RF_MAXRECT];
    int band_h[RF_MAXRECT], nstrips[RF_MAXRECT], nbands[RF_MAXRECT];
};
template <bool NINE, bool REV>
__global__ void __launch_bounds__(RF_NT, 4) k_relax_fused_rects(const __grid_constant__ RectSet R)
{
    pdl_wait();
    __shared__ double ring[RF_RING][RF_W + 2];
    const int z = blockIdx.z;
    if ((int)blockIdx.x >= R.nstrips[z] || (int)blockIdx.y >= R.nbands[z]) return;
    if (NINE) relax9_body<REV>(R.g[z], R.qin[z], R.qout[z], R.band_h[z], (int)blockIdx.x, (int)blockIdx.y, ring);
    else relax5_body<REV>(R.g[z], R.qin[z], R.qout[z], R.band_h[z], (int)blockIdx.x, (int)blockIdx.y, ring);
}

// strips x bands of one (sub-)array: one resident wave (148 SMs x 4 CTAs) when the array is large; short bands on
// small grids (the row loop is a latency chain: one global load round trip per step)
void tile(unsigned ii, unsigned jj, unsigned share, unsigned &nstrips, unsigned &band_h, unsigned &nbands)
{
    nstrips = ii / RF_TW + 1;
    unsigned nb_target = (592 / share) / nstrips; if (nb_target < 1) nb_target = 1;
    band_h = (jj + nb_target - 1) / nb_target;
    if (band_h < 4) band_h = 4;
    band_h += band_h & 1;
    nbands = (jj + band_h - 1) / band_h;
}

}  // namespace

extern "C" int bmgx2_relax_GS_oop_rects(int k, const double *so, const double *qf, const double *q_in, double *q_out,
                                        const double *sor, bmgx_len ii, bmgx_len jj, int kf, int ifd, int nstncl, int nsorv,
                                        int irelax_sym, int updown, int jpn, const bmgx_rect *rects, int nrects,
                                        bmgx_stream_t s)
{
    (void)nstncl; (void)nsorv;
    if (q_in == q_out) return bmgx_set_error(BMGX_ERR_ARG, "relax_GS_oop: q_in aliases q_out");
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_GS_oop: periodic BCs not supported");
    if (nrects < 1 || nrects > RF_MAXRECT) return bmgx_set_error(BMGX_ERR_ARG, "relax_GS_oop_rects: 1..5 rectangles");
    if (ii < 1 || jj < 1) return BMGX_OK;
    const bool fwd = (irelax_sym == BMGX_RELAX_NONSYM) || (updown == BMGX_DOWN && irelax_sym == BMGX_RELAX_SYM);
    const bool nine = (k < kf || ifd != 1);
    const size_t pl = (size_t)ii * jj;
    RectSet R{};
    unsigned gx = 1, gy = 1;
    for (int z = 0; z < nrects; z++) {
        const bmgx_rect &r = rects[z];
        // the origin must keep the parity of the full array's indices: the colouring is by index parity
        if ((r.i0 & 1) || (r.j0 & 1) || r.ni < 1 || r.nj < 1 || r.i0 + r.ni > ii || r.j0 + r.nj > jj || r.wi0 < 1 || r.wj0 < 1
            || r.wi1 > r.ni || r.wj1 > r.nj)
            return bmgx_set_error(BMGX_ERR_ARG, "relax_GS_oop_rects: rectangle outside the array, odd origin or bad window");
        const size_t off = (size_t)r.j0 * ii + r.i0;
        Grid &g = R.g[z];
        g.kw = so + (KW - 1) * pl + off; g.ks = so + (KS - 1) * pl + off;
        g.ksw = nine ? so + (KSW - 1) * pl + off : nullptr; g.knw = nine ? so + (KNW - 1) * pl + off : nullptr;
        g.sor = sor + (MSOR - 1) * pl + off; g.qf = qf + off;
        g.II = (int)r.ni; g.JJ = (int)r.nj; g.P = (long long)ii;
        g.wi0 = (int)r.wi0; g.wi1 = (int)r.wi1; g.wj0 = (int)r.wj0; g.wj1 = (int)r.wj1;
        R.qin[z] = q_in + off; R.qout[z] = q_out + off;
        unsigned ns, bh, nb;
        tile(r.ni, r.nj, (unsigned)nrects, ns, bh, nb);
        R.nstrips[z] = (int)ns; R.band_h[z] = (int)bh; R.nbands[z] = (int)nb;
        gx = ns > gx ? ns : gx; gy = nb > gy ? nb : gy;
    }
    dim3 grid(gx, gy, nrects);
    cudaStream_t st = (cudaStream_t)s;
    if (nine) {
        if (fwd) launch_pdl(k_relax_fused_rects<true, false>, grid, dim3(RF_NT), 0, st, R);
        else launch_pdl(k_relax_fused_rects<true, true>, grid, dim3(RF_NT), 0, st, R);
    } else {
        if (fwd) launch_pdl(k_relax_fused_rects<false, false>, grid, dim3(RF_NT), 0, st, R);
        else launch_pdl(k_relax_fused_rects<false, true>, grid, dim3(RF_NT), 0, st, R);
    }
    BMGX_LAUNCH_CHECK("relax_GS_oop_rects");
    return BMGX_OK;
}

extern "C" int bmgx2_relax_GS_oop(int k, const double *so, const double *qf, const double *q_in, double *q_out,
                                  const double *sor, bmgx_len ii, bmgx_len jj, int kf, int ifd, int nstncl,
                                  int nsorv, int irelax_sym, int updown, int jpn, bmgx_stream_t s)
{
    (void)nstncl; (void)nsorv;
    if (q_in == q_out) return bmgx_set_error(BMGX_ERR_ARG, "relax_GS_oop: q_in aliases q_out");
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_GS_oop: periodic BCs not supported");
    if (ii < 1 || jj < 1) return BMGX_OK;
    const bool fwd = (irelax_sym == BMGX_RELAX_NONSYM) || (updown == BMGX_DOWN && irelax_sym == BMGX_RELAX_SYM);
    const bool nine = (k < kf || ifd != 1);
    const size_t pl = (size_t)ii * jj;
    Grid g;
    g.kw = so + (KW - 1) * pl; g.ks = so + (KS - 1) * pl;
    g.ksw = nine ? so + (KSW - 1) * pl : nullptr; g.knw = nine ? so + (KNW - 1) * pl : nullptr;
    g.sor = sor + (MSOR - 1) * pl; g.qf = qf; g.II = (int)ii; g.JJ = (int)jj;
    g.P = (long long)ii; g.wi0 = 1; g.wi1 = (int)ii; g.wj0 = 1; g.wj1 = (int)jj;
    unsigned nstrips, band_h, nbands;
    tile(ii, jj, 1, nstrips, band_h, nbands);
    dim3 grid(nstrips, nbands);
    cudaStream_t st = (cudaStream_t)s;
    if (nine) {
        if (fwd) launch_pdl(k_relax9_fused<false>, grid, dim3(RF_NT), 0, st, g, q_in, q_out, (int)band_h);
        else launch_pdl(k_relax9_fused<true>, grid, dim3(RF_NT), 0, st, g, q_in, q_out, (int)band_h);
    } else {
        if (fwd) launch_pdl(k_relax5_fused<false>, grid, dim3(RF_NT), 0, st, g, q_in, q_out, (int)band_h);
        else launch_pdl(k_relax5_fused<true>, grid, dim3(RF_NT), 0, st, g, q_in, q_out, (int)band_h);
    }
    BMGX_LAUNCH_CHECK("relax_GS_oop");
    return BMGX_OK;
}
