// bmgx3_relax_fused.cu -- 3D 7-point red-black Gauss-Seidel, BOTH colours in ONE pass over
// memory (out of place: q_in -> q_out).
//
// Same arithmetic as k3_relax_colour<true> (bmgx3_stencil.cu; reference
// src/3d/ftn/BMG3_SymStd_relax_GS.f90:142-177), same term order, -fmad=false => bitwise the
// oracle's result.  Points of one colour are independent, so the schedule is free:
//
//   * a CTA owns a tile of R7_OX x R7_OY columns (plus a halo of 2 on each side whose values
//     are recomputed, never stored) and a band of planes, and walks up the planes;
//   * the iterate lives in a shared-memory ring of 5 planes; per step p the first colour is
//     relaxed on plane p (it needs only old values of planes p-1, p, p+1), then the second
//     colour on plane p-1 (whose six neighbours are now final), then plane p-1 is stored;
//   * a thread owns two (column pair, row) items.  In plane p one column of a pair has the
//     first colour, the other the second, and the roles swap from plane to plane, so the
//     thread loads the coefficients of BOTH columns of plane p together (full sectors),
//     uses one column's set at once and carries the other column's set in registers to the
//     next step, where that column gets its second-colour update;
//   * tiles overlap by the halo and bands by two planes, so CTAs are independent: no
//     inter-CTA ordering, no atomics.
//
// Algorithmic traffic 56 B/point (PW, PS, B, SOR, QF, Q read + Q write); the per-colour
// kernels move 2x that (every array is pulled once per colour, half of each sector unused).
#include "bmgx3_common.cuh"

#define R7_SX 64
// tile 64 x 32 with 512 threads, one CTA per SM (128 registers): 385^3 0.926 ms, 513^3 2.08 ms per sweep; 64 x 16 with 256
// threads and two CTAs per SM: 0.935 / 2.2 ms (fewer recomputed halo rows win); 80 registers for three CTAs spill: 1.67 ms
#ifndef R7_SY
#define R7_SY 32
#endif
#ifndef R7_NT
#define R7_NT 512      // R7_SX * R7_SY / 4: four iterate elements and two (column pair, row) items per thread
#endif
#ifndef R7_MINB
#define R7_MINB 1
#endif
#define R7_WR (R7_NT / 32)   // rows between the two items of a thread
#define R7_OX (R7_SX - 4)
#define R7_OY (R7_SY - 4)
#define R7_RING 5

namespace {

struct G7 {
    const double *pw, *ps, *kb, *sor, *qf;
    int II, JJ, KK;
};

// coefficients of one column of one plane that its update needs (the coupling to the plane
// above is stored in that plane and joins later)
struct Col7 { double pw, pwe, ps, psn, kbd, sor, qf; };

// FIRST_ODD: the first colour is (i+j+k) odd (DOWN sweep); else even (UP)
// INNER: the whole tile, halo included, lies inside the array in x and y (and its update window inside the interior), so
// every in-plane bounds test is true at compile time; two thirds of the CTAs of a 385^3 sweep
template <bool FIRST_ODD, bool INNER>
__device__ __forceinline__ void relax7_body(const G7 &g, const double *__restrict__ qin, double *__restrict__ qout, int band_h)
{
    extern __shared__ double ring_raw[];
    double (*ring)[R7_SY][R7_SX] = reinterpret_cast<double (*)[R7_SY][R7_SX]>(ring_raw);
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int II = g.II, JJ = g.JJ, KK = g.KK;
    const int gi0 = 1 + (int)blockIdx.x * R7_OX - 2, gj0 = 1 + (int)blockIdx.y * R7_OY - 2;
    const int kb = 1 + (int)blockIdx.z * band_h, ke = min(KK, kb + band_h - 1);
    constexpr int pF = FIRST_ODD ? 1 : 0;

    // addressing is hoisted out of the plane loop: per thread four iterate elements and two
    // (column pair, row) items, each an in-plane offset plus validity masks; a plane adds PJ
    const long long PJ = (long long)II * JJ;
    long long eoff[4];
    bool eok[4];
#pragma unroll
    for (int n = 0; n < 4; n++) {
        const int e = t + n * R7_NT, gi = gi0 + (e & 63), gj = gj0 + (e >> 6);
        eok[n] = INNER || (gi >= 1 && gi <= II && gj >= 1 && gj <= JJ);
        eoff[n] = eok[n] ? (long long)(gj - 1) * II + (gi - 1) : 0;
    }
    auto load_plane = [&](int p, double (&v)[4]) {
        const bool pok = p >= 1 && p <= KK;
        const double *src = qin + (long long)(p - 1) * PJ;
#pragma unroll
        for (int n = 0; n < 4; n++) v[n] = (pok && eok[n]) ? __ldg(src + eoff[n]) : 0.0;
    };
    auto put_plane = [&](int slot, const double (&v)[4]) {
#pragma unroll
        for (int n = 0; n < 4; n++) {
            const int e = t + n * R7_NT;
            ring[slot][e >> 6][e & 63] = v[n];
        }
    };

    // ring slots of planes p-2, p-1, p, p+1, p+2, rotated once per step
    int s_m2 = 0, s_m1 = 1, s_0 = 2, s_p1 = 3, s_p2 = 4;
    {
        // step p = kb-1 needs planes kb-2, kb-1, kb (kb-3 is never read)
        double v[4];
        load_plane(kb - 2, v); put_plane(s_m1, v);
        load_plane(kb - 1, v); put_plane(s_0, v);
        load_plane(kb, v);     put_plane(s_p1, v);
    }
    // items: rows ly = w and w + 8, column pair (2 lane, 2 lane + 1)
    const int c0 = gi0 + 2 * lane;
    const bool k0 = INNER || (c0 >= 1 && c0 <= II), k1 = INNER || (c0 + 1 >= 1 && c0 + 1 <= II),
               k2 = INNER || (c0 + 2 >= 1 && c0 + 2 <= II);   // (the pair of the last lane reaches one column past the tile)
    long long ioff[2];
    bool rok[2], rnok[2];
    Col7 sv[2];          // second-colour column set of plane p-1
    double kbc[2][2];    // KB(c0,p), KB(c1,p)
#pragma unroll
    for (int n = 0; n < 2; n++) {
        const int gj = gj0 + w + R7_WR * n;
        rok[n] = INNER || (gj >= 1 && gj <= JJ);
        rnok[n] = INNER || (gj + 1 >= 1 && gj + 1 <= JJ);
        ioff[n] = (long long)(gj - 1) * II + (c0 - 1);
        sv[n] = Col7{0, 0, 0, 0, 0, 0, 0};
        const bool pok = kb - 1 >= 1 && kb - 1 <= KK;
        const double *kp = g.kb + (long long)(kb - 2) * PJ + ioff[n];
        kbc[n][0] = (pok && rok[n] && k0) ? __ldg(kp) : 0.0;
        kbc[n][1] = (pok && rok[n] && k1) ? __ldg(kp + 1) : 0.0;
    }
    __syncthreads();

    for (int p = kb - 1; p <= ke + 1; p++) {
        // ---- loads of this step: coefficients of plane p (both columns), KB of plane p+1, iterate plane p+2
        double pw0[2], pw1[2], pw2[2], ps0[2], ps1[2], pn0[2], pn1[2], kn0[2], kn1[2], so0[2], so1[2], qf0[2], qf1[2];
        const bool pok = p >= 1 && p <= KK, pnok = p + 1 >= 1 && p + 1 <= KK;
        const long long po = (long long)(p - 1) * PJ;
#pragma unroll
        for (int n = 0; n < 2; n++) {
            const long long o = po + ioff[n];
            const bool r = pok && rok[n], rn = pok && rnok[n], ru = pnok && rok[n];
            pw0[n] = (r && k0) ? __ldg(g.pw + o) : 0.0;
            pw1[n] = (r && k1) ? __ldg(g.pw + o + 1) : 0.0;
            pw2[n] = (r && k2) ? __ldg(g.pw + o + 2) : 0.0;
            ps0[n] = (r && k0) ? __ldg(g.ps + o) : 0.0;
            ps1[n] = (r && k1) ? __ldg(g.ps + o + 1) : 0.0;
            pn0[n] = (rn && k0) ? __ldg(g.ps + o + II) : 0.0;
            pn1[n] = (rn && k1) ? __ldg(g.ps + o + II + 1) : 0.0;
            kn0[n] = (ru && k0) ? __ldg(g.kb + o + PJ) : 0.0;
            kn1[n] = (ru && k1) ? __ldg(g.kb + o + PJ + 1) : 0.0;
            so0[n] = (r && k0) ? __ldg(g.sor + o) : 0.0;
            so1[n] = (r && k1) ? __ldg(g.sor + o + 1) : 0.0;
            qf0[n] = (r && k0) ? __ldg(g.qf + o) : 0.0;
            qf1[n] = (r && k1) ? __ldg(g.qf + o + 1) : 0.0;
        }
        double nxt[4];
        load_plane(p + 2, nxt);
        // the next step's coefficient sectors are requested now (no register, never waited on), so that
        // its loads -- consumed straight away -- find them in L2 instead of paying the DRAM latency
        if (p + 1 <= KK) {
#pragma unroll
            for (int n = 0; n < 2; n++) {
                if (rok[n] && k0) {
                    const long long o = po + PJ + ioff[n];
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(g.pw + o));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(g.ps + o));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(g.sor + o));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(g.qf + o));
                    if (p + 2 <= KK) asm volatile("prefetch.global.L2 [%0];" ::"l"(g.kb + o + PJ));
                }
            }
        }

        const bool pin = p >= 2 && p <= KK - 1;
        bool a0[2];
        // ---- first colour on plane p
#pragma unroll
        for (int n = 0; n < 2; n++) {
            const int ly = w + R7_WR * n, gj = gj0 + ly;
            a0[n] = (((c0 + gj + p) & 1) == pF);
            const int lx = 2 * lane + (a0[n] ? 0 : 1), ca = c0 + (a0[n] ? 0 : 1);
            if (pin && (INNER || (gj >= 2 && gj <= JJ - 1 && ca >= 2 && ca <= II - 1)) && lx >= 1 && lx <= R7_SX - 2 && ly >= 1
                && ly <= R7_SY - 2) {
                const double cw = a0[n] ? pw0[n] : pw1[n], ce = a0[n] ? pw1[n] : pw2[n];
                const double cs = a0[n] ? ps0[n] : ps1[n], cn = a0[n] ? pn0[n] : pn1[n];
                const double cd = a0[n] ? kbc[n][0] : kbc[n][1], cu = a0[n] ? kn0[n] : kn1[n];
                const double f = a0[n] ? qf0[n] : qf1[n], r = a0[n] ? so0[n] : so1[n];
                const double v = f + cw * ring[s_0][ly][lx - 1] + cn * ring[s_0][ly + 1][lx] + ce * ring[s_0][ly][lx + 1]
                                 + cs * ring[s_0][ly - 1][lx] + cd * ring[s_m1][ly][lx] + cu * ring[s_p1][ly][lx];
                ring[s_0][ly][lx] = v * r;
            }
        }
        __syncthreads();
        // ---- second colour on plane p-1 (same columns), coefficients carried from the previous step
        const bool bin = p - 1 >= max(2, kb) && p - 1 <= min(KK - 1, ke);
#pragma unroll
        for (int n = 0; n < 2; n++) {
            const int ly = w + R7_WR * n, gj = gj0 + ly;
            const int lx = 2 * lane + (a0[n] ? 0 : 1), ca = c0 + (a0[n] ? 0 : 1);
            if (bin && (INNER || (gj >= 2 && gj <= JJ - 1 && ca >= 2 && ca <= II - 1)) && lx >= 2 && lx <= R7_SX - 3 && ly >= 2
                && ly <= R7_SY - 3) {
                const double cu = a0[n] ? kbc[n][0] : kbc[n][1];
                const double v = sv[n].qf + sv[n].pw * ring[s_m1][ly][lx - 1] + sv[n].psn * ring[s_m1][ly + 1][lx]
                                 + sv[n].pwe * ring[s_m1][ly][lx + 1] + sv[n].ps * ring[s_m1][ly - 1][lx]
                                 + sv[n].kbd * ring[s_m2][ly][lx] + cu * ring[s_0][ly][lx];
                ring[s_m1][ly][lx] = v * sv[n].sor;
            }
            // the other column of plane p gets its second-colour update in the next step
            if (a0[n]) sv[n] = Col7{pw1[n], pw2[n], ps1[n], pn1[n], kbc[n][1], so1[n], qf1[n]};
            else sv[n] = Col7{pw0[n], pw1[n], ps0[n], pn0[n], kbc[n][0], so0[n], qf0[n]};
            kbc[n][0] = kn0[n];
            kbc[n][1] = kn1[n];
        }
        put_plane(s_p2, nxt);   // slot of plane p-3: last read by the previous step
        __syncthreads();
        // ---- plane p-1 is final: store the tile's own columns (ghost ring included as loaded)
        if (p - 1 >= kb && p - 1 <= ke) {
#pragma unroll
            for (int n = 0; n < 4; n++) {
                const int e = t + n * R7_NT, ly = e >> 6, lx = e & 63;
                if (lx >= 2 && lx <= R7_SX - 3 && ly >= 2 && ly <= R7_SY - 3 && eok[n])
                    qout[(long long)(p - 2) * PJ + eoff[n]] = ring[s_m1][ly][lx];
            }
        }
        // rotate: plane p+1 becomes plane p
        const int freed = s_m2;
        s_m2 = s_m1; s_m1 = s_0; s_0 = s_p1; s_p1 = s_p2; s_p2 = freed;
    }
}

template <bool FIRST_ODD>
__global__ void __launch_bounds__(R7_NT, R7_MINB)
k3_relax7_fused(G7 g, const double *__restrict__ qin, double *__restrict__ qout, int band_h)
{
    const int gi0 = 1 + (int)blockIdx.x * R7_OX - 2, gj0 = 1 + (int)blockIdx.y * R7_OY - 2;
    // halo inside the array (one column more for the last lane's pair) and the update window inside the interior
    const bool inner = gi0 >= 1 && gi0 + R7_SX <= g.II && gj0 >= 1 && gj0 + R7_SY <= g.JJ;
    if (inner) relax7_body<FIRST_ODD, true>(g, qin, qout, band_h);
    else relax7_body<FIRST_ODD, false>(g, qin, qout, band_h);
}

}  // namespace

// Out-of-place form of bmgx3_relax_GS for the 7-point operator (ifd == 1): reads q_in, writes
// the relaxed iterate (interior + copied ghost layer) to q_out.
extern "C" int bmgx3_relax_GS_oop(int kg, const double *so, const double *qf, const double *q_in, double *q_out,
                                  const double *sor, bmgx_len ii, bmgx_len jj, bmgx_len kk, int ifd, int nstncl,
                                  int nsorv, int irelax_sym, int updown, int jpn, bmgx_stream_t s)
{
    (void)kg; (void)nstncl; (void)nsorv;
    if (q_in == q_out) return bmgx_set_error(BMGX_ERR_ARG, "relax_GS3_oop: q_in aliases q_out");
    if (ifd != 1) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_GS3_oop: 7-point operators only");
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "relax_GS3_oop: periodic BCs not supported");
    if (ii < 1 || jj < 1 || kk < 1) return BMGX_OK;
    const bool up = (updown == BMGX_UP) || (irelax_sym == BMGX_RELAX_NONSYM);
    const size_t pl = (size_t)ii * jj * kk;
    G7 g;
    g.pw = so + (KPW - 1) * pl; g.ps = so + (KPS - 1) * pl; g.kb = so + (KB - 1) * pl;
    g.sor = sor + (MSOR - 1) * pl; g.qf = qf; g.II = (int)ii; g.JJ = (int)jj; g.KK = (int)kk;
    const unsigned ntx = cdiv(ii, R7_OX), nty = cdiv(jj, R7_OY);
    // about four resident waves (148 SMs x R7_MINB CTAs) of equal bands
    unsigned nb = (4 * 148 * R7_MINB + ntx * nty - 1) / (ntx * nty);
    if (nb < 1) nb = 1;
    unsigned band_h = (kk + nb - 1) / nb;
    if (band_h < 8) band_h = 8;
    dim3 grid(ntx, nty, cdiv(kk, band_h));
    // DOWN: pts = 1 first, i.e. (i+j+k) odd (relax_GS.f90:144-154); UP: even first
    const size_t smem = sizeof(double) * R7_RING * R7_SY * R7_SX;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k3_relax7_fused<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k3_relax7_fused<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return bmgx_set_cuda_error(e, "relax_GS3_oop/attr");
        attr_set = true;
    }
    if (!up) k3_relax7_fused<true><<<grid, R7_NT, smem, (cudaStream_t)s>>>(g, q_in, q_out, (int)band_h);
    else k3_relax7_fused<false><<<grid, R7_NT, smem, (cudaStream_t)s>>>(g, q_in, q_out, (int)band_h);
    BMGX_LAUNCH_CHECK("relax_GS3_oop");
    return BMGX_OK;
}
