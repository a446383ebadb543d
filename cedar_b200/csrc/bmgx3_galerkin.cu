// bmgx3_galerkin.cu -- 3D Galerkin coarse operator A_c = P^T A P.
// Reference: src/3d/ftn/BMG3_SymStd_SETUP_ITLI07_ex.f90 (7-point fine operator),
//            src/3d/ftn/BMG3_SymStd_SETUP_ITLI27_ex.f90 (27-point fine operator):
// 14 loop nests, one per coarse stencil entry, each a hand-expanded sum of products
// CI * SO * CI over the 3x3x3 fine neighbourhoods of the two coarse points.
//
// Here: SOC(S, e) is the coupling between the coarse points c = S - s_e and
// c2 = c + d_e (table below, read off BMG3_SymStd_residual.f90:74-100):
//     sum over f1 in N(c), f2 in N(c2), |f1 - f2|_inf <= 1 of  P(f1,c) (-A)(f1,f2) P(f2,c2)
// with off-diagonals stored positive and the diagonal entry negated back, evaluated
// as P^T (A P) in two stages through a slab of scratch (see "two-stage product"
// below).  Same terms as the reference, different summation order: parity to
// rounding (tests: 1e-12 of the diagonal scale).
#include "bmgx3_common.cuh"

namespace {

struct RowT { int e, sx, sy, sz, dx, dy, dz; };
// for every stored entry e: storage shift s (S = c + s) and neighbour offset d (c2 = c + d)
__device__ __forceinline__ constexpr RowT row_of(int e)
{
    return e == KPW ? RowT{KPW, 0, 0, 0, -1, 0, 0}
         : e == KPS ? RowT{KPS, 0, 0, 0, 0, -1, 0}
         : e == KB ? RowT{KB, 0, 0, 0, 0, 0, -1}
         : e == KPSW ? RowT{KPSW, 0, 0, 0, -1, -1, 0}
         : e == KPNW ? RowT{KPNW, 0, 1, 0, -1, 1, 0}
         : e == KBW ? RowT{KBW, 0, 0, 0, -1, 0, -1}
         : e == KBNW ? RowT{KBNW, 0, 1, 0, -1, 1, -1}
         : e == KBN ? RowT{KBN, 0, 1, 0, 0, 1, -1}
         : e == KBNE ? RowT{KBNE, 1, 1, 0, 1, 1, -1}
         : e == KBE ? RowT{KBE, 1, 0, 0, 1, 0, -1}
         : e == KBSE ? RowT{KBSE, 1, 0, 0, 1, -1, -1}
         : e == KBS ? RowT{KBS, 0, 0, 0, 0, -1, -1}
         : e == KBSW ? RowT{KBSW, 0, 0, 0, -1, -1, -1}
                     : RowT{KP, 0, 0, 0, 0, 0, 0};
}

// name of the weight with which a fine point of parity type (px,py,pz) takes the coarse
// point U - (a,b,c), U its upper coarse point (runtime twin of Wname<> in bmgx3_common.cuh;
// folds to a constant once the loops around it are unrolled)
__device__ __forceinline__ constexpr int wname(int px, int py, int pz, int a, int b, int c)
{
    const int t = px + 2 * py + 4 * pz, s = a + 2 * b + 4 * c;
    return t == 0 ? 0
         : t == 1 ? (a ? LXYL : LXYR)
         : t == 2 ? (b ? LXYB : LXYA)
         : t == 3 ? (s == 0 ? LXYNE : s == 1 ? LXYNW : s == 2 ? LXYSE : LXYSW)
         : t == 4 ? (c ? LXZB : LXZA)
         : t == 5 ? (s == 0 ? LXZNE : s == 1 ? LXZNW : s == 4 ? LXZSE : LXZSW)
         : t == 6 ? (s == 0 ? LYZNW : s == 2 ? LYZNE : s == 4 ? LYZSW : LYZSE)
                  : (s == 0 ? LTNE : s == 1 ? LTNW : s == 2 ? LTSE : s == 3 ? LTSW : s == 4 ? LBNE : s == 5 ? LBNW
                     : s == 6 ? LBSE : LBSW);
}

// entry and storage shift of the coupling p ~ p + (dx,dy,dz) (runtime twin of Cpl<>)
__device__ __forceinline__ double cplr(const Op3 &A, bool seven, int i, int j, int k, int dx, int dy, int dz)
{
    if (seven && (dx != 0) + (dy != 0) + (dz != 0) != 1) return 0.0;
    const int e0 = (dx == -1 && dy == 0) ? KPW : (dx == -1 && dy == 1) ? KPNW : (dx == 0 && dy == 1) ? KPS
                 : (dx == 1 && dy == 1) ? KPSW : (dx == 1 && dy == 0) ? KPW : (dx == 1 && dy == -1) ? KPNW
                 : (dx == 0 && dy == -1) ? KPS : KPSW;
    const int em = (dx == 0 && dy == 0) ? KB : (dx == -1 && dy == 0) ? KBW : (dx == -1 && dy == 1) ? KBNW
                 : (dx == 0 && dy == 1) ? KBN : (dx == 1 && dy == 1) ? KBNE : (dx == 1 && dy == 0) ? KBE
                 : (dx == 1 && dy == -1) ? KBSE : (dx == 0 && dy == -1) ? KBS : KBSW;
    const int ep = (dx == 0 && dy == 0) ? KB : (dx == -1 && dy == 0) ? KBE : (dx == -1 && dy == 1) ? KBSE
                 : (dx == 0 && dy == 1) ? KBS : (dx == 1 && dy == 1) ? KBSW : (dx == 1 && dy == 0) ? KBW
                 : (dx == 1 && dy == -1) ? KBNW : (dx == 0 && dy == -1) ? KBN : KBNE;
    const int e = dz == 0 ? e0 : (dz < 0 ? em : ep);
    return A.at(e, i + (dx > 0), j + (dy > 0), k + (dz > 0));
}

// ---- two-stage product -----------------------------------------------------------
// Stage 1 (k3_ap): T = (-A) P, one row per fine point f.  Fine points are grouped in
// cells: cell (ic,jc,kc) holds the 8 points 2(c-1) - (px,py,pz), p in {0,1}^3 -- the
// C-point c and the 7 F-points whose upper coarse point is c (the convention CI uses).
// Row f of type p reaches the coarse points c + o, o_d in {-1,0} where p_d = 1 and
// {-1,0,1} where p_d = 0: 27 + 3*18 + 3*12 + 8 = 125 values per cell, stored
// structure-of-arrays by (type, o) over the coarse grid, in slabs of coarse planes.
// Stage 2 (k3_ptap): SOC(S,e) = sum over the 27 fine points f of supp(c) of P(f,c) T(f,c2),
// c = S - s_e, c2 = c + d_e; one thread per storage point, 14 entries.
// Every index is a compile-time constant after unrolling; 4x fewer multiplications and
// ~6x fewer loads than evaluating the triple sum per entry.

// number of T values of type (px,py,pz) and offset of the type in the 125-table
__host__ __device__ constexpr int tcount(int px, int py, int pz) { return (3 - px) * (3 - py) * (3 - pz); }
__host__ __device__ constexpr int tbase(int px, int py, int pz)
{
    int b = 0;
    for (int t = 0; t < px + 2 * py + 4 * pz; t++) b += tcount(t & 1, (t >> 1) & 1, (t >> 2) & 1);
    return b;
}
// index of offset o (o_d in -1..1-p_d) inside its type
__host__ __device__ constexpr int tslot(int px, int py, int pz, int ox, int oy, int oz)
{
    return (ox + 1) + (3 - px) * ((oy + 1) + (3 - py) * (oz + 1));
}
enum { NT = 125 };

struct Slab {
    double *t;       // NT * plane * nk doubles
    size_t plane;    // IIC * JJC
    int k0, nk;      // first coarse plane (1-based cell index) and number of planes held
    __device__ __forceinline__ size_t at(int slot, int ic, int jc, int kc, int IIC) const
    {
        return ((size_t)slot * nk + (size_t)(kc - k0)) * plane + (size_t)(jc - 1) * IIC + (size_t)(ic - 1);
    }
};

template <bool SEVEN, int PX, int PY, int PZ>
__device__ __forceinline__ void ap_point(const Op3 &A, const Ci3 &W, const Slab &S, int ic, int jc, int kc)
{
    const int fi = 2 * (ic - 1) - PX, fj = 2 * (jc - 1) - PY, fk = 2 * (kc - 1) - PZ;
    const bool valid = fi >= 2 && fi <= A.II - 1 && fj >= 2 && fj <= A.JJ - 1 && fk >= 2 && fk <= A.KK - 1;
    double acc[27];
#pragma unroll
    for (int n = 0; n < 27; n++) acc[n] = 0.0;
    if (valid) {
#pragma unroll
        for (int dz = -1; dz <= 1; dz++) {
#pragma unroll
            for (int dy = -1; dy <= 1; dy++) {
#pragma unroll
                for (int dx = -1; dx <= 1; dx++) {
                    if (SEVEN && (dx != 0) + (dy != 0) + (dz != 0) > 1) continue;
                    const double a = (dx || dy || dz) ? cplr(A, SEVEN, fi, fj, fk, dx, dy, dz) : -A.diag(fi, fj, fk);
                    // f2 = f + d: type q, upper coarse point U2 = cell + u
                    const int qx = PX ^ (dx != 0), qy = PY ^ (dy != 0), qz = PZ ^ (dz != 0);
                    const int ux = PX ? (dx < 0 ? -1 : 0) : (dx > 0 ? 1 : 0);
                    const int uy = PY ? (dy < 0 ? -1 : 0) : (dy > 0 ? 1 : 0);
                    const int uz = PZ ? (dz < 0 ? -1 : 0) : (dz > 0 ? 1 : 0);
#pragma unroll
                    for (int c = 0; c <= qz; c++) {
#pragma unroll
                        for (int b = 0; b <= qy; b++) {
#pragma unroll
                            for (int a2 = 0; a2 <= qx; a2++) {
                                const int l = wname(qx, qy, qz, a2, b, c);
                                const double w = l ? W.at(l, ic + ux, jc + uy, kc + uz) : 1.0;
                                acc[tslot(PX, PY, PZ, ux - a2, uy - b, uz - c)] += a * w;
                            }
                        }
                    }
                }
            }
        }
    }
#pragma unroll
    for (int n = 0; n < tcount(PX, PY, PZ); n++)
        S.t[S.at(tbase(PX, PY, PZ) + n, ic, jc, kc, W.IIC)] = acc[n];
}

// cells (1..IIC, 1..JJC, slab planes); one thread handles the two x-types of one (py,pz)
template <bool SEVEN, int PY, int PZ>
__global__ void __launch_bounds__(128)
k3_ap(Op3 A, Ci3 W, Slab S)
{
    const int ic = blockIdx.x * blockDim.x + threadIdx.x + 1, jc = blockIdx.y * blockDim.y + threadIdx.y + 1,
              kc = S.k0 + blockIdx.z;
    if (ic > W.IIC || jc > W.JJC) return;
    ap_point<SEVEN, 0, PY, PZ>(A, W, S, ic, jc, kc);
    ap_point<SEVEN, 1, PY, PZ>(A, W, S, ic, jc, kc);
}

template <int E>
__device__ __forceinline__ double ptap_entry(const Ci3 &W, const Slab &S, int ic, int jc, int kc)
{
    constexpr RowT r = row_of(E);
    const int c0 = ic - r.sx, c1 = jc - r.sy, c2 = kc - r.sz;   // c
    double acc = 0.0;
#pragma unroll
    for (int ez = -1; ez <= 1; ez++) {
#pragma unroll
        for (int ey = -1; ey <= 1; ey++) {
#pragma unroll
            for (int ex = -1; ex <= 1; ex++) {
                // f = fine(c) + e lives in cell c + u with type p; T(f, c + d) sits at o = d - u
                const int px = ex != 0, py = ey != 0, pz = ez != 0;
                const int ux = ex > 0, uy = ey > 0, uz = ez > 0;
                const int ox = r.dx - ux, oy = r.dy - uy, oz = r.dz - uz;
                if (ox < -1 || ox > 1 - px || oy < -1 || oy > 1 - py || oz < -1 || oz > 1 - pz) continue;
                const int l = wname(px, py, pz, ux, uy, uz);
                const double w = l ? W.at(l, c0 + ux, c1 + uy, c2 + uz) : 1.0;
                acc += w * S.t[S.at(tbase(px, py, pz) + tslot(px, py, pz, ox, oy, oz), c0 + ux, c1 + uy, c2 + uz, W.IIC)];
            }
        }
    }
    return E == KP ? -acc : acc;
}

// storage points (2..IIC-1, 2..JJC-1, planes ka..kb)
__global__ void __launch_bounds__(128, 3)
k3_ptap(Ci3 W, Slab S, double *__restrict__ soc, int ka)
{
    const int ic = blockIdx.x * blockDim.x + threadIdx.x + 2, jc = blockIdx.y * blockDim.y + threadIdx.y + 2,
              kc = ka + blockIdx.z;
    if (ic > W.IIC - 1 || jc > W.JJC - 1) return;
    const size_t cc = ((size_t)(kc - 1) * W.JJC + (jc - 1)) * W.IIC + (ic - 1);
#define ENTRY(E) soc[(size_t)(E - 1) * W.pl + cc] = ptap_entry<E>(W, S, ic, jc, kc);
    ENTRY(1) ENTRY(2) ENTRY(3) ENTRY(4) ENTRY(5) ENTRY(6) ENTRY(7) ENTRY(8) ENTRY(9) ENTRY(10) ENTRY(11) ENTRY(12)
    ENTRY(13) ENTRY(14)
#undef ENTRY
}

template <bool SEVEN>
int launch(const Op3 &A, const Ci3 &W, double *soc, cudaStream_t st)
{
    // slabs of coarse planes bound the scratch: storage planes ka..kb need the cells of planes ka-1..kb+1
    const int SLAB = 32;
    const size_t plane = (size_t)W.IIC * W.JJC;
    const int nkmax = min(SLAB + 2, W.KKC);
    Slab S;
    S.plane = plane;
    // stream-ordered scratch from the library's own pool (bmgx_util.cu): cached across setups, the application's
    // default pool is left alone
    cudaError_t e = bmgx_scratch_alloc((void **)&S.t, (size_t)NT * plane * nkmax * sizeof(double), st);
    if (e != cudaSuccess) return bmgx_set_cuda_error(e, "SETUP_ITLI3/scratch");
    dim3 b(32, 4);
    for (int ka = 2; ka <= W.KKC - 1; ka += SLAB) {
        const int kb = min(ka + SLAB - 1, W.KKC - 1);
        S.k0 = ka - 1;
        S.nk = kb + 1 - S.k0 + 1;
        dim3 g1(cdiv(W.IIC, 32), cdiv(W.JJC, 4), S.nk);
        k3_ap<SEVEN, 0, 0><<<g1, b, 0, st>>>(A, W, S);
        k3_ap<SEVEN, 1, 0><<<g1, b, 0, st>>>(A, W, S);
        k3_ap<SEVEN, 0, 1><<<g1, b, 0, st>>>(A, W, S);
        k3_ap<SEVEN, 1, 1><<<g1, b, 0, st>>>(A, W, S);
        dim3 g2(cdiv(W.IIC - 2, 32), cdiv(W.JJC - 2, 4), kb - ka + 1);
        k3_ptap<<<g2, b, 0, st>>>(W, S, soc, ka);
        cudaError_t le = cudaGetLastError();
        if (le != cudaSuccess) { bmgx_scratch_free(S.t, st); return bmgx_set_cuda_error(le, "SETUP_ITLI3"); }
        bmgx_count_launch(5);
    }
    e = bmgx_scratch_free(S.t, st);
    if (e != cudaSuccess) return bmgx_set_cuda_error(e, "SETUP_ITLI3/free");
    return BMGX_OK;
}

int galerkin3(bool seven, const double *so, double *soc, const double *ci, bmgx_len iif, bmgx_len jjf, bmgx_len kkf,
              bmgx_len iic, bmgx_len jjc, bmgx_len kkc, int ipn, bmgx_stream_t s)
{
    if (!bmgx_nonperiodic(ipn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "SETUP_ITLI3: periodic BCs not supported");
    if (iic < 3 || jjc < 3 || kkc < 3) return BMGX_OK;
    Op3 A(so, iif, jjf, kkf);
    Ci3 W(ci, iic, jjc, kkc);
    return seven ? launch<true>(A, W, soc, (cudaStream_t)s) : launch<false>(A, W, soc, (cudaStream_t)s);
}

}  // namespace

// BMG3_SymStd_SETUP_ITLI07_ex(so, soc, ci, iif, jjf, kkf, iic, jjc, kkc, ipn), include/cedar/3d/coarsen.h:12-14
extern "C" int bmgx3_SETUP_ITLI07_ex(const double *so, double *soc, const double *ci, bmgx_len iif, bmgx_len jjf,
                                     bmgx_len kkf, bmgx_len iic, bmgx_len jjc, bmgx_len kkc, int ipn, bmgx_stream_t s)
{
    return galerkin3(true, so, soc, ci, iif, jjf, kkf, iic, jjc, kkc, ipn, s);
}

// BMG3_SymStd_SETUP_ITLI27_ex(...), include/cedar/3d/coarsen.h:15-16
extern "C" int bmgx3_SETUP_ITLI27_ex(const double *so, double *soc, const double *ci, bmgx_len iif, bmgx_len jjf,
                                     bmgx_len kkf, bmgx_len iic, bmgx_len jjc, bmgx_len kkc, int ipn, bmgx_stream_t s)
{
    return galerkin3(false, so, soc, ci, iif, jjf, kkf, iic, jjc, kkc, ipn, s);
}
