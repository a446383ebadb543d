// bmgx3_galerkin.cu -- 3D Galerkin coarse operator A_c = P^T A P.
// Reference: src/3d/ftn/BMG3_SymStd_SETUP_ITLI07_ex.f90 (7-point fine operator),
//            src/3d/ftn/BMG3_SymStd_SETUP_ITLI27_ex.f90 (27-point fine operator):
// 14 loop nests, one per coarse stencil entry, each a hand-expanded sum of products
// CI * SO * CI over the 3x3x3 fine neighbourhoods of the two coarse points.
//
// Here: one thread per coarse storage point S computes its 14 entries from the
// definition.  SOC(S, e) is the coupling between the coarse points c = S - s_e and
// c2 = c + d_e (table below, read off BMG3_SymStd_residual.f90:74-100):
//     sum over f1 in N(c), f2 in N(c2), |f1 - f2|_inf <= 1 of  P(f1,c) (-A)(f1,f2) P(f2,c2)
// with off-diagonals stored positive and the diagonal entry negated back.  All
// offsets are compile-time after unrolling, so each entry is a straight-line sum
// of exactly the non-vanishing products.  Same terms as the reference, different
// summation order: parity to rounding (tests: 1e-12 of the diagonal scale).
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

// P(f, c) for f = fine(c) + (x,y,z), x,y,z in {-1,0,1}; (ic,jc,kc) = c
__device__ __forceinline__ double pw(const Ci3 &W, int ic, int jc, int kc, int x, int y, int z)
{
    if (x == 0 && y == 0 && z == 0) return 1.0;
    const int l = wname(x != 0, y != 0, z != 0, x > 0, y > 0, z > 0);
    return W.at(l, ic + (x > 0), jc + (y > 0), kc + (z > 0));
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

template <bool SEVEN, int E>
__device__ __forceinline__ double entry(const Op3 &A, const Ci3 &W, int ic, int jc, int kc)
{
    constexpr RowT r = row_of(E);
    const int c0 = ic - r.sx, c1 = jc - r.sy, c2 = kc - r.sz;               // c
    const int g0 = 2 * (c0 - 1), g1 = 2 * (c1 - 1), g2 = 2 * (c2 - 1);      // fine(c)
    double acc = 0.0;
#pragma unroll
    for (int z = -1; z <= 1; z++) {
#pragma unroll
        for (int y = -1; y <= 1; y++) {
#pragma unroll
            for (int x = -1; x <= 1; x++) {
                double row = 0.0;
#pragma unroll
                for (int dz = -1; dz <= 1; dz++) {
#pragma unroll
                    for (int dy = -1; dy <= 1; dy++) {
#pragma unroll
                        for (int dx = -1; dx <= 1; dx++) {
                            // f2 = f1 + (dx,dy,dz) relative to fine(c2) = fine(c) + 2 d
                            const int rx = x + dx - 2 * r.dx, ry = y + dy - 2 * r.dy, rz = z + dz - 2 * r.dz;
                            if (rx < -1 || rx > 1 || ry < -1 || ry > 1 || rz < -1 || rz > 1) continue;
                            if (SEVEN && (dx != 0) + (dy != 0) + (dz != 0) > 1) continue;
                            const double a = (dx || dy || dz) ? cplr(A, SEVEN, g0 + x, g1 + y, g2 + z, dx, dy, dz)
                                                              : -A.diag(g0 + x, g1 + y, g2 + z);
                            row += a * pw(W, c0 + r.dx, c1 + r.dy, c2 + r.dz, rx, ry, rz);
                        }
                    }
                }
                acc += pw(W, c0, c1, c2, x, y, z) * row;
            }
        }
    }
    return E == KP ? -acc : acc;
}

template <bool SEVEN, int E0, int E1>
__global__ void __launch_bounds__(128)
k3_galerkin(Op3 A, Ci3 W, double *__restrict__ soc)
{
    const int ic = blockIdx.x * blockDim.x + threadIdx.x + 2, jc = blockIdx.y * blockDim.y + threadIdx.y + 2,
              kc = blockIdx.z + 2;
    if (ic > W.IIC - 1 || jc > W.JJC - 1) return;
    const size_t cc = ((size_t)(kc - 1) * W.JJC + (jc - 1)) * W.IIC + (ic - 1);
#define ENTRY(E) if (E >= E0 && E <= E1) soc[(size_t)(E - 1) * W.pl + cc] = entry<SEVEN, E>(A, W, ic, jc, kc);
    ENTRY(1) ENTRY(2) ENTRY(3) ENTRY(4) ENTRY(5) ENTRY(6) ENTRY(7) ENTRY(8) ENTRY(9) ENTRY(10) ENTRY(11) ENTRY(12)
    ENTRY(13) ENTRY(14)
#undef ENTRY
}

template <bool SEVEN>
int launch(const Op3 &A, const Ci3 &W, double *soc, cudaStream_t st)
{
    dim3 b(32, 4), g(cdiv(W.IIC - 2, 32), cdiv(W.JJC - 2, 4), W.KKC - 2);
    // split over the entries to bound register pressure / code size per kernel
    k3_galerkin<SEVEN, 1, 4><<<g, b, 0, st>>>(A, W, soc);
    BMGX_LAUNCH_CHECK("SETUP_ITLI3/a");
    k3_galerkin<SEVEN, 5, 9><<<g, b, 0, st>>>(A, W, soc);
    BMGX_LAUNCH_CHECK("SETUP_ITLI3/b");
    k3_galerkin<SEVEN, 10, 14><<<g, b, 0, st>>>(A, W, soc);
    BMGX_LAUNCH_CHECK("SETUP_ITLI3/c");
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
