// bmgx3_common.cuh -- 3D stencil storage conventions shared by the 3D kernels.
//
// Reference storage (src/3d/ftn/BMG_stencils_f90.h:43-64): a symmetric 7/27-point
// operator keeps, per grid point S=(I,J,K), the diagonal and the 3 (13) couplings
// to the "lower" half of its neighbourhood, OFF-DIAGONALS STORED POSITIVE:
//   kp  S            kpw  S~S-x          kps  S~S-y           kb   S~S-z
//   kpsw S~S-x-y     kpnw (S-y)~(S-x)
//   kbw S~S-x-z      kbnw (S-y)~(S-x-z)  kbn (S-y)~(S-z)      kbne (S-x-y)~(S-z)
//   kbe (S-x)~(S-z)  kbse (S-x)~(S-y-z)  kbs S~S-y-z          kbsw S~S-x-y-z
// (read off BMG3_SymStd_residual.f90:67-100).  coupling<DX,DY,DZ>() below turns
// "the coupling between point p and p+(DX,DY,DZ)" into the entry and the storage
// point, so that the setup kernels can be written in neighbour-offset form.
//
// Interpolation weights (BMG_stencils_f90.h:52-64; read off
// BMG3_SymStd_interp_add.f90:100-230): CI(ic,jc,kc,.) holds the weights of the
// fine points whose "upper" coarse point (x+, y+, z+ side) is (ic,jc,kc).
#pragma once
#include "bmgx_common.cuh"

#define IX4(n1, n2, n3, i, j, k, l) \
    ((((size_t)((l)-1) * (size_t)(n3) + (size_t)((k)-1)) * (size_t)(n2) + (size_t)((j)-1)) * (size_t)(n1) + (size_t)((i)-1))

enum { KP = 1, KPW = 2, KPS = 3, KB = 4, KPSW = 5, KPNW = 6, KBW = 7, KBNW = 8, KBN = 9, KBNE = 10, KBE = 11,
       KBSE = 12, KBS = 13, KBSW = 14 };
enum { LXYL = 1, LXYR = 2, LXYA = 3, LXYB = 4, LXZA = 5, LXZB = 6, LXYNE = 7, LXYSE = 8, LXYSW = 9, LXYNW = 10,
       LXZSW = 11, LXZNW = 12, LXZNE = 13, LXZSE = 14, LYZSW = 15, LYZNW = 16, LYZNE = 17, LYZSE = 18, LBSW = 19,
       LBNW = 20, LBNE = 21, LBSE = 22, LTSW = 23, LTNW = 24, LTNE = 25, LTSE = 26 };

// fine operator view: SO(II,JJ,KK,NSTN), NSTN = 4 (7-point) or 14 (27-point)
struct Op3 {
    const double *so;
    int II, JJ, KK;
    size_t pl;  // II*JJ*KK
    __host__ __device__ Op3(const double *s, int ii, int jj, int kk) : so(s), II(ii), JJ(jj), KK(kk), pl((size_t)ii * jj * kk) {}
    // entry e at storage point (i,j,k); 0 outside the array (reads past the ghost layer)
    __device__ __forceinline__ double at(int e, int i, int j, int k) const
    {
        if (i < 1 || i > II || j < 1 || j > JJ || k < 1 || k > KK) return 0.0;
        return __ldg(so + (size_t)(e - 1) * pl + ((size_t)(k - 1) * JJ + (j - 1)) * II + (i - 1));
    }
    __device__ __forceinline__ double diag(int i, int j, int k) const { return at(KP, i, j, k); }
};

// entry and storage offset of the coupling p ~ p+(DX,DY,DZ)
template <int DX, int DY, int DZ> struct Cpl {
    // within the plane (DZ == 0)
    static constexpr int e0 = (DX == -1 && DY == 0) ? KPW : (DX == -1 && DY == 1) ? KPNW : (DX == 0 && DY == 1) ? KPS
                            : (DX == 1 && DY == 1) ? KPSW : (DX == 1 && DY == 0) ? KPW : (DX == 1 && DY == -1) ? KPNW
                            : (DX == 0 && DY == -1) ? KPS : KPSW;
    // towards the plane below (DZ == -1): stored in p's plane
    static constexpr int em = (DX == 0 && DY == 0) ? KB : (DX == -1 && DY == 0) ? KBW : (DX == -1 && DY == 1) ? KBNW
                            : (DX == 0 && DY == 1) ? KBN : (DX == 1 && DY == 1) ? KBNE : (DX == 1 && DY == 0) ? KBE
                            : (DX == 1 && DY == -1) ? KBSE : (DX == 0 && DY == -1) ? KBS : KBSW;
    // towards the plane above (DZ == +1): stored in the plane above, mirrored names
    static constexpr int ep = (DX == 0 && DY == 0) ? KB : (DX == -1 && DY == 0) ? KBE : (DX == -1 && DY == 1) ? KBSE
                            : (DX == 0 && DY == 1) ? KBS : (DX == 1 && DY == 1) ? KBSW : (DX == 1 && DY == 0) ? KBW
                            : (DX == 1 && DY == -1) ? KBNW : (DX == 0 && DY == -1) ? KBN : KBNE;
    static constexpr int entry = DZ == 0 ? e0 : (DZ < 0 ? em : ep);
    static constexpr int sx = DX > 0 ? 1 : 0;
    static constexpr int sy = DY > 0 ? 1 : 0;
    static constexpr int sz = DZ > 0 ? 1 : 0;
    static constexpr bool seven = (DX != 0) + (DY != 0) + (DZ != 0) == 1;   // present in a 7-point operator
};

// coupling (stored positive) between p=(i,j,k) and p+(DX,DY,DZ); SEVEN: 7-point storage
template <bool SEVEN, int DX, int DY, int DZ>
__device__ __forceinline__ double cpl(const Op3 &A, int i, int j, int k)
{
    static_assert(DX != 0 || DY != 0 || DZ != 0, "off-diagonal only");
    using C = Cpl<DX, DY, DZ>;
    if (SEVEN && !C::seven) return 0.0;
    return A.at(C::entry, i + C::sx, j + C::sy, k + C::sz);
}

// interpolation weight view CI(IIC,JJC,KKC,26)
struct Ci3 {
    const double *ci;
    int IIC, JJC, KKC;
    size_t pl;
    __host__ __device__ Ci3(const double *c, int ii, int jj, int kk) : ci(c), IIC(ii), JJC(jj), KKC(kk), pl((size_t)ii * jj * kk) {}
    __device__ __forceinline__ double at(int l, int ic, int jc, int kc) const
    {
        if (ic < 1 || ic > IIC || jc < 1 || jc > JJC || kc < 1 || kc > KKC) return 0.0;
        return ci[(size_t)(l - 1) * pl + ((size_t)(kc - 1) * JJC + (jc - 1)) * IIC + (ic - 1)];
    }
};

// name of the weight with which a fine point of parity type (PX,PY,PZ) (1 = between
// two coarse points along that axis) takes the coarse point U-(A,B,C), U its upper
// coarse point.  0 = injection (weight 1), -1 = no such weight.
template <int PX, int PY, int PZ, int A, int B, int C> struct Wname {
    static constexpr int t = PX + 2 * PY + 4 * PZ, s = A + 2 * B + 4 * C;
    static constexpr int value =
        (A > PX || B > PY || C > PZ) ? -1
        : t == 0 ? 0
        : t == 1 ? (A ? LXYL : LXYR)
        : t == 2 ? (B ? LXYB : LXYA)
        : t == 3 ? (s == 0 ? LXYNE : s == 1 ? LXYNW : s == 2 ? LXYSE : LXYSW)
        : t == 4 ? (C ? LXZB : LXZA)
        : t == 5 ? (s == 0 ? LXZNE : s == 1 ? LXZNW : s == 4 ? LXZSE : LXZSW)
        : t == 6 ? (s == 0 ? LYZNW : s == 2 ? LYZNE : s == 4 ? LYZSW : LYZSE)
        : (s == 0 ? LTNE : s == 1 ? LTNW : s == 2 ? LTSE : s == 3 ? LTSW : s == 4 ? LBNE : s == 5 ? LBNW
           : s == 6 ? LBSE : LBSW);
};
