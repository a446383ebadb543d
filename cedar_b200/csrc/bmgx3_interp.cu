// bmgx3_interp.cu -- 3D operator-induced interpolation setup.
// Reference: src/3d/ftn/BMG3_SymStd_SETUP_interp_OI.f90:82-805 (non-periodic; 27-point
// branch 120-537, 7-point branch 539-805).
//
// Written in neighbour-offset form (C(dx,dy,dz) = coupling between the F-point p and
// p+(dx,dy,dz), see bmgx3_common.cuh) instead of the reference's hand-expanded index
// lists; the sums contain the same terms but are added in a different order, so
// parity with the reference is to rounding (tests: 1e-13 relative), not bitwise.
//
// Three dependent stages, one launch each:
//   edges   F-points between two C-points on a coarse line: the operator is collapsed
//           onto the line (f90:128-226);
//   faces   F-points in the middle of a coarse-plane cell: collapsed onto the plane,
//           corner weights use the edge weights (f90:228-381);
//   centres F-points in the middle of a coarse cell: use edge and face weights (f90:383-537).
// The reference's quirks are kept on purpose (f90:419-446: in the 27-point cell-centre
// `ep` the north- and bottom-face sums are not divided by the diagonal, and the
// south-face sum counts the (+,-,-) coupling twice instead of (-,-,-)).
#include "bmgx3_common.cuh"

namespace {

__device__ __forceinline__ double oi_fix3(double base, double d, double c, double ep)
{
    // base + (d - c) * max(d - (1+ep) c, 0) / (|d - (1+ep) c| + eMACH),  eMACH = 1d-13 (f90:78)
    const double t = d - (1.0 + ep) * c;
    return base + (d - c) * fmax(t, 0.0) / (fabs(t) + 1.0e-13);
}

struct CiW {
    double *ci;
    int IIC, JJC, KKC;
    size_t pl;
    __device__ __forceinline__ double &w(int l, int ic, int jc, int kc) const
    {
        return ci[(size_t)(l - 1) * pl + ((size_t)(kc - 1) * JJC + (jc - 1)) * IIC + (ic - 1)];
    }
};

#define C(dx, dy, dz) cpl<SEVEN, dx, dy, dz>(A, pi, pj, pk)
// sum over the 3x3 face dx = X (resp. dy = Y, dz = Z)
#define FACE_X(X) (C(X, -1, -1) + C(X, 0, -1) + C(X, 1, -1) + C(X, -1, 0) + C(X, 0, 0) + C(X, 1, 0) + C(X, -1, 1) + C(X, 0, 1) + C(X, 1, 1))
#define FACE_Y(Y) (C(-1, Y, -1) + C(0, Y, -1) + C(1, Y, -1) + C(-1, Y, 0) + C(0, Y, 0) + C(1, Y, 0) + C(-1, Y, 1) + C(0, Y, 1) + C(1, Y, 1))
#define FACE_Z(Z) (C(-1, -1, Z) + C(0, -1, Z) + C(1, -1, Z) + C(-1, 0, Z) + C(0, 0, Z) + C(1, 0, Z) + C(-1, 1, Z) + C(0, 1, Z) + C(1, 1, Z))
// the 8 couplings inside the plane dx = 0 (resp. dy, dz), centre excluded
#define RING_X (C(0, -1, -1) + C(0, 0, -1) + C(0, 1, -1) + C(0, -1, 0) + C(0, 1, 0) + C(0, -1, 1) + C(0, 0, 1) + C(0, 1, 1))
#define RING_Y (C(-1, 0, -1) + C(0, 0, -1) + C(1, 0, -1) + C(-1, 0, 0) + C(1, 0, 0) + C(-1, 0, 1) + C(0, 0, 1) + C(1, 0, 1))
#define RING_Z (C(-1, -1, 0) + C(0, -1, 0) + C(1, -1, 0) + C(-1, 0, 0) + C(1, 0, 0) + C(-1, 1, 0) + C(0, 1, 0) + C(1, 1, 0))

template <bool SEVEN>
__global__ void __launch_bounds__(128)
k3_oi_edges(Op3 A, CiW W, int iicf1, int jjcf1, int kkcf1)
{
    const int ic = blockIdx.x * blockDim.x + threadIdx.x + 2, jc = blockIdx.y * blockDim.y + threadIdx.y + 2,
              kc = blockIdx.z + 2;
    const int i = 2 * (ic - 1), j = 2 * (jc - 1), k = 2 * (kc - 1);
    const int iic1 = W.IIC - 1, jjc1 = W.JJC - 1, kkc1 = W.KKC - 1;
    if (ic >= 3 && ic <= iicf1 && jc <= jjc1 && kc <= kkc1) {   // p = (i-1, j, k)
        const int pi = i - 1, pj = j, pk = k;
        const double a = FACE_X(-1), b = FACE_X(1), d = A.diag(pi, pj, pk);
        const double ep = fmin(fabs(a / d), fabs(b / d));
        const double c = oi_fix3(a + b, d, a + b + RING_X, ep);
        W.w(LXYL, ic, jc, kc) = a / c;
        W.w(LXYR, ic, jc, kc) = b / c;
    }
    if (jc >= 3 && jc <= jjcf1 && ic <= iic1 && kc <= kkc1) {   // p = (i, j-1, k)
        const int pi = i, pj = j - 1, pk = k;
        const double a = FACE_Y(1), b = FACE_Y(-1), d = A.diag(pi, pj, pk);
        const double ep = fmin(fabs(a / d), fabs(b / d));
        const double c = oi_fix3(a + b, d, a + b + RING_Y, ep);
        W.w(LXYA, ic, jc, kc) = a / c;
        W.w(LXYB, ic, jc, kc) = b / c;
    }
    if (kc >= 3 && kc <= kkcf1 && ic <= iic1 && jc <= jjc1) {   // p = (i, j, k-1)
        const int pi = i, pj = j, pk = k - 1;
        const double a = FACE_Z(1), b = FACE_Z(-1), d = A.diag(pi, pj, pk);
        const double ep = fmin(fabs(a / d), fabs(b / d));
        const double c = oi_fix3(a + b, d, a + b + RING_Z, ep);
        W.w(LXZA, ic, jc, kc) = a / c;
        W.w(LXZB, ic, jc, kc) = b / c;
    }
}

// 2D operator-induced weights of a face-centre point from the collapsed couplings
// (dw..dsw), the collapsed diagonal `sum`, and the four edge weights on each side
struct Face {
    double dw, dnw, dn, dne, de, dse, ds, dsw, sum, d;
    __device__ __forceinline__ double inv() const
    {
        const double ep = fmin(fmin(fabs((dsw + dw + dnw) / d), fabs((dnw + dn + dne) / d)),
                               fmin(fabs((dne + de + dse) / d), fabs((dse + ds + dsw) / d)));
        const double dp = dw + dnw + dn + dne + de + dse + ds + dsw;
        return 1.0 / oi_fix3(dp, sum, dp, ep);
    }
};

template <bool SEVEN>
__global__ void __launch_bounds__(128)
k3_oi_faces(Op3 A, CiW W, int iicf1, int jjcf1, int kkcf1)
{
    const int ic = blockIdx.x * blockDim.x + threadIdx.x + 2, jc = blockIdx.y * blockDim.y + threadIdx.y + 2,
              kc = blockIdx.z + 2;
    const int i = 2 * (ic - 1), j = 2 * (jc - 1), k = 2 * (kc - 1);
    const int iic1 = W.IIC - 1, jjc1 = W.JJC - 1, kkc1 = W.KKC - 1;
#define COL_Z(dx, dy) (C(dx, dy, 0) + C(dx, dy, -1) + C(dx, dy, 1))
#define COL_Y(dx, dz) (C(dx, 1, dz) + C(dx, 0, dz) + C(dx, -1, dz))
#define COL_X(dy, dz) (C(-1, dy, dz) + C(0, dy, dz) + C(1, dy, dz))
    if (ic >= 3 && ic <= iicf1 && jc >= 3 && jc <= jjcf1 && kc <= kkc1) {   // p = (i-1, j-1, k): w/e = -/+x, s/n = -/+y
        const int pi = i - 1, pj = j - 1, pk = k;
        Face f;
        f.dnw = COL_Z(-1, 1); f.dn = COL_Z(0, 1); f.dne = COL_Z(1, 1); f.dw = COL_Z(-1, 0); f.de = COL_Z(1, 0);
        f.dsw = COL_Z(-1, -1); f.ds = COL_Z(0, -1); f.dse = COL_Z(1, -1);
        f.d = A.diag(pi, pj, pk);
        f.sum = f.d - C(0, 0, -1) - C(0, 0, 1);
        const double dp = f.inv();
        W.w(LXYNW, ic, jc, kc) = dp * (f.dnw + W.w(LXYA, ic - 1, jc, kc) * f.dw + W.w(LXYL, ic, jc, kc) * f.dn);
        W.w(LXYNE, ic, jc, kc) = dp * (f.dne + W.w(LXYR, ic, jc, kc) * f.dn + W.w(LXYA, ic, jc, kc) * f.de);
        W.w(LXYSE, ic, jc, kc) = dp * (f.dse + W.w(LXYB, ic, jc, kc) * f.de + W.w(LXYR, ic, jc - 1, kc) * f.ds);
        W.w(LXYSW, ic, jc, kc) = dp * (f.dsw + W.w(LXYL, ic, jc - 1, kc) * f.ds + W.w(LXYB, ic - 1, jc, kc) * f.dw);
    }
    if (ic >= 3 && ic <= iicf1 && kc >= 3 && kc <= kkcf1 && jc <= jjc1) {   // p = (i-1, j, k-1): w/e = -/+x, s/n = -/+z
        const int pi = i - 1, pj = j, pk = k - 1;
        Face f;
        f.dnw = COL_Y(-1, 1); f.dn = COL_Y(0, 1); f.dne = COL_Y(1, 1); f.dw = COL_Y(-1, 0); f.de = COL_Y(1, 0);
        f.dsw = COL_Y(-1, -1); f.ds = COL_Y(0, -1); f.dse = COL_Y(1, -1);
        f.d = A.diag(pi, pj, pk);
        f.sum = f.d - C(0, 1, 0) - C(0, -1, 0);
        const double dp = f.inv();
        W.w(LXZNW, ic, jc, kc) = dp * (f.dnw + W.w(LXZA, ic - 1, jc, kc) * f.dw + W.w(LXYL, ic, jc, kc) * f.dn);
        W.w(LXZNE, ic, jc, kc) = dp * (f.dne + W.w(LXYR, ic, jc, kc) * f.dn + W.w(LXZA, ic, jc, kc) * f.de);
        W.w(LXZSE, ic, jc, kc) = dp * (f.dse + W.w(LXZB, ic, jc, kc) * f.de + W.w(LXYR, ic, jc, kc - 1) * f.ds);
        W.w(LXZSW, ic, jc, kc) = dp * (f.dsw + W.w(LXYL, ic, jc, kc - 1) * f.ds + W.w(LXZB, ic - 1, jc, kc) * f.dw);
    }
    if (jc >= 3 && jc <= jjcf1 && kc >= 3 && kc <= kkcf1 && ic <= iic1) {   // p = (i, j-1, k-1): w/e = +/-y (!), s/n = -/+z
        const int pi = i, pj = j - 1, pk = k - 1;
        Face f;
        f.dnw = COL_X(1, 1); f.dn = COL_X(0, 1); f.dne = COL_X(-1, 1); f.dw = COL_X(1, 0); f.de = COL_X(-1, 0);
        f.dsw = COL_X(1, -1); f.ds = COL_X(0, -1); f.dse = COL_X(-1, -1);
        f.d = A.diag(pi, pj, pk);
        f.sum = f.d - C(-1, 0, 0) - C(1, 0, 0);
        const double dp = f.inv();
        W.w(LYZNW, ic, jc, kc) = dp * (f.dnw + W.w(LXZA, ic, jc, kc) * f.dw + W.w(LXYA, ic, jc, kc) * f.dn);
        W.w(LYZNE, ic, jc, kc) = dp * (f.dne + W.w(LXYB, ic, jc, kc) * f.dn + W.w(LXZA, ic, jc - 1, kc) * f.de);
        W.w(LYZSE, ic, jc, kc) = dp * (f.dse + W.w(LXZB, ic, jc - 1, kc) * f.de + W.w(LXYB, ic, jc, kc - 1) * f.ds);
        W.w(LYZSW, ic, jc, kc) = dp * (f.dsw + W.w(LXYA, ic, jc, kc - 1) * f.ds + W.w(LXZB, ic, jc, kc) * f.dw);
    }
}

// ---- cell centres ------------------------------------------------------------
// weight with which the neighbour n = p + (OX,OY,OZ) of the cell-centre point p
// (upper coarse point (ic,jc,kc)) takes the corner coarse point p + (SX,SY,SZ);
// OX in {0,SX} etc.  n lies on a coarse line/plane along every axis where O != 0.
template <int SX, int SY, int SZ, int OX, int OY, int OZ>
__device__ __forceinline__ double corner_weight(const CiW &W, int ic, int jc, int kc)
{
    constexpr int TX = OX == 0, TY = OY == 0, TZ = OZ == 0;                     // parity type of n
    constexpr int AX = TX && SX < 0, AY = TY && SY < 0, AZ = TZ && SZ < 0;        // corner = U(n) - (AX,AY,AZ)
    constexpr int name = Wname<TX, TY, TZ, AX, AY, AZ>::value;
    static_assert(name >= 0, "no such weight");
    if (name == 0) return 1.0;
    const int ui = TX ? ic : (SX > 0 ? ic : ic - 1), uj = TY ? jc : (SY > 0 ? jc : jc - 1),
              uk = TZ ? kc : (SZ > 0 ? kc : kc - 1);
    return W.w(name, ui, uj, uk);
}

template <bool SEVEN, int SX, int SY, int SZ>
__device__ __forceinline__ double corner_sum(const Op3 &A, const CiW &W, int pi, int pj, int pk, int ic, int jc, int kc)
{
    // the 7 neighbours of p inside the 2x2x2 cube towards the corner (f90:447-533)
    return C(SX, SY, SZ)
           + corner_weight<SX, SY, SZ, SX, 0, 0>(W, ic, jc, kc) * C(SX, 0, 0)
           + corner_weight<SX, SY, SZ, 0, SY, 0>(W, ic, jc, kc) * C(0, SY, 0)
           + corner_weight<SX, SY, SZ, 0, 0, SZ>(W, ic, jc, kc) * C(0, 0, SZ)
           + corner_weight<SX, SY, SZ, SX, SY, 0>(W, ic, jc, kc) * C(SX, SY, 0)
           + corner_weight<SX, SY, SZ, SX, 0, SZ>(W, ic, jc, kc) * C(SX, 0, SZ)
           + corner_weight<SX, SY, SZ, 0, SY, SZ>(W, ic, jc, kc) * C(0, SY, SZ);
}

template <bool SEVEN>
__global__ void __launch_bounds__(128)
k3_oi_centres(Op3 A, CiW W, int iicf1, int jjcf1, int kkcf1)
{
    const int ic = blockIdx.x * blockDim.x + threadIdx.x + 3, jc = blockIdx.y * blockDim.y + threadIdx.y + 3,
              kc = blockIdx.z + 3;
    if (ic > iicf1 || jc > jjcf1 || kc > kkcf1) return;
    const int pi = 2 * (ic - 1) - 1, pj = 2 * (jc - 1) - 1, pk = 2 * (kc - 1) - 1;
    const double d = A.diag(pi, pj, pk);
    double tot, ep;
    if (SEVEN) {
        tot = C(-1, 0, 0) + C(0, 1, 0) + C(1, 0, 0) + C(0, -1, 0) + C(0, 0, -1) + C(0, 0, 1);
        ep = fmin(fmin(fmin(fabs(C(-1, 0, 0) / d), fabs(C(0, 1, 0) / d)), fmin(fabs(C(1, 0, 0) / d), fabs(C(0, -1, 0) / d))),
                  fmin(fabs(C(0, 0, -1) / d), fabs(C(0, 0, 1) / d)));
    } else {
        const double fw = FACE_X(-1), fe = FACE_X(1), fn = FACE_Y(1), fb = FACE_Z(-1), ft = FACE_Z(1);
        // reference quirk: (+,-,-) twice, (-,-,-) missing
        const double fs = C(0, -1, 0) + C(-1, -1, 0) + C(-1, -1, 1) + C(0, -1, 1) + C(1, -1, 1) + C(1, -1, 0)
                          + C(1, -1, -1) + C(0, -1, -1) + C(1, -1, -1);
        tot = fw + fe + RING_X;
        // reference quirk: the north and bottom sums are not divided by the diagonal
        ep = fmin(fmin(fabs(fw) / d, fabs(fe) / d), fmin(fabs(fn), fabs(fs) / d));
        ep = fmin(ep, fmin(fabs(fb), fabs(ft) / d));
    }
    const double dp = 1.0 / oi_fix3(tot, d, tot, ep);
    W.w(LTNW, ic, jc, kc) = dp * corner_sum<SEVEN, -1, 1, 1>(A, W, pi, pj, pk, ic, jc, kc);
    W.w(LTNE, ic, jc, kc) = dp * corner_sum<SEVEN, 1, 1, 1>(A, W, pi, pj, pk, ic, jc, kc);
    W.w(LBNW, ic, jc, kc) = dp * corner_sum<SEVEN, -1, 1, -1>(A, W, pi, pj, pk, ic, jc, kc);
    W.w(LBNE, ic, jc, kc) = dp * corner_sum<SEVEN, 1, 1, -1>(A, W, pi, pj, pk, ic, jc, kc);
    W.w(LBSW, ic, jc, kc) = dp * corner_sum<SEVEN, -1, -1, -1>(A, W, pi, pj, pk, ic, jc, kc);
    W.w(LTSW, ic, jc, kc) = dp * corner_sum<SEVEN, -1, -1, 1>(A, W, pi, pj, pk, ic, jc, kc);
    W.w(LTSE, ic, jc, kc) = dp * corner_sum<SEVEN, 1, -1, 1>(A, W, pi, pj, pk, ic, jc, kc);
    W.w(LBSE, ic, jc, kc) = dp * corner_sum<SEVEN, 1, -1, -1>(A, W, pi, pj, pk, ic, jc, kc);
}

}  // namespace

// BMG3_SymStd_SETUP_interp_OI(so, soc, ci, iif, jjf, kkf, iic, jjc, kkc, ifd, nstncl, irelax, jpn, yo)
extern "C" int bmgx3_SETUP_interp_OI(const double *so, const double *soc, double *ci, bmgx_len iif, bmgx_len jjf,
                                     bmgx_len kkf, bmgx_len iic, bmgx_len jjc, bmgx_len kkc, int ifd, int nstncl,
                                     int irelax, int jpn, double *yo, bmgx_stream_t s)
{
    (void)soc; (void)irelax; (void)yo; (void)nstncl;
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "SETUP_interp_OI3: periodic BCs not supported");
    if (iic < 3 || jjc < 3 || kkc < 3) return BMGX_OK;
    const int iicf1 = (iif - 2) / 2 + 2, jjcf1 = (jjf - 2) / 2 + 2, kkcf1 = (kkf - 2) / 2 + 2;
    Op3 A(so, iif, jjf, kkf);
    CiW W{ci, (int)iic, (int)jjc, (int)kkc, (size_t)iic * jjc * kkc};
    cudaStream_t st = (cudaStream_t)s;
    // index ranges reach max(cf1, c1) = c-1 or c along each axis
    const unsigned mi = max(iicf1, (int)iic - 1), mj = max(jjcf1, (int)jjc - 1), mk = max(kkcf1, (int)kkc - 1);
    dim3 b(32, 4), g(cdiv(mi - 1, 32), cdiv(mj - 1, 4), mk - 1);
    if (ifd != 1) k3_oi_edges<false><<<g, b, 0, st>>>(A, W, iicf1, jjcf1, kkcf1);
    else k3_oi_edges<true><<<g, b, 0, st>>>(A, W, iicf1, jjcf1, kkcf1);
    BMGX_LAUNCH_CHECK("SETUP_interp_OI3/edges");
    if (ifd != 1) k3_oi_faces<false><<<g, b, 0, st>>>(A, W, iicf1, jjcf1, kkcf1);
    else k3_oi_faces<true><<<g, b, 0, st>>>(A, W, iicf1, jjcf1, kkcf1);
    BMGX_LAUNCH_CHECK("SETUP_interp_OI3/faces");
    if (iicf1 >= 3 && jjcf1 >= 3 && kkcf1 >= 3) {
        dim3 g2(cdiv(iicf1 - 2, 32), cdiv(jjcf1 - 2, 4), kkcf1 - 2);
        if (ifd != 1) k3_oi_centres<false><<<g2, b, 0, st>>>(A, W, iicf1, jjcf1, kkcf1);
        else k3_oi_centres<true><<<g2, b, 0, st>>>(A, W, iicf1, jjcf1, kkcf1);
        BMGX_LAUNCH_CHECK("SETUP_interp_OI3/centres");
    }
    return BMGX_OK;
}
