// bmgx3_transfer.cu -- 3D restriction (R = P^T) and interpolate-and-add.
// Reference: src/3d/ftn/BMG3_SymStd_restrict.f90:115-160,
//            src/3d/ftn/BMG3_SymStd_interp_add.f90:88-230 (non-periodic).
// Reference term order, -fmad=false => bitwise.
#include "bmgx3_common.cuh"

namespace {

__global__ void __launch_bounds__(128)
k3_restrict(const double *__restrict__ q, double *__restrict__ qc, const double *__restrict__ ci, int Nx, int Ny,
            int Nxc, int Nyc, int Nzc, size_t pc)
{
    // pc: stride between the 26 weight arrays of ci (Nxc*Nyc*Nzc for a plain array; larger when ci is a window of
    // planes of a thicker array, bmgx3_restrict_v)
    int ic = blockIdx.x * blockDim.x + threadIdx.x + 2, jc = blockIdx.y * blockDim.y + threadIdx.y + 2, kc = blockIdx.z + 2;
    if (ic > Nxc - 1 || jc > Nyc - 1) return;
    const size_t PJc = (size_t)Nxc * Nyc, PJ = (size_t)Nx * Ny;
    const size_t cc = ((size_t)(kc - 1) * Nyc + (jc - 1)) * Nxc + (ic - 1);
    const size_t f = ((size_t)(2 * (kc - 1) - 1) * Ny + (2 * (jc - 1) - 1)) * Nx + (2 * (ic - 1) - 1);
#define CI(l, off) ci[(size_t)((l)-1) * pc + cc + (off)]
    double r = CI(LXYNE, 0) * q[f - Nx - 1] + CI(LXYA, 0) * q[f - Nx] + CI(LXYNW, 1) * q[f - Nx + 1]
               + CI(LXYR, 0) * q[f - 1] + q[f] + CI(LXYL, 1) * q[f + 1]
               + CI(LXYSE, Nxc) * q[f + Nx - 1] + CI(LXYB, Nxc) * q[f + Nx] + CI(LXYSW, Nxc + 1) * q[f + Nx + 1];
    r = r + CI(LTNE, 0) * q[f - PJ - Nx - 1] + CI(LYZNW, 0) * q[f - PJ - Nx] + CI(LTNW, 1) * q[f - PJ - Nx + 1]
        + CI(LXZNE, 0) * q[f - PJ - 1] + CI(LXZA, 0) * q[f - PJ] + CI(LXZNW, 1) * q[f - PJ + 1]
        + CI(LTSE, Nxc) * q[f - PJ + Nx - 1] + CI(LYZNE, Nxc) * q[f - PJ + Nx] + CI(LTSW, Nxc + 1) * q[f - PJ + Nx + 1];
    r = r + CI(LBNE, PJc) * q[f + PJ - Nx - 1] + CI(LYZSW, PJc) * q[f + PJ - Nx] + CI(LBNW, PJc + 1) * q[f + PJ - Nx + 1]
        + CI(LXZSE, PJc) * q[f + PJ - 1] + CI(LXZB, PJc) * q[f + PJ] + CI(LXZSW, PJc + 1) * q[f + PJ + 1]
        + CI(LBSE, PJc + Nxc) * q[f + PJ + Nx - 1] + CI(LYZSE, PJc + Nxc) * q[f + PJ + Nx]
        + CI(LBSW, PJc + Nxc + 1) * q[f + PJ + Nx + 1];
#undef CI
    qc[cc] = r;
}

// One thread per fine interior point (i,j,k); (ic,jc,kc) is its upper coarse point.
// RES is scaled by 1/diag in place first (interp_add.f90:88-94), then
//   coarse-plane points (k even) add `a + RES` with the weighted sum `a` formed
//   first (f90:100-150); points between coarse planes (k odd) add term by term
//   (f90:153-230).
// One thread per x-pair of fine points (i0 = 2ic-3 odd, i1 = 2ic-2 even) of a fine row (j,k): both share
// the coarse point (ic,jc,kc) their weights are stored at; the (j,k) parities are uniform per warp, so
// there is no divergence and twice the loads are in flight.
__global__ void __launch_bounds__(128)
k3_interp_add(double *__restrict__ q, const double *__restrict__ qc, const double *__restrict__ so,
              double *__restrict__ res, const double *__restrict__ ci, int IIC, int JJC, int KKC, int IIF, int JJF, size_t pc)
{
    const int ic = blockIdx.x * blockDim.x + threadIdx.x + 2, j = blockIdx.y * blockDim.y + threadIdx.y + 2,
              k = blockIdx.z + 2;
    if (ic > IIF / 2 + 1 || j > JJF - 1) return;
    const int i0 = 2 * ic - 3, i1 = i0 + 1;
    const bool v0 = i0 >= 2, v1 = i1 <= IIF - 1;
    const size_t f0 = ((size_t)(k - 1) * JJF + (j - 1)) * IIF + (i0 - 1), f1 = f0 + 1;
    const size_t X = 1, Y = IIC, Z = (size_t)IIC * JJC;   // pc: stride between the weight arrays (see k3_restrict)
    const int jc = (j + 1) / 2 + 1, kc = (k + 1) / 2 + 1;
    const size_t cc = ((size_t)(kc - 1) * JJC + (jc - 1)) * IIC + (ic - 1);
    const double n0 = v0 ? res[f0] : 0.0, n1 = v1 ? res[f1] : 0.0;
    const double d0 = v0 ? __ldg(so + f0) : 1.0, d1 = v1 ? __ldg(so + f1) : 1.0;   // so(i,j,k,kp): plane 0
    double w0 = v0 ? q[f0] : 0.0, w1 = v1 ? q[f1] : 0.0;
    const double r0 = n0 / d0, r1 = n1 / d1;
    const bool py = j & 1, pz = k & 1;
#define CI(l) __ldg(ci + (size_t)((l)-1) * pc + cc)
#define QC(o) __ldg(qc + cc - (o))
    if (!pz) {
        if (!py) {
            { double a = CI(LXYR) * QC(0) + CI(LXYL) * QC(X); w0 = w0 + a + r0; }
            w1 = w1 + QC(0);
        } else {
            { double a = CI(LXYSW) * QC(X + Y) + CI(LXYNW) * QC(X) + CI(LXYNE) * QC(0) + CI(LXYSE) * QC(Y); w0 = w0 + a + r0; }
            { double a = CI(LXYA) * QC(0) + CI(LXYB) * QC(Y); w1 = w1 + a + r1; }
        }
    } else {
        if (!py) {
            w0 = w0 + CI(LXZNW) * QC(X) + CI(LXZNE) * QC(0) + CI(LXZSW) * QC(X + Z) + CI(LXZSE) * QC(Z) + r0;
            w1 = w1 + CI(LXZA) * QC(0) + CI(LXZB) * QC(Z) + r1;
        } else {
            w0 = w0 + CI(LTNW) * QC(X) + CI(LTNE) * QC(0) + CI(LTSW) * QC(X + Y) + CI(LTSE) * QC(Y)
                 + CI(LBNW) * QC(X + Z) + CI(LBNE) * QC(Z) + CI(LBSW) * QC(X + Y + Z) + CI(LBSE) * QC(Y + Z) + r0;
            w1 = w1 + CI(LYZNW) * QC(0) + CI(LYZNE) * QC(Y) + CI(LYZSW) * QC(Z) + CI(LYZSE) * QC(Y + Z) + r1;
        }
    }
#undef CI
#undef QC
    if (v0) { res[f0] = r0; q[f0] = w0; }
    if (v1) { res[f1] = r1; q[f1] = w1; }
}

}  // namespace

extern "C" int bmgx3_restrict(const double *q, double *qc, const double *ci, bmgx_len nx, bmgx_len ny, bmgx_len nz,
                              bmgx_len nxc, bmgx_len nyc, bmgx_len nzc, int jpn, bmgx_stream_t s)
{
    (void)nz;
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "restrict3: periodic BCs not supported");
    if (nxc < 3 || nyc < 3 || nzc < 3) return BMGX_OK;
    dim3 b(32, 4), g(cdiv(nxc - 2, 32), cdiv(nyc - 2, 4), nzc - 2);
    k3_restrict<<<g, b, 0, (cudaStream_t)s>>>(q, qc, ci, nx, ny, nxc, nyc, nzc, (size_t)nxc * nyc * nzc);
    BMGX_LAUNCH_CHECK("restrict3");
    return BMGX_OK;
}

extern "C" int bmgx3_restrict_v(const double *q, double *qc, const double *ci, bmgx_len nx, bmgx_len ny, bmgx_len nz,
                                bmgx_len nxc, bmgx_len nyc, bmgx_len nzc, size_t cstride, bmgx_stream_t s)
{
    (void)nz;
    if (nxc < 3 || nyc < 3 || nzc < 3) return BMGX_OK;
    if (cstride < (size_t)nxc * nyc * nzc) return bmgx_set_error(BMGX_ERR_ARG, "restrict3_v: weight stride smaller than the window");
    dim3 b(32, 4), g(cdiv(nxc - 2, 32), cdiv(nyc - 2, 4), nzc - 2);
    k3_restrict<<<g, b, 0, (cudaStream_t)s>>>(q, qc, ci, nx, ny, nxc, nyc, nzc, cstride);
    BMGX_LAUNCH_CHECK("restrict3_v");
    return BMGX_OK;
}

extern "C" int bmgx3_interp_add(double *q, const double *qc, const double *so, double *res, const double *ci,
                                bmgx_len iic, bmgx_len jjc, bmgx_len kkc, bmgx_len iif, bmgx_len jjf, bmgx_len kkf,
                                int nstncl, int jpn, bmgx_stream_t s)
{
    (void)nstncl;
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "interp_add3: periodic BCs not supported");
    if (iif < 3 || jjf < 3 || kkf < 3) return BMGX_OK;
    dim3 b(32, 4), g(cdiv(iif / 2, 32), cdiv(jjf - 2, 4), kkf - 2);
    k3_interp_add<<<g, b, 0, (cudaStream_t)s>>>(q, qc, so, res, ci, iic, jjc, kkc, iif, jjf, (size_t)iic * jjc * kkc);
    BMGX_LAUNCH_CHECK("interp_add3");
    return BMGX_OK;
}

extern "C" int bmgx3_interp_add_v(double *q, const double *qc, const double *so, double *res, const double *ci,
                                  bmgx_len iic, bmgx_len jjc, bmgx_len kkc, bmgx_len iif, bmgx_len jjf, bmgx_len kkf,
                                  size_t cstride, bmgx_stream_t s)
{
    if (iif < 3 || jjf < 3 || kkf < 3) return BMGX_OK;
    if (cstride < (size_t)iic * jjc * kkc) return bmgx_set_error(BMGX_ERR_ARG, "interp_add3_v: weight stride smaller than the window");
    dim3 b(32, 4), g(cdiv(iif / 2, 32), cdiv(jjf - 2, 4), kkf - 2);
    k3_interp_add<<<g, b, 0, (cudaStream_t)s>>>(q, qc, so, res, ci, iic, jjc, kkc, iif, jjf, cstride);
    BMGX_LAUNCH_CHECK("interp_add3_v");
    return BMGX_OK;
}
