// bmgx3_transfer.cu -- 3D restriction (R = P^T) and interpolate-and-add.
// Reference: src/3d/ftn/BMG3_SymStd_restrict.f90:115-160,
//            src/3d/ftn/BMG3_SymStd_interp_add.f90:88-230 (non-periodic).
// Reference term order, -fmad=false => bitwise.
#include "bmgx3_common.cuh"

namespace {

__global__ void __launch_bounds__(128)
k3_restrict(const double *__restrict__ q, double *__restrict__ qc, const double *__restrict__ ci, int Nx, int Ny,
            int Nxc, int Nyc, int Nzc)
{
    int ic = blockIdx.x * blockDim.x + threadIdx.x + 2, jc = blockIdx.y * blockDim.y + threadIdx.y + 2, kc = blockIdx.z + 2;
    if (ic > Nxc - 1 || jc > Nyc - 1) return;
    const size_t pc = (size_t)Nxc * Nyc * Nzc, PJc = (size_t)Nxc * Nyc, PJ = (size_t)Nx * Ny;
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
__global__ void __launch_bounds__(128)
k3_interp_add(double *__restrict__ q, const double *__restrict__ qc, const double *__restrict__ so,
              double *__restrict__ res, const double *__restrict__ ci, int IIC, int JJC, int KKC, int IIF, int JJF)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x + 2, j = blockIdx.y * blockDim.y + threadIdx.y + 2, k = blockIdx.z + 2;
    if (i > IIF - 1 || j > JJF - 1) return;
    const size_t f = ((size_t)(k - 1) * JJF + (j - 1)) * IIF + (i - 1);
    const size_t pc = (size_t)IIC * JJC * KKC, X = 1, Y = IIC, Z = (size_t)IIC * JJC;
    const int ic = (i + 1) / 2 + 1, jc = (j + 1) / 2 + 1, kc = (k + 1) / 2 + 1;
    const size_t cc = ((size_t)(kc - 1) * JJC + (jc - 1)) * IIC + (ic - 1);
    const double r = res[f] / so[f];   // so(i,j,k,kp): plane 0
    res[f] = r;
    const bool px = i & 1, py = j & 1, pz = k & 1;
#define CI(l) ci[(size_t)((l)-1) * pc + cc]
    double v = q[f];
    if (!pz) {
        if (!px && !py) v = v + qc[cc];
        else if (px && !py) { double a = CI(LXYR) * qc[cc] + CI(LXYL) * qc[cc - X]; v = v + a + r; }
        else if (!px && py) { double a = CI(LXYA) * qc[cc] + CI(LXYB) * qc[cc - Y]; v = v + a + r; }
        else {
            double a = CI(LXYSW) * qc[cc - X - Y] + CI(LXYNW) * qc[cc - X] + CI(LXYNE) * qc[cc] + CI(LXYSE) * qc[cc - Y];
            v = v + a + r;
        }
    } else {
        if (!px && !py) v = v + CI(LXZA) * qc[cc] + CI(LXZB) * qc[cc - Z] + r;
        else if (!px && py)
            v = v + CI(LYZNW) * qc[cc] + CI(LYZNE) * qc[cc - Y] + CI(LYZSW) * qc[cc - Z] + CI(LYZSE) * qc[cc - Y - Z] + r;
        else if (px && !py)
            v = v + CI(LXZNW) * qc[cc - X] + CI(LXZNE) * qc[cc] + CI(LXZSW) * qc[cc - X - Z] + CI(LXZSE) * qc[cc - Z] + r;
        else
            v = v + CI(LTNW) * qc[cc - X] + CI(LTNE) * qc[cc] + CI(LTSW) * qc[cc - X - Y] + CI(LTSE) * qc[cc - Y]
                + CI(LBNW) * qc[cc - X - Z] + CI(LBNE) * qc[cc - Z] + CI(LBSW) * qc[cc - X - Y - Z]
                + CI(LBSE) * qc[cc - Y - Z] + r;
    }
#undef CI
    q[f] = v;
}

}  // namespace

extern "C" int bmgx3_restrict(const double *q, double *qc, const double *ci, bmgx_len nx, bmgx_len ny, bmgx_len nz,
                              bmgx_len nxc, bmgx_len nyc, bmgx_len nzc, int jpn, bmgx_stream_t s)
{
    (void)nz;
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "restrict3: periodic BCs not supported");
    if (nxc < 3 || nyc < 3 || nzc < 3) return BMGX_OK;
    dim3 b(32, 4), g(cdiv(nxc - 2, 32), cdiv(nyc - 2, 4), nzc - 2);
    k3_restrict<<<g, b, 0, (cudaStream_t)s>>>(q, qc, ci, nx, ny, nxc, nyc, nzc);
    BMGX_LAUNCH_CHECK("restrict3");
    return BMGX_OK;
}

extern "C" int bmgx3_interp_add(double *q, const double *qc, const double *so, double *res, const double *ci,
                                bmgx_len iic, bmgx_len jjc, bmgx_len kkc, bmgx_len iif, bmgx_len jjf, bmgx_len kkf,
                                int nstncl, int jpn, bmgx_stream_t s)
{
    (void)nstncl;
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "interp_add3: periodic BCs not supported");
    if (iif < 3 || jjf < 3 || kkf < 3) return BMGX_OK;
    dim3 b(64, 2), g(cdiv(iif - 2, 64), cdiv(jjf - 2, 2), kkf - 2);
    k3_interp_add<<<g, b, 0, (cudaStream_t)s>>>(q, qc, so, res, ci, iic, jjc, kkc, iif, jjf);
    BMGX_LAUNCH_CHECK("interp_add3");
    return BMGX_OK;
}
