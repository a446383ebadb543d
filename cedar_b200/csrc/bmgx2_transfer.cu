// bmgx2_transfer.cu -- 2D restriction (R = P^T) and interpolate-and-add.
// Reference: src/2d/ftn/BMG2_SymStd_restrict.f90:73-94,
//            src/2d/ftn/BMG2_SymStd_interp_add.f90:88-141 (non-periodic).
#include "bmgx_common.cuh"

__global__ void __launch_bounds__(256)
k_restrict(const double *__restrict__ q, double *__restrict__ qc, const double *__restrict__ ci,
           unsigned Nx, unsigned Nxc, unsigned Nyc)
{
    unsigned ic = blockIdx.x * blockDim.x + threadIdx.x + 2;
    unsigned jc = blockIdx.y * blockDim.y + threadIdx.y + 2;
    if (ic > Nxc - 1 || jc > Nyc - 1) return;
    unsigned i = 2 * (ic - 1), j = 2 * (jc - 1);
    size_t pc = (size_t)Nxc * Nyc;
    size_t cc = IX2(Nxc, ic, jc);
    size_t f = IX2(Nx, i, j);
    double r = ci[(LNE - 1) * pc + cc] * q[f - Nx - 1]
               + ci[(LA - 1) * pc + cc] * q[f - Nx]
               + ci[(LNW - 1) * pc + cc + 1] * q[f - Nx + 1]
               + ci[(LR - 1) * pc + cc] * q[f - 1]
               + q[f]
               + ci[(LL - 1) * pc + cc + 1] * q[f + 1]
               + ci[(LSE - 1) * pc + cc + Nxc] * q[f + Nx - 1]
               + ci[(LB - 1) * pc + cc + Nxc] * q[f + Nx]
               + ci[(LSW - 1) * pc + cc + Nxc + 1] * q[f + Nx + 1];
    qc[cc] = r;
}

extern "C" int bmgx2_restrict(const double *q, double *qc, const double *ci, bmgx_len nx, bmgx_len ny,
                              bmgx_len nxc, bmgx_len nyc, int jpn, bmgx_stream_t s)
{
    (void)ny;
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "restrict: periodic BCs not supported");
    if (nxc < 3 || nyc < 3) return BMGX_OK;
    dim3 b(64, 4), g(cdiv(nxc - 2, 64), cdiv(nyc - 2, 4));
    k_restrict<<<g, b, 0, (cudaStream_t)s>>>(q, qc, ci, nx, nxc, nyc);
    BMGX_LAUNCH_CHECK("restrict");
    return BMGX_OK;
}

// One thread per fine interior point.  The reference walks coarse cells and
// touches the 4 fine points of each; per fine point that is:
//   C-point (i even, j even):  Q += QC
//   x-edge  (i odd,  j even):  Q = (Q + (LR*QC(ic,jc) + LL*QC(ic-1,jc))) + RES
//   y-edge  (i even, j odd ):  Q = (Q + (LA*QC(ic,jc) + LB*QC(ic,jc-1))) + RES
//   centre  (i odd,  j odd ):  Q = (Q + (LSW*QC(ic-1,jc-1)+LNW*QC(ic-1,jc)+LNE*QC(ic,jc)+LSE*QC(ic,jc-1))) + RES
// with (ic,jc) the coarse point up-right of the fine point, and RES already
// divided by the diagonal (the division is stored back: persistent side effect).
__global__ void __launch_bounds__(256)
k_interp_add(double *__restrict__ q, const double *__restrict__ qc, double *__restrict__ res,
             const double *__restrict__ so, const double *__restrict__ ci,
             unsigned IIC, unsigned JJC, unsigned IIF, unsigned JJF)
{
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x + 2;
    unsigned j = blockIdx.y * blockDim.y + threadIdx.y + 2;
    if (i > IIF - 1 || j > JJF - 1) return;
    size_t f = IX2(IIF, i, j);
    size_t pc = (size_t)IIC * JJC;
    double r = res[f] / so[f];  // SO(i,j,KO): KO is plane 0
    res[f] = r;
    unsigned ic = (i + 1) / 2 + 1, jc = (j + 1) / 2 + 1;  // i even: i/2+1 ; i odd: (i+1)/2+1
    size_t cc = IX2(IIC, ic, jc);
    bool io = i & 1, jo = j & 1;
    double qv = q[f];
    if (!io && !jo) {
        qv = qv + qc[cc];
    } else if (io && !jo) {
        double a = ci[(LR - 1) * pc + cc] * qc[cc] + ci[(LL - 1) * pc + cc] * qc[cc - 1];
        qv = qv + a + r;
    } else if (!io && jo) {
        double a = ci[(LA - 1) * pc + cc] * qc[cc] + ci[(LB - 1) * pc + cc] * qc[cc - IIC];
        qv = qv + a + r;
    } else {
        double a = ci[(LSW - 1) * pc + cc] * qc[cc - IIC - 1] + ci[(LNW - 1) * pc + cc] * qc[cc - 1]
                   + ci[(LNE - 1) * pc + cc] * qc[cc] + ci[(LSE - 1) * pc + cc] * qc[cc - IIC];
        qv = qv + a + r;
    }
    q[f] = qv;
}

extern "C" int bmgx2_interp_add(double *q, const double *qc, double *res, const double *so, const double *ci,
                                bmgx_len iic, bmgx_len jjc, bmgx_len iif, bmgx_len jjf, int nstncl, int jpn,
                                bmgx_stream_t s)
{
    (void)nstncl;
    if (!bmgx_nonperiodic(jpn)) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "interp_add: periodic BCs not supported");
    if (iif < 3 || jjf < 3) return BMGX_OK;
    dim3 b(128, 2), g(cdiv(iif - 2, 128), cdiv(jjf - 2, 2));
    k_interp_add<<<g, b, 0, (cudaStream_t)s>>>(q, qc, res, so, ci, iic, jjc, iif, jjf);
    BMGX_LAUNCH_CHECK("interp_add");
    return BMGX_OK;
}
