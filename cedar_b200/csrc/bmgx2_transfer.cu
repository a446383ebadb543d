// bmgx2_transfer.cu -- 2D restriction (R = P^T) and interpolate-and-add.
// Reference: src/2d/ftn/BMG2_SymStd_restrict.f90:73-94,
//            src/2d/ftn/BMG2_SymStd_interp_add.f90:88-141 (non-periodic).
#include "bmgx_common.cuh"

__global__ void __launch_bounds__(256)
k_restrict(const double *__restrict__ q, double *__restrict__ qc, const double *__restrict__ ci,
           unsigned Nx, unsigned Ny, unsigned Nxc, unsigned Nyc, size_t cpitch, size_t pc)
{
    // coarse arrays: row pitch cpitch, plane stride pc (= Nxc, Nxc*Nyc for plain arrays; a window of a
    // larger array otherwise, bmgx2_restrict_v)
    pdl_wait();
    unsigned ic = blockIdx.x * blockDim.x + threadIdx.x + 2;
    unsigned jc = blockIdx.y * blockDim.y + threadIdx.y + 2;
    if (ic > Nxc - 1 || jc > Nyc - 1) return;
    unsigned i = 2 * (ic - 1), j = 2 * (jc - 1);
    q += (size_t)blockIdx.z * Nx * Ny; qc += blockIdx.z * pc;   // batch entry
    size_t cc = IX2(cpitch, ic, jc);
    size_t f = IX2(Nx, i, j);
    double r = ci[(LNE - 1) * pc + cc] * q[f - Nx - 1]
               + ci[(LA - 1) * pc + cc] * q[f - Nx]
               + ci[(LNW - 1) * pc + cc + 1] * q[f - Nx + 1]
               + ci[(LR - 1) * pc + cc] * q[f - 1]
               + q[f]
               + ci[(LL - 1) * pc + cc + 1] * q[f + 1]
               + ci[(LSE - 1) * pc + cc + cpitch] * q[f + Nx - 1]
               + ci[(LB - 1) * pc + cc + cpitch] * q[f + Nx]
               + ci[(LSW - 1) * pc + cc + cpitch + 1] * q[f + Nx + 1];
    qc[cc] = r;
}

extern "C" int bmgx2_restrict(const double *q, double *qc, const double *ci, bmgx_len nx, bmgx_len ny,
                              bmgx_len nxc, bmgx_len nyc, int jpn, bmgx_stream_t s)
{
    if (!bmgx_nonperiodic(jpn)) {
        // BMG2_SymStd_restrict.f90:96-111: refresh the fine vector's ghost layer first (the "const" argument is
        // mutated, as in the reference): y rows, then x columns, each only for the N/2+1 coarsening
        const int ipn = jpn < 0 ? -jpn : jpn;
        if (ipn < 1 || ipn > 3) return bmgx_set_error(BMGX_ERR_ARG, "restrict: unknown boundary code");
        if (bmgx_batch() != 1) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "restrict: periodic BCs not supported in a batch");
        if (nx < 3 || ny < 3) return BMGX_OK;
        double *qm = const_cast<double *>(q);
        if ((ipn == 1 || ipn == 3) && ny / 2 + 1 == nyc && bmgx_wrap_y(qm, nx, ny, (cudaStream_t)s)) return BMGX_ERR_CUDA;
        if ((ipn == 2 || ipn == 3) && nx / 2 + 1 == nxc && bmgx_wrap_x(qm, nx, 1, 1, ny, (cudaStream_t)s)) return BMGX_ERR_CUDA;
    }
    if (nxc < 3 || nyc < 3) return BMGX_OK;
    dim3 b(64, 4), g(cdiv(nxc - 2, 64), cdiv(nyc - 2, 4), bmgx_batch());
    launch_pdl(k_restrict, g, b, 0, (cudaStream_t)s, q, qc, ci, nx, ny, nxc, nyc, (size_t)nxc, (size_t)nxc * nyc);
    BMGX_LAUNCH_CHECK("restrict");
    return BMGX_OK;
}

extern "C" int bmgx2_restrict_v(const double *q, double *qc, const double *ci, bmgx_len nx, bmgx_len ny, bmgx_len nxc,
                                bmgx_len nyc, size_t cpitch, size_t cplane, bmgx_stream_t s)
{
    if (nxc < 3 || nyc < 3) return BMGX_OK;
    if (cpitch < nxc) return bmgx_set_error(BMGX_ERR_ARG, "restrict_v: coarse pitch smaller than the window");
    dim3 b(64, 4), g(cdiv(nxc - 2, 64), cdiv(nyc - 2, 4), 1);
    launch_pdl(k_restrict, g, b, 0, (cudaStream_t)s, q, qc, ci, nx, ny, nxc, nyc, cpitch, cplane);
    BMGX_LAUNCH_CHECK("restrict_v");
    return BMGX_OK;
}

// The reference walks coarse cells and touches the 4 fine points of each; per fine point that is:
//   C-point (i even, j even):  Q += QC
//   x-edge  (i odd,  j even):  Q = (Q + (LR*QC(ic,jc) + LL*QC(ic-1,jc))) + RES
//   y-edge  (i even, j odd ):  Q = (Q + (LA*QC(ic,jc) + LB*QC(ic,jc-1))) + RES
//   centre  (i odd,  j odd ):  Q = (Q + (LSW*QC(ic-1,jc-1)+LNW*QC(ic-1,jc)+LNE*QC(ic,jc)+LSE*QC(ic,jc-1))) + RES
// with (ic,jc) the coarse point up-right of the fine point, and RES already
// divided by the diagonal (the division is stored back: persistent side effect).
// One thread per coarse cell (ic,jc) = the fine points i in {2ic-3, 2ic-2}, j in {2jc-3, 2jc-2}
// (centre, y-edge / x-edge, C-point): no divergence, 24 independent loads in flight per thread.
__global__ void __launch_bounds__(256)
k_interp_add(double *__restrict__ q, const double *__restrict__ qc, double *__restrict__ res,
             const double *__restrict__ so, const double *__restrict__ ci,
             unsigned IIC, unsigned JJC, unsigned IIF, unsigned JJF, size_t cpitch, size_t pc)
{
    // coarse arrays: row pitch cpitch, plane stride pc (bmgx2_interp_add_v: a window of a larger array)
    pdl_wait();
    const unsigned ic = blockIdx.x * blockDim.x + threadIdx.x + 2;
    const unsigned jc = blockIdx.y * blockDim.y + threadIdx.y + 2;
    if (ic > IIF / 2 + 1 || jc > JJF / 2 + 1) return;
    const unsigned i0 = 2 * ic - 3, i1 = i0 + 1, j0 = 2 * jc - 3, j1 = j0 + 1;   // i0, j0 odd
    const bool xi0 = i0 >= 2, xi1 = i1 <= IIF - 1, yj0 = j0 >= 2, yj1 = j1 <= JJF - 1;
    const size_t cc = IX2(cpitch, ic, jc);
    q += (size_t)blockIdx.z * IIF * JJF; res += (size_t)blockIdx.z * IIF * JJF; qc += blockIdx.z * pc;   // batch entry
    const size_t f00 = IX2(IIF, i0, j0), f10 = f00 + 1, f01 = f00 + IIF, f11 = f01 + 1;
    const bool p00 = xi0 && yj0, p10 = xi1 && yj0, p01 = xi0 && yj1, p11 = xi1 && yj1;
    // loads first (all independent)
    const double r00 = p00 ? res[f00] : 0.0, r10 = p10 ? res[f10] : 0.0, r01 = p01 ? res[f01] : 0.0,
                 r11 = p11 ? res[f11] : 0.0;
    const double d00 = p00 ? __ldg(so + f00) : 1.0, d10 = p10 ? __ldg(so + f10) : 1.0, d01 = p01 ? __ldg(so + f01) : 1.0,
                 d11 = p11 ? __ldg(so + f11) : 1.0;
    const double q00 = p00 ? q[f00] : 0.0, q10 = p10 ? q[f10] : 0.0, q01 = p01 ? q[f01] : 0.0, q11 = p11 ? q[f11] : 0.0;
    const double c11 = __ldg(qc + cc), c01 = __ldg(qc + cc - 1), c10 = __ldg(qc + cc - cpitch), c00 = __ldg(qc + cc - cpitch - 1);
    const double lr = __ldg(ci + (LR - 1) * pc + cc), ll = __ldg(ci + (LL - 1) * pc + cc);
    const double la = __ldg(ci + (LA - 1) * pc + cc), lb = __ldg(ci + (LB - 1) * pc + cc);
    const double lsw = __ldg(ci + (LSW - 1) * pc + cc), lnw = __ldg(ci + (LNW - 1) * pc + cc);
    const double lne = __ldg(ci + (LNE - 1) * pc + cc), lse = __ldg(ci + (LSE - 1) * pc + cc);
    if (p00) {   // cell centre (i odd, j odd)
        const double r = r00 / d00;
        res[f00] = r;
        const double a = lsw * c00 + lnw * c01 + lne * c11 + lse * c10;
        q[f00] = q00 + a + r;
    }
    if (p10) {   // y-edge (i even, j odd)
        const double r = r10 / d10;
        res[f10] = r;
        const double a = la * c11 + lb * c10;
        q[f10] = q10 + a + r;
    }
    if (p01) {   // x-edge (i odd, j even)
        const double r = r01 / d01;
        res[f01] = r;
        const double a = lr * c11 + ll * c01;
        q[f01] = q01 + a + r;
    }
    if (p11) {   // C-point
        res[f11] = r11 / d11;
        q[f11] = q11 + c11;
    }
}

extern "C" int bmgx2_interp_add(double *q, const double *qc, double *res, const double *so, const double *ci,
                                bmgx_len iic, bmgx_len jjc, bmgx_len iif, bmgx_len jjf, int nstncl, int jpn,
                                bmgx_stream_t s)
{
    (void)nstncl;
    const bool periodic = !bmgx_nonperiodic(jpn);
    const int ipn = jpn < 0 ? -jpn : jpn;
    if (periodic && (ipn < 1 || ipn > 3)) return bmgx_set_error(BMGX_ERR_ARG, "interp_add: unknown boundary code");
    if (periodic && bmgx_batch() != 1)
        return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "interp_add: periodic BCs not supported in a batch");
    if (iif < 3 || jjf < 3) return BMGX_OK;
    dim3 b(64, 4), g(cdiv(iif / 2, 64), cdiv(jjf / 2, 4), bmgx_batch());
    launch_pdl(k_interp_add, g, b, 0, (cudaStream_t)s, q, qc, res, so, ci, iic, jjc, iif, jjf, (size_t)iic, (size_t)iic * jjc);
    BMGX_LAUNCH_CHECK("interp_add");
    if (periodic) {
        // BMG2_SymStd_interp_add.f90:143-165: ghost layer of the corrected vector, y rows first, then x columns
        if ((ipn == 1 || ipn == 3) && bmgx_wrap_y(q, iif, jjf, (cudaStream_t)s)) return BMGX_ERR_CUDA;
        if ((ipn == 2 || ipn == 3) && bmgx_wrap_x(q, iif, 1, 1, jjf, (cudaStream_t)s)) return BMGX_ERR_CUDA;
    }
    return BMGX_OK;
}

extern "C" int bmgx2_interp_add_v(double *q, const double *qc, double *res, const double *so, const double *ci,
                                  bmgx_len iic, bmgx_len jjc, bmgx_len iif, bmgx_len jjf, size_t cpitch, size_t cplane,
                                  bmgx_stream_t s)
{
    if (iif < 3 || jjf < 3) return BMGX_OK;
    if (cpitch < iic) return bmgx_set_error(BMGX_ERR_ARG, "interp_add_v: coarse pitch smaller than the window");
    dim3 b(64, 4), g(cdiv(iif / 2, 64), cdiv(jjf / 2, 4), 1);
    launch_pdl(k_interp_add, g, b, 0, (cudaStream_t)s, q, qc, res, so, ci, iic, jjc, iif, jjf, cpitch, cplane);
    BMGX_LAUNCH_CHECK("interp_add_v");
    return BMGX_OK;
}
