// bmgx2_tail.cu -- the coarse tail of a 2D V-cycle in ONE launch.
//
// On the smallest levels a cycle is a chain of ~30 dependent kernels with a few thousand points each: every node of
// the captured graph costs ~3 us of launch/dependency latency and does ~0.1 us of work (profiles/r01_launches_*.csv:
// levels <= 513^2 take 0.19 ms of the 1.165 ms 4097^2 cycle).  This kernel runs the whole sub-cycle of those levels --
// relax, residual, restrict, zero, direct solve on the coarsest grid, interpolate-and-add, relax
// (include/cedar/cycle/vcycle.h, reference cycle/vcycle.h:57-115) -- in one CTA of 1024 threads with __syncthreads()
// between the stages; the level arrays (a few hundred KB) stay in L1/L2.
// Measured (profiles/r02_tailsweep.log, 4097^2 V(1,1)): levels <= 33^2 in the tail: 1.168 -> 1.142 ms per cycle (44
// instead of 68 graph nodes); including 65^2 gives the gain back (1.161 ms) and 129^2 loses (1.243 ms): one SM draws
// ~120 GB/s from L2, so a 65^2 colour pass (118 KB) costs 1 us against 1.7 us for a whole graph node.
//
// Same arithmetic as the per-level kernels, term by term (compiled -fmad=false): relaxation is the in-place 4-colour
// form of bmgx2_relax.cu (reference src/2d/ftn/BMG2_SymStd_relax_GS.f90:78-114: a colour's points are independent, so
// the order inside a colour is free), residual / restrict / interp_add are the expressions of bmgx2_residual.cu and
// bmgx2_transfer.cu, the direct solve is k_solve_cg of bmgx2_cg.cu.  Coarse levels are always 9-point.
#include "bmgx_common.cuh"

namespace {

struct Tail
{
    bmgx_tail_level lv[BMGX_TAIL_MAX];
    int n, npre, npost;
    const double *abd;
    double *bbd;
    unsigned nabd1;
};

#define TAIL_NT 1024

__device__ __forceinline__ void t_relax(const bmgx_tail_level &L, bool fwd)
{
    const unsigned II = L.ii, JJ = L.jj, I1 = II - 1, J1 = JJ - 1;
    const size_t pl = (size_t)II * JJ;
    const double *__restrict__ so = L.so;
    const double *kw = so + (KW - 1) * pl, *ks = so + (KS - 1) * pl, *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
    const double *sor = L.sor + (MSOR - 1) * pl;
    const double *qf = L.qf;
    double *q = L.q;
    for (int a = 0; a < 2; a++) {
        const unsigned jbeg = fwd ? 2 + a : 3 - a;
        for (int b = 0; b < 2; b++) {
            const unsigned ibeg = fwd ? 2 + b : 3 - b;
            if (jbeg <= J1 && ibeg <= I1) {
                const unsigned nrow = (J1 - jbeg) / 2 + 1, ncol = (I1 - ibeg) / 2 + 1;
                for (unsigned t = threadIdx.x; t < nrow * ncol; t += TAIL_NT) {
                    const unsigned j = jbeg + 2 * (t / ncol), i = ibeg + 2 * (t % ncol);
                    const size_t c = IX2(II, i, j);
                    double acc = qf[c] + __ldg(kw + c) * q[c - 1] + __ldg(kw + c + 1) * q[c + 1] + __ldg(ks + c) * q[c - II]
                                 + __ldg(ks + c + II) * q[c + II];
                    acc = acc + __ldg(ksw + c) * q[c - II - 1] + __ldg(knw + c + 1) * q[c - II + 1]
                          + __ldg(knw + c + II) * q[c + II - 1] + __ldg(ksw + c + II + 1) * q[c + II + 1];
                    q[c] = acc * __ldg(sor + c);
                }
            }
            __syncthreads();
        }
    }
}

__device__ __forceinline__ void t_residual(const bmgx_tail_level &L)
{
    const unsigned II = L.ii, JJ = L.jj, nx = II - 2, ny = JJ - 2;
    const size_t pl = (size_t)II * JJ;
    const double *__restrict__ so = L.so;
    const double *ko = so, *kw = so + (KW - 1) * pl, *ks = so + (KS - 1) * pl, *ksw = so + (KSW - 1) * pl, *knw = so + (KNW - 1) * pl;
    const double *qf = L.qf, *q = L.q;
    double *res = L.res;
    for (unsigned t = threadIdx.x; t < nx * ny; t += TAIL_NT) {
        const unsigned j = 2 + t / nx, i = 2 + t % nx;
        const size_t c = IX2(II, i, j);
        double acc = qf[c] + __ldg(kw + c) * q[c - 1] + __ldg(kw + c + 1) * q[c + 1] + __ldg(ks + c) * q[c - II]
                     + __ldg(ks + c + II) * q[c + II];
        acc = acc + __ldg(ksw + c) * q[c - II - 1] + __ldg(knw + c + 1) * q[c - II + 1] + __ldg(knw + c + II) * q[c + II - 1]
              + __ldg(ksw + c + II + 1) * q[c + II + 1];
        res[c] = acc - __ldg(ko + c) * q[c];
    }
}

// restriction of L.res to C.qf, and C.q = 0 (whole array, ghosts included, like grid_func::set(0))
__device__ __forceinline__ void t_restrict_zero(const bmgx_tail_level &L, const bmgx_tail_level &C)
{
    const unsigned Nx = L.ii, Nxc = C.ii, Nyc = C.jj, nxc = Nxc - 2, nyc = Nyc - 2;
    const size_t pc = (size_t)Nxc * Nyc;
    const double *q = L.res;
    const double *__restrict__ ci = C.ci;
    double *qc = C.qf;
    for (unsigned t = threadIdx.x; t < nxc * nyc; t += TAIL_NT) {
        const unsigned jc = 2 + t / nxc, ic = 2 + t % nxc;
        const unsigned i = 2 * (ic - 1), j = 2 * (jc - 1);
        const size_t cc = IX2(Nxc, ic, jc), f = IX2(Nx, i, j);
        double r = __ldg(ci + (LNE - 1) * pc + cc) * q[f - Nx - 1]
                   + __ldg(ci + (LA - 1) * pc + cc) * q[f - Nx]
                   + __ldg(ci + (LNW - 1) * pc + cc + 1) * q[f - Nx + 1]
                   + __ldg(ci + (LR - 1) * pc + cc) * q[f - 1]
                   + q[f]
                   + __ldg(ci + (LL - 1) * pc + cc + 1) * q[f + 1]
                   + __ldg(ci + (LSE - 1) * pc + cc + Nxc) * q[f + Nx - 1]
                   + __ldg(ci + (LB - 1) * pc + cc + Nxc) * q[f + Nx]
                   + __ldg(ci + (LSW - 1) * pc + cc + Nxc + 1) * q[f + Nx + 1];
        qc[cc] = r;
    }
    double *xc = C.q;
    for (size_t t = threadIdx.x; t < pc; t += TAIL_NT) xc[t] = 0.0;
}

// bmgx2_transfer.cu k_interp_add: one coarse cell = its four fine points
__device__ __forceinline__ void t_interp_add(const bmgx_tail_level &L, const bmgx_tail_level &C)
{
    const unsigned IIF = L.ii, JJF = L.jj, IIC = C.ii, JJC = C.jj;
    const size_t pc = (size_t)IIC * JJC;
    double *q = L.q, *res = L.res;
    const double *qc = C.q;
    const double *__restrict__ so = L.so, *__restrict__ ci = C.ci;
    const unsigned ncx = IIF / 2, ncy = JJF / 2;     // cells ic = 2..IIF/2+1, jc = 2..JJF/2+1
    for (unsigned t = threadIdx.x; t < ncx * ncy; t += TAIL_NT) {
        const unsigned jc = 2 + t / ncx, ic = 2 + t % ncx;
        const unsigned i0 = 2 * ic - 3, i1 = i0 + 1, j0 = 2 * jc - 3, j1 = j0 + 1;
        const bool xi0 = i0 >= 2, xi1 = i1 <= IIF - 1, yj0 = j0 >= 2, yj1 = j1 <= JJF - 1;
        const size_t cc = IX2(IIC, ic, jc);
        const size_t f00 = IX2(IIF, i0, j0), f10 = f00 + 1, f01 = f00 + IIF, f11 = f01 + 1;
        const bool p00 = xi0 && yj0, p10 = xi1 && yj0, p01 = xi0 && yj1, p11 = xi1 && yj1;
        const double r00 = p00 ? res[f00] : 0.0, r10 = p10 ? res[f10] : 0.0, r01 = p01 ? res[f01] : 0.0, r11 = p11 ? res[f11] : 0.0;
        const double d00 = p00 ? __ldg(so + f00) : 1.0, d10 = p10 ? __ldg(so + f10) : 1.0, d01 = p01 ? __ldg(so + f01) : 1.0,
                     d11 = p11 ? __ldg(so + f11) : 1.0;
        const double q00 = p00 ? q[f00] : 0.0, q10 = p10 ? q[f10] : 0.0, q01 = p01 ? q[f01] : 0.0, q11 = p11 ? q[f11] : 0.0;
        const double c11 = qc[cc], c01 = qc[cc - 1], c10 = qc[cc - IIC], c00 = qc[cc - IIC - 1];
        const double lr = __ldg(ci + (LR - 1) * pc + cc), ll = __ldg(ci + (LL - 1) * pc + cc);
        const double la = __ldg(ci + (LA - 1) * pc + cc), lb = __ldg(ci + (LB - 1) * pc + cc);
        const double lsw = __ldg(ci + (LSW - 1) * pc + cc), lnw = __ldg(ci + (LNW - 1) * pc + cc);
        const double lne = __ldg(ci + (LNE - 1) * pc + cc), lse = __ldg(ci + (LSE - 1) * pc + cc);
        if (p00) {
            const double r = r00 / d00;
            res[f00] = r;
            const double a = lsw * c00 + lnw * c01 + lne * c11 + lse * c10;
            q[f00] = q00 + a + r;
        }
        if (p10) {
            const double r = r10 / d10;
            res[f10] = r;
            const double a = la * c11 + lb * c10;
            q[f10] = q10 + a + r;
        }
        if (p01) {
            const double r = r01 / d01;
            res[f01] = r;
            const double a = lr * c11 + ll * c01;
            q[f01] = q01 + a + r;
        }
        if (p11) {
            res[f11] = r11 / d11;
            q[f11] = q11 + c11;
        }
    }
}

#define ABD(i, j) abd[(size_t)((j)-1) * nabd1 + ((i)-1)]
// bmgx2_cg.cu k_solve_cg: DTBSV(U,T,N), DTBSV(U,N,N) in LAPACK's sequential order, one thread
__device__ void t_solve_cg(const bmgx_tail_level &C, const double *__restrict__ abd, double *bbd, unsigned nabd1)
{
    if (threadIdx.x != 0) return;
    const unsigned II = C.ii, JJ = C.jj;
    const int I1 = II - 1, J1 = JJ - 1;
    double *q = C.q;
    const double *qf = C.qf;
    const int kd = I1, kplus1 = kd + 1;
    int n = 0;
    for (int j = 2; j <= J1; j++)
        for (int i = 2; i <= I1; i++) bbd[n++] = qf[IX2(II, i, j)];
    for (int j = 1; j <= n; j++) {
        double temp = bbd[j - 1];
        const int l = kplus1 - j;
        const int i0 = j - kd > 1 ? j - kd : 1;
        for (int i = i0; i <= j - 1; i++) temp = temp - ABD(l + i, j) * bbd[i - 1];
        temp = temp / ABD(kplus1, j);
        bbd[j - 1] = temp;
    }
    for (int j = n; j >= 1; j--) {
        if (bbd[j - 1] != 0.0) {
            const int l = kplus1 - j;
            bbd[j - 1] = bbd[j - 1] / ABD(kplus1, j);
            const double temp = bbd[j - 1];
            const int i0 = j - kd > 1 ? j - kd : 1;
            for (int i = j - 1; i >= i0; i--) bbd[i - 1] = bbd[i - 1] - temp * ABD(l + i, j);
        }
    }
    n = 0;
    for (int j = 2; j <= J1; j++)
        for (int i = 2; i <= I1; i++) q[IX2(II, i, j)] = bbd[n++];
}
#undef ABD

__global__ void __launch_bounds__(TAIL_NT, 1) k_vcycle_tail(const __grid_constant__ Tail T)
{
    pdl_wait();
    const int n = T.n;
    for (int l = 0; l + 1 < n; l++) {
        for (int s = 0; s < T.npre; s++) t_relax(T.lv[l], true);
        t_residual(T.lv[l]);
        __syncthreads();
        t_restrict_zero(T.lv[l], T.lv[l + 1]);
        __syncthreads();
    }
    t_solve_cg(T.lv[n - 1], T.abd, T.bbd, T.nabd1);
    __syncthreads();
    for (int l = n - 2; l >= 0; l--) {
        t_interp_add(T.lv[l], T.lv[l + 1]);
        __syncthreads();
        for (int s = 0; s < T.npost; s++) t_relax(T.lv[l], false);
    }
}

}  // namespace

extern "C" int bmgx2_vcycle_tail(const bmgx_tail_level *lv, int nlevels, const double *abd, double *bbd, bmgx_len nabd1,
                                 int nrelax_pre, int nrelax_post, bmgx_stream_t s)
{
    if (nlevels < 1 || nlevels > BMGX_TAIL_MAX) return bmgx_set_error(BMGX_ERR_ARG, "vcycle_tail: 1..BMGX_TAIL_MAX levels");
    if (bmgx_batch() != 1) return bmgx_set_error(BMGX_ERR_UNSUPPORTED, "vcycle_tail: not batch aware");
    Tail T{};
    for (int l = 0; l < nlevels; l++) {
        T.lv[l] = lv[l];
        if (lv[l].ii < 3 || lv[l].jj < 3 || (size_t)lv[l].ii * lv[l].jj > (size_t)BMGX_TAIL_MAX_POINTS)
            return bmgx_set_error(BMGX_ERR_ARG, "vcycle_tail: level too large for the one-CTA tail (or empty)");
        if (l > 0 && (lv[l].ii != (lv[l - 1].ii - 3) / 2 + 3 || lv[l].jj != (lv[l - 1].jj - 3) / 2 + 3))
            return bmgx_set_error(BMGX_ERR_ARG, "vcycle_tail: levels are not successive coarsenings");
    }
    T.n = nlevels; T.npre = nrelax_pre; T.npost = nrelax_post; T.abd = abd; T.bbd = bbd; T.nabd1 = nabd1;
    launch_pdl(k_vcycle_tail, dim3(1), dim3(TAIL_NT), 0, (cudaStream_t)s, T);
    BMGX_LAUNCH_CHECK("vcycle_tail");
    return BMGX_OK;
}
