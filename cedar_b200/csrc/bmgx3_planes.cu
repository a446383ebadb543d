// bmgx3_planes.cu -- 3D <-> 2D staging of plane relaxation.
// Reference: include/cedar/3d/relax_planes.h:77-161 (copy_coeff), src/3d/relax_planes.cc:26-238
// (copy_rhs, copy32, copy23).  dir: 0 = xy planes (index k), 1 = xz (index j), 2 = yz (index i).
// The in-plane coordinates (u,v) of the 2D arrays are (i,j), (i,k), (j,k) respectively.
// copy_rhs keeps the reference's term order (-fmad=false => bitwise).
#include "bmgx3_common.cuh"

namespace {

struct Map {
    int dir, ipl, II, JJ, KK;
    // 3D linear index of the in-plane point (u,v), 1-based
    __device__ __forceinline__ void ijk(int u, int v, int &i, int &j, int &k) const
    {
        if (dir == 0) { i = u; j = v; k = ipl; }
        else if (dir == 1) { i = u; j = ipl; k = v; }
        else { i = ipl; j = u; k = v; }
    }
    __host__ __device__ int n1() const { return dir == 2 ? JJ : II; }
    __host__ __device__ int n2() const { return dir == 0 ? JJ : KK; }
};

// x2(u,v) = x3(plane) or back, all points including ghosts
__global__ void __launch_bounds__(256) k3_plane_copy(Map m, double *x3, double *x2, int to2d)
{
    const int u = blockIdx.x * blockDim.x + threadIdx.x + 1, v = blockIdx.y + 1;
    if (u > m.n1()) return;
    m.ipl += 2 * blockIdx.z; x2 += (size_t)blockIdx.z * m.n1() * m.n2();   // batch entry: every other plane
    int i, j, k;
    m.ijk(u, v, i, j, k);
    const size_t c3 = IX3(m.II, m.JJ, i, j, k), c2 = IX2(m.n1(), u, v);
    if (to2d) x2[c2] = x3[c3];
    else x3[c3] = x2[c2];
}

// so2(u,v,0..n2d-1) = the in-plane entries of so3 at the given plane, all points including ghosts
__global__ void __launch_bounds__(256) k3_plane_coeff(Map m, const double *__restrict__ so3, double *__restrict__ so2, int full)
{
    const int u = blockIdx.x * blockDim.x + threadIdx.x + 1, v = blockIdx.y + 1;
    if (u > m.n1()) return;
    int i, j, k;
    m.ijk(u, v, i, j, k);
    const size_t pl3 = (size_t)m.II * m.JJ * m.KK, pl2 = (size_t)m.n1() * m.n2();
    const size_t c3 = IX3(m.II, m.JJ, i, j, k), c2 = IX2(m.n1(), u, v);
    // 2D slots c,w,s,sw,nw <- 3D entries (relax_planes.h:84-161)
    const int e_xy[5] = {KP, KPW, KPS, KPSW, KPNW}, e_xz[5] = {KP, KPW, KB, KBW, KBE}, e_yz[5] = {KP, KPS, KB, KBS, KBN};
    const int *e = m.dir == 0 ? e_xy : (m.dir == 1 ? e_xz : e_yz);
    const int n = full ? 5 : 3;
    for (int s = 0; s < n; s++) so2[(size_t)s * pl2 + c2] = so3[(size_t)(e[s] - 1) * pl3 + c3];
}

template <bool SEVEN, int DX, int DY, int DZ>
__device__ __forceinline__ double term(const Op3 &A, const double *__restrict__ x, int i, int j, int k)
{
    return cpl<SEVEN, DX, DY, DZ>(A, i, j, k) * x[IX3(A.II, A.JJ, i + DX, j + DY, k + DZ)];
}

// b2 = b + off-plane couplings * x, interior of the plane (src/3d/relax_planes.cc:26-178)
template <bool SEVEN, int DIR>
__global__ void __launch_bounds__(256)
k3_plane_rhs(Map m, Op3 A, const double *__restrict__ x, const double *__restrict__ b, double *__restrict__ b2)
{
    const int u = blockIdx.x * blockDim.x + threadIdx.x + 2, v = blockIdx.y + 2;
    if (u > m.n1() - 1) return;
    m.ipl += 2 * blockIdx.z; b2 += (size_t)blockIdx.z * m.n1() * m.n2();   // batch entry: every other plane
    int i, j, k;
    m.ijk(u, v, i, j, k);
    double r = b[IX3(m.II, m.JJ, i, j, k)];
#define T(dx, dy, dz) r = r + term<SEVEN, dx, dy, dz>(A, x, i, j, k);
    if (SEVEN) {
        if (DIR == 0) { T(0, 0, -1) T(0, 0, 1) }
        else if (DIR == 1) { T(0, -1, 0) T(0, 1, 0) }
        else { T(-1, 0, 0) T(1, 0, 0) }
    } else if (DIR == 0) {
        T(0, 0, -1) T(-1, 0, -1) T(-1, 1, -1) T(0, 1, -1) T(1, 1, -1) T(1, 0, -1) T(1, -1, -1) T(0, -1, -1) T(-1, -1, -1)
        T(-1, 0, 1) T(-1, 1, 1) T(0, 1, 1) T(1, 1, 1) T(1, 0, 1) T(0, 0, 1) T(1, -1, 1) T(0, -1, 1) T(-1, -1, 1)
    } else if (DIR == 1) {
        T(-1, 1, 0) T(0, 1, 0) T(1, 1, 0) T(-1, 1, -1) T(0, 1, -1) T(1, 1, -1) T(-1, 1, 1) T(0, 1, 1) T(1, 1, 1)
        T(-1, -1, 0) T(0, -1, 0) T(1, -1, 0) T(-1, -1, -1) T(0, -1, -1) T(1, -1, -1) T(-1, -1, 1) T(0, -1, 1) T(1, -1, 1)
    } else {
        T(-1, 1, 0) T(-1, 0, 0) T(-1, -1, 0) T(-1, 1, -1) T(-1, 0, -1) T(-1, -1, -1) T(-1, 1, 1) T(-1, 0, 1) T(-1, -1, 1)
        T(1, 1, 0) T(1, 0, 0) T(1, -1, 0) T(1, 1, -1) T(1, 0, -1) T(1, -1, -1) T(1, 1, 1) T(1, 0, 1) T(1, -1, 1)
    }
#undef T
    b2[IX2(m.n1(), u, v)] = r;
}

// nb > 1 (batch context): planes ipl, ipl+2, ... must all exist
int check_plane(const Map &m, const char *what, unsigned nb = 1)
{
    const int n = m.dir == 0 ? m.KK : (m.dir == 1 ? m.JJ : m.II);
    if (m.dir < 0 || m.dir > 2 || m.ipl < 1 || m.ipl + 2 * ((int)nb - 1) > n) return bmgx_set_error(BMGX_ERR_ARG, what);
    return BMGX_OK;
}

}  // namespace

// copy32 / copy23 (src/3d/relax_planes.cc:181-238): whole plane including ghosts
extern "C" int bmgx3_plane_copy(double *x3, double *x2, bmgx_len ii, bmgx_len jj, bmgx_len kk, int dir, bmgx_len ipl,
                                int to2d, bmgx_stream_t s)
{
    Map m{dir, (int)ipl, (int)ii, (int)jj, (int)kk};
    if (check_plane(m, "plane_copy: bad plane", bmgx_batch())) return BMGX_ERR_ARG;
    dim3 g(cdiv(m.n1(), 256), m.n2(), bmgx_batch());
    k3_plane_copy<<<g, 256, 0, (cudaStream_t)s>>>(m, x3, x2, to2d);
    BMGX_LAUNCH_CHECK("plane_copy");
    return BMGX_OK;
}

// copy_coeff (include/cedar/3d/relax_planes.h:77-161) for ONE plane
extern "C" int bmgx3_plane_coeff(const double *so3, double *so2, bmgx_len ii, bmgx_len jj, bmgx_len kk, int nstncl,
                                 int dir, bmgx_len ipl, bmgx_stream_t s)
{
    Map m{dir, (int)ipl, (int)ii, (int)jj, (int)kk};
    if (check_plane(m, "plane_coeff: bad plane")) return BMGX_ERR_ARG;
    dim3 g(cdiv(m.n1(), 256), m.n2());
    k3_plane_coeff<<<g, 256, 0, (cudaStream_t)s>>>(m, so3, so2, nstncl == 14);
    BMGX_LAUNCH_CHECK("plane_coeff");
    return BMGX_OK;
}

// copy_rhs (src/3d/relax_planes.cc:26-178)
extern "C" int bmgx3_plane_rhs(const double *so, const double *x, const double *b, double *b2, bmgx_len ii, bmgx_len jj,
                               bmgx_len kk, int nstncl, int dir, bmgx_len ipl, bmgx_stream_t s)
{
    Map m{dir, (int)ipl, (int)ii, (int)jj, (int)kk};
    if (check_plane(m, "plane_rhs: bad plane", bmgx_batch())) return BMGX_ERR_ARG;
    if (m.n1() < 3 || m.n2() < 3) return BMGX_OK;
    Op3 A(so, ii, jj, kk);
    dim3 g(cdiv(m.n1() - 2, 256), m.n2() - 2, bmgx_batch());
    cudaStream_t st = (cudaStream_t)s;
    const bool seven = nstncl != 14;
#define LAUNCH(SV, D) k3_plane_rhs<SV, D><<<g, 256, 0, st>>>(m, A, x, b, b2)
    if (seven) { if (dir == 0) LAUNCH(true, 0); else if (dir == 1) LAUNCH(true, 1); else LAUNCH(true, 2); }
    else { if (dir == 0) LAUNCH(false, 0); else if (dir == 1) LAUNCH(false, 1); else LAUNCH(false, 2); }
#undef LAUNCH
    BMGX_LAUNCH_CHECK("plane_rhs");
    return BMGX_OK;
}
