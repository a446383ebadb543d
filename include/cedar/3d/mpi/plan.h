// cedar/3d/mpi/plan.h -- slab decomposition of the multi-GPU 3D solver (host-only).
//
// The reference cuts a 3D grid into px x py x pz blocks (include/cedar/3d/mpi/solver.h:104 setup_space,
// include/cedar/decomp.h:59-86) and exchanges one ghost layer after every colour of every sweep
// (src/3d/ftn/mpi/BMG3_SymStd_relax_GS.f90:145,190).  Here the grid is cut into SLABS of xy planes, one per GPU:
//   * the colour orderings of both point smoothers have depth 2 across planes (7-point red-black: black needs the new
//     red of the planes next to it; 27-point 8 colours: the four k-odd colours need the new k-even planes next to
//     them, never further), so a slab with W = 5 halo planes runs a whole sweep, the residual and the restriction on
//     one exchange -- the deep-halo schedule of include/cedar/2d/mpi/plan.h with colour depth 2 on every level;
//   * a halo is a few CONTIGUOUS planes, a window of planes of a coarse array is contiguous too, so the serial
//     kernels run unchanged on the slabs (only the stride between the 26 weight arrays is passed,
//     bmgx3_restrict_v / bmgx3_interp_add_v);
//   * two neighbours per rank instead of 26.
// Levels whose global plane count is at most `agglomerate` are solved replicated on every rank by the serial
// cdr3::solver.  The hierarchy is the serial one, so the result is the serial solver's (the 3D setup kernels are not
// bitwise, see tests/test_gpu_kernels3d.py, the cycle kernels are).
#ifndef CEDAR_3D_MPI_PLAN_H
#define CEDAR_3D_MPI_PLAN_H

#include <cedar/2d/mpi/plan.h>

namespace cedar { namespace cdr3 { namespace mpi {

using cdr2::mpi::coarse_n;

inline int compute_num_levels(int nx, int ny, int nz, int min_coarse = 3)   // include/cedar/3d/solver.h:54-72
{
	for (int ng = 1;; ng++)
		if (std::min({(nx - 1) / (1 << ng) + 1, (ny - 1) / (1 << ng) + 1, (nz - 1) / (1 << ng) + 1}) < min_coarse) return ng;
}

struct slab_geom
{
	int nx, ny;          // global interior extents of the level (every slab holds whole planes)
	int z0, nz;          // first own plane (0-based global interior index) and own planes
	int wb, wt;          // planes beyond the own ones below / above: W towards a neighbour, 1 at the boundary
	int ii, jj, kk;      // array extents: nx + 2, ny + 2, nz + wb + wt (own planes start at array plane wb)
};

struct slab_decomp
{
	int gnx, gny, gnz, np, ndist, nlevels, W;
	std::vector<std::array<int, 3>> gn;
	std::vector<std::vector<int>> lo;    // lo[lvl][slab], with the end as sentinel

	slab_decomp(int gnx, int gny, int gnz, int np, int ndist, int W)
	    : gnx(gnx), gny(gny), gnz(gnz), np(np), ndist(ndist), nlevels(compute_num_levels(gnx, gny, gnz)), W(W)
	{
		if (W < 3 || W % 2 == 0) throw std::invalid_argument("halo width must be odd and at least 3");
		if (ndist < 1 || ndist > nlevels - 1) throw std::invalid_argument("need 1 <= distributed levels < levels");
		gn.push_back({gnx, gny, gnz});
		for (int l = 0; l < ndist; l++) gn.push_back({coarse_n(gn.back()[0]), coarse_n(gn.back()[1]), coarse_n(gn.back()[2])});
		const int al = 1 << ndist;   // slab boundaries on coarse planes of every distributed level
		lo.resize(ndist + 1);
		for (int l = 0; l <= ndist; l++) {
			for (int c = 0; c < np; c++) lo[l].push_back(((int)((long)c * gnz / np) / al * al) >> l);
			lo[l].push_back(gn[l][2]);
		}
		for (int c = 0; c < np; c++)
			if (lo[ndist][c + 1] - lo[ndist][c] < std::max(W, 2) && np > 1)
				throw std::invalid_argument("slabs thinner than the halo on the coarsest distributed level");
	}
	slab_geom geom(int lvl, int b) const
	{
		slab_geom g;
		g.nx = gn[lvl][0]; g.ny = gn[lvl][1];
		g.z0 = lo[lvl][b]; g.nz = lo[lvl][b + 1] - g.z0;
		g.wb = b > 0 ? W : 1; g.wt = b + 1 < np ? W : 1;
		g.ii = g.nx + 2; g.jj = g.ny + 2; g.kk = g.nz + g.wb + g.wt;
		return g;
	}
	// window of planes of the level lvl+1 array that the serial transfer kernels see as the coarse grid of the whole
	// level-lvl array
	void coarse_window(int lvl, int b, int &oz, int &kkc) const
	{
		const slab_geom f = geom(lvl, b);
		oz = (f.wb - 1) / 2; kkc = (f.kk - 3) / 2 + 3;
	}
};

}}}
#endif
