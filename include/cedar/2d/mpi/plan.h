// cedar/2d/mpi/plan.h -- block decomposition and exchange schedule of the multi-GPU 2D solver.
//
// Host-only (no device code): the decomposition of the reference's cdr2::mpi::solver
// (include/cedar/2d/mpi/solver.h:104 setup_space, include/cedar/decomp.h:59-86 grid_decomp,
// include/cedar/mpi/block_partition.h:14-17) and the ORDER of kernels and ghost exchanges of one
// setup / one V-cycle / one residual norm, as a flat list of operations.  The CUDA executor
// (cedar_b200/host/dist2.cc) runs the list with the C-ABI kernels and peer-memory exchanges; the CPU
// tests run the very same list with the oracle's kernels over gloo (tests/dist_plan_sim.py).
//
// Deep halos instead of an exchange per colour.  The reference exchanges one ghost line after every
// colour of a sweep, after the residual and after the interpolation
// (src/2d/ftn/mpi/BMG2_SymStd_relax_GS.f90:126,171, BMG2_SymStd_residual.f90, BMG2_SymStd_interp_add.f90).
// Here every array of a block carries W lines of its neighbours (W odd, so the global colouring is
// kept; the outermost line plays the role of the ghost ring) and every kernel simply runs over the
// whole extended array, recomputing the neighbours' values it needs.  Each kernel consumes a known
// number of lines of validity (a 4-colour sweep 4, a residual 1, a restriction halves it, ...); the
// planner tracks, per array, how many halo lines still hold exactly the values the neighbour holds and
// inserts an exchange only when a kernel would otherwise produce a wrong OWN value.  With W = 15 a
// V(1,1) cycle needs ONE exchange per level (the restricted right-hand side) instead of twelve; the
// result is bit for bit the serial solver's (every recomputed value is computed from the same operands
// in the same order as on its owner).
#ifndef CEDAR_2D_MPI_PLAN_H
#define CEDAR_2D_MPI_PLAN_H

#include <algorithm>
#include <array>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace cedar { namespace cdr2 { namespace mpi {

// include/cedar/decomp.h:59-86: prime factors, largest first, onto the currently longest dimension
inline std::array<int, 2> grid_decomp(std::array<long, 2> n, int nprocs)
{
	std::vector<int> fac;
	for (int p = 2, m = nprocs; m > 1; p++)
		while (m % p == 0) { fac.push_back(p); m /= p; }
	std::sort(fac.rbegin(), fac.rend());
	std::array<int, 2> dec{1, 1};
	for (int f : fac) {
		int ind = n[1] > n[0] ? 1 : 0;
		dec[ind] *= f;
		n[ind] /= f;
	}
	return dec;
}

inline int coarse_n(int n) { return (n - 1) / 2 + 1; }   // include/cedar/2d/solver.h:86-87

inline int compute_num_levels(int nx, int ny, int min_coarse = 3)   // include/cedar/2d/solver.h:57-73
{
	for (int ng = 1;; ng++)
		if (std::min((nx - 1) / (1 << ng) + 1, (ny - 1) / (1 << ng) + 1) < min_coarse) return ng;
}

// array slots of a level
enum arr { X0 = 0, X1 = 1, B = 2, RES = 3, A = 4, SOR = 5, P = 6, NARR = 7 };

// Lines that cross block cuts.  The reference reduces every cut line to an interface system (LineSolve_A ->
// gather -> LineSolve_B -> scatter -> LineSolve_C, src/2d/ftn/mpi/BMG2_SymStd_relax_lines_x.f90:156-215), three
// latency-bound stages per colour whose arithmetic differs from the serial solve.  With every GPU one NVSwitch hop
// from every other the lines are TRANSPOSED instead: the blocks of one process row (x-lines; column for y-lines)
// share their lines out evenly; each block builds the right-hand side of its own points and stores it straight into
// the line buffer of the line's solver (peer memory), the solver runs the serial factor/solve on whole lines and
// stores the solution back into the owners' iterate.  Same operands, same order as the serial kernels: the
// decomposed line relaxation equals the serial GPU solver bit for bit.
struct line_group
{
	int G = 1, me = 0;               // blocks in the group and this block's place in it
	std::vector<int> members;        // block ids, in line direction order
	int nlines = 0, len = 0;         // lines the group owns (own extent across the lines), global line length
	std::vector<int> first;          // first[m] .. first[m+1]: lines (0-based within the group) member m solves; even
	int off0 = 0;                    // global interior index (0-based) of the group's first line
	int mine() const { return first[me + 1] - first[me]; }
};

// geometry of one block on one level: own points and the halo-extended array that holds them
struct block_geom
{
	int x0, y0, nx, ny;       // first own point (0-based global interior index) and own counts
	int wl, wr, wb, wt;       // lines beyond the own points on each side: W towards a neighbour, 1 at the boundary
	int ii, jj;               // array extents: nx + wl + wr, ny + wb + wt (the own points start at [wb][wl], 0-based)
};

struct decomp
{
	int gnx, gny, px, py, ndist, nlevels, W;
	std::vector<std::array<int, 2>> gn;                 // global size per level
	std::vector<std::array<std::vector<int>, 2>> lo;    // lo[lvl][dim][coord], with the end as sentinel

	decomp(int gnx, int gny, int px, int py, int ndist, int W)
	    : gnx(gnx), gny(gny), px(px), py(py), ndist(ndist), nlevels(compute_num_levels(gnx, gny)), W(W)
	{
		if (W < 1 || W % 2 == 0) throw std::invalid_argument("halo width must be odd");
		if (ndist < 1 || ndist > nlevels - 1) throw std::invalid_argument("need 1 <= distributed levels < levels");
		gn.push_back({gnx, gny});
		for (int l = 0; l < ndist; l++) gn.push_back({coarse_n(gn.back()[0]), coarse_n(gn.back()[1])});
		// block boundaries on coarse points of every distributed level (multiples of 2^ndist)
		const int al = 1 << ndist;
		lo.resize(ndist + 1);
		for (int l = 0; l <= ndist; l++)
			for (int d = 0; d < 2; d++) {
				const int n0 = d == 0 ? gnx : gny, p = d == 0 ? px : py;
				for (int c = 0; c < p; c++) lo[l][d].push_back(((int)((long)c * n0 / p) / al * al) >> l);
				lo[l][d].push_back(gn[l][d]);
			}
		for (int d = 0; d < 2; d++)
			for (int c = 0; c < (d == 0 ? px : py); c++)
				if (lo[ndist][d][c + 1] - lo[ndist][d][c] < std::max(W, 2) && (px * py > 1))
					throw std::invalid_argument("blocks smaller than the halo width on the coarsest distributed level");
	}
	int nblocks() const { return px * py; }
	int neighbour(int b, int dx, int dy) const
	{
		const int cx = b % px + dx, cy = b / px + dy;   // x fastest, like the reference's topology
		return (cx < 0 || cx >= px || cy < 0 || cy >= py) ? -1 : cy * px + cx;
	}
	block_geom geom(int lvl, int b) const
	{
		const int cx = b % px, cy = b / px;
		block_geom g;
		g.x0 = lo[lvl][0][cx]; g.nx = lo[lvl][0][cx + 1] - g.x0;
		g.y0 = lo[lvl][1][cy]; g.ny = lo[lvl][1][cy + 1] - g.y0;
		g.wl = cx > 0 ? W : 1; g.wr = cx + 1 < px ? W : 1;
		g.wb = cy > 0 ? W : 1; g.wt = cy + 1 < py ? W : 1;
		g.ii = g.nx + g.wl + g.wr; g.jj = g.ny + g.wb + g.wt;
		return g;
	}
	// the group of blocks that shares the lines of direction dirn (0 = x-lines: the process row, 1 = y-lines: the
	// process column) through block b on level lvl, and how the group's lines are shared out
	line_group lines(int lvl, int b, int dirn) const
	{
		const int cx = b % px, cy = b / px;
		line_group g;
		g.G = dirn == 0 ? px : py;
		g.me = dirn == 0 ? cx : cy;
		for (int m = 0; m < g.G; m++) g.members.push_back(dirn == 0 ? cy * px + m : m * px + cx);
		const block_geom k = geom(lvl, b);
		g.nlines = dirn == 0 ? k.ny : k.nx;
		g.off0 = dirn == 0 ? k.y0 : k.x0;
		g.len = gn[lvl][dirn];
		if (g.off0 % 2) throw std::logic_error("block boundaries of a relaxed level must be even");
		for (int m = 0; m <= g.G; m++) g.first.push_back(m == g.G ? g.nlines : (int)((long)m * g.nlines / g.G) / 2 * 2);
		return g;
	}
	// leading levels on which every block has at least `split_min` own points per dimension (sweeps are split there)
	int split_levels(int split_min) const
	{
		if (split_min <= 0 || px * py == 1) return 0;
		int n = 0;
		for (int l = 0; l < ndist; l++) {
			for (int b = 0; b < px * py; b++) {
				const block_geom g = geom(l, b);
				if (std::min(g.nx, g.ny) < std::max(split_min, 4 * W)) return n;
			}
			n = l + 1;
		}
		return n;
	}
	// window of the level lvl+1 array that the serial transfer kernels see as "the coarse grid" of the whole
	// level-lvl array: first entry (0-based) and extents
	void coarse_window(int lvl, int b, int &ox, int &oy, int &iic, int &jjc) const
	{
		const block_geom f = geom(lvl, b);
		ox = (f.wl - 1) / 2; oy = (f.wb - 1) / 2;
		iic = (f.ii - 3) / 2 + 3; jjc = (f.jj - 3) / 2 + 3;
	}
};

// Rectangles of a split sweep (bmgx2_relax_GS_oop_rects): sub-array [i0, i0+ni) x [j0, j0+nj) of the block's array
// (0-based, even origin) and the window of it that is stored (1-based, inclusive, sub-array coordinates).
struct sweep_rect { int i0, j0, ni, nj, wi0, wi1, wj0, wj1; };

// part 1 = interior: reads only OWN points of the iterate and the right-hand side (its sub-array starts one line
// inside the own region towards every neighbour), so it can run while the halos are being pulled; stores the points at
// least nc lines inside that sub-array.  part 2 = frames: everything else (the halo lines and the own points next to
// them), one rectangle per neighbour side, each with nc + 1 recomputed lines towards the interior.  Together the
// stored windows tile the whole array and every stored value equals the whole-array sweep's: a point's value depends
// on values at most nc lines away (nc colours), and each window keeps that distance from every sub-array edge that is
// not an edge of the array itself.
inline std::vector<sweep_rect> sweep_rects(const block_geom &g, int nc, int part)
{
	const bool L = g.wl > 1, R = g.wr > 1, B = g.wb > 1, T = g.wt > 1;
	// interior sub-array (array coordinates, inclusive) and its stored window
	const int s0 = L ? g.wl + 1 : 0, e0 = R ? g.wl + g.nx - 1 : g.ii - 1;
	const int t0 = B ? g.wb + 1 : 0, f0 = T ? g.wb + g.ny - 1 : g.jj - 1;
	const int cl = L ? s0 + nc : 0, cr = R ? e0 - nc : g.ii - 1;
	const int rb = B ? t0 + nc : 0, rt = T ? f0 - nc : g.jj - 1;
	if (cl > cr || rb > rt) throw std::invalid_argument("block too small to split a sweep");
	auto rect = [](int a0, int a1, int b0, int b1, int wa0, int wa1, int wb0, int wb1) {
		// sub-array cols [a0, a1], rows [b0, b1]; stored cols [wa0, wa1], rows [wb0, wb1] (array coordinates, inclusive)
		sweep_rect r;
		r.i0 = a0; r.j0 = b0; r.ni = a1 - a0 + 1; r.nj = b1 - b0 + 1;
		r.wi0 = wa0 - a0 + 1; r.wi1 = wa1 - a0 + 1; r.wj0 = wb0 - b0 + 1; r.wj1 = wb1 - b0 + 1;
		return r;
	};
	auto even_down = [](int v) { return v < 0 ? 0 : v - (v & 1); };
	std::vector<sweep_rect> out;
	if (part == 1) {
		out.push_back(rect(s0, e0, t0, f0, cl, cr, rb, rt));
		return out;
	}
	if (L) out.push_back(rect(0, std::min(g.ii - 1, cl + nc), 0, g.jj - 1, 0, cl - 1, 0, g.jj - 1));
	if (R) out.push_back(rect(even_down(cr - nc), g.ii - 1, 0, g.jj - 1, cr + 1, g.ii - 1, 0, g.jj - 1));
	const int a0 = even_down(cl - (nc + 1)), a1 = std::min(g.ii - 1, cr + nc + 1);
	if (B) out.push_back(rect(a0, a1, 0, std::min(g.jj - 1, rb + nc), cl, cr, 0, rb - 1));
	if (T) out.push_back(rect(a0, a1, even_down(rt - nc), g.jj - 1, cl, cr, rt + 1, g.jj - 1));
	return out;
}

enum class op_kind { exchange, sync, sweep, residual, restrict_, zero, gather, coarse, scatter, interp, copy, sumsq, norm,
                     recip, setup_interp, galerkin, gather_op, setup_coarse, join, lines_rhs, lines_solve, setup_lines };

// relaxation of the distributed levels (config "solver.relaxation": point, line-x, line-y, line-xy)
enum class relax_kind { point = 0, line_x = 1, line_y = 2, line_xy = 3 };
inline relax_kind relax_from_name(const std::string &n)
{
	if (n == "point") return relax_kind::point;
	if (n == "line-x") return relax_kind::line_x;
	if (n == "line-y") return relax_kind::line_y;
	if (n == "line-xy") return relax_kind::line_xy;
	throw std::invalid_argument("relaxation of the decomposed 2D solver: point, line-x, line-y or line-xy");
}
inline const char *relax_name(relax_kind k)
{
	static const char *n[] = {"point", "line-x", "line-y", "line-xy"};
	return n[(int)k];
}


struct plan_op
{
	op_kind kind;
	int lvl = 0;
	int a = 0, b = 0;           // array slots (meaning depends on kind)
	int dir = 0;                // sweep direction (0 = down, 1 = up)
	std::vector<int> arrays;    // exchange: the slots that travel
	bool wait_all = false;
	bool async = false;         // exchange: runs beside the following operations until the matching `join`
	int part = 0;               // sweep: 0 = whole array, 1 = interior, 2 = frames (sweep_rects)
	int id = 0;                 // async exchange / join: running number of the pair within the list
	// lines_rhs / lines_solve / setup_lines: dir = 0 x-lines, 1 y-lines; part = F-index parity class of the colour:
	// the global F-index (2 = first interior line) of the colour's lines is part, part + 2, ...
};

inline const char *op_name(op_kind k)
{
	static const char *n[] = {"exchange", "sync", "sweep", "residual", "restrict", "zero", "gather", "coarse", "scatter",
	                          "interp", "copy", "sumsq", "norm", "recip", "setup_interp", "galerkin", "gather_op",
	                          "setup_coarse", "join", "lines_rhs", "lines_solve", "setup_lines"};
	return n[(int)k];
}

// Builds operation lists.  `ncol0` = colours of the finest operator's sweep (2 five-point, 4 nine-point);
// coarse operators are nine-point.
class planner
{
public:
	// split_levels: on levels 0 .. split_levels-1 a sweep that directly follows the exchange of its operands is split:
	// the exchange starts, the interior is relaxed meanwhile, the frames after the halos have arrived
	// ncol_coarse: lines of validity a sweep of a coarse level consumes (2D nine-point: its 4 colours; the slab-
	// decomposed 3D path: 2, the depth of the colour ordering across planes).  inplace0 / inplace_coarse: the sweep
	// kernel of that level works in place (3D 27-point) instead of from one array into the other.
	planner(int ndist, int W, int ncol0, int npre, int npost, int split_levels = 0, int ncol_coarse = 4, bool inplace0 = false,
	        bool inplace_coarse = false, relax_kind relax = relax_kind::point)
	    : ndist(ndist), W(W), ncol0(ncol0), npre(npre), npost(npost), split_levels(split_levels), ncol_coarse(ncol_coarse),
	      inplace0(inplace0), inplace_coarse(inplace_coarse), relax(relax)
	{
		if (npre + npost < 1) throw std::invalid_argument("a cycle needs at least one sweep");
		// a whole 4-colour sweep over the extended array must leave the own points exact: colour c is exact out to line
		// W - 1 - (c - 1), so W - 4 >= 0 and W odd
		if (W < std::max(ncol0, ncol_coarse) + 1 || W % 2 == 0)
			throw std::invalid_argument("halo width must be odd and at least the number of colours + 1 (2D: 5)");
	}

	// interpolation + Galerkin operators of the distributed levels, agglomerated operator
	std::vector<plan_op> setup()
	{
		begin();
		exchange(0, {A}, true);
		for (int l = 0; l < ndist; l++) {
			push(op_kind::recip, l);
			push(op_kind::setup_interp, l);          // writes P of level l+1 (own points + what the halo allows)
			exchange(l + 1, {P}, false);
			push(op_kind::galerkin, l);              // writes A of level l+1
			exchange(l + 1, {A}, false);
		}
		// line relaxation: every block pulls the diagonals of the lines it solves out of their owners' operators and
		// factors them (a gather: every rank waits for every rank)
		for (int l = 0; l < ndist && relax != relax_kind::point; l++)
			for (int dirn = 0; dirn < 2; dirn++)
				if ((int)relax & (1 << dirn)) {
					plan_op o; o.kind = op_kind::setup_lines; o.lvl = l; o.dir = dirn; o.a = A; o.wait_all = true;
					ops.push_back(o);
					step++; last_all = step;
				}
		gather(op_kind::gather_op, ndist, A);
		push(op_kind::setup_coarse, ndist);
		return take();
	}

	// one V-cycle on (X0, B) of level 0; own values of both must be current, halos need not be
	std::vector<plan_op> cycle()
	{
		begin();
		cur.assign(ndist + 1, X0);
		val.assign(ndist + 1, std::array<int, 4>{0, 0, 0, 0});
		exchange(0, {X0, B}, true);
		ncycle(0);
		if (cur[0] != X0) { before_write(0, X0); push(op_kind::copy, 0, cur[0], X0); }
		overlap();
		return take();
	}

	// res = b - A x on level 0 and the global sum of squares of its own points (every rank gets the same bits)
	std::vector<plan_op> norm()
	{
		begin();
		val.assign(ndist + 1, std::array<int, 4>{0, 0, 0, 0});
		exchange(0, {X0, B}, true);
		push(op_kind::residual, 0, X0);
		push(op_kind::sumsq, 0, RES);
		plan_op o; o.kind = op_kind::norm; o.wait_all = true;
		ops.push_back(o); step++; last_all = step;
		return take();
	}

	int exchanges(const std::vector<plan_op> &p) const
	{
		int n = 0;
		for (auto &o : p) n += o.kind == op_kind::exchange || o.kind == op_kind::sync || o.kind == op_kind::gather
		                       || o.kind == op_kind::gather_op || o.kind == op_kind::norm;
		return n;
	}

	static std::string to_json(const std::vector<plan_op> &p)
	{
		std::ostringstream s;
		s << "[";
		for (std::size_t i = 0; i < p.size(); i++) {
			const plan_op &o = p[i];
			s << (i ? "," : "") << "{\"op\":\"" << op_name(o.kind) << "\",\"lvl\":" << o.lvl << ",\"a\":" << o.a << ",\"b\":" << o.b
			  << ",\"dir\":" << o.dir << ",\"wait_all\":" << (o.wait_all ? "true" : "false") << ",\"async\":"
			  << (o.async ? "true" : "false") << ",\"part\":" << o.part << ",\"id\":" << o.id << ",\"arrays\":[";
			for (std::size_t k = 0; k < o.arrays.size(); k++) s << (k ? "," : "") << o.arrays[k];
			s << "]}";
		}
		s << "]";
		return s.str();
	}

private:
	static constexpr int INF = 1 << 20;
	int ndist, W, ncol0, npre, npost, split_levels, ncol_coarse;
	bool inplace0, inplace_coarse;
	relax_kind relax;
	std::vector<plan_op> ops;
	std::vector<std::array<int, 4>> val;   // X0, X1, B, RES: halo lines that hold the neighbours' exact values
	std::vector<int> cur;                  // slot that holds the iterate of each level
	long step = 0, last_all = 0;
	std::map<std::pair<int, int>, std::pair<long, bool>> pulled;   // (lvl, slot) -> exchange step that read it, by all ranks?

	void begin() { ops.clear(); pulled.clear(); step = 0; last_all = 0; }
	// exchange(l, operands of the sweep) [zero(l, x)] sweep(l)  ->  [zero] exchange(async) sweep(interior) join sweep(frames).
	// The interior part reads own points only, the pull writes halo lines only, the zero is moved ahead of both.
	void overlap()
	{
		std::vector<plan_op> out;
		int pairs = 0;
		for (std::size_t i = 0; i < ops.size(); i++) {
			const plan_op &e = ops[i];
			std::size_t k = i + 1;
			const bool z = k < ops.size() && ops[k].kind == op_kind::zero && ops[k].lvl == e.lvl;
			if (z) k++;
			bool ok = e.kind == op_kind::exchange && e.lvl < split_levels && k < ops.size() && ops[k].kind == op_kind::sweep
			          && ops[k].lvl == e.lvl && (!z || ops[i + 1].a == ops[k].a);
			if (ok)
				for (int a : e.arrays) ok = ok && (a == ops[k].a || a == B);
			if (!ok) { out.push_back(e); continue; }
			if (z) out.push_back(ops[i + 1]);
			plan_op ex = e; ex.async = true; ex.id = pairs;
			plan_op in = ops[k]; in.part = 1;
			plan_op fr = ops[k]; fr.part = 2;
			plan_op jn; jn.kind = op_kind::join; jn.lvl = e.lvl; jn.id = pairs++;
			out.push_back(ex); out.push_back(in); out.push_back(jn); out.push_back(fr);
			i = k;
		}
		ops.swap(out);
	}
	std::vector<plan_op> take() { std::vector<plan_op> r; r.swap(ops); return r; }
	void push(op_kind k, int lvl, int a = 0, int b = 0, int dir = 0)
	{
		plan_op o; o.kind = k; o.lvl = lvl; o.a = a; o.b = b; o.dir = dir;
		ops.push_back(o);
	}
	void exchange(int lvl, std::vector<int> arrays, bool wait_all)
	{
		plan_op o; o.kind = op_kind::exchange; o.lvl = lvl; o.arrays = arrays; o.wait_all = wait_all;
		ops.push_back(o);
		step++;
		if (wait_all) last_all = step;
		for (int a : arrays) {
			pulled[{lvl, a}] = {step, false};
			if (a < 4 && !val.empty()) val[lvl][a] = W;
		}
	}
	void gather(op_kind k, int lvl, int a)
	{
		plan_op o; o.kind = k; o.lvl = lvl; o.a = a; o.wait_all = true;
		ops.push_back(o);
		step++; last_all = step;
		pulled[{lvl, a}] = {step, true};
	}
	// A region other ranks read in exchange step s may be overwritten only after a later step has waited for them:
	// a rank publishes step s+1 after its reads of step s (cedar_b200/csrc/bmgx2_halo.cu).
	void before_write(int lvl, int a)
	{
		auto it = pulled.find({lvl, a});
		if (it == pulled.end()) return;
		const bool ok = it->second.second ? last_all > it->second.first : step > it->second.first;
		pulled.erase(it);
		if (ok) return;
		plan_op o; o.kind = op_kind::sync; o.wait_all = true;
		ops.push_back(o);
		step++; last_all = step;
	}
	int ncol(int lvl) const { return lvl == 0 ? ncol0 : ncol_coarse; }
	bool inplace(int lvl) const { return lvl == 0 ? inplace0 : inplace_coarse; }

	// zebra line relaxation of the iterate, in place (src/2d/ftn/BMG2_SymStd_relax_lines_x.f90:83-99: on the way down
	// the odd lines first; include/cedar/multilevel.h:165-180: x then y on the way down, y then x on the way up).  Per
	// colour: right-hand sides of the own points into the solvers' line buffers; every rank waits for every rank; whole
	// lines are solved and stored into their owners' iterate; the ghost exchange that follows waits for every rank too,
	// because a block's own lines were written by the other blocks of its group.
	void line_sweep(int l, int dir)
	{
		const int x = cur[l];
		for (int k = 0; k < 2; k++) {
			const int dirn = dir == 0 ? k : 1 - k;
			if (!((int)relax & (1 << dirn))) continue;
			for (int a = 0; a < 2; a++) {
				const int first = dir == 0 ? 3 - a : 2 + a;
				if (val[l][x] < 1) exchange(l, {x}, false);
				plan_op r; r.kind = op_kind::lines_rhs; r.lvl = l; r.a = x; r.dir = dirn; r.part = first;
				ops.push_back(r);
				plan_op sy; sy.kind = op_kind::sync; sy.wait_all = true;
				ops.push_back(sy);
				step++; last_all = step;
				before_write(l, x);
				plan_op sv = r; sv.kind = op_kind::lines_solve;
				ops.push_back(sv);
				exchange(l, {x}, true);
			}
		}
	}

	void sweep(int l, int dir)
	{
		if (relax != relax_kind::point) { line_sweep(l, dir); return; }
		const int nc = ncol(l), src = cur[l], dst = inplace(l) ? src : (src == X0 ? X1 : X0);
		std::vector<int> need;
		if (std::min(val[l][src] - 1, W - 1) - (nc - 1) < 0) need.push_back(src);
		if (val[l][B] - (nc - 1) < 0) need.push_back(B);
		if (!need.empty()) exchange(l, need, false);
		before_write(l, dst);
		push(op_kind::sweep, l, src, dst, dir);
		// first colour: needs the old iterate one line further out and the right-hand side; never beyond line W - 1 (the
		// outermost line is the frozen ring); every further colour needs the previous colours one line further out
		val[l][dst] = std::min(std::min(val[l][src] - 1, val[l][B]), W - 1) - (nc - 1);
		cur[l] = dst;
	}
	void residual(int l)
	{
		const int x = cur[l];
		std::vector<int> need;
		if (val[l][x] - 1 < 1) need.push_back(x);
		if (val[l][B] < 1) need.push_back(B);
		if (!need.empty()) exchange(l, need, false);
		before_write(l, RES);
		push(op_kind::residual, l, x);
		val[l][RES] = std::min(std::min(val[l][x] - 1, val[l][B]), W - 1);
	}
	void ncycle(int l)
	{
		for (int i = 0; i < npre; i++) sweep(l, 0);
		residual(l);
		before_write(l + 1, B);
		push(op_kind::restrict_, l);
		if (l + 1 == ndist) {
			gather(op_kind::gather, ndist, B);
			push(op_kind::coarse, ndist);
			before_write(ndist, X0);
			push(op_kind::scatter, ndist, X0);
			val[ndist][X0] = W;
			cur[ndist] = X0;
		} else {
			exchange(l + 1, {B}, false);
			before_write(l + 1, X0);
			push(op_kind::zero, l + 1, X0);
			val[l + 1][X0] = INF;
			cur[l + 1] = X0;
			ncycle(l + 1);
		}
		// interpolate and add: fine halo line k needs the coarse iterate out to line (k + 1) / 2 + 1 at most
		const int x = cur[l], xc = cur[l + 1];
		if (val[l + 1][xc] < 1) exchange(l + 1, {xc}, false);
		before_write(l, x);
		before_write(l, RES);
		push(op_kind::interp, l, x, xc);
		val[l][x] = std::min(std::min(val[l][x], val[l][RES]), std::min(2 * val[l + 1][xc] - 1, W));
		for (int i = 0; i < npost; i++) sweep(l, 1);
	}
};

}}}
#endif
