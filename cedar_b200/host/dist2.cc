// dist2.cc -- the block-decomposed (multi-GPU) 2D BoxMG solver: CUDA executor of the operation lists of
// include/cedar/2d/mpi/plan.h, and its C API (include/cedar_capi.h, cdr2_dist_*).
//
// Counterpart of the reference's cdr2::mpi::solver (include/cedar/2d/mpi/solver.h) with its MPI halo exchange
// (src/2d/mpi/halo.cc) replaced by pulls out of the neighbours' device memory over NVLink
// (cedar_b200/csrc/bmgx2_halo.cu).  One process per GPU; a process may also own several blocks (they then all
// live on its GPU: the single-GPU tests of the decomposed path).  The library does no inter-process
// communication of its own: the caller moves the 64-byte IPC handles between the processes once
// (cdr2_dist_export_handles / cdr2_dist_connect), like an application would with its MPI communicator.
//
// Levels whose global size is at most `agglomerate` are solved REPLICATED on every rank by the serial
// cdr2::solver: the ranks pull each other's blocks of the restricted right-hand side into a global array, run
// the same serial cycle, and take their block (halo included) of the result -- no scatter, no idle GPUs.
#include <cmath>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include <bmgx.h>
#include <cedar_capi.h>
#include <cedar/gpu.h>
#include <cedar/halo_exchanger_base.h>
#include <cedar/2d/mpi/plan.h>

using namespace cedar::cdr2::mpi;
using cedar::gpu::check;

void cdr_set_error(const std::string &msg);

namespace {

constexpr size_t ALIGN_D = 32;   // array offsets in the arena: multiples of 32 doubles (256 B)
inline size_t up(size_t n) { return (n + ALIGN_D - 1) / ALIGN_D * ALIGN_D; }

struct seg_table { bmgx_seg *dev = nullptr; int n = 0; int chunks = 1; };

// lines of one direction through a block on one level (line relaxation across block cuts, plan.h line_group)
struct line_store
{
	line_group grp;
	double *sor = nullptr;                    // factor of the lines this block solves: 2 planes (mine + 2) x (len + 2)
	double *strip = nullptr;                  // setup only: gathered diagonals, laid out like a serial operator
	bmgx_line_group dst{};                    // where the right-hand sides go
	std::map<int, bmgx_line_owners> own;      // where the solutions go, per iterate slot
};

struct dist_block
{
	int id = -1;
	char *arena = nullptr;
	std::vector<block_geom> g;
	cdr2_operator *cop = nullptr;
	cdr2_solver *csol = nullptr;
	double *GB = nullptr, *GX = nullptr;   // the replicated solver's own b and (current) x: global arrays of the agglomerated level
	double *GA = nullptr;                  // setup: the gathered operator of that level
	double *tmpP = nullptr, *tmpA = nullptr;              // setup: contiguous coarse windows
	double *work = nullptr, *red_all = nullptr, *norm_out = nullptr;
	std::map<std::string, seg_table> segs;
	std::map<std::pair<int, int>, line_store> lines;   // (level, direction)
};

}  // namespace

struct cdr2_dist
{
	std::unique_ptr<decomp> dec;
	std::unique_ptr<planner> plan;
	int nst = 5, npre = 1, npost = 1, W = 15;
	relax_kind relax = relax_kind::point;
	bool use_graph = true;
	double timeout_s = 10.0;
	std::vector<dist_block> blk;                 // owned blocks
	std::vector<char *> base;                    // arena of every block as mapped in this process
	std::vector<bool> opened;                    // base[i] came from bmgx_ipc_open
	void **peer_sync_dev = nullptr;
	bmgx_stream_t stream = nullptr, side = nullptr;   // side: exchanges that run beside the interior sweeps
	std::vector<void *> fork_ev, join_ev;             // one pair per asynchronous exchange of the cycle
	bool overlap = true;
	std::vector<plan_op> p_setup, p_cycle, p_norm;
	void *graph = nullptr;
	unsigned long long graph_nodes = 0;
	int warm = 0;
	bool connected = false, is_setup = false;
	size_t sync_words = 0;
	std::unique_ptr<cedar::halo_exchanger_base> exchanger;   // the reference's seam over the peer-memory exchange
	// e2e staging
	double *db[2] = {nullptr, nullptr}, *dx[2] = {nullptr, nullptr};
	bmgx_stream_t s_up = nullptr, s_down = nullptr;
	void *ev[8] = {};

	int nblocks() const { return dec->nblocks(); }
	int nstl(int lvl) const { return lvl == 0 ? nst : 5; }
	int planes(int lvl, int a) const { return a == A ? nstl(lvl) : a == SOR ? 2 : a == P ? 8 : 1; }

	// arena layout of ANY block (every rank can compute every other rank's offsets), in doubles
	size_t off_sync() const { return 0; }
	size_t off_red() const { return up(sync_words); }
	size_t off_arr(int b, int lvl, int a) const
	{
		size_t o = off_red() + ALIGN_D;
		for (int l = 0; l <= dec->ndist; l++) {
			const block_geom g = dec->geom(l, b);
			const size_t pl = (size_t)g.ii * g.jj;   // planes of an array are contiguous, as the serial kernels expect
			for (int s = 0; s < NARR; s++) {
				if (l == lvl && s == a) return o;
				if (s == P && l == 0) continue;
				o += up((size_t)planes(l, s) * pl);
			}
		}
		return o;
	}
	// line buffers (line relaxation) follow the arrays: one per relaxed level and line direction
	bool relaxes(int dirn) const { return ((int)relax >> dirn) & 1; }
	size_t off_lb(int b, int lvl, int dirn) const
	{
		size_t o = off_arr(b, dec->ndist + 1, 0);
		for (int l = 0; l < dec->ndist; l++)
			for (int dd = 0; dd < 2; dd++) {
				if (!relaxes(dd)) continue;
				if (l == lvl && dd == dirn) return o;
				const line_group g = dec->lines(l, b, dd);
				if (g.G > 1) o += up((size_t)(g.mine() + 2) * (g.len + 2));
			}
		return o;
	}
	double *lb(int b, int lvl, int dirn) const { return reinterpret_cast<double *>(base[b]) + off_lb(b, lvl, dirn); }
	size_t arena_doubles(int b) const { return off_lb(b, dec->ndist, 0); }
	size_t plane_stride(int b, int lvl) const { const block_geom g = dec->geom(lvl, b); return (size_t)g.ii * g.jj; }
	double *ptr(int b, int lvl, int a, int plane = 0) const
	{
		return reinterpret_cast<double *>(base[b]) + off_arr(b, lvl, a) + (size_t)plane * plane_stride(b, lvl);
	}
};

namespace {

template <class F> int guard(F &&f)
{
	try { f(); return 0; }
	catch (std::exception &e) { cdr_set_error(e.what()); return 1; }
	catch (...) { cdr_set_error("unknown error"); return 1; }
}

std::string key_of(const plan_op &o)
{
	std::string k = std::string(op_name(o.kind)) + ":" + std::to_string(o.lvl) + ":" + std::to_string(o.a) + ":" + std::to_string(o.dir);
	for (int a : o.arrays) k += "," + std::to_string(a);
	return k;
}

void upload_segs(seg_table &t, const std::vector<bmgx_seg> &v, bmgx_stream_t s)
{
	t.n = (int)v.size();
	size_t big = 0;
	for (auto &sg : v) big = std::max(big, (size_t)sg.rows * sg.cols);
	t.chunks = (int)std::min<size_t>(64, std::max<size_t>(1, (big + 1023) / 1024));   // 256 threads x 4 loads in flight
	if (t.n == 0) return;
	check(bmgx_malloc((void **)&t.dev, v.size() * sizeof(bmgx_seg)), "segment table");
	check(bmgx_memcpy_h2d(t.dev, v.data(), v.size() * sizeof(bmgx_seg), s), "segment table upload");
	check(bmgx_stream_sync(s), "segment table sync");
}

// rectangles of one ghost exchange of `arrays` on level lvl for block b: the W own lines of every neighbour next
// to the common side (a W x W corner for the diagonal neighbours) -> this block's halo
std::vector<bmgx_seg> exchange_segs(const cdr2_dist &d, int lvl, int b, const std::vector<int> &arrays)
{
	std::vector<bmgx_seg> v;
	const decomp &dec = *d.dec;
	const block_geom g = dec.geom(lvl, b);
	const int W = dec.W;
	for (int dy = -1; dy <= 1; dy++)
		for (int dx = -1; dx <= 1; dx++) {
			if (!dx && !dy) continue;
			const int nb = dec.neighbour(b, dx, dy);
			if (nb < 0) continue;
			const block_geom h = dec.geom(lvl, nb);
			if ((dx && h.nx < W) || (dy && h.ny < W)) throw std::runtime_error("neighbour block smaller than the halo width");
			// columns: mine (destination) and the neighbour's (source)
			int dc, sc, nc, dr, sr, nr;
			if (dx < 0) { dc = 0; nc = W; sc = h.wl + h.nx - W; }
			else if (dx > 0) { dc = g.wl + g.nx; nc = W; sc = h.wl; }
			else { dc = g.wl; nc = g.nx; sc = h.wl; }
			if (dy < 0) { dr = 0; nr = W; sr = h.wb + h.ny - W; }
			else if (dy > 0) { dr = g.wb + g.ny; nr = W; sr = h.wb; }
			else { dr = g.wb; nr = g.ny; sr = h.wb; }
			for (int a : arrays)
				for (int p = 0; p < d.planes(lvl, a); p++) {
					bmgx_seg s{};
					s.src = d.ptr(nb, lvl, a, p) + (size_t)sr * h.ii + sc;
					s.dst = d.ptr(b, lvl, a, p) + (size_t)dr * g.ii + dc;
					s.rows = nr; s.cols = nc; s.src_pitch = h.ii; s.dst_pitch = g.ii; s.src_rank = nb;
					v.push_back(s);
				}
		}
	return v;
}

// own points of every block of array `a` on the agglomerated level -> a global array (ghost ring untouched)
std::vector<bmgx_seg> gather_segs(const cdr2_dist &d, int lvl, int a, double *global)
{
	std::vector<bmgx_seg> v;
	const decomp &dec = *d.dec;
	const size_t gii = dec.gn[lvl][0] + 2, gjj = dec.gn[lvl][1] + 2;
	for (int q = 0; q < dec.nblocks(); q++) {
		const block_geom h = dec.geom(lvl, q);
		for (int p = 0; p < d.planes(lvl, a); p++) {
			bmgx_seg s{};
			s.src = d.ptr(q, lvl, a, p) + (size_t)h.wb * h.ii + h.wl;
			s.dst = global + (size_t)p * gii * gjj + (size_t)(1 + h.y0) * gii + (1 + h.x0);
			s.rows = h.ny; s.cols = h.nx; s.src_pitch = h.ii; s.dst_pitch = gii; s.src_rank = q;
			v.push_back(s);
		}
	}
	return v;
}

line_store &lines_of(cdr2_dist &d, dist_block &K, int lvl, int dirn)
{
	auto it = K.lines.find({lvl, dirn});
	if (it != K.lines.end()) return it->second;
	line_store ls;
	ls.grp = d.dec->lines(lvl, K.id, dirn);
	const line_group &g = ls.grp;
	if (g.G > BMGX_LINE_GROUP_MAX) throw std::runtime_error("line relaxation: more than BMGX_LINE_GROUP_MAX blocks along a line");
	ls.dst.nmembers = g.G;
	ls.dst.pitch = (size_t)g.len + 2;
	for (int m = 0; m <= g.G; m++) ls.dst.first[m] = g.first[m];
	for (int m = 0; m < g.G; m++) ls.dst.lb[m] = d.lb(g.members[m], lvl, dirn);
	return K.lines.emplace(std::make_pair(lvl, dirn), ls).first->second;
}

const bmgx_line_owners &owners_of(cdr2_dist &d, line_store &ls, int lvl, int dirn, int slot)
{
	auto it = ls.own.find(slot);
	if (it != ls.own.end()) return it->second;
	bmgx_line_owners o{};
	const line_group &g = ls.grp;
	o.nmembers = g.G;
	for (int m = 0; m < g.G; m++) {
		const block_geom h = d.dec->geom(lvl, g.members[m]);
		o.start[m] = dirn == 0 ? h.x0 : h.y0;
		o.w_along[m] = dirn == 0 ? h.wl : h.wb;
		o.w_across = dirn == 0 ? h.wb : h.wl;
		o.pitch[m] = h.ii;
		o.q[m] = d.ptr(g.members[m], lvl, slot);
	}
	o.start[g.G] = g.len;
	return ls.own.emplace(slot, o).first->second;
}

// diagonals (KO and the coupling along the line) of the lines this block solves, out of their owners' operators into
// a strip laid out like a serial operator: x-lines (mine + 2) rows of len + 2, y-lines len + 2 rows of (mine + 2)
std::vector<bmgx_seg> line_setup_segs(const cdr2_dist &d, int lvl, int dirn, const line_store &ls)
{
	std::vector<bmgx_seg> v;
	const line_group &g = ls.grp;
	const int mine = g.mine(), t0 = g.first[g.me];
	if (mine == 0 || g.G == 1) return v;      // nothing to gather: no lines here, or whole lines in this block
	const size_t sii = dirn == 0 ? g.len + 2 : mine + 2, sjj = dirn == 0 ? mine + 2 : g.len + 2;
	const int planes[2] = {0, dirn == 0 ? 1 : 2};   // KO, KW | KS
	for (int m = 0; m < g.G; m++) {
		const int q = g.members[m];
		const block_geom h = d.dec->geom(lvl, q);
		for (int p : planes) {
			bmgx_seg s{};
			if (dirn == 0) {
				s.src = d.ptr(q, lvl, A, p) + (size_t)(h.wb + t0) * h.ii + h.wl;
				s.dst = ls.strip + (size_t)p * sii * sjj + sii + 1 + h.x0;
				s.rows = mine; s.cols = h.nx;
			} else {
				s.src = d.ptr(q, lvl, A, p) + (size_t)h.wb * h.ii + h.wl + t0;
				s.dst = ls.strip + (size_t)p * sii * sjj + (size_t)(1 + h.y0) * sii + 1;
				s.rows = h.ny; s.cols = mine;
			}
			s.src_pitch = h.ii; s.dst_pitch = sii; s.src_rank = q;
			v.push_back(s);
		}
	}
	return v;
}

seg_table &table(cdr2_dist &d, dist_block &K, const plan_op &o)
{
	const std::string k = key_of(o);
	auto it = K.segs.find(k);
	if (it != K.segs.end()) return it->second;
	std::vector<bmgx_seg> v;
	const int nd = d.dec->ndist;
	switch (o.kind) {
	case op_kind::exchange: v = exchange_segs(d, o.lvl, K.id, o.arrays); break;
	case op_kind::gather: v = gather_segs(d, nd, o.a, K.GB); break;   // GB = the replicated solver's own right-hand side
	case op_kind::gather_op: v = gather_segs(d, nd, A, K.GA); break;
	case op_kind::setup_lines: v = line_setup_segs(d, o.lvl, o.dir, lines_of(d, K, o.lvl, o.dir)); break;
	case op_kind::norm:
		for (int q = 0; q < d.nblocks(); q++) {
			bmgx_seg s{};
			s.src = reinterpret_cast<double *>(d.base[q]) + d.off_red();
			s.dst = K.red_all + q;
			s.rows = s.cols = 1; s.src_pitch = s.dst_pitch = 1; s.src_rank = q;
			v.push_back(s);
		}
		break;
	default: break;
	}
	seg_table t;
	upload_segs(t, v, d.stream);
	return K.segs.emplace(k, t).first->second;
}

void step(cdr2_dist &d, const plan_op &o, bmgx_stream_t st)
{
	const int world = d.nblocks();
	const bool local = d.blk.size() > 1;
	if (local)
		for (auto &K : d.blk)
			check(bmgx_halo_step(nullptr, 0, 1, d.peer_sync_dev, K.arena, K.id, world, 1, 0, d.timeout_s, st), "halo signal");
	for (auto &K : d.blk) {
		seg_table &t = table(d, K, o);
		check(bmgx_halo_step(t.dev, t.n, t.chunks, d.peer_sync_dev, K.arena, K.id, world, local ? 2 : 3, o.wait_all ? 1 : 0,
		                     d.timeout_s, st), "halo step");
	}
}

// The reference's ghost-exchange seam (include/cedar/halo_exchanger_base.h; reference kernels/halo_exchange.h:22-31,
// called from inside its MPI kernels as halof->exchange_func(k, q)) over the peer-memory exchange: one pull of the W
// halo lines of the array `gf` / `so` on level k out of the neighbours' arrays.  k is the reference's level number
// (k = number of distributed levels + 1 is the finest grid, k = 1 the agglomerated one).  The pointer names the array
// (iterate, right-hand side, residual or operator of an owned block); the call is collective: every process makes it,
// and it exchanges that array for all the blocks the process owns.
struct peer_exchanger : cedar::halo_exchanger_base
{
	cdr2_dist *d;
	explicit peer_exchanger(cdr2_dist *d) : d(d) {}
	void exchange(int k, const double *p, bool sten)
	{
		const int lvl = d->dec->ndist + 1 - k;
		if (lvl < 0 || lvl > d->dec->ndist) throw std::runtime_error("halo exchange: level out of range");
		if (!d->connected) throw std::runtime_error("halo exchange: cdr2_dist_connect has not run");
		for (auto &K : d->blk)
			for (int a : {X0, X1, B, RES, A}) {
				if ((a == A) != sten || d->ptr(K.id, lvl, a) != p) continue;
				plan_op o; o.kind = op_kind::exchange; o.lvl = lvl; o.arrays = {a};
				step(*d, o, d->stream);
				return;
			}
		throw std::runtime_error("halo exchange: the pointer is not a level array of a block owned by this process");
	}
	void exchange_func(int k, cedar::real_t *gf) override { exchange(k, gf, false); }
	void exchange_sten(int k, cedar::real_t *so) override { exchange(k, so, true); }
};

void run_op(cdr2_dist &d, const plan_op &o)
{
	const decomp &dec = *d.dec;
	const int l = o.lvl, nd = dec.ndist;
	bmgx_stream_t s = d.stream;
	switch (o.kind) {
	case op_kind::join:
		// the exchange started by the matching asynchronous operation has delivered the halos
		if (d.overlap) check(bmgx_stream_wait_event(s, d.join_ev[o.id]), "join");
		return;
	case op_kind::setup_lines:
		// gather the diagonals of my lines (every rank waits for every rank: the operators are complete), factor them
		for (auto &K : d.blk) {
			line_store &ls = lines_of(d, K, l, o.dir);
			const int mine = ls.grp.mine();
			if (ls.grp.G == 1) {
				// no cut across these lines: the serial kernels on the block's own (halo-extended) arrays
				check(bmgx_malloc((void **)&ls.sor, 2 * d.plane_stride(K.id, l) * 8), "line factor");
				check(bmgx_memset0(ls.sor, 2 * d.plane_stride(K.id, l) * 8, s), "line factor");
				continue;
			}
			if (mine == 0) continue;
			const size_t n = (size_t)(mine + 2) * (ls.grp.len + 2);
			check(bmgx_malloc((void **)&ls.strip, 3 * n * 8), "line strip");
			check(bmgx_malloc((void **)&ls.sor, 2 * n * 8), "line factor");
			check(bmgx_memset0(ls.strip, 3 * n * 8, s), "line strip");
			check(bmgx_memset0(ls.sor, 2 * n * 8, s), "line factor");
		}
		step(d, o, s);
		for (auto &K : d.blk) {
			line_store &ls = lines_of(d, K, l, o.dir);
			const int mine = ls.grp.mine();
			if (ls.grp.G == 1) {
				const block_geom g = K.g[l];
				if (o.dir == 0) check(bmgx2_SETUP_lines_x(d.ptr(K.id, l, A), ls.sor, g.ii, g.jj, d.nstl(l), 0, s), "SETUP_lines_x");
				else check(bmgx2_SETUP_lines_y(d.ptr(K.id, l, A), ls.sor, g.ii, g.jj, d.nstl(l), 0, s), "SETUP_lines_y");
				continue;
			}
			if (mine == 0) continue;
			if (o.dir == 0) check(bmgx2_SETUP_lines_x(ls.strip, ls.sor, ls.grp.len + 2, mine + 2, 3, 0, s), "SETUP_lines_x");
			else check(bmgx2_SETUP_lines_y(ls.strip, ls.sor, mine + 2, ls.grp.len + 2, 3, 0, s), "SETUP_lines_y");
		}
		return;
	case op_kind::exchange: case op_kind::sync: case op_kind::gather: case op_kind::gather_op: case op_kind::norm:
		if (o.async && d.overlap) {
			// fork: the exchange kernel goes to the side stream behind everything enqueued so far; the operations up to
			// the join (the interior of the next sweep) run beside it
			check(bmgx_event_record(d.fork_ev[o.id], s), "fork");
			check(bmgx_stream_wait_event(d.side, d.fork_ev[o.id]), "fork");
			step(d, o, d.side);
			check(bmgx_event_record(d.join_ev[o.id], d.side), "fork");
			return;
		}
		step(d, o, s);
		if (o.kind == op_kind::norm)
			for (auto &K : d.blk) check(bmgx_sum_ordered(K.red_all, d.nblocks(), K.norm_out, s), "norm fold");
		return;
	default: break;
	}
	for (auto &K : d.blk) {
		const int b = K.id;
		const block_geom g = K.g[l];
		const int ifd = (l == 0 && d.nst == 3) ? 1 : 0, nst = d.nstl(l);
		// the coarse window the transfer kernels see (level l + 1 arrays)
		int ox = 0, oy = 0, iic = 0, jjc = 0;
		size_t cpitch = 0, cplane = 0, coff = 0;
		if (l < nd) {
			dec.coarse_window(l, b, ox, oy, iic, jjc);
			cpitch = K.g[l + 1].ii; cplane = d.plane_stride(b, l + 1); coff = (size_t)oy * cpitch + ox;
		}
		switch (o.kind) {
		case op_kind::sweep:
			if (o.part == 0) {
				check(bmgx2_relax_GS_oop(1, d.ptr(b, l, A), d.ptr(b, l, B), d.ptr(b, l, o.a), d.ptr(b, l, o.b), d.ptr(b, l, SOR),
				                         g.ii, g.jj, 1, ifd, nst, 2, BMGX_RELAX_SYM, o.dir ? BMGX_UP : BMGX_DOWN, 0, s), "relax_GS_oop");
			} else {
				// a split sweep: interior (own points only, beside the exchange) or the frames around it
				const std::vector<sweep_rect> rs = sweep_rects(g, ifd ? 2 : 4, o.part);
				bmgx_rect br[5];
				for (std::size_t k = 0; k < rs.size(); k++)
					br[k] = bmgx_rect{(bmgx_len)rs[k].i0, (bmgx_len)rs[k].j0, (bmgx_len)rs[k].ni, (bmgx_len)rs[k].nj,
					                  (bmgx_len)rs[k].wi0, (bmgx_len)rs[k].wi1, (bmgx_len)rs[k].wj0, (bmgx_len)rs[k].wj1};
				if (!rs.empty())
					check(bmgx2_relax_GS_oop_rects(1, d.ptr(b, l, A), d.ptr(b, l, B), d.ptr(b, l, o.a), d.ptr(b, l, o.b),
					                               d.ptr(b, l, SOR), g.ii, g.jj, 1, ifd, nst, 2, BMGX_RELAX_SYM,
					                               o.dir ? BMGX_UP : BMGX_DOWN, 0, br, (int)rs.size(), s), "relax_GS_oop_rects");
			}
			break;
		case op_kind::lines_rhs: {
			line_store &ls = lines_of(d, K, l, o.dir);
			if (ls.grp.G == 1) break;       // fused with the solve below
			check(bmgx2_lines_rhs_push(o.dir, ifd ? 0 : 1, d.ptr(b, l, A), d.ptr(b, l, B), d.ptr(b, l, o.a), g.ii, g.jj, g.wl, g.wb,
			                           g.nx, g.ny, o.part, 1 + (o.dir == 0 ? g.x0 : g.y0), &ls.dst, s), "lines_rhs_push");
			break;
		}
		case op_kind::lines_solve: {
			line_store &ls = lines_of(d, K, l, o.dir);
			const int mine = ls.grp.mine();
			if (ls.grp.G == 1) {
				// whole lines live in this block: one colour of the serial sweep over the extended array (the halo lines are
				// recomputed from stale neighbours and replaced by the exchange that follows)
				if (o.dir == 0)
					check(bmgx2_relax_lines_x_colour(d.ptr(b, l, A), d.ptr(b, l, B), d.ptr(b, l, o.a), ls.sor, g.ii, g.jj, ifd ? 0 : 1,
					                                 o.part, s), "relax_lines_x");
				else
					check(bmgx2_relax_lines_y_colour(d.ptr(b, l, A), d.ptr(b, l, B), d.ptr(b, l, o.a), ls.sor, d.ptr(b, l, RES), g.ii,
					                                 g.jj, ifd ? 0 : 1, o.part, s), "relax_lines_y");
				break;
			}
			if (mine == 0) break;
			double *buf = d.lb(b, l, o.dir);
			check(bmgx2_lines_solve_T(buf, ls.sor, mine + 2, ls.grp.len + 2, o.part, s), "lines_solve_T");
			check(bmgx2_lines_scatter_push(o.dir, buf, mine + 2, ls.grp.len + 2, o.part, ls.grp.first[ls.grp.me],
			                               &owners_of(d, ls, l, o.dir, o.a), s), "lines_scatter_push");
			break;
		}
		case op_kind::residual:
			check(bmgx2_residual(0, d.ptr(b, l, A), d.ptr(b, l, B), d.ptr(b, l, o.a), d.ptr(b, l, RES), g.ii, g.jj, 0, ifd, nst,
			                     0, 0, 0, 0, s), "residual");
			break;
		case op_kind::restrict_:
			check(bmgx2_restrict_v(d.ptr(b, l, RES), d.ptr(b, l + 1, B) + coff, d.ptr(b, l + 1, P) + coff, g.ii, g.jj, iic, jjc,
			                       cpitch, cplane, s), "restrict");
			break;
		case op_kind::interp:
			check(bmgx2_interp_add_v(d.ptr(b, l, o.a), d.ptr(b, l + 1, o.b) + coff, d.ptr(b, l, RES), d.ptr(b, l, A),
			                         d.ptr(b, l + 1, P) + coff, iic, jjc, g.ii, g.jj, cpitch, cplane, s), "interp_add");
			break;
		case op_kind::zero:
			check(bmgx_memset0(d.ptr(b, l, o.a), (size_t)g.ii * g.jj * 8, s), "zero");
			break;
		case op_kind::copy:
			check(bmgx_memcpy_d2d(d.ptr(b, l, o.b), d.ptr(b, l, o.a), (size_t)g.ii * g.jj * 8, s), "copy");
			break;
		case op_kind::coarse: {
			// the replicated serial solver works on its own device buffers: the gather wrote its b; clear its x, cycle,
			// and remember where its iterate ended up (the fused relaxation ping-pongs x with a scratch array)
			const size_t n = (size_t)(dec.gn[nd][0] + 2) * (dec.gn[nd][1] + 2);
			double *cb = nullptr, *cx = nullptr;
			if (cdr2_solver_resident_ptrs(K.csol, &cb, &cx)) throw std::runtime_error(cdr_last_error());
			if (cb != K.GB) throw std::logic_error("replicated solver moved its right-hand side");
			check(bmgx_memset0(cx, n * 8, s), "coarse x0");
			if (cdr2_solver_vcycle_resident(K.csol, s)) throw std::runtime_error(std::string("agglomerated cycle: ") + cdr_last_error());
			if (cdr2_solver_resident_ptrs(K.csol, &cb, &K.GX)) throw std::runtime_error(cdr_last_error());
			break;
		}
		case op_kind::scatter: {
			// this block's window of the replicated solution, halo included: array entry (r, c) is the global
			// array's (y0 + 1 - wb + r, x0 + 1 - wl + c)
			const size_t gii = dec.gn[nd][0] + 2;
			const double *src = K.GX + (size_t)(g.y0 + 1 - g.wb) * gii + (g.x0 + 1 - g.wl);
			check(bmgx_copy2d(d.ptr(b, l, o.a), g.ii, src, gii, g.ii, g.jj, s), "scatter");
			break;
		}
		case op_kind::sumsq:
			check(bmgx2_sumsq_win(d.ptr(b, l, o.a) + (size_t)g.wb * g.ii + g.wl, g.ii, g.nx, g.ny, K.work,
			                      reinterpret_cast<double *>(K.arena) + d.off_red(), s), "sumsq");
			break;
		case op_kind::recip:
			check(bmgx2_SETUP_recip(d.ptr(b, l, A), d.ptr(b, l, SOR), g.ii, g.jj, nst, 2, s), "SETUP_recip");
			break;
		case op_kind::setup_interp: {
			// on the contiguous window, then into the (wider) level l + 1 array
			const size_t pc = (size_t)iic * jjc;
			check(bmgx_memset0(K.tmpP, 8 * pc * 8, s), "P window");
			check(bmgx2_SETUP_interp_OI(d.ptr(b, l, A), nullptr, K.tmpP, g.ii, g.jj, iic, jjc, ifd, nst, 0, 0, s), "SETUP_interp_OI");
			for (int p = 0; p < 8; p++)
				check(bmgx_copy2d(d.ptr(b, l + 1, P, p) + coff, cpitch, K.tmpP + p * pc, iic, iic, jjc, s), "P window copy");
			break;
		}
		case op_kind::galerkin: {
			const size_t pc = (size_t)iic * jjc;
			check(bmgx_memset0(K.tmpA, 5 * pc * 8, s), "A window");
			check(bmgx2_SETUP_ITLI_ex(d.ptr(b, l, A), K.tmpA, K.tmpP, g.ii, g.jj, iic, jjc, ifd, nst, 0, s), "SETUP_ITLI_ex");
			for (int p = 0; p < 5; p++)
				check(bmgx_copy2d(d.ptr(b, l + 1, A, p) + coff, cpitch, K.tmpA + p * pc, iic, iic, jjc, s), "A window copy");
			break;
		}
		case op_kind::setup_coarse: {
			const unsigned gx = dec.gn[nd][0], gy = dec.gn[nd][1];
			const size_t n = (size_t)(gx + 2) * (gy + 2) * 5;
			if (cdr2_operator_create(5, gx, gy, &K.cop)) throw std::runtime_error(cdr_last_error());
			double *h = cdr2_operator_data(K.cop);
			if (!h) throw std::runtime_error(cdr_last_error());
			check(bmgx_memcpy_d2h(h, K.GA, n * 8, s), "agglomerated operator download");
			check(bmgx_stream_sync(s), "sync");
			const int nl = dec.nlevels - nd;
			const std::string cfg = "{\"log\": [\"error\"], \"solver\": {\"relaxation\": \"" + std::string(relax_name(d.relax)) + "\", \"cycle\": {\"type\": \"v\", "
			                        "\"nrelax-pre\": " + std::to_string(d.npre) + ", \"nrelax-post\": " + std::to_string(d.npost) +
			                        "}, \"max-iter\": 1, \"tol\": 0, \"cuda-graph\": false, \"num-levels\": " + std::to_string(nl) + "}}";
			if (cdr2_solver_create(K.cop, cfg.c_str(), &K.csol)) throw std::runtime_error(cdr_last_error());
			if (cdr2_solver_nlevels(K.csol) != nl) throw std::runtime_error("agglomerated solver: unexpected number of levels");
			if (cdr2_solver_resident_ptrs(K.csol, &K.GB, &K.GX)) throw std::runtime_error(cdr_last_error());
			break;
		}
		default: throw std::logic_error("unknown plan operation");
		}
	}
}

void run(cdr2_dist &d, const std::vector<plan_op> &p)
{
	for (auto &o : p) run_op(d, o);
}

void check_device_error(cdr2_dist &d)
{
	for (auto &K : d.blk) {
		unsigned long long e = 0;
		check(bmgx_memcpy_d2h(&e, reinterpret_cast<unsigned long long *>(K.arena) + d.nblocks() + 2, 8, d.stream), "error word");
		check(bmgx_stream_sync(d.stream), "sync");
		if (e) throw std::runtime_error("ghost exchange timed out waiting for block " + std::to_string((long long)e - 1) +
		                                " (a rank is missing, or the ranks did not issue the same sequence of exchanges)");
	}
}

void cycle(cdr2_dist &d)
{
	if (!d.is_setup) throw std::runtime_error("cdr2_dist_setup has not run");
	if (!d.use_graph || d.warm < 2) { run(d, d.p_cycle); d.warm++; return; }
	if (!d.graph) {
		const unsigned long long before = bmgx_launch_count();
		check(bmgx_graph_begin(d.stream), "cycle graph capture");
		try { run(d, d.p_cycle); }
		catch (...) { void *g = nullptr; bmgx_graph_end(d.stream, &g); if (g) bmgx_graph_destroy(g); throw; }
		check(bmgx_graph_end(d.stream, &d.graph), "cycle graph instantiate");
		d.graph_nodes = bmgx_launch_count() - before;
	} else {
		cedar::gpu::add_launches(d.graph_nodes);
	}
	check(bmgx_graph_launch(d.graph, d.stream), "cycle graph launch");
}

dist_block &find(cdr2_dist &d, int b)
{
	for (auto &K : d.blk) if (K.id == b) return K;
	throw std::runtime_error("block " + std::to_string(b) + " is not owned by this process");
}

}  // namespace

extern "C" {

int cdr2_dist_create(unsigned gnx, unsigned gny, int px, int py, int nstencil, int nowned, const int *owned, int nrelax_pre,
                     int nrelax_post, const char *relaxation, int agglomerate, int halo_width, int split_min, int use_graph,
                     cdr2_dist **out)
{
	return guard([&] {
		if (nstencil != 3 && nstencil != 5) throw std::runtime_error("nstencil must be 3 or 5");
		if (!cedar::gpu::available()) throw std::runtime_error("no CUDA device: the BoxMG kernels have no CPU fallback");
		auto d = std::make_unique<cdr2_dist>();
		d->relax = relax_from_name(relaxation ? relaxation : "point");
		d->nst = nstencil; d->npre = nrelax_pre; d->npost = nrelax_post; d->W = halo_width; d->use_graph = use_graph != 0;
		if (const char *t = getenv("CDR_HALO_TIMEOUT_S")) d->timeout_s = atof(t);
		const int nl = compute_num_levels(gnx, gny);
		int ndist = 0, nx = gnx, ny = gny;
		while (ndist < nl - 1 && std::min(nx, ny) > agglomerate) { ndist++; nx = coarse_n(nx); ny = coarse_n(ny); }
		if (ndist < 1) throw std::runtime_error("grid not larger than the agglomeration size: use the serial solver");
		for (;; ndist--) {   // fewer distributed levels if the blocks get thinner than the halo
			try { d->dec = std::make_unique<decomp>(gnx, gny, px, py, ndist, halo_width); break; }
			catch (std::invalid_argument &) { if (ndist <= 1) throw; }
		}
		d->plan = std::make_unique<planner>(d->dec->ndist, halo_width, nstencil == 3 ? 2 : 4, nrelax_pre, nrelax_post,
		                                    d->relax == relax_kind::point ? d->dec->split_levels(split_min) : 0, 4, false, false,
		                                    d->relax);
		d->p_setup = d->plan->setup(); d->p_cycle = d->plan->cycle(); d->p_norm = d->plan->norm();
		if (const char *e = getenv("CDR_DIST_OVERLAP")) d->overlap = atoi(e) != 0;
		for (auto &o : d->p_cycle)
			if (o.kind == op_kind::join) {
				void *e0 = nullptr, *e1 = nullptr;
				check(bmgx_event_create(&e0), "event"); check(bmgx_event_create(&e1), "event");
				d->fork_ev.push_back(e0); d->join_ev.push_back(e1);
			}
		const int nb = d->dec->nblocks();
		d->sync_words = nb + 3;
		d->base.assign(nb, nullptr); d->opened.assign(nb, false);
		check(bmgx_stream_create(&d->stream), "stream");
		check(bmgx_stream_create(&d->side), "stream");
		for (int i = 0; i < nowned; i++) {
			dist_block K;
			K.id = owned[i];
			if (K.id < 0 || K.id >= nb) throw std::runtime_error("block id out of range");
			for (int l = 0; l <= d->dec->ndist; l++) K.g.push_back(d->dec->geom(l, K.id));
			const size_t bytes = d->arena_doubles(K.id) * 8;
			check(bmgx_malloc((void **)&K.arena, bytes), "arena");
			check(bmgx_memset0(K.arena, bytes, d->stream), "arena clear");
			d->base[K.id] = K.arena;
			const int ndl = d->dec->ndist;
			const size_t gn = (size_t)(d->dec->gn[ndl][0] + 2) * (d->dec->gn[ndl][1] + 2);
			check(bmgx_malloc((void **)&K.GA, gn * 5 * 8), "GA");
			check(bmgx_memset0(K.GA, gn * 5 * 8, d->stream), "clear");
			size_t wmax = 0;
			for (int l = 0; l < ndl; l++) {
				int ox, oy, iic, jjc;
				d->dec->coarse_window(l, K.id, ox, oy, iic, jjc);
				wmax = std::max(wmax, (size_t)iic * jjc);
			}
			check(bmgx_malloc((void **)&K.tmpP, wmax * 8 * 8), "tmpP"); check(bmgx_malloc((void **)&K.tmpA, wmax * 5 * 8), "tmpA");
			check(bmgx_malloc((void **)&K.work, (bmgx_sumsq_win_work_len() + nb + 1) * 8), "work");
			K.red_all = K.work + bmgx_sumsq_win_work_len(); K.norm_out = K.red_all + nb;
			d->blk.push_back(std::move(K));
		}
		check(bmgx_stream_sync(d->stream), "sync");
		*out = d.release();
	});
}

void cdr2_dist_destroy(cdr2_dist *d)
{
	if (!d) return;
	bmgx_device_sync();
	if (d->graph) bmgx_graph_destroy(d->graph);
	for (size_t i = 0; i < d->base.size(); i++) if (d->opened[i] && d->base[i]) bmgx_ipc_close(d->base[i]);
	for (auto &K : d->blk) {
		for (auto &kv : K.segs) if (kv.second.dev) bmgx_free(kv.second.dev);
		if (K.csol) cdr2_solver_destroy(K.csol);
		if (K.cop) cdr2_operator_destroy(K.cop);
		for (auto &kv : K.lines) { if (kv.second.sor) bmgx_free(kv.second.sor); if (kv.second.strip) bmgx_free(kv.second.strip); }
		bmgx_free(K.GA); bmgx_free(K.tmpP); bmgx_free(K.tmpA); bmgx_free(K.work);
		bmgx_free(K.arena);
	}
	for (int q = 0; q < 2; q++) { if (d->db[q]) bmgx_free(d->db[q]); if (d->dx[q]) bmgx_free(d->dx[q]); }
	for (auto e : d->ev) if (e) bmgx_event_destroy(e);
	if (d->s_up) bmgx_stream_destroy(d->s_up);
	if (d->s_down) bmgx_stream_destroy(d->s_down);
	if (d->peer_sync_dev) bmgx_free(d->peer_sync_dev);
	for (auto e : d->fork_ev) bmgx_event_destroy(e);
	for (auto e : d->join_ev) bmgx_event_destroy(e);
	if (d->side) bmgx_stream_destroy(d->side);
	if (d->stream) bmgx_stream_destroy(d->stream);
	delete d;
}

int cdr2_dist_info(cdr2_dist *d, int *ndist, int *nlevels, int *halo_width, int *exchanges_per_cycle, int *ops_per_cycle)
{
	return guard([&] {
		if (ndist) *ndist = d->dec->ndist;
		if (nlevels) *nlevels = d->dec->nlevels;
		if (halo_width) *halo_width = d->W;
		if (exchanges_per_cycle) *exchanges_per_cycle = d->plan->exchanges(d->p_cycle);
		if (ops_per_cycle) *ops_per_cycle = (int)d->p_cycle.size();
	});
}

int cdr2_dist_block_extent(cdr2_dist *d, int lvl, int block, int *x0, int *y0, int *nx, int *ny)
{
	return guard([&] {
		if (lvl < 0 || lvl > d->dec->ndist || block < 0 || block >= d->nblocks()) throw std::runtime_error("level or block out of range");
		const block_geom g = d->dec->geom(lvl, block);
		*x0 = g.x0; *y0 = g.y0; *nx = g.nx; *ny = g.ny;
	});
}

int cdr2_dist_export_handles(cdr2_dist *d, void *handles)
{
	return guard([&] {
		char *h = static_cast<char *>(handles);
		for (auto &K : d->blk) check(bmgx_ipc_export(K.arena, h + (size_t)K.id * BMGX_IPC_HANDLE_BYTES), "ipc export");
	});
}

int cdr2_dist_connect(cdr2_dist *d, const void *handles)
{
	return guard([&] {
		const char *h = static_cast<const char *>(handles);
		const int nb = d->nblocks();
		for (int q = 0; q < nb; q++) {
			if (d->base[q]) continue;
			if (!h) throw std::runtime_error("connect: blocks of other processes need their IPC handles");
			void *p = nullptr;
			check(bmgx_ipc_open(h + (size_t)q * BMGX_IPC_HANDLE_BYTES, &p), "ipc open");
			d->base[q] = static_cast<char *>(p); d->opened[q] = true;
		}
		check(bmgx_malloc((void **)&d->peer_sync_dev, nb * sizeof(void *)), "peer table");
		check(bmgx_memcpy_h2d(d->peer_sync_dev, d->base.data(), nb * sizeof(void *), d->stream), "peer table upload");
		check(bmgx_stream_sync(d->stream), "sync");
		d->connected = true;
	});
}

/* own entries of the fine operator of one block: host array (nstencil, ny+2, nx+2), ghost ring ignored */
int cdr2_dist_set_operator(cdr2_dist *d, int block, const double *so)
{
	return guard([&] {
		dist_block &K = find(*d, block);
		const block_geom g = K.g[0];
		const size_t hp = g.nx + 2, hpl = hp * (g.ny + 2);
		for (int p = 0; p < d->nst; p++)
			check(bmgx_memcpy2d_h2d(d->ptr(block, 0, A, p) + (size_t)g.wb * g.ii + g.wl, g.ii, so + p * hpl + hp + 1, hp, g.nx, g.ny,
			                        d->stream), "operator upload");
		check(bmgx_stream_sync(d->stream), "sync");
	});
}

int cdr2_dist_setup(cdr2_dist *d)
{
	return guard([&] {
		if (!d->connected) throw std::runtime_error("cdr2_dist_connect has not run");
		run(*d, d->p_setup);
		check(bmgx_stream_sync(d->stream), "sync");
		check_device_error(*d);
		for (auto &K : d->blk) {
			bmgx_free(K.tmpP); bmgx_free(K.tmpA); bmgx_free(K.GA); K.tmpP = K.tmpA = K.GA = nullptr;
			for (auto &kv : K.lines) if (kv.second.strip) { bmgx_free(kv.second.strip); kv.second.strip = nullptr; }
		}
		d->is_setup = true;
	});
}

static void move_gf(cdr2_dist *d, int block, int slot, double *host, bool to_device)
{
	dist_block &K = find(*d, block);
	const block_geom g = K.g[0];
	const size_t hp = g.nx + 2;
	double *dev = d->ptr(block, 0, slot) + (size_t)g.wb * g.ii + g.wl;
	if (to_device) check(bmgx_memcpy2d_h2d(dev, g.ii, host + hp + 1, hp, g.nx, g.ny, d->stream), "upload");
	else check(bmgx_memcpy2d_d2h(host + hp + 1, hp, dev, g.ii, g.nx, g.ny, d->stream), "download");
	check(bmgx_stream_sync(d->stream), "sync");
}

/* grid functions of a block: host arrays (ny+2, nx+2); only the own points travel.  which: "x" or "b" */
int cdr2_dist_upload(cdr2_dist *d, int block, const char *which, const double *host)
{
	return guard([&] { move_gf(d, block, which[0] == 'b' ? B : X0, const_cast<double *>(host), true); });
}
int cdr2_dist_download(cdr2_dist *d, int block, const char *which, double *host)
{
	return guard([&] { move_gf(d, block, which[0] == 'b' ? B : which[0] == 'r' ? RES : X0, host, false); });
}

/* own points of a level array -> host (planes, ny, nx); which: "A", "P", "SOR", "x", "b" */
int cdr2_dist_level_array(cdr2_dist *d, int block, int lvl, const char *which, double *out)
{
	return guard([&] {
		dist_block &K = find(*d, block);
		if (lvl < 0 || lvl > d->dec->ndist) throw std::runtime_error("level out of range");
		const std::string w(which);
		const int a = w == "A" ? A : w == "P" ? P : w == "SOR" ? SOR : w == "b" ? B : w == "res" ? RES : X0;
		if (a == P && lvl == 0) throw std::runtime_error("level 0 has no interpolation operator");
		const block_geom g = K.g[lvl];
		for (int p = 0; p < d->planes(lvl, a); p++)
			check(bmgx_memcpy2d_d2h(out + (size_t)p * g.nx * g.ny, g.nx, d->ptr(block, lvl, a, p) + (size_t)g.wb * g.ii + g.wl, g.ii,
			                        g.nx, g.ny, d->stream), "level array download");
		check(bmgx_stream_sync(d->stream), "sync");
	});
}

/* own points of a grid function of any distributed level <- host (nyl, nxl); which = "x" | "b" | "res"; planes of the
 * operator: which = "A" with host (planes, nyl, nxl) */
int cdr2_dist_upload_level(cdr2_dist *d, int block, int lvl, const char *which, const double *host)
{
	return guard([&] {
		dist_block &K = find(*d, block);
		if (lvl < 0 || lvl > d->dec->ndist) throw std::runtime_error("level out of range");
		const std::string w(which);
		const int a = w == "A" ? A : w == "b" ? B : w == "res" ? RES : X0;
		const block_geom g = K.g[lvl];
		for (int p = 0; p < d->planes(lvl, a); p++)
			check(bmgx_memcpy2d_h2d(d->ptr(block, lvl, a, p) + (size_t)g.wb * g.ii + g.wl, g.ii, host + (size_t)p * g.nx * g.ny, g.nx,
			                        g.nx, g.ny, d->stream), "level array upload");
		check(bmgx_stream_sync(d->stream), "sync");
	});
}

/* the whole halo-extended array of a level -> host (planes, jj, ii) with the extents of cdr2_dist_geometry */
int cdr2_dist_extended_array(cdr2_dist *d, int block, int lvl, const char *which, double *out)
{
	return guard([&] {
		dist_block &K = find(*d, block);
		if (lvl < 0 || lvl > d->dec->ndist) throw std::runtime_error("level out of range");
		const std::string w(which);
		const int a = w == "A" ? A : w == "b" ? B : w == "res" ? RES : X0;
		const block_geom g = K.g[lvl];
		check(bmgx_memcpy_d2h(out, d->ptr(block, lvl, a), (size_t)d->planes(lvl, a) * g.ii * g.jj * 8, d->stream), "extended array");
		check(bmgx_stream_sync(d->stream), "sync");
	});
}

/* device pointer of a level array of an owned block (what the reference's kernels hand to halof->exchange_func) */
double *cdr2_dist_level_ptr(cdr2_dist *d, int block, int lvl, const char *which)
{
	double *p = nullptr;
	guard([&] {
		find(*d, block);
		if (lvl < 0 || lvl > d->dec->ndist) throw std::runtime_error("level out of range");
		const std::string w(which);
		p = d->ptr(block, lvl, w == "A" ? A : w == "b" ? B : w == "res" ? RES : X0);
	});
	return p;
}

/* the solver's ghost exchanger as the reference's halo_exchanger_base (C++ callers: static_cast to
 * cedar::halo_exchanger_base * and hand it to kernel_manager::add_halo) */
void *cdr2_dist_halo_exchanger(cdr2_dist *d)
{
	if (!d->exchanger) d->exchanger = std::make_unique<peer_exchanger>(d);
	return d->exchanger.get();
}

/* one ghost exchange through that interface: exchange_func(k, ptr) (stencil == 0) or exchange_sten(k, ptr); collective */
int cdr2_dist_exchange(cdr2_dist *d, int k, int stencil, double *ptr)
{
	return guard([&] {
		auto *h = static_cast<cedar::halo_exchanger_base *>(cdr2_dist_halo_exchanger(d));
		if (stencil) h->exchange_sten(k, ptr); else h->exchange_func(k, ptr);
	});
}

int cdr2_dist_vcycle(cdr2_dist *d)
{
	return guard([&] { cycle(*d); });
}

int cdr2_dist_sync(cdr2_dist *d)
{
	return guard([&] { check(bmgx_stream_sync(d->stream), "sync"); check_device_error(*d); });
}

static double global_norm(cdr2_dist *d)
{
	run(*d, d->p_norm);
	double v = 0;
	check(bmgx_memcpy_d2h(&v, d->blk[0].norm_out, 8, d->stream), "norm download");
	check(bmgx_stream_sync(d->stream), "sync");
	return std::sqrt(v);
}

/* multilevel::solve (include/cedar/multilevel.h:278-298) on the resident x, b: relative residual after each cycle */
int cdr2_dist_solve(cdr2_dist *d, int maxiter, double tol, double *hist, int hist_cap, int *niter, double *res0)
{
	return guard([&] {
		const double r0 = global_norm(d);
		if (res0) *res0 = r0;
		int it = 0;
		for (; it < maxiter;) {
			cycle(*d);
			const double r = global_norm(d) / r0;
			if (hist && it < hist_cap) hist[it] = r;
			it++;
			if (r < tol) break;
		}
		if (niter) *niter = it;
		check_device_error(*d);
	});
}

/* device-resident timing: `warmup` + `n` cycles from x = 0 on the resident b; mean device time per cycle (CUDA events
 * on the launching stream) */
int cdr2_dist_time_vcycles(cdr2_dist *d, int warmup, int n, float *ms_per_cycle)
{
	return guard([&] {
		void *e0, *e1;
		check(bmgx_event_create(&e0), "event"); check(bmgx_event_create(&e1), "event");
		float total = 0;
		for (int i = 0; i < warmup + n; i++) {
			for (auto &K : d->blk) check(bmgx_memset0(d->ptr(K.id, 0, X0), (size_t)K.g[0].ii * K.g[0].jj * 8, d->stream), "x0");
			check(bmgx_event_record(e0, d->stream), "event record");
			cycle(*d);
			check(bmgx_event_record(e1, d->stream), "event record");
			check(bmgx_event_sync(e1), "event sync");
			float ms = 0;
			check(bmgx_event_elapsed_ms(e0, e1, &ms), "event elapsed");
			if (i >= warmup) total += ms;
		}
		bmgx_event_destroy(e0); bmgx_event_destroy(e1);
		*ms_per_cycle = total / (n > 0 ? n : 1);
		check_device_error(*d);
	});
}

/* A stream of right-hand sides for this process's (single) block: for k < nrhs one cycle from x = 0 on b[k], result in
 * x[k]; host arrays (ny+2, nx+2), pinned for overlap.  Upload of b[k+1], cycle k and download of x[k-1] overlap.
 * Every rank calls it with the same nrhs. */
int cdr2_dist_vcycle_stream(cdr2_dist *d, int nrhs, const double *const *bh, double *const *xh)
{
	return guard([&] {
		if (d->blk.size() != 1) throw std::runtime_error("vcycle_stream: one block per process");
		dist_block &K = d->blk[0];
		const block_geom g = K.g[0];
		const size_t n = (size_t)g.nx * g.ny, hp = g.nx + 2;
		if (!d->db[0]) {
			for (int q = 0; q < 2; q++) { check(bmgx_malloc((void **)&d->db[q], n * 8), "staging"); check(bmgx_malloc((void **)&d->dx[q], n * 8), "staging"); }
			for (auto &e : d->ev) check(bmgx_event_create(&e), "event");
			check(bmgx_stream_create(&d->s_up), "stream"); check(bmgx_stream_create(&d->s_down), "stream");
		}
		void **b_ready = d->ev, **b_free = d->ev + 2, **x_ready = d->ev + 4, **x_free = d->ev + 6;
		bmgx_stream_t cs = d->stream;
		double *bo = d->ptr(K.id, 0, B) + (size_t)g.wb * g.ii + g.wl, *xo = d->ptr(K.id, 0, X0) + (size_t)g.wb * g.ii + g.wl;
		for (int k = 0; k < nrhs; k++) {
			const int q = k & 1;
			if (k >= 2) check(bmgx_stream_wait_event(d->s_up, b_free[q]), "wait");
			check(bmgx_memcpy2d_h2d(d->db[q], g.nx, bh[k] + hp + 1, hp, g.nx, g.ny, d->s_up), "upload");
			check(bmgx_event_record(b_ready[q], d->s_up), "record");
			check(bmgx_stream_wait_event(cs, b_ready[q]), "wait");
			check(bmgx_copy2d(bo, g.ii, d->db[q], g.nx, g.nx, g.ny, cs), "b in");
			check(bmgx_event_record(b_free[q], cs), "record");
			check(bmgx_memset0(d->ptr(K.id, 0, X0), (size_t)g.ii * g.jj * 8, cs), "x0");
			cycle(*d);
			if (k >= 2) check(bmgx_stream_wait_event(cs, x_free[q]), "wait");
			check(bmgx_copy2d(d->dx[q], g.nx, xo, g.ii, g.nx, g.ny, cs), "x out");
			check(bmgx_event_record(x_ready[q], cs), "record");
			check(bmgx_stream_wait_event(d->s_down, x_ready[q]), "wait");
			check(bmgx_memcpy2d_d2h(xh[k] + hp + 1, hp, d->dx[q], g.nx, g.nx, g.ny, d->s_down), "download");
			check(bmgx_event_record(x_free[q], d->s_down), "record");
		}
		check(bmgx_stream_sync(d->s_down), "sync");
		check(bmgx_stream_sync(cs), "sync");
		check_device_error(*d);
	});
}

/* per-operation device times of one cycle (eager, one host synchronisation per operation, CUDA events around each):
 * out[k] = ms of operation k of the cycle list (cdr2_dist_plan_json "cycle"), averaged over n cycles.  Exchange
 * operations include the wait for the neighbours.  A diagnostic: the sum exceeds the graph-replayed cycle. */
int cdr2_dist_profile_ops(cdr2_dist *d, int n, float *out, int cap, int *nops)
{
	return guard([&] {
		if (!d->is_setup) throw std::runtime_error("cdr2_dist_setup has not run");
		const int m = (int)d->p_cycle.size();
		if (nops) *nops = m;
		if (cap < m) throw std::runtime_error("profile buffer too small");
		void *e0, *e1;
		check(bmgx_event_create(&e0), "event"); check(bmgx_event_create(&e1), "event");
		for (int k = 0; k < m; k++) out[k] = 0;
		const bool was = d->overlap;
		d->overlap = false;      // one operation at a time
		for (int it = 0; it < n + 1; it++) {
			for (auto &K : d->blk) check(bmgx_memset0(d->ptr(K.id, 0, X0), (size_t)K.g[0].ii * K.g[0].jj * 8, d->stream), "x0");
			for (int k = 0; k < m; k++) {
				check(bmgx_stream_sync(d->stream), "sync");
				check(bmgx_event_record(e0, d->stream), "record");
				run_op(*d, d->p_cycle[k]);
				check(bmgx_event_record(e1, d->stream), "record");
				check(bmgx_event_sync(e1), "sync");
				float ms = 0;
				check(bmgx_event_elapsed_ms(e0, e1, &ms), "elapsed");
				if (it > 0) out[k] += ms / n;
			}
		}
		bmgx_event_destroy(e0); bmgx_event_destroy(e1);
		d->overlap = was;
		check_device_error(*d);
	});
}

/* the operation lists as JSON (host only, no device needed): what = "setup", "cycle" or "norm" */
int cdr2_dist_plan_json(unsigned gnx, unsigned gny, int px, int py, int nstencil, int nrelax_pre, int nrelax_post,
                        const char *relaxation, int agglomerate, int halo_width, int split_min, const char *what, char *out,
                        size_t cap, int *ndist_out)
{
	return guard([&] {
		const int nl = compute_num_levels(gnx, gny);
		int ndist = 0, nx = gnx, ny = gny;
		while (ndist < nl - 1 && std::min(nx, ny) > agglomerate) { ndist++; nx = coarse_n(nx); ny = coarse_n(ny); }
		if (ndist < 1) throw std::runtime_error("grid not larger than the agglomeration size");
		std::unique_ptr<decomp> dec;
		for (;; ndist--) {
			try { dec = std::make_unique<decomp>(gnx, gny, px, py, ndist, halo_width); break; }
			catch (std::invalid_argument &) { if (ndist <= 1) throw; }
		}
		const relax_kind rk = relax_from_name(relaxation ? relaxation : "point");
		planner pl(dec->ndist, halo_width, nstencil == 3 ? 2 : 4, nrelax_pre, nrelax_post,
		           rk == relax_kind::point ? dec->split_levels(split_min) : 0, 4, false, false, rk);
		const std::string w(what);
		const std::string js = planner::to_json(w == "setup" ? pl.setup() : w == "norm" ? pl.norm() : pl.cycle());
		if (js.size() + 1 > cap) throw std::runtime_error("plan buffer too small");
		std::memcpy(out, js.c_str(), js.size() + 1);
		if (ndist_out) *ndist_out = dec->ndist;
	});
}

/* block geometry without a device: extents of a block's halo-extended array on a level and its coarse window */
int cdr2_dist_geometry(unsigned gnx, unsigned gny, int px, int py, int ndist, int halo_width, int lvl, int block, int *out12)
{
	return guard([&] {
		decomp dec(gnx, gny, px, py, ndist, halo_width);
		const block_geom g = dec.geom(lvl, block);
		int v[12] = {g.x0, g.y0, g.nx, g.ny, g.wl, g.wr, g.wb, g.wt, 0, 0, 0, 0};
		if (lvl < ndist) dec.coarse_window(lvl, block, v[8], v[9], v[10], v[11]);
		std::memcpy(out12, v, sizeof(v));
	});
}

}

extern "C" int cdr2_dist_sweep_rects(unsigned gnx, unsigned gny, int px, int py, int ndist, int halo_width, int lvl, int block,
                                     int ncolours, int part, int *out8, int cap, int *nrects)
{
	return guard([&] {
		decomp dec(gnx, gny, px, py, ndist, halo_width);
		const std::vector<sweep_rect> rs = sweep_rects(dec.geom(lvl, block), ncolours, part);
		if ((int)rs.size() > cap) throw std::runtime_error("rectangle buffer too small");
		for (std::size_t k = 0; k < rs.size(); k++) {
			const int v[8] = {rs[k].i0, rs[k].j0, rs[k].ni, rs[k].nj, rs[k].wi0, rs[k].wi1, rs[k].wj0, rs[k].wj1};
			std::memcpy(out8 + 8 * k, v, sizeof(v));
		}
		*nrects = (int)rs.size();
	});
}
