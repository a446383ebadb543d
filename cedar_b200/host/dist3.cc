// dist3.cc -- the slab-decomposed (multi-GPU) 3D BoxMG solver: CUDA executor of the operation lists of
// include/cedar/2d/mpi/plan.h on the slabs of include/cedar/3d/mpi/plan.h, and its C API (cdr3_dist_*,
// include/cedar_capi.h).  Counterpart of the reference's cdr3::mpi::solver (include/cedar/3d/mpi/solver.h) with the
// MPI halo exchange (src/3d/mpi/halo.cc; per colour, src/3d/ftn/mpi/BMG3_SymStd_relax_GS.f90:145,190) replaced by
// pulls of whole planes out of the two neighbours' device memory (cedar_b200/csrc/bmgx2_halo.cu).  Same structure as
// dist2.cc: one process per GPU (or several slabs in one process), IPC handles moved by the caller, levels of at most
// `agglomerate` planes solved replicated by the serial cdr3::solver.
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
#include <cedar/3d/mpi/plan.h>

using namespace cedar::cdr3::mpi;
using cedar::cdr2::mpi::planner;
using cedar::cdr2::mpi::plan_op;
using cedar::cdr2::mpi::op_kind;
using cedar::cdr2::mpi::op_name;
using cedar::gpu::check;
namespace a2 = cedar::cdr2::mpi;   // array slots X0, X1, B, RES, A, SOR, P

void cdr_set_error(const std::string &msg);

namespace {

constexpr size_t ALIGN_D = 32;
inline size_t up(size_t n) { return (n + ALIGN_D - 1) / ALIGN_D * ALIGN_D; }
struct seg_table { bmgx_seg *dev = nullptr; int n = 0; int chunks = 1; };

struct slab
{
	int id = -1;
	char *arena = nullptr;
	std::vector<slab_geom> g;
	cdr3_operator *cop = nullptr;
	cdr3_solver *csol = nullptr;
	double *GB = nullptr, *GX = nullptr, *GA = nullptr;
	double *tmpP = nullptr, *tmpA = nullptr;
	double *work = nullptr, *red_all = nullptr, *norm_out = nullptr;
	std::map<std::string, seg_table> segs;
};

}  // namespace

struct cdr3_dist
{
	std::unique_ptr<slab_decomp> dec;
	std::unique_ptr<planner> plan;
	int nst = 4, npre = 1, npost = 1, W = 5;
	bool use_graph = true;
	double timeout_s = 10.0;
	std::vector<slab> blk;
	std::vector<char *> base;
	std::vector<bool> opened;
	void **peer_sync_dev = nullptr;
	bmgx_stream_t stream = nullptr;
	std::vector<plan_op> p_setup, p_cycle, p_norm;
	void *graph = nullptr;
	unsigned long long graph_nodes = 0;
	int warm = 0;
	bool connected = false, is_setup = false;
	size_t sync_words = 0;
	std::unique_ptr<cedar::halo_exchanger_base> exchanger;   // the reference's seam over the peer-memory exchange

	int nblocks() const { return dec->np; }
	int nstl(int lvl) const { return lvl == 0 ? nst : 14; }
	int planes(int lvl, int a) const { return a == a2::A ? nstl(lvl) : a == a2::SOR ? 2 : a == a2::P ? 26 : 1; }
	size_t off_red() const { return up(sync_words); }
	size_t vol(int b, int lvl) const { const slab_geom g = dec->geom(lvl, b); return (size_t)g.ii * g.jj * g.kk; }
	size_t off_arr(int b, int lvl, int a) const
	{
		size_t o = off_red() + ALIGN_D;
		for (int l = 0; l <= dec->ndist; l++) {
			const size_t v = vol(b, l);
			for (int s = 0; s < a2::NARR; s++) {
				if (l == lvl && s == a) return o;
				if (s == a2::P && l == 0) continue;
				o += up((size_t)planes(l, s) * v);
			}
		}
		return o;
	}
	size_t arena_doubles(int b) const { return off_arr(b, dec->ndist + 1, 0); }
	double *ptr(int b, int lvl, int a, int entry = 0) const
	{
		return reinterpret_cast<double *>(base[b]) + off_arr(b, lvl, a) + (size_t)entry * vol(b, lvl);
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
	std::string k = std::string(op_name(o.kind)) + ":" + std::to_string(o.lvl) + ":" + std::to_string(o.a);
	for (int a : o.arrays) k += "," + std::to_string(a);
	return k;
}

void upload_segs(seg_table &t, const std::vector<bmgx_seg> &v, bmgx_stream_t s)
{
	t.n = (int)v.size();
	size_t big = 0;
	for (auto &sg : v) big = std::max(big, (size_t)sg.rows * sg.cols);
	t.chunks = (int)std::min<size_t>(64, std::max<size_t>(1, (big + 1023) / 1024));
	if (t.n == 0) return;
	check(bmgx_malloc((void **)&t.dev, v.size() * sizeof(bmgx_seg)), "segment table");
	check(bmgx_memcpy_h2d(t.dev, v.data(), v.size() * sizeof(bmgx_seg), s), "segment table upload");
	check(bmgx_stream_sync(s), "segment table sync");
}

// a run of whole planes is one contiguous rectangle: rows = planes, cols = ii * jj
bmgx_seg plane_seg(const double *src, double *dst, int nplanes, size_t pl, int rank)
{
	bmgx_seg s{};
	s.src = src; s.dst = dst; s.rows = nplanes; s.cols = (uint32_t)pl; s.src_pitch = pl; s.dst_pitch = pl; s.src_rank = rank;
	return s;
}

seg_table &table(cdr3_dist &d, slab &K, const plan_op &o)
{
	const std::string k = key_of(o);
	auto it = K.segs.find(k);
	if (it != K.segs.end()) return it->second;
	std::vector<bmgx_seg> v;
	const slab_decomp &dec = *d.dec;
	const int nd = dec.ndist, W = dec.W, b = K.id;
	if (o.kind == op_kind::exchange) {
		const slab_geom g = dec.geom(o.lvl, b);
		const size_t pl = (size_t)g.ii * g.jj;
		for (int dz = -1; dz <= 1; dz += 2) {
			const int nb = b + dz;
			if (nb < 0 || nb >= dec.np) continue;
			const slab_geom h = dec.geom(o.lvl, nb);
			if (h.nz < W) throw std::runtime_error("neighbour slab thinner than the halo");
			const int dp = dz < 0 ? 0 : g.wb + g.nz, sp = dz < 0 ? h.wb + h.nz - W : h.wb;
			for (int a : o.arrays)
				for (int e = 0; e < d.planes(o.lvl, a); e++)
					v.push_back(plane_seg(d.ptr(nb, o.lvl, a, e) + (size_t)sp * pl, d.ptr(b, o.lvl, a, e) + (size_t)dp * pl, W, pl, nb));
		}
	} else if (o.kind == op_kind::gather || o.kind == op_kind::gather_op) {
		const int a = o.kind == op_kind::gather ? o.a : a2::A;
		double *global = o.kind == op_kind::gather ? K.GB : K.GA;
		const size_t pl = (size_t)(dec.gn[nd][0] + 2) * (dec.gn[nd][1] + 2), gvol = pl * (dec.gn[nd][2] + 2);
		for (int q = 0; q < dec.np; q++) {
			const slab_geom h = dec.geom(nd, q);
			for (int e = 0; e < d.planes(nd, a); e++)
				v.push_back(plane_seg(d.ptr(q, nd, a, e) + (size_t)h.wb * pl, global + (size_t)e * gvol + (size_t)(1 + h.z0) * pl, h.nz, pl, q));
		}
	} else if (o.kind == op_kind::norm) {
		for (int q = 0; q < dec.np; q++) {
			bmgx_seg s{};
			s.src = reinterpret_cast<double *>(d.base[q]) + d.off_red();
			s.dst = K.red_all + q;
			s.rows = s.cols = 1; s.src_pitch = s.dst_pitch = 1; s.src_rank = q;
			v.push_back(s);
		}
	}
	seg_table t;
	upload_segs(t, v, d.stream);
	return K.segs.emplace(k, t).first->second;
}

void step(cdr3_dist &d, const plan_op &o)
{
	const int world = d.nblocks();
	const bool local = d.blk.size() > 1;
	if (local)
		for (auto &K : d.blk)
			check(bmgx_halo_step(nullptr, 0, 1, d.peer_sync_dev, K.arena, K.id, world, 1, 0, d.timeout_s, d.stream), "halo signal");
	for (auto &K : d.blk) {
		seg_table &t = table(d, K, o);
		check(bmgx_halo_step(t.dev, t.n, t.chunks, d.peer_sync_dev, K.arena, K.id, world, local ? 2 : 3, o.wait_all ? 1 : 0,
		                     d.timeout_s, d.stream), "halo step");
	}
}

// halo_exchanger_base over the slab exchange: see peer_exchanger in dist2.cc (k = distributed levels + 1 is the finest grid)
struct slab_exchanger : cedar::halo_exchanger_base
{
	cdr3_dist *d;
	explicit slab_exchanger(cdr3_dist *d) : d(d) {}
	void exchange(int k, const double *p, bool sten)
	{
		const int lvl = d->dec->ndist + 1 - k;
		if (lvl < 0 || lvl > d->dec->ndist) throw std::runtime_error("halo exchange: level out of range");
		if (!d->connected) throw std::runtime_error("halo exchange: cdr3_dist_connect has not run");
		for (auto &K : d->blk)
			for (int a : {(int)a2::X0, (int)a2::X1, (int)a2::B, (int)a2::RES, (int)a2::A}) {
				if ((a == a2::A) != sten || d->ptr(K.id, lvl, a) != p) continue;
				plan_op o; o.kind = op_kind::exchange; o.lvl = lvl; o.arrays = {a};
				step(*d, o);
				return;
			}
		throw std::runtime_error("halo exchange: the pointer is not a level array of a slab owned by this process");
	}
	void exchange_func(int k, cedar::real_t *gf) override { exchange(k, gf, false); }
	void exchange_sten(int k, cedar::real_t *so) override { exchange(k, so, true); }
};

void run_op(cdr3_dist &d, const plan_op &o)
{
	const slab_decomp &dec = *d.dec;
	const int l = o.lvl, nd = dec.ndist;
	bmgx_stream_t s = d.stream;
	switch (o.kind) {
	case op_kind::join: return;
	case op_kind::exchange: case op_kind::sync: case op_kind::gather: case op_kind::gather_op: case op_kind::norm:
		step(d, o);
		if (o.kind == op_kind::norm)
			for (auto &K : d.blk) check(bmgx_sum_ordered(K.red_all, d.nblocks(), K.norm_out, s), "norm fold");
		return;
	default: break;
	}
	for (auto &K : d.blk) {
		const int b = K.id;
		const slab_geom g = K.g[l];
		const int ifd = (l == 0 && d.nst == 4) ? 1 : 0, nst = d.nstl(l);
		const size_t pl = (size_t)g.ii * g.jj;
		int oz = 0, kkc = 0, iic = 0, jjc = 0;
		size_t cstride = 0, coff = 0;
		if (l < nd) {
			dec.coarse_window(l, b, oz, kkc);
			const slab_geom c = K.g[l + 1];
			iic = c.ii; jjc = c.jj;
			cstride = d.vol(b, l + 1); coff = (size_t)oz * iic * jjc;
		}
		switch (o.kind) {
		case op_kind::sweep:
			if (o.a != o.b)
				check(bmgx3_relax_GS_oop(1, d.ptr(b, l, a2::A), d.ptr(b, l, a2::B), d.ptr(b, l, o.a), d.ptr(b, l, o.b), d.ptr(b, l, a2::SOR),
				                         g.ii, g.jj, g.kk, ifd, nst, 2, BMGX_RELAX_SYM, o.dir ? BMGX_UP : BMGX_DOWN, 0, s), "relax_GS3_oop");
			else
				check(bmgx3_relax_GS(1, d.ptr(b, l, a2::A), d.ptr(b, l, a2::B), d.ptr(b, l, o.a), d.ptr(b, l, a2::SOR), g.ii, g.jj, g.kk,
				                     ifd, nst, 2, BMGX_RELAX_SYM, o.dir ? BMGX_UP : BMGX_DOWN, 0, s), "relax_GS3");
			break;
		case op_kind::residual:
			check(bmgx3_residual(1, 1, ifd, d.ptr(b, l, o.a), d.ptr(b, l, a2::B), d.ptr(b, l, a2::A), d.ptr(b, l, a2::RES), g.ii, g.jj,
			                     g.kk, nst, s), "residual3");
			break;
		case op_kind::restrict_:
			check(bmgx3_restrict_v(d.ptr(b, l, a2::RES), d.ptr(b, l + 1, a2::B) + coff, d.ptr(b, l + 1, a2::P) + coff, g.ii, g.jj, g.kk,
			                       iic, jjc, kkc, cstride, s), "restrict3");
			break;
		case op_kind::interp:
			check(bmgx3_interp_add_v(d.ptr(b, l, o.a), d.ptr(b, l + 1, o.b) + coff, d.ptr(b, l, a2::A), d.ptr(b, l, a2::RES),
			                         d.ptr(b, l + 1, a2::P) + coff, iic, jjc, kkc, g.ii, g.jj, g.kk, cstride, s), "interp_add3");
			break;
		case op_kind::zero:
			check(bmgx_memset0(d.ptr(b, l, o.a), pl * g.kk * 8, s), "zero");
			break;
		case op_kind::copy:
			check(bmgx_memcpy_d2d(d.ptr(b, l, o.b), d.ptr(b, l, o.a), pl * g.kk * 8, s), "copy");
			break;
		case op_kind::coarse: {
			const size_t n = (size_t)(dec.gn[nd][0] + 2) * (dec.gn[nd][1] + 2) * (dec.gn[nd][2] + 2);
			double *cb = nullptr, *cx = nullptr;
			if (cdr3_solver_resident_ptrs(K.csol, &cb, &cx)) throw std::runtime_error(cdr_last_error());
			if (cb != K.GB) throw std::logic_error("replicated solver moved its right-hand side");
			check(bmgx_memset0(cx, n * 8, s), "coarse x0");
			if (cdr3_solver_vcycle_resident(K.csol, s)) throw std::runtime_error(std::string("agglomerated cycle: ") + cdr_last_error());
			if (cdr3_solver_resident_ptrs(K.csol, &cb, &K.GX)) throw std::runtime_error(cdr_last_error());
			break;
		}
		case op_kind::scatter:
			// this slab's planes of the replicated solution, halo included: array plane p is global plane z0 + 1 - wb + p
			check(bmgx_memcpy_d2d(d.ptr(b, l, o.a), K.GX + (size_t)(g.z0 + 1 - g.wb) * pl, pl * g.kk * 8, s), "scatter");
			break;
		case op_kind::sumsq:
			// own planes = interior planes of the array that starts one plane below them
			check(bmgx3_sumsq(d.ptr(b, l, o.a) + (size_t)(g.wb - 1) * pl, g.ii, g.jj, g.nz + 2, K.work,
			                  reinterpret_cast<double *>(K.arena) + d.off_red(), s), "sumsq3");
			break;
		case op_kind::recip:
			check(bmgx3_SETUP_recip(d.ptr(b, l, a2::A), d.ptr(b, l, a2::SOR), g.ii, g.jj, g.kk, nst, 2, s), "SETUP_recip3");
			break;
		case op_kind::setup_interp: {
			// on a contiguous window of planes, then entry by entry into the (thicker) level l + 1 array
			const size_t wv = (size_t)iic * jjc * kkc;
			check(bmgx_memset0(K.tmpP, 26 * wv * 8, s), "P window");
			check(bmgx3_SETUP_interp_OI(d.ptr(b, l, a2::A), nullptr, K.tmpP, g.ii, g.jj, g.kk, iic, jjc, kkc, ifd, nst, 0, 0, nullptr, s),
			      "SETUP_interp_OI3");
			for (int e = 0; e < 26; e++)
				check(bmgx_memcpy_d2d(d.ptr(b, l + 1, a2::P, e) + coff, K.tmpP + e * wv, wv * 8, s), "P window copy");
			break;
		}
		case op_kind::galerkin: {
			const size_t wv = (size_t)iic * jjc * kkc;
			check(bmgx_memset0(K.tmpA, 14 * wv * 8, s), "A window");
			if (ifd)
				check(bmgx3_SETUP_ITLI07_ex(d.ptr(b, l, a2::A), K.tmpA, K.tmpP, g.ii, g.jj, g.kk, iic, jjc, kkc, 0, s), "SETUP_ITLI07_ex");
			else
				check(bmgx3_SETUP_ITLI27_ex(d.ptr(b, l, a2::A), K.tmpA, K.tmpP, g.ii, g.jj, g.kk, iic, jjc, kkc, 0, s), "SETUP_ITLI27_ex");
			for (int e = 0; e < 14; e++)
				check(bmgx_memcpy_d2d(d.ptr(b, l + 1, a2::A, e) + coff, K.tmpA + e * wv, wv * 8, s), "A window copy");
			break;
		}
		case op_kind::setup_coarse: {
			const unsigned gx = dec.gn[nd][0], gy = dec.gn[nd][1], gz = dec.gn[nd][2];
			const size_t n = (size_t)(gx + 2) * (gy + 2) * (gz + 2) * 14;
			if (cdr3_operator_create(14, gx, gy, gz, &K.cop)) throw std::runtime_error(cdr_last_error());
			double *h = cdr3_operator_data(K.cop);
			if (!h) throw std::runtime_error(cdr_last_error());
			check(bmgx_memcpy_d2h(h, K.GA, n * 8, s), "agglomerated operator download");
			check(bmgx_stream_sync(s), "sync");
			const int nl = dec.nlevels - nd;
			const std::string cfg = "{\"log\": [\"error\"], \"solver\": {\"relaxation\": \"point\", \"cycle\": {\"type\": \"v\", "
			                        "\"nrelax-pre\": " + std::to_string(d.npre) + ", \"nrelax-post\": " + std::to_string(d.npost) +
			                        "}, \"max-iter\": 1, \"tol\": 0, \"cuda-graph\": false, \"num-levels\": " + std::to_string(nl) + "}}";
			if (cdr3_solver_create(K.cop, cfg.c_str(), &K.csol)) throw std::runtime_error(cdr_last_error());
			if (cdr3_solver_nlevels(K.csol) != nl) throw std::runtime_error("agglomerated solver: unexpected number of levels");
			if (cdr3_solver_resident_ptrs(K.csol, &K.GB, &K.GX)) throw std::runtime_error(cdr_last_error());
			break;
		}
		default: throw std::logic_error("unknown plan operation");
		}
	}
}

void run(cdr3_dist &d, const std::vector<plan_op> &p) { for (auto &o : p) run_op(d, o); }

void check_device_error(cdr3_dist &d)
{
	for (auto &K : d.blk) {
		unsigned long long e = 0;
		check(bmgx_memcpy_d2h(&e, reinterpret_cast<unsigned long long *>(K.arena) + d.nblocks() + 2, 8, d.stream), "error word");
		check(bmgx_stream_sync(d.stream), "sync");
		if (e) throw std::runtime_error("ghost exchange timed out waiting for slab " + std::to_string((long long)e - 1));
	}
}

void cycle(cdr3_dist &d)
{
	if (!d.is_setup) throw std::runtime_error("cdr3_dist_setup has not run");
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

slab &find(cdr3_dist &d, int b)
{
	for (auto &K : d.blk) if (K.id == b) return K;
	throw std::runtime_error("slab " + std::to_string(b) + " is not owned by this process");
}

std::unique_ptr<slab_decomp> make_decomp(unsigned gnx, unsigned gny, unsigned gnz, int np, int agglomerate, int W)
{
	const int nl = compute_num_levels(gnx, gny, gnz);
	int ndist = 0, nz = gnz;
	while (ndist < nl - 1 && nz > agglomerate) { ndist++; nz = coarse_n(nz); }
	if (ndist < 1) throw std::runtime_error("grid not thicker than the agglomeration size: use the serial solver");
	for (;; ndist--) {
		try { return std::make_unique<slab_decomp>(gnx, gny, gnz, np, ndist, W); }
		catch (std::invalid_argument &) { if (ndist <= 1) throw; }
	}
}

std::unique_ptr<planner> make_planner(const slab_decomp &dec, int nst, int W, int npre, int npost)
{
	// colour depth across planes is 2 for both smoothers; the 7-point sweep is out of place (fused), the 27-point
	// sweeps (every coarse level) work in place
	return std::make_unique<planner>(dec.ndist, W, 2, npre, npost, 0, 2, nst != 4, true);
}

}  // namespace

extern "C" {

int cdr3_dist_create(unsigned gnx, unsigned gny, unsigned gnz, int nslabs, int nstencil, int nowned, const int *owned,
                     int nrelax_pre, int nrelax_post, int agglomerate, int halo_width, int use_graph, cdr3_dist **out)
{
	return guard([&] {
		if (nstencil != 4 && nstencil != 14) throw std::runtime_error("nstencil must be 4 or 14");
		if (!cedar::gpu::available()) throw std::runtime_error("no CUDA device: the BoxMG kernels have no CPU fallback");
		auto d = std::make_unique<cdr3_dist>();
		d->nst = nstencil; d->npre = nrelax_pre; d->npost = nrelax_post; d->W = halo_width; d->use_graph = use_graph != 0;
		if (const char *t = getenv("CDR_HALO_TIMEOUT_S")) d->timeout_s = atof(t);
		d->dec = make_decomp(gnx, gny, gnz, nslabs, agglomerate, halo_width);
		d->plan = make_planner(*d->dec, nstencil, halo_width, nrelax_pre, nrelax_post);
		d->p_setup = d->plan->setup(); d->p_cycle = d->plan->cycle(); d->p_norm = d->plan->norm();
		const int nb = nslabs;
		d->sync_words = nb + 3;
		d->base.assign(nb, nullptr); d->opened.assign(nb, false);
		check(bmgx_stream_create(&d->stream), "stream");
		const int ndl = d->dec->ndist;
		const size_t gv = (size_t)(d->dec->gn[ndl][0] + 2) * (d->dec->gn[ndl][1] + 2) * (d->dec->gn[ndl][2] + 2);
		for (int i = 0; i < nowned; i++) {
			slab K;
			K.id = owned[i];
			if (K.id < 0 || K.id >= nb) throw std::runtime_error("slab id out of range");
			for (int l = 0; l <= ndl; l++) K.g.push_back(d->dec->geom(l, K.id));
			const size_t bytes = d->arena_doubles(K.id) * 8;
			check(bmgx_malloc((void **)&K.arena, bytes), "arena");
			check(bmgx_memset0(K.arena, bytes, d->stream), "arena clear");
			d->base[K.id] = K.arena;
			check(bmgx_malloc((void **)&K.GA, gv * 14 * 8), "GA");
			check(bmgx_memset0(K.GA, gv * 14 * 8, d->stream), "clear");
			size_t wmax = 0;
			for (int l = 0; l < ndl; l++) {
				int oz, kkc;
				d->dec->coarse_window(l, K.id, oz, kkc);
				wmax = std::max(wmax, (size_t)K.g[l + 1].ii * K.g[l + 1].jj * kkc);
			}
			check(bmgx_malloc((void **)&K.tmpP, wmax * 26 * 8), "tmpP"); check(bmgx_malloc((void **)&K.tmpA, wmax * 14 * 8), "tmpA");
			check(bmgx_malloc((void **)&K.work, (bmgx_norm_work_len() + nb + 1) * 8), "work");
			K.red_all = K.work + bmgx_norm_work_len(); K.norm_out = K.red_all + nb;
			d->blk.push_back(std::move(K));
		}
		check(bmgx_stream_sync(d->stream), "sync");
		*out = d.release();
	});
}

void cdr3_dist_destroy(cdr3_dist *d)
{
	if (!d) return;
	bmgx_device_sync();
	if (d->graph) bmgx_graph_destroy(d->graph);
	for (size_t i = 0; i < d->base.size(); i++) if (d->opened[i] && d->base[i]) bmgx_ipc_close(d->base[i]);
	for (auto &K : d->blk) {
		for (auto &kv : K.segs) if (kv.second.dev) bmgx_free(kv.second.dev);
		if (K.csol) cdr3_solver_destroy(K.csol);
		if (K.cop) cdr3_operator_destroy(K.cop);
		bmgx_free(K.GA); bmgx_free(K.tmpP); bmgx_free(K.tmpA); bmgx_free(K.work); bmgx_free(K.arena);
	}
	if (d->peer_sync_dev) bmgx_free(d->peer_sync_dev);
	if (d->stream) bmgx_stream_destroy(d->stream);
	delete d;
}

int cdr3_dist_info(cdr3_dist *d, int *ndist, int *nlevels, int *halo_width, int *exchanges_per_cycle, int *ops_per_cycle)
{
	return guard([&] {
		if (ndist) *ndist = d->dec->ndist;
		if (nlevels) *nlevels = d->dec->nlevels;
		if (halo_width) *halo_width = d->W;
		if (exchanges_per_cycle) *exchanges_per_cycle = d->plan->exchanges(d->p_cycle);
		if (ops_per_cycle) *ops_per_cycle = (int)d->p_cycle.size();
	});
}

int cdr3_dist_slab_extent(cdr3_dist *d, int lvl, int slab_id, int *nx, int *ny, int *z0, int *nz)
{
	return guard([&] {
		if (lvl < 0 || lvl > d->dec->ndist || slab_id < 0 || slab_id >= d->nblocks()) throw std::runtime_error("level or slab out of range");
		const slab_geom g = d->dec->geom(lvl, slab_id);
		*nx = g.nx; *ny = g.ny; *z0 = g.z0; *nz = g.nz;
	});
}

int cdr3_dist_export_handles(cdr3_dist *d, void *handles)
{
	return guard([&] {
		char *h = static_cast<char *>(handles);
		for (auto &K : d->blk) check(bmgx_ipc_export(K.arena, h + (size_t)K.id * BMGX_IPC_HANDLE_BYTES), "ipc export");
	});
}

int cdr3_dist_connect(cdr3_dist *d, const void *handles)
{
	return guard([&] {
		const char *h = static_cast<const char *>(handles);
		const int nb = d->nblocks();
		for (int q = 0; q < nb; q++) {
			if (d->base[q]) continue;
			if (!h) throw std::runtime_error("connect: slabs of other processes need their IPC handles");
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

/* own planes of the fine operator of one slab: host array (nstencil, nzl+2, ny+2, nx+2); ghost planes ignored */
int cdr3_dist_set_operator(cdr3_dist *d, int slab_id, const double *so)
{
	return guard([&] {
		slab &K = find(*d, slab_id);
		const slab_geom g = K.g[0];
		const size_t pl = (size_t)g.ii * g.jj, hv = pl * (g.nz + 2);
		for (int e = 0; e < d->nst; e++)
			check(bmgx_memcpy_h2d(d->ptr(slab_id, 0, a2::A, e) + (size_t)g.wb * pl, so + e * hv + pl, pl * g.nz * 8, d->stream),
			      "operator upload");
		check(bmgx_stream_sync(d->stream), "sync");
	});
}

int cdr3_dist_setup(cdr3_dist *d)
{
	return guard([&] {
		if (!d->connected) throw std::runtime_error("cdr3_dist_connect has not run");
		run(*d, d->p_setup);
		check(bmgx_stream_sync(d->stream), "sync");
		check_device_error(*d);
		for (auto &K : d->blk) { bmgx_free(K.tmpP); bmgx_free(K.tmpA); bmgx_free(K.GA); K.tmpP = K.tmpA = K.GA = nullptr; }
		d->is_setup = true;
	});
}

static void move_gf3(cdr3_dist *d, int slab_id, int slot, int lvl, double *host, bool to_device)
{
	slab &K = find(*d, slab_id);
	const slab_geom g = K.g[lvl];
	const size_t pl = (size_t)g.ii * g.jj;
	double *dev = d->ptr(slab_id, lvl, slot) + (size_t)g.wb * pl;
	if (to_device) check(bmgx_memcpy_h2d(dev, host + pl, pl * g.nz * 8, d->stream), "upload");
	else check(bmgx_memcpy_d2h(host + pl, dev, pl * g.nz * 8, d->stream), "download");
	check(bmgx_stream_sync(d->stream), "sync");
}

/* grid functions of a slab: host arrays (nzl+2, ny+2, nx+2), own planes only; which = "x" | "b" (download also "res") */
int cdr3_dist_upload(cdr3_dist *d, int slab_id, const char *which, const double *host)
{
	return guard([&] { move_gf3(d, slab_id, which[0] == 'b' ? a2::B : a2::X0, 0, const_cast<double *>(host), true); });
}
int cdr3_dist_download(cdr3_dist *d, int slab_id, const char *which, double *host)
{
	return guard([&] { move_gf3(d, slab_id, which[0] == 'b' ? a2::B : which[0] == 'r' ? a2::RES : a2::X0, 0, host, false); });
}

/* own planes of a level array -> host (entries, nzl, ny+2, nx+2); which = "A" | "P" */
int cdr3_dist_level_array(cdr3_dist *d, int slab_id, int lvl, const char *which, double *out)
{
	return guard([&] {
		slab &K = find(*d, slab_id);
		if (lvl < 0 || lvl > d->dec->ndist) throw std::runtime_error("level out of range");
		const int a = which[0] == 'A' ? a2::A : a2::P;
		if (a == a2::P && lvl == 0) throw std::runtime_error("level 0 has no interpolation operator");
		const slab_geom g = K.g[lvl];
		const size_t pl = (size_t)g.ii * g.jj;
		for (int e = 0; e < d->planes(lvl, a); e++)
			check(bmgx_memcpy_d2h(out + (size_t)e * pl * g.nz, d->ptr(slab_id, lvl, a, e) + (size_t)g.wb * pl, pl * g.nz * 8, d->stream),
			      "level array download");
		check(bmgx_stream_sync(d->stream), "sync");
	});
}

static int slot3(const char *which)
{
	const std::string w(which);
	return w == "A" ? a2::A : w == "b" ? a2::B : w == "res" ? a2::RES : a2::X0;
}
/* own planes of an array of any distributed level <- host (entries, nzl, ny+2, nx+2); which = "x" | "b" | "res" | "A" */
int cdr3_dist_upload_level(cdr3_dist *d, int slab_id, int lvl, const char *which, const double *host)
{
	return guard([&] {
		slab &K = find(*d, slab_id);
		if (lvl < 0 || lvl > d->dec->ndist) throw std::runtime_error("level out of range");
		const int a = slot3(which);
		const slab_geom g = K.g[lvl];
		const size_t pl = (size_t)g.ii * g.jj;
		for (int e = 0; e < d->planes(lvl, a); e++)
			check(bmgx_memcpy_h2d(d->ptr(slab_id, lvl, a, e) + (size_t)g.wb * pl, host + (size_t)e * pl * g.nz, pl * g.nz * 8, d->stream),
			      "level array upload");
		check(bmgx_stream_sync(d->stream), "sync");
	});
}
/* the whole halo-extended array of a level -> host (entries, kk, ny+2, nx+2) */
int cdr3_dist_extended_array(cdr3_dist *d, int slab_id, int lvl, const char *which, double *out)
{
	return guard([&] {
		find(*d, slab_id);
		if (lvl < 0 || lvl > d->dec->ndist) throw std::runtime_error("level out of range");
		const int a = slot3(which);
		check(bmgx_memcpy_d2h(out, d->ptr(slab_id, lvl, a), (size_t)d->planes(lvl, a) * d->vol(slab_id, lvl) * 8, d->stream), "extended array");
		check(bmgx_stream_sync(d->stream), "sync");
	});
}
double *cdr3_dist_level_ptr(cdr3_dist *d, int slab_id, int lvl, const char *which)
{
	double *p = nullptr;
	guard([&] {
		find(*d, slab_id);
		if (lvl < 0 || lvl > d->dec->ndist) throw std::runtime_error("level out of range");
		p = d->ptr(slab_id, lvl, slot3(which));
	});
	return p;
}
void *cdr3_dist_halo_exchanger(cdr3_dist *d)
{
	if (!d->exchanger) d->exchanger = std::make_unique<slab_exchanger>(d);
	return d->exchanger.get();
}
int cdr3_dist_exchange(cdr3_dist *d, int k, int stencil, double *ptr)
{
	return guard([&] {
		auto *h = static_cast<cedar::halo_exchanger_base *>(cdr3_dist_halo_exchanger(d));
		if (stencil) h->exchange_sten(k, ptr); else h->exchange_func(k, ptr);
	});
}

int cdr3_dist_vcycle(cdr3_dist *d) { return guard([&] { cycle(*d); }); }
int cdr3_dist_sync(cdr3_dist *d) { return guard([&] { check(bmgx_stream_sync(d->stream), "sync"); check_device_error(*d); }); }

static double global_norm3(cdr3_dist *d)
{
	run(*d, d->p_norm);
	double v = 0;
	check(bmgx_memcpy_d2h(&v, d->blk[0].norm_out, 8, d->stream), "norm download");
	check(bmgx_stream_sync(d->stream), "sync");
	return std::sqrt(v);
}

int cdr3_dist_solve(cdr3_dist *d, int maxiter, double tol, double *hist, int hist_cap, int *niter, double *res0)
{
	return guard([&] {
		const double r0 = global_norm3(d);
		if (res0) *res0 = r0;
		int it = 0;
		for (; it < maxiter;) {
			cycle(*d);
			const double r = global_norm3(d) / r0;
			if (hist && it < hist_cap) hist[it] = r;
			it++;
			if (r < tol) break;
		}
		if (niter) *niter = it;
		check_device_error(*d);
	});
}

int cdr3_dist_time_vcycles(cdr3_dist *d, int warmup, int n, float *ms_per_cycle)
{
	return guard([&] {
		void *e0, *e1;
		check(bmgx_event_create(&e0), "event"); check(bmgx_event_create(&e1), "event");
		float total = 0;
		for (int i = 0; i < warmup + n; i++) {
			for (auto &K : d->blk)
				check(bmgx_memset0(d->ptr(K.id, 0, a2::X0), (size_t)K.g[0].ii * K.g[0].jj * K.g[0].kk * 8, d->stream), "x0");
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

int cdr3_dist_plan_json(unsigned gnx, unsigned gny, unsigned gnz, int nslabs, int nstencil, int nrelax_pre, int nrelax_post,
                        int agglomerate, int halo_width, const char *what, char *out, size_t cap, int *ndist_out)
{
	return guard([&] {
		auto dec = make_decomp(gnx, gny, gnz, nslabs, agglomerate, halo_width);
		auto pl = make_planner(*dec, nstencil, halo_width, nrelax_pre, nrelax_post);
		const std::string w(what);
		const std::string js = planner::to_json(w == "setup" ? pl->setup() : w == "norm" ? pl->norm() : pl->cycle());
		if (js.size() + 1 > cap) throw std::runtime_error("plan buffer too small");
		std::memcpy(out, js.c_str(), js.size() + 1);
		if (ndist_out) *ndist_out = dec->ndist;
	});
}

/* slab geometry without a device: nx, ny, z0, nz, wb, wt, coarse window oz, kkc */
int cdr3_dist_geometry(unsigned gnx, unsigned gny, unsigned gnz, int nslabs, int ndist, int halo_width, int lvl, int slab_id, int *out8)
{
	return guard([&] {
		slab_decomp dec(gnx, gny, gnz, nslabs, ndist, halo_width);
		const slab_geom g = dec.geom(lvl, slab_id);
		int v[8] = {g.nx, g.ny, g.z0, g.nz, g.wb, g.wt, 0, 0};
		if (lvl < ndist) dec.coarse_window(lvl, slab_id, v[6], v[7]);
		std::memcpy(out8, v, sizeof(v));
	});
}

}
