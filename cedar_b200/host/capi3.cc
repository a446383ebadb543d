// capi3.cc -- 3D part of the C API over the C++ host layer (include/cedar_capi.h).
#include <cstring>
#include <memory>
#include <string>

#include <cedar_capi.h>
#include <cedar/3d/solver.h>
#include <cedar/3d/gallery.h>
#include "batch_stage.h"

using namespace cedar;
using namespace cedar::cdr3;

void cdr_set_error(const std::string &msg);   // capi.cc

struct cdr3_operator
{
	int nstencil;
	std::unique_ptr<stencil_op<seven_pt>> seven;
	std::unique_ptr<stencil_op<xxvii_pt>> full;
	unsigned nx, ny, nz;
};

struct cdr3_solver
{
	cdr3_operator *op;
	std::unique_ptr<solver<seven_pt>> s7;
	std::unique_ptr<solver<xxvii_pt>> s27;
	grid_func b, x;
	batch_stage stage;
	template <class F> auto each(F &&f) { return s7 ? f(*s7) : f(*s27); }
};

template <class F> static int guard(F &&f)
{
	try { f(); return 0; }
	catch (std::exception &e) { cdr_set_error(e.what()); return 1; }
	catch (...) { cdr_set_error("unknown error"); return 1; }
}

static std::shared_ptr<config> make_config(const char *json)
{
	if (!json || !*json) return std::make_shared<config>("");
	auto conf = std::make_shared<config>(jvalue::parse(json));
	auto lv = conf->getvec<std::string>("log");
	if (!lv.empty()) log::set_levels(lv);
	return conf;
}

template <class ARR> static void copy_out(const ARR &a, double *out, size_t n, size_t *written)
{
	size_t m = a.size();
	if (m > n) throw std::runtime_error("output buffer too small");
	if (m) std::memcpy(out, a.cdata(), m * sizeof(double));
	*written = m;
}

template <class LVL> static void level_array(LVL &L, const std::string &w, double *out, size_t n, size_t *written)
{
	if (w == "A") copy_out(L.A, out, n, written);
	else if (w == "P") copy_out(L.P, out, n, written);
	else if (w == "x") copy_out(L.x, out, n, written);
	else if (w == "res") copy_out(L.res, out, n, written);
	else if (w == "b") copy_out(L.b, out, n, written);
	else if (w == "SOR0") copy_out(L.SOR[0], out, n, written);
	else throw std::runtime_error("unknown level array " + w);
}

static void upload(grid_func &g, const double *src) { std::memcpy(g.data(), src, g.size() * sizeof(double)); }

template <class F> static float time_loop(cdr3_solver *s, int warmup, int n, F &&body)
{
	void *e0, *e1;
	gpu::check(bmgx_event_create(&e0), "event"); gpu::check(bmgx_event_create(&e1), "event");
	float total = 0;
	for (int i = 0; i < warmup + n; i++) {
		s->x.set(0.0); s->x.dev();
		gpu::check(bmgx_event_record(e0, gpu::stream()), "event record");
		body();
		gpu::check(bmgx_event_record(e1, gpu::stream()), "event record");
		gpu::check(bmgx_event_sync(e1), "event sync");
		float ms = 0;
		gpu::check(bmgx_event_elapsed_ms(e0, e1, &ms), "event elapsed");
		if (i >= warmup) total += ms;
	}
	bmgx_event_destroy(e0); bmgx_event_destroy(e1);
	return total / (n > 0 ? n : 1);
}

extern "C" {

int cdr3_operator_create(int nstencil, unsigned nx, unsigned ny, unsigned nz, cdr3_operator **op)
{
	return guard([&] {
		if (nstencil != 4 && nstencil != 14) throw std::runtime_error("nstencil must be 4 or 14");
		auto o = std::make_unique<cdr3_operator>();
		o->nstencil = nstencil; o->nx = nx; o->ny = ny; o->nz = nz;
		if (nstencil == 4) o->seven = std::make_unique<stencil_op<seven_pt>>(nx, ny, nz);
		else o->full = std::make_unique<stencil_op<xxvii_pt>>(nx, ny, nz);
		*op = o.release();
	});
}

double *cdr3_operator_data(cdr3_operator *op)
{
	try { return op->nstencil == 4 ? op->seven->data() : op->full->data(); }
	catch (std::exception &e) { cdr_set_error(e.what()); return nullptr; }
}

int cdr3_operator_gallery(const char *name, unsigned nx, unsigned ny, unsigned nz, cdr3_operator **op)
{
	return guard([&] {
		auto o = std::make_unique<cdr3_operator>();
		o->nx = nx; o->ny = ny; o->nz = nz;
		std::string n(name);
		if (n == "poisson") { o->nstencil = 4; o->seven = std::make_unique<stencil_op<seven_pt>>(gallery::poisson(nx, ny, nz)); }
		else if (n == "fe") { o->nstencil = 14; o->full = std::make_unique<stencil_op<xxvii_pt>>(gallery::fe(nx, ny, nz)); }
		else throw std::runtime_error("unknown gallery operator " + n);
		*op = o.release();
	});
}

void cdr3_operator_destroy(cdr3_operator *op) { delete op; }

int cdr3_solver_create(cdr3_operator *op, const char *config_json, cdr3_solver **solver_out)
{
	return guard([&] {
		auto conf = make_config(config_json);
		auto s = std::make_unique<cdr3_solver>();
		s->op = op;
		if (op->nstencil == 4) s->s7 = std::make_unique<solver<seven_pt>>(*op->seven, conf);
		else s->s27 = std::make_unique<solver<xxvii_pt>>(*op->full, conf);
		s->b = grid_func(op->nx, op->ny, op->nz);
		s->x = grid_func(op->nx, op->ny, op->nz);
		gpu::sync();
		*solver_out = s.release();
	});
}

void cdr3_solver_destroy(cdr3_solver *s) { delete s; }

int cdr3_solver_nlevels(cdr3_solver *s) { return static_cast<int>(s->each([](auto &sv) { return sv.nlevels(); })); }

int cdr3_solver_level_shape(cdr3_solver *s, int lvl, unsigned *nx, unsigned *ny, unsigned *nz)
{
	return guard([&] {
		if (lvl < 0 || lvl >= cdr3_solver_nlevels(s)) throw std::runtime_error("level out of range");
		if (lvl == 0) { *nx = s->op->nx; *ny = s->op->ny; *nz = s->op->nz; return; }
		auto &A = s->s7 ? s->s7->levels.get(lvl).A : s->s27->levels.get(lvl).A;
		*nx = A.shape(0); *ny = A.shape(1); *nz = A.shape(2);
	});
}

int cdr3_solver_level_array(cdr3_solver *s, int lvl, const char *which, double *out, size_t n, size_t *written)
{
	return guard([&] {
		std::string w(which);
		if (lvl < 0 || lvl >= cdr3_solver_nlevels(s)) throw std::runtime_error("level out of range");
		if (lvl == 0) {
			if (s->s7) level_array(s->s7->levels.get<seven_pt>(0), w, out, n, written);
			else level_array(s->s27->levels.get<xxvii_pt>(0), w, out, n, written);
		} else {
			if (s->s7) level_array(s->s7->levels.get(lvl), w, out, n, written);
			else level_array(s->s27->levels.get(lvl), w, out, n, written);
		}
	});
}

int cdr3_solver_solve(cdr3_solver *s, const double *b, double *x, double *hist, int hist_cap, int *niter, double *res0)
{
	return guard([&] {
		upload(s->b, b);
		upload(s->x, x);
		s->each([&](auto &sv) {
			sv.solve(s->b, s->x);
			auto &h = sv.history();
			if (res0) *res0 = sv.res0();
			if (niter) *niter = static_cast<int>(h.size());
			for (int i = 0; hist && i < hist_cap && i < static_cast<int>(h.size()); i++) hist[i] = h[i];
			return 0;
		});
		std::memcpy(x, s->x.cdata(), s->x.size() * sizeof(double));
	});
}

int cdr3_solver_vcycle(cdr3_solver *s, const double *b, double *x)
{
	return guard([&] {
		upload(s->b, b);
		upload(s->x, x);
		s->each([&](auto &sv) { sv.vcycle(s->x, s->b); return 0; });
		std::memcpy(x, s->x.cdata(), s->x.size() * sizeof(double));
	});
}

int cdr3_solver_matvec(cdr3_solver *s, const double *x, double *y)
{
	return guard([&] {
		grid_func gx(s->op->nx, s->op->ny, s->op->nz), gy(s->op->nx, s->op->ny, s->op->nz);
		std::memcpy(gx.data(), x, gx.size() * sizeof(double));
		if (s->s7) s->s7->get_kernels()->run<kernels::matvec<stypes>>(s->s7->levels.get<seven_pt>(0).A, gx, gy);
		else s->s27->get_kernels()->run<kernels::matvec<stypes>>(s->s27->levels.get<xxvii_pt>(0).A, gx, gy);
		std::memcpy(y, gy.cdata(), gy.size() * sizeof(double));
	});
}

int cdr3_solver_buffers(cdr3_solver *s, double **b, double **x)
{
	return guard([&] { *b = s->b.data(); *x = s->x.data(); });
}

int cdr3_solver_vcycle_pinned(cdr3_solver *s)
{
	return guard([&] {
		s->b.data(); s->x.data();
		s->each([&](auto &sv) { sv.vcycle(s->x, s->b); return 0; });
		s->x.cdata();
	});
}

int cdr3_solver_vcycle0_pinned(cdr3_solver *s)
{
	return guard([&] {
		s->b.data();
		s->x.zero_device();
		s->each([&](auto &sv) { sv.vcycle(s->x, s->b); return 0; });
		s->x.cdata();
	});
}

int cdr3_solver_vcycle_batch(cdr3_solver *s, int nrhs, const double *const *b, double *const *x)
{
	return guard([&] {
		run_batch(s->stage, s->b, s->x, nrhs, b, x, [&] { s->each([&](auto &sv) { sv.vcycle(s->x, s->b); return 0; }); });
	});
}

int cdr3_solver_time_vcycles(cdr3_solver *s, const double *b, int warmup, int n, float *ms_per_cycle)
{
	return guard([&] {
		upload(s->b, b);
		s->b.cdev();
		*ms_per_cycle = time_loop(s, warmup, n, [&] { s->each([&](auto &sv) { sv.vcycle(s->x, s->b); return 0; }); });
	});
}

int cdr3_solver_time_solves(cdr3_solver *s, const double *b, int warmup, int n, float *ms_per_solve, int *niter)
{
	return guard([&] {
		upload(s->b, b);
		s->b.cdev();
		*ms_per_solve = time_loop(s, warmup, n, [&] { s->each([&](auto &sv) { sv.solve(s->b, s->x); return 0; }); });
		if (niter) *niter = static_cast<int>(s->each([](auto &sv) { return sv.history().size(); }));
	});
}

}

// One plane-relaxation sweep through the kernel registry, exactly the call sequence of the reference's own test
// (test/3d/test_planes.cc:13-52): build_kernel_manager(conf); kman->setup<plane_relax<dir>>(so);
// kman->run<plane_relax<dir>>(so, x, b, cdir).  Host arrays in, x out.
template <relax_dir rdir, class OP>
static void plane_sweep(OP &so, config &conf, grid_func &x, const grid_func &b, cycle::Dir cdir)
{
	auto kman = build_kernel_manager(conf);
	kman->template setup<kernels::plane_relax<stypes, rdir>>(so);
	kman->template run<kernels::plane_relax<stypes, rdir>>(so, x, b, cdir);
}

extern "C" int cdr3_plane_relax(cdr3_operator *op, const char *config_json, const char *dir, int updown, const double *b, double *x)
{
	return guard([&] {
		auto conf = make_config(config_json);
		grid_func gx(op->nx, op->ny, op->nz), gb(op->nx, op->ny, op->nz);
		upload(gx, x); upload(gb, b);
		const std::string d(dir);
		const cycle::Dir cdir = updown ? cycle::Dir::UP : cycle::Dir::DOWN;
		auto go = [&](auto &so) {
			if (d == "xy") plane_sweep<relax_dir::xy>(so, *conf, gx, gb, cdir);
			else if (d == "xz") plane_sweep<relax_dir::xz>(so, *conf, gx, gb, cdir);
			else if (d == "yz") plane_sweep<relax_dir::yz>(so, *conf, gx, gb, cdir);
			else throw std::runtime_error("plane direction must be xy, xz or yz");
		};
		if (op->seven) go(*op->seven); else go(*op->full);
		std::memcpy(x, gx.cdata(), gx.size() * sizeof(double));
	});
}

extern "C" int cdr3_solver_resident_ptrs(cdr3_solver *s, double **b_dev, double **x_dev)
{
	return guard([&] { *b_dev = s->b.dev(); *x_dev = s->x.dev(); });
}

extern "C" int cdr3_solver_vcycle_resident(cdr3_solver *s, void *stream)
{
	return guard([&] {
		gpu::stream_guard on_callers_stream(stream);
		s->b.dev(); s->x.dev();
		s->each([&](auto &slv) { slv.vcycle(s->x, s->b); });
	});
}
