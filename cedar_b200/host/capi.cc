// capi.cc -- C API over the C++ host layer (include/cedar_capi.h).
#include <cstring>
#include <memory>
#include <sstream>
#include <string>

#include <cedar_capi.h>
#include <cedar/2d/solver.h>
#include <cedar/2d/gallery.h>
#include "batch_stage.h"

using namespace cedar;
using namespace cedar::cdr2;

static thread_local std::string g_err;
void cdr_set_error(const std::string &msg) { g_err = msg; }

struct cdr2_operator
{
	int nstencil;
	std::unique_ptr<stencil_op<five_pt>> five;
	std::unique_ptr<stencil_op<nine_pt>> nine;
	unsigned nx, ny;
};

struct cdr2_solver
{
	cdr2_operator *op;
	std::unique_ptr<solver<five_pt>> s5;
	std::unique_ptr<solver<nine_pt>> s9;
	grid_func b, x;
	batch_stage stage;
};

template <class F> static int guard(F &&f)
{
	try { f(); return 0; }
	catch (std::exception &e) { g_err = e.what(); return 1; }
	catch (...) { g_err = "unknown error"; return 1; }
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
	else if (w == "SOR1") copy_out(L.SOR[1], out, n, written);
	else throw std::runtime_error("unknown level array " + w);
}

static void upload(grid_func &g, const double *src)
{
	std::memcpy(g.data(), src, g.size() * sizeof(double));
}

extern "C" {

const char *cdr_last_error(void) { return g_err.c_str(); }
unsigned long long cdr_gpu_launches(void) { return gpu::launches(); }

void cdr_set_log(const char *levels)
{
	std::vector<std::string> names;
	std::stringstream ss(levels ? levels : "");
	std::string tok;
	while (std::getline(ss, tok, ','))
		if (!tok.empty()) names.push_back(tok);
	log::set_levels(names);
}

int cdr2_operator_create(int nstencil, unsigned nx, unsigned ny, cdr2_operator **op)
{
	return guard([&] {
		if (nstencil != 3 && nstencil != 5) throw std::runtime_error("nstencil must be 3 or 5");
		auto o = std::make_unique<cdr2_operator>();
		o->nstencil = nstencil; o->nx = nx; o->ny = ny;
		if (nstencil == 3) o->five = std::make_unique<stencil_op<five_pt>>(nx, ny);
		else o->nine = std::make_unique<stencil_op<nine_pt>>(nx, ny);
		*op = o.release();
	});
}

double *cdr2_operator_data(cdr2_operator *op)
{
	try { return op->nstencil == 3 ? op->five->data() : op->nine->data(); }
	catch (std::exception &e) { g_err = e.what(); return nullptr; }
}

int cdr2_operator_gallery(const char *name, unsigned nx, unsigned ny, double dx, double dy, cdr2_operator **op)
{
	return guard([&] {
		auto o = std::make_unique<cdr2_operator>();
		o->nx = nx; o->ny = ny;
		std::string n(name);
		if (n == "poisson") { o->nstencil = 3; o->five = std::make_unique<stencil_op<five_pt>>(gallery::poisson(nx, ny)); }
		else if (n == "diag_diffusion") { o->nstencil = 3; o->five = std::make_unique<stencil_op<five_pt>>(gallery::diag_diffusion(nx, ny, dx, dy)); }
		else if (n == "fe") { o->nstencil = 5; o->nine = std::make_unique<stencil_op<nine_pt>>(gallery::fe(nx, ny)); }
		else throw std::runtime_error("unknown gallery operator " + n);
		*op = o.release();
	});
}

void cdr2_operator_destroy(cdr2_operator *op) { delete op; }

int cdr2_solver_create(cdr2_operator *op, const char *config_json, cdr2_solver **solver_out)
{
	return guard([&] {
		auto conf = make_config(config_json);
		auto s = std::make_unique<cdr2_solver>();
		s->op = op;
		if (op->nstencil == 3) s->s5 = std::make_unique<solver<five_pt>>(*op->five, conf);
		else s->s9 = std::make_unique<solver<nine_pt>>(*op->nine, conf);
		s->b = grid_func(op->nx, op->ny);
		s->x = grid_func(op->nx, op->ny);
		gpu::sync();
		*solver_out = s.release();
	});
}

void cdr2_solver_destroy(cdr2_solver *s) { delete s; }

int cdr2_solver_nlevels(cdr2_solver *s) { return static_cast<int>(s->s5 ? s->s5->nlevels() : s->s9->nlevels()); }

int cdr2_solver_level_shape(cdr2_solver *s, int lvl, unsigned *nx, unsigned *ny)
{
	return guard([&] {
		if (lvl < 0 || lvl >= cdr2_solver_nlevels(s)) throw std::runtime_error("level out of range");
		if (lvl == 0) { *nx = s->op->nx; *ny = s->op->ny; return; }
		auto &A = s->s5 ? s->s5->levels.get(lvl).A : s->s9->levels.get(lvl).A;
		*nx = A.shape(0); *ny = A.shape(1);
	});
}

int cdr2_solver_level_array(cdr2_solver *s, int lvl, const char *which, double *out, size_t n, size_t *written)
{
	return guard([&] {
		std::string w(which);
		if (lvl < 0 || lvl >= cdr2_solver_nlevels(s)) throw std::runtime_error("level out of range");
		if (lvl == 0) {
			if (s->s5) level_array(s->s5->levels.get<five_pt>(0), w, out, n, written);
			else level_array(s->s9->levels.get<nine_pt>(0), w, out, n, written);
		} else {
			if (s->s5) level_array(s->s5->levels.get(lvl), w, out, n, written);
			else level_array(s->s9->levels.get(lvl), w, out, n, written);
		}
	});
}

int cdr2_solver_solve(cdr2_solver *s, const double *b, double *x, double *hist, int hist_cap, int *niter, double *res0)
{
	return guard([&] {
		upload(s->b, b);
		upload(s->x, x);
		const std::vector<real_t> *h;
		if (s->s5) { s->s5->solve(s->b, s->x); h = &s->s5->history(); if (res0) *res0 = s->s5->res0(); }
		else { s->s9->solve(s->b, s->x); h = &s->s9->history(); if (res0) *res0 = s->s9->res0(); }
		std::memcpy(x, s->x.cdata(), s->x.size() * sizeof(double));
		if (niter) *niter = static_cast<int>(h->size());
		for (int i = 0; hist && i < hist_cap && i < static_cast<int>(h->size()); i++) hist[i] = (*h)[i];
	});
}

int cdr2_solver_vcycle(cdr2_solver *s, const double *b, double *x)
{
	return guard([&] {
		upload(s->b, b);
		upload(s->x, x);
		if (s->s5) s->s5->vcycle(s->x, s->b); else s->s9->vcycle(s->x, s->b);
		std::memcpy(x, s->x.cdata(), s->x.size() * sizeof(double));
	});
}

int cdr2_solver_matvec(cdr2_solver *s, const double *x, double *y)
{
	return guard([&] {
		// y = A x with the solver's fine-level operator (what the reference's C interface calls the operator's
		// apply, src/2d/interface/c/operator.cc:76-90); ghost entries of y are zero
		grid_func gx(s->op->nx, s->op->ny), gy(s->op->nx, s->op->ny);
		upload(gx, x);
		if (s->s5) s->s5->get_kernels()->run<kernels::matvec<stypes>>(s->s5->levels.get<five_pt>(0).A, gx, gy);
		else s->s9->get_kernels()->run<kernels::matvec<stypes>>(s->s9->levels.get<nine_pt>(0).A, gx, gy);
		std::memcpy(y, gy.cdata(), gy.size() * sizeof(double));
	});
}

int cdr2_solver_vcycle_dev(cdr2_solver *s, const double *b_dev, double *x_dev, void *stream)
{
	return guard([&] {
		// device-resident caller (the block-decomposed solver's agglomerated coarse
		// levels): no host traffic; everything is enqueued on the caller's stream
		gpu::stream_guard on_callers_stream(stream);
		size_t bytes = s->b.size() * sizeof(double);
		gpu::check(bmgx_memcpy_d2d(s->b.dev(), b_dev, bytes, gpu::stream()), "vcycle_dev b");
		gpu::check(bmgx_memcpy_d2d(s->x.dev(), x_dev, bytes, gpu::stream()), "vcycle_dev x");
		if (s->s5) s->s5->vcycle(s->x, s->b); else s->s9->vcycle(s->x, s->b);
		gpu::check(bmgx_memcpy_d2d(x_dev, s->x.cdev(), bytes, gpu::stream()), "vcycle_dev x out");
	});
}

int cdr2_solver_resident_ptrs(cdr2_solver *s, double **b_dev, double **x_dev)
{
	return guard([&] { *b_dev = s->b.dev(); *x_dev = s->x.dev(); });
}

int cdr2_solver_vcycle_resident(cdr2_solver *s, void *stream)
{
	return guard([&] {
		gpu::stream_guard on_callers_stream(stream);
		s->b.dev(); s->x.dev();    // the caller wrote the device buffers
		if (s->s5) s->s5->vcycle(s->x, s->b); else s->s9->vcycle(s->x, s->b);
	});
}

int cdr2_solver_buffers(cdr2_solver *s, double **b, double **x)
{
	return guard([&] { *b = s->b.data(); *x = s->x.data(); });
}

int cdr2_solver_vcycle_pinned(cdr2_solver *s)
{
	return guard([&] {
		s->b.data(); s->x.data();  // the caller wrote the host side: device copies are stale
		if (s->s5) s->s5->vcycle(s->x, s->b); else s->s9->vcycle(s->x, s->b);
		s->x.cdata();              // download (synchronises)
	});
}

int cdr2_solver_vcycle0_pinned(cdr2_solver *s)
{
	return guard([&] {
		s->b.data();               // the caller wrote b on the host
		s->x.zero_device();        // x0 = 0: nothing to upload
		if (s->s5) s->s5->vcycle(s->x, s->b); else s->s9->vcycle(s->x, s->b);
		s->x.cdata();              // download (synchronises)
	});
}

int cdr2_solver_vcycle_batch(cdr2_solver *s, int nrhs, const double *const *b, double *const *x)
{
	return guard([&] {
		run_batch(s->stage, s->b, s->x, nrhs, b, x,
		          [&] { if (s->s5) s->s5->vcycle(s->x, s->b); else s->s9->vcycle(s->x, s->b); });
	});
}

int cdr_pinned_alloc(void **p, size_t bytes) { return guard([&] { gpu::check(bmgx_host_alloc(p, bytes), "pinned alloc"); }); }
int cdr_pinned_free(void *p) { return guard([&] { gpu::check(bmgx_host_free(p), "pinned free"); }); }

int cdr2_solver_solve_pinned(cdr2_solver *s, double *hist, int hist_cap, int *niter, double *res0)
{
	return guard([&] {
		s->b.data(); s->x.data();
		const std::vector<real_t> *h;
		if (s->s5) { s->s5->solve(s->b, s->x); h = &s->s5->history(); if (res0) *res0 = s->s5->res0(); }
		else { s->s9->solve(s->b, s->x); h = &s->s9->history(); if (res0) *res0 = s->s9->res0(); }
		s->x.cdata();
		if (niter) *niter = static_cast<int>(h->size());
		for (int i = 0; hist && i < hist_cap && i < static_cast<int>(h->size()); i++) hist[i] = (*h)[i];
	});
}

int cdr2_solver_time_vcycles(cdr2_solver *s, const double *b, int warmup, int n, float *ms_per_cycle)
{
	return guard([&] {
		upload(s->b, b);
		s->b.cdev();
		void *e0, *e1;
		gpu::check(bmgx_event_create(&e0), "event"); gpu::check(bmgx_event_create(&e1), "event");
		float total = 0;
		for (int i = 0; i < warmup + n; i++) {
			s->x.set(0.0); s->x.dev();
			gpu::check(bmgx_event_record(e0, gpu::stream()), "event record");
			if (s->s5) s->s5->vcycle(s->x, s->b); else s->s9->vcycle(s->x, s->b);
			gpu::check(bmgx_event_record(e1, gpu::stream()), "event record");
			gpu::check(bmgx_event_sync(e1), "event sync");
			float ms = 0;
			gpu::check(bmgx_event_elapsed_ms(e0, e1, &ms), "event elapsed");
			if (i >= warmup) total += ms;
		}
		bmgx_event_destroy(e0); bmgx_event_destroy(e1);
		*ms_per_cycle = total / (n > 0 ? n : 1);
	});
}

int cdr2_solver_time_solves(cdr2_solver *s, const double *b, int warmup, int n, float *ms_per_solve, int *niter)
{
	return guard([&] {
		upload(s->b, b);
		s->b.cdev();
		void *e0, *e1;
		gpu::check(bmgx_event_create(&e0), "event"); gpu::check(bmgx_event_create(&e1), "event");
		float total = 0;
		for (int i = 0; i < warmup + n; i++) {
			s->x.set(0.0); s->x.dev();
			gpu::check(bmgx_event_record(e0, gpu::stream()), "event record");
			if (s->s5) s->s5->solve(s->b, s->x); else s->s9->solve(s->b, s->x);
			gpu::check(bmgx_event_record(e1, gpu::stream()), "event record");
			gpu::check(bmgx_event_sync(e1), "event sync");
			float ms = 0;
			gpu::check(bmgx_event_elapsed_ms(e0, e1, &ms), "event elapsed");
			if (i >= warmup) total += ms;
		}
		bmgx_event_destroy(e0); bmgx_event_destroy(e1);
		*ms_per_solve = total / (n > 0 ? n : 1);
		if (niter) *niter = static_cast<int>(s->s5 ? s->s5->history().size() : s->s9->history().size());
	});
}

int cdr2_solve_host(int nstencil, unsigned nx, unsigned ny, const double *so, const double *b, double *x,
                    const char *config_json, double *hist, int hist_cap, int *niter)
{
	cdr2_operator *op = nullptr;
	cdr2_solver *sol = nullptr;
	int rc = cdr2_operator_create(nstencil, nx, ny, &op);
	if (rc) return rc;
	rc = guard([&] {
		size_t n = static_cast<size_t>(nx + 2) * (ny + 2) * nstencil;
		std::memcpy(cdr2_operator_data(op), so, n * sizeof(double));
	});
	if (!rc) rc = cdr2_solver_create(op, config_json, &sol);
	if (!rc) {
		std::memset(x, 0, static_cast<size_t>(nx + 2) * (ny + 2) * sizeof(double));
		rc = cdr2_solver_solve(sol, b, x, hist, hist_cap, niter, nullptr);
	}
	cdr2_solver_destroy(sol);
	cdr2_operator_destroy(op);
	return rc;
}

}
