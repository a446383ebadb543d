// capi_bmg.cc -- the reference's C interface by name (include/cedar/capi.h: bmg2_* / bmg3_* / bmg_timer_save) over
// the GPU solver.  Reference: src/2d/interface/c/{topo,operator,solver}.cc, src/3d/interface/c/*.cc.  One process,
// one GPU: the communicator is ignored and the process grid must be 1 x 1 (x 1); the block-decomposed multi-GPU
// solver is cdr2_dist_* (include/cedar_capi.h).  The functions have no status channel in the reference either:
// failures are logged on stderr.
#include <cstdio>
#include <fstream>
#include <memory>
#include <string>

#include <cedar/capi.h>
#include <cedar/2d/solver.h>
#include <cedar/3d/solver.h>

using namespace cedar;

namespace {

std::shared_ptr<config> load_config()
{
	// the reference reads ./config.json (operator.cc:22); defaults when it is absent
	std::ifstream probe("config.json");
	return probe ? std::make_shared<config>("config.json") : std::make_shared<config>("");
}

template <class F> bool guarded(const char *where, F &&f)
{
	try { f(); return true; }
	catch (std::exception &e) { std::fprintf(stderr, "cedar %s: %s\n", where, e.what()); }
	catch (...) { std::fprintf(stderr, "cedar %s: unknown error\n", where); }
	return false;
}

}  // namespace

struct bmg2_topology { unsigned nx, ny; };
struct bmg3_topology { unsigned nx, ny, nz; };

struct bmg2_op
{
	bmg2_op(unsigned nx, unsigned ny, std::shared_ptr<config> c) : conf(c), op(nx, ny), xgf(nx, ny), bgf(nx, ny)
	{
		kman = cdr2::build_kernel_manager(*conf);
	}
	std::shared_ptr<config> conf;
	cdr2::stencil_op<cdr2::nine_pt> op;
	cdr2::grid_func xgf, bgf;
	cdr2::kman_ptr kman;
};
struct bmg3_op
{
	bmg3_op(unsigned nx, unsigned ny, unsigned nz, std::shared_ptr<config> c)
	    : conf(c), op(nx, ny, nz), xgf(nx, ny, nz), bgf(nx, ny, nz)
	{
		kman = cdr3::build_kernel_manager(*conf);
	}
	std::shared_ptr<config> conf;
	cdr3::stencil_op<cdr3::xxvii_pt> op;
	cdr3::grid_func xgf, bgf;
	cdr3::kman_ptr kman;
};
struct bmg2_slv
{
	std::unique_ptr<cdr2::solver<cdr2::nine_pt>> s;
	cdr2::grid_func x, b;
};
struct bmg3_slv
{
	std::unique_ptr<cdr3::solver<cdr3::xxvii_pt>> s;
	cdr3::grid_func x, b;
};

extern "C" {

void bmg_timer_save(const char *fname)
{
	guarded("bmg_timer_save", [&] {
		std::ofstream o(fname);
		o << "{\"device_launches\": " << gpu::launches() << "}\n";
	});
}

// ---- 2D -----------------------------------------------------------------------------------------------------
bmg2_topo bmg2_topo_create(bmg_comm_t, unsigned int ngx, unsigned int ngy, unsigned int lnx[], unsigned int lny[],
                           int nprocx, int nprocy)
{
	if (nprocx != 1 || nprocy != 1 || (lnx && lnx[0] != ngx) || (lny && lny[0] != ngy)) {
		std::fprintf(stderr, "cedar bmg2_topo_create: one process owns the whole grid here (process grid 1 x 1); the "
		                     "multi-GPU solver is cdr2_dist_* (include/cedar_capi.h)\n");
		return nullptr;
	}
	return new bmg2_topology{ngx, ngy};
}
void bmg2_topo_destroy(bmg2_topo t) { delete t; }

bmg2_operator bmg2_operator_create(bmg2_topo topo)
{
	bmg2_operator r = nullptr;
	if (!topo) return nullptr;
	guarded("bmg2_operator_create", [&] { r = new bmg2_op(topo->nx, topo->ny, load_config()); });
	return r;
}

void bmg2_operator_set(bmg2_operator op, unsigned int nvals, grid_coord_2d coords[], double vals[])
{
	guarded("bmg2_operator_set", [&] {
		auto &so = op->op;
		for (unsigned int n = 0; n < nvals; n++) {
			// global 0-based vertex -> array index with the ghost layer (is = 1: operator.cc:36-37)
			len_t ci = static_cast<len_t>(coords[n].i - 1 + 2);
			len_t cj = static_cast<len_t>(coords[n].j - 1 + 2);
			bmg2_dir dir = coords[n].dir;
			// BoxMG stores the off-diagonals positive; the caller's array is negated in place, as in the reference
			if (dir != 0) vals[n] = -1 * vals[n];
			// vertex-based input -> symmetric storage: E, N, NE, SE live at the neighbour that sees them as W, S, SW, NW
			if (dir == BMG2_SE) { ci++; dir = BMG2_NW; }
			else if (dir == BMG2_N) { cj++; dir = BMG2_S; }
			else if (dir == BMG2_NE) { ci++; cj++; dir = BMG2_SW; }
			else if (dir == BMG2_E) { ci++; dir = BMG2_W; }
			else if (dir == BMG2_NW) cj++;
			so(ci, cj, static_cast<cdr2::nine_pt>(dir)) = vals[n];
		}
	});
}

void bmg2_operator_apply(bmg2_operator op, const double *x, double *b)
{
	guarded("bmg2_operator_apply", [&] {
		auto &xgf = op->xgf;
		auto &bgf = op->bgf;
		int idx = 0;
		for (auto j : xgf.range(1))
			for (auto i : xgf.range(0)) xgf(i, j) = x[idx++];
		op->kman->run<kernels::matvec<cdr2::stypes>>(op->op, xgf, bgf);
		idx = 0;
		for (auto j : bgf.range(1))
			for (auto i : bgf.range(0)) b[idx++] = bgf(i, j);
	});
}

void bmg2_operator_dump(bmg2_operator op)
{
	guarded("bmg2_operator_dump", [&] {
		std::ofstream o("op0-0.txt", std::ios::out | std::ios::trunc);
		auto &so = op->op;
		o.precision(17);
		for (auto j : so.range(1))
			for (auto i : so.range(0)) {
				o << i << ", " << j;
				for (int d = 0; d < 5; d++) o << " " << so(i, j, static_cast<cdr2::nine_pt>(d));
				o << '\n';
			}
	});
}

void bmg2_operator_destroy(bmg2_operator op) { delete op; }

bmg2_solver bmg2_solver_create(bmg2_operator *op)
{
	bmg2_solver r = nullptr;
	if (!op || !*op) return nullptr;
	guarded("bmg2_solver_create", [&] {
		auto s = std::make_unique<bmg2_slv>();
		auto &so = (*op)->op;
		s->s = std::make_unique<cdr2::solver<cdr2::nine_pt>>(so, (*op)->conf);
		s->x = cdr2::grid_func(so.shape(0), so.shape(1));
		s->b = cdr2::grid_func(so.shape(0), so.shape(1));
		r = s.release();
	});
	return r;
}

void bmg2_solver_run(bmg2_solver slv, double *x, const double *b)
{
	guarded("bmg2_solver_run", [&] {
		int idx = 0;
		for (auto j : slv->b.range(1))
			for (auto i : slv->b.range(0)) slv->b(i, j) = b[idx++];
		slv->x.set(0.0);
		slv->s->solve(slv->b, slv->x);
		idx = 0;
		for (auto j : slv->x.range(1))
			for (auto i : slv->x.range(0)) x[idx++] = slv->x(i, j);
	});
}

void bmg2_solver_destroy(bmg2_solver s) { delete s; }

// ---- 3D -----------------------------------------------------------------------------------------------------
bmg3_topo bmg3_topo_create(bmg_comm_t, unsigned int ngx, unsigned int ngy, unsigned int ngz, unsigned int lnx[],
                           unsigned int lny[], unsigned int lnz[], int nprocx, int nprocy, int nprocz)
{
	if (nprocx != 1 || nprocy != 1 || nprocz != 1 || (lnx && lnx[0] != ngx) || (lny && lny[0] != ngy) || (lnz && lnz[0] != ngz)) {
		std::fprintf(stderr, "cedar bmg3_topo_create: one process owns the whole grid here (process grid 1 x 1 x 1)\n");
		return nullptr;
	}
	return new bmg3_topology{ngx, ngy, ngz};
}
void bmg3_topo_destroy(bmg3_topo t) { delete t; }

bmg3_operator bmg3_operator_create(bmg3_topo topo)
{
	bmg3_operator r = nullptr;
	if (!topo) return nullptr;
	guarded("bmg3_operator_create", [&] { r = new bmg3_op(topo->nx, topo->ny, topo->nz, load_config()); });
	return r;
}

void bmg3_operator_set(bmg3_operator op, unsigned int nvals, grid_coord_3d coords[], double vals[])
{
	guarded("bmg3_operator_set", [&] {
		auto &so = op->op;
		for (unsigned int n = 0; n < nvals; n++) {
			len_t ci = static_cast<len_t>(coords[n].i - 1 + 2);
			len_t cj = static_cast<len_t>(coords[n].j - 1 + 2);
			len_t ck = static_cast<len_t>(coords[n].k - 1 + 2);
			if (coords[n].dir != 0) vals[n] = -1 * vals[n];      // src/3d/interface/c/operator.cc:40
			so(ci, cj, ck, static_cast<cdr3::xxvii_pt>(coords[n].dir)) = vals[n];
		}
	});
}

void bmg3_operator_apply(bmg3_operator op, const double *x, double *b)
{
	guarded("bmg3_operator_apply", [&] {
		auto &xgf = op->xgf;
		auto &bgf = op->bgf;
		int idx = 0;
		for (auto k : xgf.range(2))
			for (auto j : xgf.range(1))
				for (auto i : xgf.range(0)) xgf(i, j, k) = x[idx++];
		op->kman->run<kernels::matvec<cdr3::stypes>>(op->op, xgf, bgf);
		idx = 0;
		for (auto k : bgf.range(2))
			for (auto j : bgf.range(1))
				for (auto i : bgf.range(0)) b[idx++] = bgf(i, j, k);
	});
}

void bmg3_operator_dump(bmg3_operator op)
{
	guarded("bmg3_operator_dump", [&] {
		std::ofstream o("op0-0-0.txt", std::ios::out | std::ios::trunc);
		auto &so = op->op;
		o.precision(17);
		for (auto k : so.range(2))
			for (auto j : so.range(1))
				for (auto i : so.range(0)) {
					o << i << ", " << j << ", " << k;
					for (int d = 0; d < 14; d++) o << " " << so(i, j, k, static_cast<cdr3::xxvii_pt>(d));
					o << '\n';
				}
	});
}

void bmg3_operator_destroy(bmg3_operator op) { delete op; }

bmg3_solver bmg3_solver_create(bmg3_operator *op)
{
	bmg3_solver r = nullptr;
	if (!op || !*op) return nullptr;
	guarded("bmg3_solver_create", [&] {
		auto s = std::make_unique<bmg3_slv>();
		auto &so = (*op)->op;
		s->s = std::make_unique<cdr3::solver<cdr3::xxvii_pt>>(so, (*op)->conf);
		s->x = cdr3::grid_func(so.shape(0), so.shape(1), so.shape(2));
		s->b = cdr3::grid_func(so.shape(0), so.shape(1), so.shape(2));
		r = s.release();
	});
	return r;
}

void bmg3_solver_run(bmg3_solver slv, double *x, const double *b)
{
	guarded("bmg3_solver_run", [&] {
		int idx = 0;
		for (auto k : slv->b.range(2))
			for (auto j : slv->b.range(1))
				for (auto i : slv->b.range(0)) slv->b(i, j, k) = b[idx++];
		slv->x.set(0.0);
		slv->s->solve(slv->b, slv->x);
		idx = 0;
		for (auto k : slv->x.range(2))
			for (auto j : slv->x.range(1))
				for (auto i : slv->x.range(0)) x[idx++] = slv->x(i, j, k);
	});
}

void bmg3_solver_destroy(bmg3_solver s) { delete s; }

}
