// grid2.cc -- 2D grid functions, problem gallery and the default kernel registry.
// Reference: src/2d/grid_func.cc, src/2d/gallery.cc:7-113, src/2d/kernel_manager.cc:10-42.
#include <cmath>
#include <iomanip>
#include <random>

#include <cedar/2d/grid_func.h>
#include <cedar/2d/gallery.h>
#include <cedar/2d/kernel_manager.h>
#include <cedar/2d/relax.h>
#include <cedar/2d/residual.h>
#include <cedar/2d/matvec.h>
#include <cedar/2d/restrict.h>
#include <cedar/2d/interp.h>
#include <cedar/2d/coarsen.h>
#include <cedar/2d/solve_cg.h>

namespace cedar {

// device scratch shared by all norm reductions of this process
real_t *norm_scratch()
{
	static array<real_t, 1> scratch(static_cast<len_t>(bmgx_norm_work_len() + 1));
	return scratch.dev();
}

namespace cdr2 {

grid_func::grid_func(len_t nx, len_t ny, unsigned int nghosts) : array<real_t, 2>(nx + 2 * nghosts, ny + 2 * nghosts)
{
	num_ghosts = nghosts;
	set_ranges();
}

static grid_func filled(len_t nx, len_t ny, real_t v)
{
	grid_func ret(nx, ny);
	ret.set(v);
	return ret;
}

grid_func grid_func::ones(len_t nx, len_t ny) { return filled(nx, ny, 1.0); }
grid_func grid_func::zeros(len_t nx, len_t ny) { return grid_func(nx, ny); }
grid_func grid_func::like(const grid_func &likeable) { return grid_func(likeable.shape(0), likeable.shape(1)); }
grid_func grid_func::zeros_like(const grid_func &like) { return grid_func(like.shape(0), like.shape(1)); }
grid_func grid_func::ones_like(const grid_func &like) { return filled(like.shape(0), like.shape(1), 1.0); }

grid_func grid_func::random(len_t nx, len_t ny)
{
	grid_func ret(nx, ny);
	std::uniform_real_distribution<double> unif(0, 1);
	std::random_device rand_dev;
	std::default_random_engine rand_engine(rand_dev());
	real_t *d = ret.data();
	for (std::size_t i = 0, n = ret.size(); i < n; i++) d[i] = unif(rand_engine);
	return ret;
}

real_t grid_func::inf_norm() const
{
	real_t cmax = 0;
	const real_t *d = cdata();
	for (auto j : this->range(1))
		for (auto i : this->range(0)) {
			real_t v = d[static_cast<std::size_t>(j) * len(0) + i];
			if (std::abs(cmax) < std::abs(v)) cmax = v;
		}
	return cmax;
}

real_t grid_func::l2_device() const
{
	real_t *scratch = norm_scratch();
	real_t *out = scratch + bmgx_norm_work_len();
	gpu::check(bmgx2_sumsq(cdev(), len(0), len(1), scratch, out, gpu::stream()), "lp_norm");
	real_t ss = 0;
	gpu::check(bmgx_memcpy_d2h(&ss, out, sizeof(real_t), gpu::stream()), "lp_norm readback");
	gpu::sync();
	return std::sqrt(ss);
}

grid_func &grid_func::operator-=(const grid_func &rhs)
{
	real_t *a = data();
	const real_t *b = rhs.cdata();
	auto jj = rhs.range(1).begin();
	for (auto j : this->range(1)) {
		auto ii = rhs.range(0).begin();
		for (auto i : this->range(0)) {
			a[static_cast<std::size_t>(j) * len(0) + i] -= b[static_cast<std::size_t>(*jj) * rhs.len(0) + *ii];
			++ii;
		}
		++jj;
	}
	return *this;
}

std::ostream &operator<<(std::ostream &os, const grid_func &obj)
{
	unsigned int width = 4;
	os << std::setprecision(7);
	for (auto j : obj.range(1))
		for (auto i : obj.range(0))
			os << std::setw(width) << i + 1 << ", " << std::setw(width) << j + 1 << ", " << std::scientific << obj(i, j)
			   << '\n';
	return os;
}

std::ostream &operator<<(std::ostream &os, const prolong_op &P)
{
	os << std::setprecision(7);
	for (auto j : P.range(1))
		for (auto i : P.range(0))
			os << i + 1 << ", " << j + 1 << ", " << std::scientific << P(i, j + 1, inter_dir::SE) << ", "
			   << P(i, j + 1, inter_dir::B) << ", " << P(i + 1, j + 1, inter_dir::SW) << ", " << P(i, j, inter_dir::R)
			   << ", " << 1.0 << ", " << P(i + 1, j, inter_dir::L) << ", " << P(i, j, inter_dir::NE) << ", "
			   << P(i, j, inter_dir::A) << ", " << P(i + 1, j, inter_dir::NW) << '\n';
	return os;
}

// ---------------------------------------------------------------- gallery ----
namespace gallery {

// 5-point finite differences on the unit square; couplings across the Dirichlet
// boundary are dropped from W/S but kept in the centre.
stencil_op<five_pt> diag_diffusion(len_t nx, len_t ny, real_t dx, real_t dy)
{
	auto so = stencil_op<five_pt>(nx, ny);
	real_t hx = 1.0 / (so.len(0) - 1);
	real_t hy = 1.0 / (so.len(1) - 1);
	real_t xh = hy / hx;
	real_t yh = hx / hy;
	real_t *d = so.data();
	std::size_t II = so.len(0), pl = static_cast<std::size_t>(so.len(0)) * so.len(1);
	for (len_t j = 1; j <= ny; j++)
		for (len_t i = 1; i <= nx; i++) {
			std::size_t c = j * II + i;
			if (j >= 2) d[static_cast<int>(five_pt::s) * pl + c] = dy * yh;
			if (i >= 2) d[static_cast<int>(five_pt::w) * pl + c] = dx * xh;
			d[static_cast<int>(five_pt::c) * pl + c] = 2 * dx * xh + 2 * dy * yh;
		}
	return so;
}

stencil_op<five_pt> poisson(len_t nx, len_t ny)
{
	auto so = stencil_op<five_pt>(nx, ny);
	real_t hx = 1.0 / (so.len(0) - 1);
	real_t hy = 1.0 / (so.len(1) - 1);
	real_t xh = hy / hx;
	real_t yh = hx / hy;
	real_t *d = so.data();
	std::size_t II = so.len(0), pl = static_cast<std::size_t>(so.len(0)) * so.len(1);
	for (len_t j = 1; j <= ny; j++)
		for (len_t i = 1; i <= nx; i++) {
			std::size_t c = j * II + i;
			if (j >= 2) d[static_cast<int>(five_pt::s) * pl + c] = 1.0 * yh;
			if (i >= 2) d[static_cast<int>(five_pt::w) * pl + c] = 1.0 * xh;
			d[static_cast<int>(five_pt::c) * pl + c] = 2 * xh + 2 * yh;
		}
	return so;
}

// 9-point bilinear finite elements: centre 8, all eight neighbours 1
stencil_op<nine_pt> fe(len_t nx, len_t ny)
{
	auto so = stencil_op<nine_pt>(nx, ny);
	real_t *d = so.data();
	std::size_t II = so.len(0), pl = static_cast<std::size_t>(so.len(0)) * so.len(1);
	for (len_t j = 1; j <= ny; j++)
		for (len_t i = 1; i <= nx; i++) {
			std::size_t c = j * II + i;
			if (j >= 2) d[static_cast<int>(nine_pt::s) * pl + c] = 1.0;
			if (i >= 2) d[static_cast<int>(nine_pt::w) * pl + c] = 1.0;
			if (i >= 2 && j >= 2) {
				d[static_cast<int>(nine_pt::sw) * pl + c] = 1.0;
				d[static_cast<int>(nine_pt::nw) * pl + c] = 1.0;
			}
			d[static_cast<int>(nine_pt::c) * pl + c] = 8.0;
		}
	return so;
}

}

// ------------------------------------------------------------- registry ----
kman_ptr build_kernel_manager(config &conf) { return build_kernel_manager(build_kernel_params(conf)); }

kman_ptr build_kernel_manager(std::shared_ptr<kernel_params> params)
{
	log::status << *params << std::endl;
	using namespace cedar::kernels;
	auto kman = std::make_shared<kernel_manager<klist<stypes, exec_mode::serial>>>(params);
	kman->add<point_relax<stypes>, rbgs>("system");
	kman->add<line_relax<stypes, relax_dir::x>, lines<relax_dir::x>>("system");
	kman->add<line_relax<stypes, relax_dir::y>, lines<relax_dir::y>>("system");
	kman->add<coarsen_op<stypes>, galerkin>("system");
	kman->add<interp_add<stypes>, interp_b200>("system");
	kman->add<restriction<stypes>, restrict_b200>("system");
	kman->add<residual<stypes>, residual_b200>("system");
	kman->add<matvec<stypes>, matvec_b200>("system");
	kman->add<setup_interp<stypes>, setup_interp_b200>("system");
	kman->add<solve_cg<stypes>, solve_cg_b200>("system");

	kman->set<point_relax<stypes>>("system");
	kman->set<line_relax<stypes, relax_dir::x>>("system");
	kman->set<line_relax<stypes, relax_dir::y>>("system");
	kman->set<coarsen_op<stypes>>("system");
	kman->set<interp_add<stypes>>("system");
	kman->set<restriction<stypes>>("system");
	kman->set<residual<stypes>>("system");
	kman->set<matvec<stypes>>("system");
	kman->set<setup_interp<stypes>>("system");
	kman->set<solve_cg<stypes>>("system");
	return kman;
}

}}
