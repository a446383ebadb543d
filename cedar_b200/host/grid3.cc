// grid3.cc -- 3D grid functions, problem gallery and the default 3D kernel registry.
// Reference: src/3d/grid_func.cc, src/3d/gallery.cc:7-190, src/3d/kernel_manager.cc.
#include <cmath>
#include <random>

#include <cedar/3d/grid_func.h>
#include <cedar/3d/gallery.h>
#include <cedar/3d/kernel_manager.h>
#include <cedar/3d/relax.h>
#include <cedar/3d/relax_planes.h>
#include <cedar/3d/residual.h>
#include <cedar/3d/matvec.h>
#include <cedar/3d/restrict.h>
#include <cedar/3d/interp.h>
#include <cedar/3d/coarsen.h>
#include <cedar/3d/solve_cg.h>

namespace cedar {

real_t *norm_scratch();   // grid2.cc

namespace cdr3 {

grid_func::grid_func(len_t nx, len_t ny, len_t nz, unsigned int nghosts)
    : array<real_t, 3>(nx + 2 * nghosts, ny + 2 * nghosts, nz + 2 * nghosts)
{
	num_ghosts = nghosts;
	set_ranges();
}

static grid_func filled(len_t nx, len_t ny, len_t nz, real_t v)
{
	grid_func ret(nx, ny, nz);
	ret.set(v);
	return ret;
}

grid_func grid_func::ones(len_t nx, len_t ny, len_t nz) { return filled(nx, ny, nz, 1.0); }
grid_func grid_func::zeros(len_t nx, len_t ny, len_t nz) { return grid_func(nx, ny, nz); }
grid_func grid_func::like(const grid_func &l) { return grid_func(l.shape(0), l.shape(1), l.shape(2)); }
grid_func grid_func::zeros_like(const grid_func &l) { return grid_func(l.shape(0), l.shape(1), l.shape(2)); }
grid_func grid_func::ones_like(const grid_func &l) { return filled(l.shape(0), l.shape(1), l.shape(2), 1.0); }

grid_func grid_func::random(len_t nx, len_t ny, len_t nz)
{
	grid_func ret(nx, ny, nz);
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
	for (auto k : this->range(2))
		for (auto j : this->range(1))
			for (auto i : this->range(0)) {
				real_t v = d[(static_cast<std::size_t>(k) * len(1) + j) * len(0) + i];
				if (std::abs(cmax) < std::abs(v)) cmax = v;
			}
	return cmax;
}

real_t grid_func::l2_device() const
{
	real_t *scratch = norm_scratch();
	real_t *out = scratch + bmgx_norm_work_len();
	gpu::check(bmgx3_sumsq(cdev(), len(0), len(1), len(2), scratch, out, gpu::stream()), "lp_norm");
	real_t ss = 0;
	gpu::check(bmgx_memcpy_d2h(&ss, out, sizeof(real_t), gpu::stream()), "lp_norm readback");
	gpu::sync();
	return std::sqrt(ss);
}

grid_func &grid_func::operator-=(const grid_func &rhs)
{
	real_t *a = data();
	const real_t *b = rhs.cdata();
	for (std::size_t i = 0, n = size(); i < n; i++) a[i] -= b[i];
	return *this;
}

// ---------------------------------------------------------------- gallery ----
namespace gallery {

// 7-point finite differences on the unit cube; couplings across the Dirichlet
// boundary are dropped from pw/ps/b but kept in the centre
stencil_op<seven_pt> poisson(len_t nx, len_t ny, len_t nz)
{
	auto so = stencil_op<seven_pt>(nx, ny, nz);
	real_t hx = 1.0 / (so.len(0) - 1), hy = 1.0 / (so.len(1) - 1), hz = 1.0 / (so.len(2) - 1);
	real_t xh = hy * hz / hx, yh = hx * hz / hy, zh = hx * hy / hz;
	real_t *d = so.data();
	std::size_t II = so.len(0), JJ = so.len(1), pl = II * JJ * so.len(2);
	for (len_t k = 1; k <= nz; k++)
		for (len_t j = 1; j <= ny; j++)
			for (len_t i = 1; i <= nx; i++) {
				std::size_t c = (k * JJ + j) * II + i;
				if (j >= 2) d[static_cast<int>(seven_pt::ps) * pl + c] = 1.0 * yh;
				if (i >= 2) d[static_cast<int>(seven_pt::pw) * pl + c] = 1.0 * xh;
				if (k >= 2) d[static_cast<int>(seven_pt::b) * pl + c] = 1.0 * zh;
				d[static_cast<int>(seven_pt::p) * pl + c] = 2.0 * xh + 2.0 * yh + 2.0 * zh;
			}
	return so;
}

// 27-point trilinear finite elements: centre 26, all 26 neighbours 1
stencil_op<xxvii_pt> fe(len_t nx, len_t ny, len_t nz)
{
	auto so = stencil_op<xxvii_pt>(nx, ny, nz);
	real_t *d = so.data();
	std::size_t II = so.len(0), JJ = so.len(1), pl = II * JJ * so.len(2);
	auto put = [&](xxvii_pt e, std::size_t c) { d[static_cast<int>(e) * pl + c] = 1.0; };
	for (len_t k = 1; k <= nz; k++)
		for (len_t j = 1; j <= ny; j++)
			for (len_t i = 1; i <= nx; i++) {
				std::size_t c = (k * JJ + j) * II + i;
				bool x = i >= 2, y = j >= 2, z = k >= 2;
				if (x) put(xxvii_pt::pw, c);
				if (y) put(xxvii_pt::ps, c);
				if (z) put(xxvii_pt::b, c);
				if (x && y) { put(xxvii_pt::pnw, c); put(xxvii_pt::psw, c); }
				if (x && z) { put(xxvii_pt::bw, c); put(xxvii_pt::be, c); }
				if (y && z) { put(xxvii_pt::bn, c); put(xxvii_pt::bs, c); }
				if (x && y && z) { put(xxvii_pt::bnw, c); put(xxvii_pt::bne, c); put(xxvii_pt::bse, c); put(xxvii_pt::bsw, c); }
				d[static_cast<int>(xxvii_pt::p) * pl + c] = 26;
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
	kman->add<plane_relax<stypes, relax_dir::xy>, planes<relax_dir::xy>>("system");
	kman->add<plane_relax<stypes, relax_dir::xz>, planes<relax_dir::xz>>("system");
	kman->add<plane_relax<stypes, relax_dir::yz>, planes<relax_dir::yz>>("system");
	kman->add<coarsen_op<stypes>, galerkin>("system");
	kman->add<interp_add<stypes>, interp_b200>("system");
	kman->add<restriction<stypes>, restrict_b200>("system");
	kman->add<residual<stypes>, residual_b200>("system");
	kman->add<matvec<stypes>, matvec_b200>("system");
	kman->add<setup_interp<stypes>, setup_interp_b200>("system");
	kman->add<solve_cg<stypes>, solve_cg_b200>("system");

	kman->set<point_relax<stypes>>("system");
	kman->set<plane_relax<stypes, relax_dir::xy>>("system");
	kman->set<plane_relax<stypes, relax_dir::xz>>("system");
	kman->set<plane_relax<stypes, relax_dir::yz>>("system");
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
