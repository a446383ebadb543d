// cedar/kernel_registry.h -- the kernel type lists (reference: kernel_registry.h:24-48).
#ifndef CEDAR_KERNEL_REGISTRY_H
#define CEDAR_KERNEL_REGISTRY_H

#include <type_traits>
#include <cedar/types.h>
#include <cedar/type_list.h>
#include <cedar/kernels/point_relax.h>
#include <cedar/kernels/line_relax.h>
#include <cedar/kernels/plane_relax.h>
#include <cedar/kernels/coarsen_op.h>
#include <cedar/kernels/interp_add.h>
#include <cedar/kernels/matvec.h>
#include <cedar/kernels/residual.h>
#include <cedar/kernels/restrict.h>
#include <cedar/kernels/setup_interp.h>
#include <cedar/kernels/solve_cg.h>

namespace cedar {

template <class solver_types>
using ser_kernels = type_list<kernels::point_relax<solver_types>,
                              kernels::line_relax<solver_types, relax_dir::x>,
                              kernels::line_relax<solver_types, relax_dir::y>,
                              kernels::plane_relax<solver_types, relax_dir::xy>,
                              kernels::plane_relax<solver_types, relax_dir::xz>,
                              kernels::plane_relax<solver_types, relax_dir::yz>,
                              kernels::coarsen_op<solver_types>,
                              kernels::restriction<solver_types>,
                              kernels::interp_add<solver_types>,
                              kernels::residual<solver_types>,
                              kernels::setup_interp<solver_types>,
                              kernels::solve_cg<solver_types>>;

// the distributed list adds the matvec kernel; ghost exchange between GPUs is
// owned by the block-decomposed solver (NVLink), not by a registry kernel
template <class solver_types> using mpi_kernels = type_list<kernels::matvec<solver_types>>;

template <class solver_types, exec_mode mode>
using klist = typename std::conditional<mode == exec_mode::serial, ser_kernels<solver_types>,
                                        typename type_cat<ser_kernels<solver_types>, mpi_kernels<solver_types>>::type>::type;

}
#endif
