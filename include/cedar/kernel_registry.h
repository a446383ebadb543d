// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
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
                              kernels::matvec<solver_types>,
                              kernels::setup_interp<solver_types>,
                              kernels::solve_cg<solver_types>>;

// The reference's distributed list adds matvec, halo_exchange and setup_nog.  Here matvec sits on the serial list
// (a single GPU applies the operator without any exchange) and ghost exchange between GPUs is owned by the
// block-decomposed solver (NCCL over NVLink), not by a registry kernel -- so both modes share one list.
template <class solver_types, exec_mode mode> using klist = ser_kernels<solver_types>;

}
#endif
