// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_HALO_EXCHANGER_BASE_H
#define CEDAR_HALO_EXCHANGER_BASE_H
#include <cedar/types.h>
namespace cedar {
// seam the reference's MPI kernels use to refresh ghost layers from inside a kernel (reference:
// halo_exchanger_base.h, kernels/halo_exchange.h:22-31; installed with kernel_manager::add_halo).  The multi-GPU solver
// implements it over its peer-memory exchange (cedar_b200/host/dist2.cc peer_exchanger, handed out by
// cdr2_dist_halo_exchanger; contract test tests/test_gpu_dist2.py::test_halo_exchange_contract); its own cycle does
// not go through it but schedules the exchanges between kernels (include/cedar/2d/mpi/plan.h).
class halo_exchanger_base
{
public:
	virtual ~halo_exchanger_base() {}
	virtual void exchange_func(int k, real_t *gf) = 0;
	virtual void exchange_sten(int k, real_t *so) = 0;
};
}
#endif
