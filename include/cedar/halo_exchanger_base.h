// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_HALO_EXCHANGER_BASE_H
#define CEDAR_HALO_EXCHANGER_BASE_H
#include <cedar/types.h>
namespace cedar {
// seam the reference's MPI kernels use to refresh ghost layers from inside a kernel (reference:
// halo_exchanger_base.h, kernels/halo_exchange.h:22-31).  Kept for source compatibility only: nothing installs an
// exchanger here.  The multi-GPU solvers (cdr2_dist_* / cdr3_dist_*, cedar_b200/host/dist2.cc, dist3.cc) schedule
// their ghost exchanges themselves (include/cedar/2d/mpi/plan.h) as peer-memory pulls between kernels.
class halo_exchanger_base
{
public:
	virtual ~halo_exchanger_base() {}
	virtual void exchange_func(int k, real_t *gf) = 0;
	virtual void exchange_sten(int k, real_t *so) = 0;
};
}
#endif
