#ifndef CEDAR_HALO_EXCHANGER_BASE_H
#define CEDAR_HALO_EXCHANGER_BASE_H
#include <cedar/types.h>
namespace cedar {
// seam the kernels use to refresh ghost layers (reference: halo_exchanger_base.h).
// Single-GPU runs leave it null; the multi-GPU block decomposition installs an
// NVLink exchanger through kernel_manager::add_halo.
class halo_exchanger_base
{
public:
	virtual ~halo_exchanger_base() {}
	virtual void exchange_func(int k, real_t *gf) = 0;
	virtual void exchange_sten(int k, real_t *so) = 0;
};
}
#endif
