/* API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root. */
/* cedar/3d/interface/c/topo.h -- bmg3_topo_create (reference include/cedar/3d/interface/c/topo.h); see the 2D twin:
 * communicator accepted and ignored, process grid 1 x 1 x 1. */
#ifndef CEDAR_3D_INTERFACE_TOPO_H
#define CEDAR_3D_INTERFACE_TOPO_H
#include <cedar/2d/interface/c/topo.h>
#ifdef __cplusplus
extern "C" {
#endif
struct bmg3_topology;
typedef struct bmg3_topology *bmg3_topo;
bmg3_topo bmg3_topo_create(bmg_comm_t comm, unsigned int ngx, unsigned int ngy, unsigned int ngz, unsigned int lnx[],
                           unsigned int lny[], unsigned int lnz[], int nprocx, int nprocy, int nprocz);
void bmg3_topo_destroy(bmg3_topo topo);
#ifdef __cplusplus
}
#endif
#endif
