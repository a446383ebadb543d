/* API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root. */
/* cedar/2d/interface/c/topo.h -- bmg2_topo_create of the reference's C interface
 * (include/cedar/2d/interface/c/topo.h:14-20, src/2d/interface/c/topo.cc:10-69).
 *
 * The reference takes an MPI_Comm and builds the process-grid topology of its MPI solver.  This library runs the
 * whole grid on ONE GPU per process (the multi-GPU path is include/cedar_capi.h, cdr2_dist_*, which needs no MPI),
 * so the communicator is accepted and ignored and the process grid must be 1 x 1: the same symbol, the same argument
 * order.  Without <mpi.h> the communicator type is an opaque pointer-sized value; include <mpi.h> first to get the
 * reference's exact prototype (an integer or pointer first argument travels in the same register either way). */
#ifndef CEDAR_2D_INTERFACE_TOPO_H
#define CEDAR_2D_INTERFACE_TOPO_H
#ifdef __cplusplus
extern "C" {
#endif
#ifdef MPI_VERSION
typedef MPI_Comm bmg_comm_t;
#else
typedef void *bmg_comm_t;
#endif
struct bmg2_topology;
typedef struct bmg2_topology *bmg2_topo;
bmg2_topo bmg2_topo_create(bmg_comm_t comm, unsigned int ngx, unsigned int ngy, unsigned int lnx[], unsigned int lny[],
                           int nprocx, int nprocy);
/* not in the reference (which leaks the topology): releases it */
void bmg2_topo_destroy(bmg2_topo topo);
#ifdef __cplusplus
}
#endif
#endif
