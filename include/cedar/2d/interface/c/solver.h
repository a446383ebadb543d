/* API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root. */
/* cedar/2d/interface/c/solver.h -- bmg2_solver_* of the reference's C interface
 * (include/cedar/2d/interface/c/solver.h:13-15, src/2d/interface/c/solver.cc): create = setup phase from the operator
 * (settings from ./config.json when it exists, defaults otherwise), run = solve from x = 0 on packed interior
 * vectors.  The solve runs on the GPU; a failure (no device, non-SPD coarse operator) is reported on stderr and
 * leaves x untouched / returns NULL -- the reference's functions have no status either. */
#ifndef CEDAR_2D_INTERFACE_SOLVER_H
#define CEDAR_2D_INTERFACE_SOLVER_H
#include <cedar/2d/interface/c/operator.h>
#ifdef __cplusplus
extern "C" {
#endif
struct bmg2_slv;
typedef struct bmg2_slv *bmg2_solver;
bmg2_solver bmg2_solver_create(bmg2_operator *op);
void bmg2_solver_run(bmg2_solver op, double *x, const double *b);
void bmg2_solver_destroy(bmg2_solver);
#ifdef __cplusplus
}
#endif
#endif
