/* API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root. */
/* cedar/2d/interface/c/operator.h -- bmg2_operator_* of the reference's C interface
 * (include/cedar/2d/interface/c/operator.h:12-29, src/2d/interface/c/operator.cc).  Same names, argument order and
 * meaning: vertex-based, signed stencil entries in (bmg2_operator_set converts them to BoxMG's symmetric storage with
 * positive off-diagonals, operator.cc:30-58), operator application on packed interior vectors (j outer, i inner). */
#ifndef CEDAR_2D_INTERFACE_OPERATOR_H
#define CEDAR_2D_INTERFACE_OPERATOR_H
#include <cedar/2d/interface/c/topo.h>
#ifdef __cplusplus
extern "C" {
#endif
#include <cedar/2d/base_types.h>
struct bmg2_op;
typedef struct bmg2_op *bmg2_operator;
typedef struct {
	unsigned int i;
	unsigned int j;
	bmg2_dir dir;
} grid_coord_2d;
bmg2_operator bmg2_operator_create(bmg2_topo topo);
void bmg2_operator_set(bmg2_operator, unsigned int nvals, grid_coord_2d coords[], double vals[]);
void bmg2_operator_apply(bmg2_operator, const double *x, double *b);
void bmg2_operator_dump(bmg2_operator);
void bmg2_operator_destroy(bmg2_operator);
#ifdef __cplusplus
}
#endif
#endif
