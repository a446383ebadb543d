/* API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root. */
/* cedar/3d/interface/c/operator.h -- bmg3_operator_* (reference include/cedar/3d/interface/c/operator.h:12-31,
 * src/3d/interface/c/operator.cc): entries are given directly in BoxMG's 14 storage slots (cdr3_dir), off-diagonals
 * signed (stored negated); apply works on packed interior vectors (k outer, then j, i inner). */
#ifndef CEDAR_3D_INTERFACE_OPERATOR_H
#define CEDAR_3D_INTERFACE_OPERATOR_H
#include <cedar/3d/interface/c/topo.h>
#ifdef __cplusplus
extern "C" {
#endif
#include <cedar/3d/base_types.h>
struct bmg3_op;
typedef struct bmg3_op *bmg3_operator;
typedef struct {
	unsigned int i;
	unsigned int j;
	unsigned int k;
	cdr3_dir dir;
} grid_coord_3d;
bmg3_operator bmg3_operator_create(bmg3_topo topo);
void bmg3_operator_set(bmg3_operator, unsigned int nvals, grid_coord_3d coords[], double vals[]);
void bmg3_operator_apply(bmg3_operator, const double *x, double *b);
void bmg3_operator_dump(bmg3_operator);
void bmg3_operator_destroy(bmg3_operator);
#ifdef __cplusplus
}
#endif
#endif
