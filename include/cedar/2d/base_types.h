// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_2D_CORE_BASE_TYPES_H
#define CEDAR_2D_CORE_BASE_TYPES_H
/* slot numbers of the 2D symmetric stencil storage (reference 2d/base_types.h:4-14;
 * Fortran ko=1,kw=2,ks=3,ksw=4,knw=5 minus one) */
typedef enum { BMG2_C = 0, BMG2_W = 1, BMG2_S = 2, BMG2_SW = 3, BMG2_NW = 4, BMG2_SE = 5, BMG2_N = 6, BMG2_NE = 7, BMG2_E = 8 } bmg2_dir;
#endif
