// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_3D_BASE_TYPES_H
#define CEDAR_3D_BASE_TYPES_H
/* slot numbers of the 3D symmetric stencil storage (reference 3d/base_types.h;
 * Fortran kp=1 ... kbsw=14 minus one, src/3d/ftn/BMG_stencils_f90.h:43-48) */
typedef enum { BMG3_P = 0, BMG3_PW = 1, BMG3_PS = 2, BMG3_B = 3, BMG3_PSW = 4, BMG3_PNW = 5, BMG3_BW = 6, BMG3_BNW = 7,
               BMG3_BN = 8, BMG3_BNE = 9, BMG3_BE = 10, BMG3_BSE = 11, BMG3_BS = 12, BMG3_BSW = 13 } cdr3_dir;
#endif
