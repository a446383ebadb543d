// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_3D_GALLERY_H
#define CEDAR_3D_GALLERY_H
#include <cedar/3d/stencil_op.h>
namespace cedar { namespace cdr3 { namespace gallery {
// problem generators of the reference (src/3d/gallery.cc:7-56, 111-190)
stencil_op<seven_pt> poisson(len_t nx, len_t ny, len_t nz);
stencil_op<xxvii_pt> fe(len_t nx, len_t ny, len_t nz);
}}}
#endif
