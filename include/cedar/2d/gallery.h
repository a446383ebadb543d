// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_2D_GALLERY_H
#define CEDAR_2D_GALLERY_H
#include <cedar/2d/stencil_op.h>
namespace cedar { namespace cdr2 { namespace gallery {
// problem generators of the reference (src/2d/gallery.cc:7-113)
stencil_op<five_pt> poisson(len_t nx, len_t ny);
stencil_op<five_pt> diag_diffusion(len_t nx, len_t ny, real_t dx, real_t dy);
stencil_op<nine_pt> fe(len_t nx, len_t ny);
}}}
#endif
