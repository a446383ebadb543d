/* API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root. */
/* cedar/capi.h -- the reference's C interface, by name (include/cedar/capi.h:1-7) */
#include <cedar/interface/c/timer.h>
#include <cedar/2d/interface/c/operator.h>
#include <cedar/2d/interface/c/solver.h>
#include <cedar/2d/interface/c/topo.h>
#include <cedar/3d/interface/c/operator.h>
#include <cedar/3d/interface/c/solver.h>
#include <cedar/3d/interface/c/topo.h>
