// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/3d/kernel_manager.h -- builds the serial 3D registry with the CUDA kernels
// registered (and selected) under "system" (reference src/3d/kernel_manager.cc).
#ifndef CEDAR_3D_KERNEL_MANAGER_H
#define CEDAR_3D_KERNEL_MANAGER_H
#include <cedar/kernel_manager.h>
#include <cedar/kernel_registry.h>
#include <cedar/3d/types.h>
namespace cedar { namespace cdr3 {
using kman_ptr = std::shared_ptr<kernel_manager<klist<stypes, exec_mode::serial>>>;
kman_ptr build_kernel_manager(config &conf);
kman_ptr build_kernel_manager(std::shared_ptr<kernel_params> params);
}}
#endif
