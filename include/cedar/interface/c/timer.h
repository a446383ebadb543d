/* API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root. */
/* cedar/interface/c/timer.h -- bmg_timer_save of the reference's C interface (include/cedar/interface/c/timer.h:10).
 * The reference's timer tree is replaced by CUDA events and ncu (out of scope): the call is kept so that
 * applications link; it writes a one-line JSON with the number of device launches so far. */
#ifndef CEDAR_INTERFACE_TIMER_H
#define CEDAR_INTERFACE_TIMER_H
#ifdef __cplusplus
extern "C" {
#endif
void bmg_timer_save(const char *fname);
#ifdef __cplusplus
}
#endif
#endif
