// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/util/timer.h -- named timing regions.  The reference keeps hierarchical
// MPI wall-clock timers (include/cedar/util/time_log.h); here a region is a pair
// of host time stamps around enqueued GPU work, recorded only when the "timer"
// log level is on (device-side timing is done with CUDA events by bench.py).
#ifndef CEDAR_TIMER_H
#define CEDAR_TIMER_H
#include <string>
namespace cedar {
void timer_begin(const std::string &label);
void timer_end(const std::string &label);
void timer_up();
void timer_down();
}
#endif
