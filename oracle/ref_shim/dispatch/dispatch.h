/* <dispatch/dispatch.h> stand-in: dispatch_apply runs the iterations in order on the calling thread (every use on the
 * reference's hot path gives each iteration its own chunk and its own histogram / counter slot, so serial execution is
 * one of the schedules GCD may produce).  The clang block literal `^(size_t t) { ... }` is rewritten to a C++ lambda by the
 * build recipe (oracle/Makefile), nothing else in the source changes. */
#pragma once
#include <cstddef>
typedef int dispatch_queue_t;
#define DISPATCH_QUEUE_PRIORITY_HIGH 2
static inline dispatch_queue_t dispatch_get_global_queue(long, unsigned long) { return 0; }
template <class F> static inline void dispatch_apply(size_t n, dispatch_queue_t, F f) { for (size_t i = 0; i < n; i++) f(i); }
