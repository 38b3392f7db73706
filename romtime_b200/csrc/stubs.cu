// Weak placeholders for entry points whose kernels are not linked yet; each is overridden by the
// strong definition in its own translation unit.  They fail loudly (no CPU fallback).
#include "common.cuh"
#define RT_WEAK extern "C" __attribute__((weak))
#define NI(h) return rt_fail((h), RT_EINVAL, "%s: not implemented in this build", __func__)

RT_WEAK int rt_project_vec_f64(rt_handle* h, const double*, int64_t, int64_t, const double*, int64_t, int64_t, int64_t, int64_t, int64_t, double*, void*) { NI(h); }
RT_WEAK int rt_project_mat_f64(rt_handle* h, const double*, int64_t, int64_t, const int64_t*, const int64_t*, int64_t, const double*, int64_t, int64_t, int64_t, int64_t, int64_t, double*, int64_t, int64_t, void*) { NI(h); }
