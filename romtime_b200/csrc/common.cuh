// Shared internals of libromtime_b200.so: handle, error convention, scratch allocator.
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>

#include "../../include/romtime_b200.h"

struct rt_handle {
    int device = 0;
    int sm_count = 0;
    int smem_optin = 0;
    char err[512] = {0};
    // grow-only scratch buffers (slots so that independent kernels do not alias)
    static constexpr int kSlots = 12;
    void* ws[kSlots] = {nullptr};
    size_t ws_bytes[kSlots] = {0};
    int64_t launches = 0;
    double fp64_peak = 0.0;
    void* tma_encode = nullptr;  // cuTensorMapEncodeTiled, resolved lazily
    void* profile = nullptr;     // online-sweep profiler (online.cu)
    void* sweep_cache = nullptr; // staged + folded online model of the last sweep / step (online.cu)
    // largest dynamic shared-memory opt-in configured so far per kernel slot ON THIS DEVICE (the attribute is per
    // device and per function: a process-wide flag would skip the opt-in on a second GPU or for a larger size)
    static constexpr int kAttrSlots = 24;
    int attr_smem[kAttrSlots] = {0};
};

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (handle, kernel slot), re-issued when a larger size is needed
template <typename F>
inline int rt_ensure_smem(rt_handle* h, int slot, F fn, int bytes) {
    if (h->attr_smem[slot] >= bytes) return RT_OK;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) {
        if (h) snprintf(h->err, sizeof(h->err), "cudaFuncSetAttribute(%d bytes, slot %d) failed: %s", bytes, slot, cudaGetErrorString(e));
        return RT_ECUDA;
    }
    h->attr_smem[slot] = bytes;
    return RT_OK;
}

inline int rt_fail(rt_handle* h, int code, const char* fmt, ...) {
    if (h) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(h->err, sizeof(h->err), fmt, ap);
        va_end(ap);
    }
    return code;
}

#define RT_CUDA(h, expr)                                                                       \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return rt_fail((h), RT_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                           __FILE__, __LINE__);                                                \
    } while (0)

#define RT_LAUNCH_CHECK(h)                                                                     \
    do {                                                                                       \
        (h)->launches++;                                                                       \
        cudaError_t _e = cudaGetLastError();                                                   \
        if (_e != cudaSuccess)                                                                 \
            return rt_fail((h), RT_ECUDA, "kernel launch failed: %s (%s:%d)",                  \
                           cudaGetErrorString(_e), __FILE__, __LINE__);                        \
    } while (0)

#define RT_REQUIRE(h, cond)                                                                    \
    do {                                                                                       \
        if (!(cond)) return rt_fail((h), RT_EINVAL, "invalid argument: %s (%s:%d)", #cond,     \
                                    __FILE__, __LINE__);                                       \
    } while (0)

// Scratch of at least `bytes` in slot `slot`; contents are not preserved across growth.
inline int rt_scratch(rt_handle* h, int slot, size_t bytes, void** out) {
    if (bytes == 0) bytes = 256;
    if (h->ws_bytes[slot] < bytes) {
        if (h->ws[slot]) {
            RT_CUDA(h, cudaDeviceSynchronize());
            RT_CUDA(h, cudaFree(h->ws[slot]));
            h->ws[slot] = nullptr;
            h->ws_bytes[slot] = 0;
        }
        size_t want = bytes + (bytes >> 3);
        want = (want + 1023) & ~size_t(1023);
        cudaError_t e = cudaMalloc(&h->ws[slot], want);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return rt_fail(h, RT_ENOMEM, "scratch allocation of %zu bytes failed: %s", want,
                           cudaGetErrorString(e));
        }
        h->ws_bytes[slot] = want;
    }
    *out = h->ws[slot];
    return RT_OK;
}

static inline cudaStream_t rt_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
