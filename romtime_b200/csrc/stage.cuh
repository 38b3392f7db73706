// Staging of operands that TMA cannot address directly (odd leading dimension, unaligned base,
// doubly-strided views): one transposing copy into a K-major scratch buffer with an even leading
// dimension.  Only unaligned views and tiny reference-sized problems take this path.
#pragma once
#include "dgemm_tma.cuh"

namespace dg {

static __global__ void k_stage_kmajor(const double* __restrict__ src, int64_t mn, int64_t k, int64_t smn, int64_t sk,
                               double* __restrict__ dst, int64_t ld) {
    __shared__ double tile[32][33];
    const int64_t k0 = (int64_t)blockIdx.x * 32, m0 = (int64_t)blockIdx.y * 32;
    const bool k_fast = (sk <= smn);  // read along the source's fast axis
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        if (k_fast) {
            const int64_t m = m0 + r, kk = k0 + threadIdx.x;
            if (m < mn && kk < k) tile[r][threadIdx.x] = src[m * smn + kk * sk];
        } else {
            const int64_t kk = k0 + r, m = m0 + threadIdx.x;
            if (m < mn && kk < k) tile[threadIdx.x][r] = src[m * smn + kk * sk];
        }
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int64_t m = m0 + r, kk = k0 + threadIdx.x;
        if (m < mn && kk < k) dst[m * ld + kk] = tile[r][threadIdx.x];
    }
}

// Returns an operand that satisfies tma_ok(): `op` itself, or a staged K-major copy in scratch `slot`.
inline int stage_if_needed(rt_handle* h, int slot, const Operand& op, Operand* out, cudaStream_t st) {
    if (op.tma_ok()) { *out = op; return RT_OK; }
    const int64_t ld = (op.k + 1) & ~int64_t(1);
    void* buf;
    int rc = rt_scratch(h, slot, (size_t)ld * op.mn * sizeof(double), &buf);
    if (rc) return rc;
    dim3 grid((unsigned)ceil_div64(op.k, 32), (unsigned)ceil_div64(op.mn, 32));
    if (grid.y > 65535) return rt_fail(h, RT_EINVAL, "staging copy: operand too large for the unaligned path");
    k_stage_kmajor<<<grid, dim3(32, 8), 0, st>>>(op.ptr, op.mn, op.k, op.stride_mn, op.stride_k, (double*)buf, ld);
    RT_LAUNCH_CHECK(h);
    out->ptr = (const double*)buf;
    out->mn = op.mn;
    out->k = op.k;
    out->stride_mn = ld;
    out->stride_k = 1;
    return RT_OK;
}

// choose the k-split count that minimises the persistent-schedule makespan (in k-tiles)
inline int choose_splits(int ntiles, int kt_total, int sms, int64_t out_elems, int max_splits = 64) {
    int best = 1;
    double best_cost = 1e300;
    for (int s = 1; s <= max_splits; ++s) {
        const int ktps = (kt_total + s - 1) / s;
        if (s > 1 && ktps < 32) break;
        if (s > 1 && (double)out_elems * s * 8 > 4e9) break;
        const int units = ntiles * ((kt_total + ktps - 1) / ktps);
        const int waves = (units + sms - 1) / sms;
        const double cost = (double)waves * (ktps + 6);  // +6: epilogue / pipeline refill per unit
        if (cost < best_cost * 0.98) { best_cost = cost; best = s; }
    }
    return best;
}

}  // namespace dg
