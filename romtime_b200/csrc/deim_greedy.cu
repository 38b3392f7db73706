// DEIM / MDEIM greedy interpolation-point selection (reference deim/deim.py:517-561).
//
// One persistent cooperative kernel runs the whole greedy: per step j every CTA
//   (1) extends a bordered LU of P^T U_{:j} by the point found in step j-1 and solves for the
//       interpolation coefficients c (warp 0, vectors in registers, LU in shared memory) --
//       redundantly in every CTA, so no second grid barrier is needed;
//   (2) streams its rows of U[:, :j+1] from HBM with 16-byte loads (column-major, coalesced)
//       and forms res_i = u_j[i] - sum_k c_k U[i,k] with ONE fixed summation order (k ascending,
//       fma) for every row, so bit-identical rows give bit-identical residuals;
//   (3) reduces (|res|, i) with numpy.argmax semantics (first maximum, NaN is maximal) through
//       warp shuffles -> CTA -> one grid barrier -> every CTA reduces the per-CTA slots.
// Algorithmic HBM traffic: 8*Nh*(j+1) bytes in step j, 8*Nh*m(m+1)/2 in total, nothing written.
//
// The DEIM points are the row pivots of an LU of U with partial pivoting, so the bordered LU
// without further pivoting has |L| <= 1 (as stable as dgesv on P^T U, deim/deim.py:545-547).
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace {

constexpr int kThreads = 1024;
constexpr unsigned kFull = 0xffffffffu;

struct GreedyArgs {
    const double* U;  // column-major, element (i,k) at U[i + k*ld]
    int64_t Nh;
    int64_t ld;
    int m;
    int ldl;  // leading dimension of the LU workspace (odd -> conflict-free column access)
    int64_t* dofs;
    double* PT_U;
    int* info;
    double* slot_val;   // [2][gridDim.x]
    int64_t* slot_idx;  // [2][gridDim.x]
};

__device__ __forceinline__ bool val_gt(double a, double b) {
    // numpy.argmax ordering: NaN beats everything, first occurrence is kept by the index rule
    if (a != a) return !(b != b);
    if (b != b) return false;
    return a > b;
}
__device__ __forceinline__ void arg_merge(double& v, int64_t& i, double ov, int64_t oi) {
    if (val_gt(ov, v) || (!val_gt(v, ov) && oi < i)) { v = ov; i = oi; }
}
__device__ __forceinline__ void warp_arg_reduce(double& v, int64_t& i) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        double ov = __shfl_xor_sync(kFull, v, off);
        int64_t oi = __shfl_xor_sync(kFull, i, off);
        arg_merge(v, i, ov, oi);
    }
}

// ---- warp-level triangular solves; vector element e lives on lane e%32, register e/32 ----
template <int R>
__device__ __forceinline__ void solve_unit_lower(const double* LU, int ldl, int n, double (&v)[R], int lane) {
    // L y = v, L unit lower (strictly-lower part of LU)
#pragma unroll
    for (int kr = 0; kr < R; ++kr) {
        for (int kl = 0; kl < 32; ++kl) {
            const int k = kr * 32 + kl;
            if (k >= n) break;
            const double yk = __shfl_sync(kFull, v[kr], kl);
#pragma unroll
            for (int er = kr; er < R; ++er) {
                const int e = er * 32 + lane;
                if (e > k && e < n) v[er] = fma(-yk, LU[(size_t)e * ldl + k], v[er]);
            }
        }
    }
}
template <int R>
__device__ __forceinline__ void solve_upper_T(const double* LU, int ldl, int n, double (&v)[R], int lane) {
    // l^T Rm = v^T  <=>  Rm^T l = v  (Rm = upper part of LU incl. diagonal), forward substitution
#pragma unroll
    for (int kr = 0; kr < R; ++kr) {
        for (int kl = 0; kl < 32; ++kl) {
            const int k = kr * 32 + kl;
            if (k >= n) break;
            double lk = __shfl_sync(kFull, v[kr], kl) / LU[(size_t)k * ldl + k];
            if (lane == kl) v[kr] = lk;
#pragma unroll
            for (int er = kr; er < R; ++er) {
                const int e = er * 32 + lane;
                if (e > k && e < n) v[er] = fma(-lk, LU[(size_t)k * ldl + e], v[er]);
            }
        }
    }
}
template <int R>
__device__ __forceinline__ void solve_upper(const double* LU, int ldl, int n, double (&v)[R], int lane) {
    // Rm c = v, backward substitution
#pragma unroll
    for (int kr = R - 1; kr >= 0; --kr) {
        for (int kl = 31; kl >= 0; --kl) {
            const int k = kr * 32 + kl;
            if (k >= n) continue;
            double ck = __shfl_sync(kFull, v[kr], kl) / LU[(size_t)k * ldl + k];
            if (lane == kl) v[kr] = ck;
#pragma unroll
            for (int er = 0; er <= kr; ++er) {
                const int e = er * 32 + lane;
                if (e < k) v[er] = fma(-ck, LU[(size_t)e * ldl + k], v[er]);
            }
        }
    }
}

template <int R, int VEC>
__global__ void __launch_bounds__(kThreads, 1) k_deim_greedy(GreedyArgs a) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* LU = reinterpret_cast<double*>(smem_raw);            // m * ldl
    double* c_s = LU + (size_t)a.m * a.ldl;                      // m
    double* red_v = c_s + a.m;                                   // 32
    int64_t* red_i = reinterpret_cast<int64_t*>(red_v + 32);     // 32
    int64_t* dofs_s = red_i + 32;                                // m
    __shared__ int stop_s;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int m = a.m, ldl = a.ldl;
    const int64_t Nh = a.Nh, ld = a.ld;
    const double* __restrict__ U = a.U;
    if (tid == 0) stop_s = 0;
    __syncthreads();

    for (int j = 0; j < m; ++j) {
        // ---------------- (1) bordered LU + coefficient solve (warp 0) ----------------
        if (j > 0 && warp == 0) {
            const int64_t p = dofs_s[j - 1];
            double v[R];
            // new row of L:  l^T = U[p, 0:j-1] Rm^-1
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int e = r * 32 + lane;
                v[r] = (e < j - 1) ? U[p + (int64_t)e * ld] : 0.0;
            }
            solve_upper_T<R>(LU, ldl, j - 1, v, lane);
            double dot = 0.0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int e = r * 32 + lane;
                if (e < j - 1) {
                    LU[(size_t)(j - 1) * ldl + e] = v[r];
                    dot = fma(v[r], LU[(size_t)e * ldl + (j - 1)], dot);
                }
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) dot += __shfl_xor_sync(kFull, dot, off);
            if (lane == 0) LU[(size_t)(j - 1) * ldl + (j - 1)] = U[p + (int64_t)(j - 1) * ld] - dot;
            __syncwarp();
            // right-hand side b = U[dofs, j];  y = L^-1 b;  store as column j of Rm;  c = Rm^-1 y
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int e = r * 32 + lane;
                v[r] = (e < j) ? U[dofs_s[e] + (int64_t)j * ld] : 0.0;
            }
            solve_unit_lower<R>(LU, ldl, j, v, lane);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int e = r * 32 + lane;
                if (e < j) LU[(size_t)e * ldl + j] = v[r];
            }
            __syncwarp();
            solve_upper<R>(LU, ldl, j, v, lane);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int e = r * 32 + lane;
                if (e < j) c_s[e] = v[r];
            }
        }
        __syncthreads();

        // ---------------- (2) stream the residual, local arg-max ----------------
        double best_v = -1.0;
        int64_t best_i = INT64_MAX;
        const double* __restrict__ uj = U + (int64_t)j * ld;
        const int64_t stride = (int64_t)gridDim.x * kThreads * VEC;
        for (int64_t i = ((int64_t)blockIdx.x * kThreads + tid) * VEC; i < Nh; i += stride) {
            if (VEC == 2 && i + 1 < Nh) {
                double2 acc = *reinterpret_cast<const double2*>(uj + i);
                const double* col = U + i;
                int k = 0;
                for (; k + 8 <= j; k += 8) {
                    double2 x[8];
#pragma unroll
                    for (int q = 0; q < 8; ++q)
                        x[q] = __ldg(reinterpret_cast<const double2*>(col + (int64_t)(k + q) * ld));
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const double ck = c_s[k + q];
                        acc.x = fma(-ck, x[q].x, acc.x);
                        acc.y = fma(-ck, x[q].y, acc.y);
                    }
                }
                for (; k < j; ++k) {
                    const double2 x = __ldg(reinterpret_cast<const double2*>(col + (int64_t)k * ld));
                    const double ck = c_s[k];
                    acc.x = fma(-ck, x.x, acc.x);
                    acc.y = fma(-ck, x.y, acc.y);
                }
                const double a0 = fabs(acc.x), a1 = fabs(acc.y);
                if (val_gt(a0, best_v)) { best_v = a0; best_i = i; }
                if (val_gt(a1, best_v)) { best_v = a1; best_i = i + 1; }
            } else {
                // scalar path (VEC == 1, or the odd last row): same arithmetic per row
                double acc = uj[i];
                const double* col = U + i;
                int k = 0;
                for (; k + 8 <= j; k += 8) {
                    double x[8];
#pragma unroll
                    for (int q = 0; q < 8; ++q) x[q] = __ldg(col + (int64_t)(k + q) * ld);
#pragma unroll
                    for (int q = 0; q < 8; ++q) acc = fma(-c_s[k + q], x[q], acc);
                }
                for (; k < j; ++k) acc = fma(-c_s[k], __ldg(col + (int64_t)k * ld), acc);
                const double a0 = fabs(acc);
                if (val_gt(a0, best_v)) { best_v = a0; best_i = i; }
            }
        }
        // ---------------- (3) CTA reduce -> slot -> grid barrier -> global reduce ----------------
        warp_arg_reduce(best_v, best_i);
        if (lane == 0) { red_v[warp] = best_v; red_i[warp] = best_i; }
        __syncthreads();
        if (warp == 0) {
            double v = red_v[lane];
            int64_t i = red_i[lane];
            warp_arg_reduce(v, i);
            if (lane == 0) {
                a.slot_val[(size_t)(j & 1) * gridDim.x + blockIdx.x] = v;
                a.slot_idx[(size_t)(j & 1) * gridDim.x + blockIdx.x] = i;
            }
        }
        grid.sync();
        if (warp == 0) {
            double v = -1.0;
            int64_t i = INT64_MAX;
            for (unsigned s = lane; s < gridDim.x; s += 32) {
                const double ov = __ldcg(a.slot_val + (size_t)(j & 1) * gridDim.x + s);
                const int64_t oi = __ldcg(a.slot_idx + (size_t)(j & 1) * gridDim.x + s);
                arg_merge(v, i, ov, oi);
            }
            warp_arg_reduce(v, i);
            if (lane == 0) {
                dofs_s[j] = i;
                if (blockIdx.x == 0) a.dofs[j] = i;
                // residual identically zero: the next interpolation matrix is exactly singular
                // (np.linalg.solve would raise LinAlgError at deim/deim.py:547)
                if (v == 0.0 && j + 1 < m) {
                    stop_s = 1;
                    if (blockIdx.x == 0) *a.info = j + 1;
                }
            }
        }
        __syncthreads();
        if (stop_s) break;
    }
    // PT_U = U[dofs, :] (gather, exact) -- deim/deim.py:159, :212
    if (!stop_s) {
        for (int64_t e = (int64_t)blockIdx.x * kThreads + tid; e < (int64_t)m * m;
             e += (int64_t)gridDim.x * kThreads) {
            const int r = (int)(e / m), col = (int)(e % m);
            a.PT_U[e] = U[dofs_s[r] + (int64_t)col * ld];
        }
    }
}

// generic strided (Nh x m) -> column-major scratch with leading dimension ld
__global__ void k_to_colmajor(const double* __restrict__ src, int64_t Nh, int64_t m, int64_t sr,
                              int64_t sc, double* __restrict__ dst, int64_t ld) {
    __shared__ double tile[32][33];
    const int64_t i0 = (int64_t)blockIdx.x * 32, k0 = (int64_t)blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int64_t i = i0 + r, k = k0 + threadIdx.x;
        if (i < Nh && k < m) tile[r][threadIdx.x] = src[i * sr + k * sc];
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int64_t k = k0 + r, i = i0 + threadIdx.x;
        if (i < Nh && k < m) dst[i + k * ld] = tile[threadIdx.x][r];
    }
}

template <int R>
int launch_greedy(rt_handle* h, GreedyArgs& a, bool vec2, int grid, size_t smem, cudaStream_t st) {
    void* params[] = {&a};
    const void* fn = vec2 ? (const void*)k_deim_greedy<R, 2> : (const void*)k_deim_greedy<R, 1>;
    RT_CUDA(h, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    RT_CUDA(h, cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kThreads), params, smem, st));
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}

}  // namespace

extern "C" int rt_deim_greedy_f64(rt_handle* h, const double* U, int64_t Nh, int64_t m,
                                  int64_t stride_row, int64_t stride_col, int64_t* dofs,
                                  double* PT_U, int* info, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, U && dofs && PT_U && info);
    RT_REQUIRE(h, Nh >= 1 && m >= 1);
    RT_REQUIRE(h, m <= 160);  // LU of P^T U lives in shared memory (m*(m|1)*8 bytes)
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    RT_CUDA(h, cudaMemsetAsync(info, 0, sizeof(int), st));

    GreedyArgs a;
    a.Nh = Nh;
    a.m = (int)m;
    a.ldl = (int)m | 1;
    a.dofs = dofs;
    a.PT_U = PT_U;
    a.info = info;
    if (stride_row == 1 && stride_col >= Nh) {
        a.U = U;
        a.ld = stride_col;
    } else if (m == 1 && stride_row == 1) {
        a.U = U;
        a.ld = Nh;
    } else {
        // C-order (or otherwise strided) basis: one transposing copy into column-major scratch
        const int64_t ld = (Nh + 1) & ~int64_t(1);
        void* buf;
        int rc = rt_scratch(h, 0, (size_t)ld * m * sizeof(double), &buf);
        if (rc) return rc;
        dim3 g((unsigned)ceil_div64(Nh, 32), (unsigned)ceil_div64(m, 32));
        k_to_colmajor<<<g, dim3(32, 8), 0, st>>>(U, Nh, m, stride_row, stride_col, (double*)buf, ld);
        RT_LAUNCH_CHECK(h);
        a.U = (const double*)buf;
        a.ld = ld;
    }
    const bool vec2 = (reinterpret_cast<uintptr_t>(a.U) % 16 == 0) && (a.ld % 2 == 0);
    const int vec = vec2 ? 2 : 1;
    int grid = (int)ceil_div64(Nh, (int64_t)kThreads * vec);
    if (grid > h->sm_count) grid = h->sm_count;
    if (grid < 1) grid = 1;
    void* slots;
    int rc = rt_scratch(h, 1, (size_t)grid * 2 * (sizeof(double) + sizeof(int64_t)), &slots);
    if (rc) return rc;
    a.slot_val = (double*)slots;
    a.slot_idx = (int64_t*)((double*)slots + 2 * (size_t)grid);
    const size_t smem = ((size_t)a.m * a.ldl + a.m + 32) * sizeof(double) + (32 + a.m) * sizeof(int64_t);
    const int R = (int)((m + 31) / 32);
    if (R <= 1) return launch_greedy<1>(h, a, vec2, grid, smem, st);
    if (R <= 2) return launch_greedy<2>(h, a, vec2, grid, smem, st);
    if (R <= 4) return launch_greedy<4>(h, a, vec2, grid, smem, st);
    return launch_greedy<5>(h, a, vec2, grid, smem, st);
}
