// DEIM / MDEIM greedy interpolation-point selection (reference deim/deim.py:517-561).
//
// The DEIM points are the row pivots of Gaussian elimination with partial pivoting on the collateral
// basis U (Nh x m): the residual of step j,  res_j = u_j - U_{:j} (P^T U_{:j})^-1 P^T u_j,  is column j
// of the Schur complement.  One persistent cooperative kernel runs all m steps as a BLOCKED
// LEFT-LOOKING elimination (block width B <= 8 columns):
//
//   per block of columns J = [j0, j0+bw):
//     (A) warp l solves  C[:, l] = (P^T U_{:j0})^-1 P^T u_{j0+l}  with the bordered LU of P^T U kept in
//         shared memory (redundantly in every CTA: no extra grid barrier);
//     (B) every CTA streams ITS rows of U[:, :j0+bw] from HBM once (coalesced, column-major) and forms
//         the panel  W[:, l] = u_{j0+l} - U_{:j0} C[:, l]  in shared memory (rows are CTA-resident);
//     (C) bw right-looking steps inside the panel, all in shared memory: arg-max of |W[:, i]| with
//         numpy.argmax semantics (first maximum, NaN is maximal) -> per-CTA slot (value, row index and
//         the candidate's panel row) -> ONE grid barrier -> every CTA picks the winner, eliminates
//         W[:, l] -= W[:, i] * (W[p, l] / W[p, i]) for l > i, and warp 0 extends the bordered LU by the
//         new point while the other warps update the panel.
//
// HBM traffic: 8*Nh*(j0+bw) bytes per block, 8*Nh*m*(m/B+1)/2 in total -- B times less than the unblocked
// left-looking form (8*Nh*m(m+1)/2, what SURVEY.md 8d calls the algorithmic bytes); nothing is written.
// Every row goes through the same fixed sequence of FMAs, so bit-identical rows (the duplicated
// symmetric entries of an MDEIM basis) give bit-identical residuals and the lower index wins, as
// np.argmax does in the reference.  The bordered LU needs no pivoting of its own: the points ARE the
// GEPP pivots, so |L| <= 1 (as stable as dgesv on P^T U, deim/deim.py:545-547).
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace {

constexpr int kThreads = 512;  // 128 registers per thread: the slot exchange keeps 20 loads in flight per lane
constexpr unsigned kFull = 0xffffffffu;

struct GreedyArgs {
    const double* U;  // column-major, element (i,k) at U[i + k*ld]
    int64_t Nh;
    int64_t ld;
    int m;
    int ldl;          // leading dimension of the LU workspace (odd -> conflict-free column access)
    int64_t rpc;      // rows per CTA (contiguous ownership)
    int64_t* dofs;
    double* PT_U;
    int* info;
    unsigned* counter;          // grid barrier counter, zeroed by the host before the launch
    unsigned long long* slots;  // [2][gridDim.x][kSlotWords] LL words: key, row index, panel row of each CTA's candidate
    double* Wglobal;    // panel in global memory (only when it does not fit in shared memory), else null
    int64_t ldw;        // leading dimension of the panel (rows per CTA when in shared memory)
};

// numpy.argmax ordering on |x| as an unsigned 64-bit key: for non-negative doubles the IEEE bit pattern is
// monotone, NaN (canonicalised) is larger than everything, and ties go to the LOWER row index.
constexpr unsigned long long kNanKey = 0x7ff8000000000000ull;
__device__ __forceinline__ unsigned long long abs_key(double x) {
    return (x != x) ? kNanKey : (unsigned long long)__double_as_longlong(fabs(x));
}
__device__ __forceinline__ void key_merge(unsigned long long& k, int64_t& i, unsigned long long ok, int64_t oi) {
    if (ok > k || (ok == k && oi < i)) { k = ok; i = oi; }
}
// warp-wide (max key, then min index) with four redux.sync instead of a shuffle tree
__device__ __forceinline__ void warp_key_reduce(unsigned long long& k, int64_t& i) {
    const unsigned hi = (unsigned)(k >> 32), lo = (unsigned)k;
    const unsigned mh = __reduce_max_sync(kFull, hi);
    const unsigned ml = __reduce_max_sync(kFull, hi == mh ? lo : 0u);
    const bool win = (hi == mh) && (lo == ml);
    const unsigned ih = win ? (unsigned)((unsigned long long)i >> 32) : 0xffffffffu;
    const unsigned mih = __reduce_min_sync(kFull, ih);
    const unsigned il = (win && ih == mih) ? (unsigned)(unsigned long long)i : 0xffffffffu;
    const unsigned mil = __reduce_min_sync(kFull, il);
    k = ((unsigned long long)mh << 32) | ml;
    i = (int64_t)(((unsigned long long)mih << 32) | mil);
}

// Grid-wide exchange of the per-CTA candidates without a separate barrier ("LL" protocol): every 64-bit word
// carries 32 bits of payload and the 32-bit step number; a reader spins on the word itself until the step
// number matches, so the data IS the flag (one L2 round trip, no fence, no atomic).  64-bit stores are
// single-copy atomic.  The slot area is zeroed by the host before the launch and steps count from 1.
constexpr int kSlotWords = 24;  // key (2) + idx (2) + panel row (16), padded
__device__ __forceinline__ void ll_store(unsigned long long* p, unsigned data, unsigned step) {
    const unsigned long long w = ((unsigned long long)data << 32) | step;
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
// grid-wide arrive / wait on a monotone counter (all CTAs are co-resident: cooperative launch).  One lane
// per CTA polls: 148 pollers on one L2 line instead of 148 x 32 lanes spinning on the slots themselves.
__device__ __forceinline__ void grid_arrive(unsigned* ctr) {
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
}
__device__ __forceinline__ void grid_wait(const unsigned* ctr, unsigned target) {
    unsigned v;
    do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
    } while ((int)(v - target) < 0);
}

// one 16-byte load of two LL words; returns false while either word still belongs to an older step
__device__ __forceinline__ bool ll_try_pair(const unsigned long long* p, unsigned step, unsigned long long& out) {
    unsigned long long w0, w1;
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(p) : "memory");
    out = (w0 & 0xffffffff00000000ull) | (w1 >> 32);
    return ((unsigned)w0 == step) && ((unsigned)w1 == step);
}

// Warp 0 of every CTA: grid barrier, then (max key, min row index) over all CTAs' candidates and the panel
// row of the winner (lane l < 8 returns W_l[p]).  All slot loads of a lane are in flight at once: one L2 round trip.
__device__ __forceinline__ void grid_exchange(unsigned* counter, const unsigned long long* all, unsigned epoch, int64_t rpc,
                                           int lane, unsigned long long& k, int64_t& idx, double& pr) {
    __syncwarp();
    if (lane == 0) {
        grid_arrive(counter);
        grid_wait(counter, epoch * gridDim.x);
    }
    __syncwarp();
    constexpr int kMaxSlotsPerLane = 5;  // grid <= 160 CTAs (host clamps)
    unsigned long long ck[kMaxSlotsPerLane], ci[kMaxSlotsPerLane];
    bool ready;
    do {
        ready = true;
#pragma unroll
        for (int q = 0; q < kMaxSlotsPerLane; ++q) {
            const unsigned s = lane + 32u * q;
            if (s < gridDim.x) {
                const bool r0k = ll_try_pair(all + (size_t)s * kSlotWords, epoch, ck[q]);
                const bool r1k = ll_try_pair(all + (size_t)s * kSlotWords + 2, epoch, ci[q]);
                ready = ready && r0k && r1k;
            }
        }
    } while (!ready);
    k = 0ull;
    idx = INT64_MAX;
#pragma unroll
    for (int q = 0; q < kMaxSlotsPerLane; ++q)
        if (lane + 32u * q < gridDim.x) key_merge(k, idx, ck[q], (int64_t)ci[q]);
    warp_key_reduce(k, idx);
    const unsigned long long* own = all + (size_t)(idx / rpc) * kSlotWords;
    pr = 0.0;
    if (lane < 8) {
        unsigned long long bits;
        while (!ll_try_pair(own + 4 + 2 * lane, epoch, bits)) {}
        pr = __longlong_as_double((long long)bits);
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
__device__ __forceinline__ void solve_upper_T(const double* LU, const double* rdiag, int ldl, int n, double (&v)[R], int lane) {
    // l^T Rm = v^T  <=>  Rm^T l = v  (Rm = upper part of LU incl. diagonal), forward substitution
#pragma unroll
    for (int kr = 0; kr < R; ++kr) {
        for (int kl = 0; kl < 32; ++kl) {
            const int k = kr * 32 + kl;
            if (k >= n) break;
            double lk = __shfl_sync(kFull, v[kr], kl) * rdiag[k];
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
__device__ __forceinline__ void solve_upper(const double* LU, const double* rdiag, int ldl, int n, double (&v)[R], int lane) {
    // Rm c = v, backward substitution
#pragma unroll
    for (int kr = R - 1; kr >= 0; --kr) {
        for (int kl = 31; kl >= 0; --kl) {
            const int k = kr * 32 + kl;
            if (k >= n) continue;
            double ck = __shfl_sync(kFull, v[kr], kl) * rdiag[k];
            if (lane == kl) v[kr] = ck;
#pragma unroll
            for (int er = 0; er <= kr; ++er) {
                const int e = er * 32 + lane;
                if (e < k) v[er] = fma(-ck, LU[(size_t)e * ldl + k], v[er]);
            }
        }
    }
}


// ---------------------------------------------------------------------------------------------
template <int R, int B>
__global__ void __launch_bounds__(kThreads, 1) k_deim_greedy(GreedyArgs a) {
        extern __shared__ __align__(16) unsigned char smem_raw[];
    double* LU = reinterpret_cast<double*>(smem_raw);            // m * ldl
    double* Cs = LU + (size_t)a.m * a.ldl;                       // m * B   coefficients of the current block
    double* rdiag_s = Cs + (size_t)a.m * B;                      // m       1 / Rm[k, k]
    unsigned long long* red_k = reinterpret_cast<unsigned long long*>(rdiag_s + a.m);  // 32
    double* prow_s = reinterpret_cast<double*>(red_k + 32);      // 8       panel row of the winning point
    int64_t* red_i = reinterpret_cast<int64_t*>(prow_s + 8);     // 32
    int64_t* dofs_s = red_i + 32;                                // m
    double* Wsm = reinterpret_cast<double*>(dofs_s + a.m);       // B * ldw (when the panel is in shared memory)
    __shared__ int stop_s;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int m = a.m, ldl = a.ldl;
    const int64_t Nh = a.Nh, ld = a.ld;
    const double* __restrict__ U = a.U;
    const int64_t r0 = (int64_t)blockIdx.x * a.rpc;
    const int nrows = (int)max((int64_t)0, min(a.rpc, Nh - r0));
    const int64_t ldw = a.ldw;
    double* W = a.Wglobal ? a.Wglobal + r0 : Wsm;  // W[l * ldw + i_local]
    unsigned epoch = 0;  // grid barriers passed so far (identical in every CTA)
    if (tid == 0) stop_s = 0;
    __syncthreads();

    for (int j0 = 0; j0 < m; j0 += B) {
        const int bw = min(B, m - j0);
        // ---------------- (A) coefficients of the block columns: warp l -> column j0 + l ----------------
        if (j0 > 0 && warp < bw) {
            double v[R];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int e = r * 32 + lane;
                v[r] = (e < j0) ? U[dofs_s[e] + (int64_t)(j0 + warp) * ld] : 0.0;
            }
            solve_unit_lower<R>(LU, ldl, j0, v, lane);  // y = L^-1 P^T u : rows 0..j0-1 of column j0+warp of Rm
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int e = r * 32 + lane;
                if (e < j0) LU[(size_t)e * ldl + (j0 + warp)] = v[r];
            }
            solve_upper<R>(LU, rdiag_s, ldl, j0, v, lane);       // c = Rm^-1 y   (reads LU[:j0, :j0] only)
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int e = r * 32 + lane;
                if (e < j0) Cs[(size_t)e * B + warp] = v[r];
            }
        }
        if (j0 > 0 && bw < B)  // ragged last block: unused coefficient columns are zero
            for (int e = tid; e < j0 * B; e += kThreads)
                if ((e % B) >= bw) Cs[e] = 0.0;
        __syncthreads();

        // ---------------- (B) panel  W[:, l] = u_{j0+l} - U_{:j0} C[:, l]  for this CTA's rows ----------------
        for (int il = tid; il < nrows; il += kThreads) {
            const double* __restrict__ row = U + r0 + il;
            double acc[B];
#pragma unroll
            for (int l = 0; l < B; ++l) acc[l] = (l < bw) ? row[(int64_t)(j0 + l) * ld] : 0.0;
            int k = 0;
            for (; k + 8 <= j0; k += 8) {
                double x[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) x[q] = __ldg(row + (int64_t)(k + q) * ld);
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const double* c = Cs + (size_t)(k + q) * B;
#pragma unroll
                    for (int l = 0; l < B; ++l) acc[l] = fma(-c[l], x[q], acc[l]);
                }
            }
            for (; k < j0; ++k) {
                const double x = __ldg(row + (int64_t)k * ld);
                const double* c = Cs + (size_t)k * B;
#pragma unroll
                for (int l = 0; l < B; ++l) acc[l] = fma(-c[l], x, acc[l]);
            }
#pragma unroll
            for (int l = 0; l < B; ++l)
                if (l < bw) W[(size_t)l * ldw + il] = acc[l];
        }
        // (each thread reads back only the panel rows it wrote itself until the CTA-level reduction below)

        // ---------------- (C) bw right-looking steps inside the panel ----------------
        for (int i = 0; i < bw; ++i) {
            const int j = j0 + i;
            unsigned long long best_k = 0ull;
            int64_t best_i = INT64_MAX;
            const double* wi = W + (size_t)i * ldw;
            for (int il = tid; il < nrows; il += kThreads) {
                const unsigned long long k = abs_key(wi[il]);
                if (k > best_k || best_i == INT64_MAX) { best_k = k; best_i = r0 + il; }
            }
            warp_key_reduce(best_k, best_i);
            if (lane == 0) { red_k[warp] = best_k; red_i[warp] = best_i; }
            __syncthreads();
            if (warp == 0) {
                unsigned long long k = lane < kThreads / 32 ? red_k[lane] : 0ull;
                int64_t idx = lane < kThreads / 32 ? red_i[lane] : INT64_MAX;
                warp_key_reduce(k, idx);
                // ---- publish this CTA's candidate (value key, row index, its panel row) as LL words ----
                ++epoch;
                unsigned long long* mine = a.slots + ((size_t)(j & 1) * gridDim.x + blockIdx.x) * kSlotWords;
                {
                    unsigned long long payload = 0ull;
                    if (lane < 2) payload = k;
                    else if (lane < 4) payload = (unsigned long long)idx;
                    else if (lane < 20) {
                        const int l = (lane - 4) >> 1;
                        const double wv = (l < bw && idx != INT64_MAX) ? W[(size_t)l * ldw + (idx - r0)] : 0.0;
                        payload = (unsigned long long)__double_as_longlong(wv);
                    }
                    if (lane < 20) ll_store(mine + lane, (lane & 1) ? (unsigned)payload : (unsigned)(payload >> 32), epoch);
                }
                double pr;
                grid_exchange(a.counter, a.slots + (size_t)(j & 1) * gridDim.x * kSlotWords, epoch, a.rpc, lane, k, idx, pr);
                if (lane < 8) prow_s[lane] = pr;
                // row j of the bordered LU inside the block: L[j, j0+l] = W_l[p] / rho_l (l < i), Rm[j, j0+l] = W_l[p] (l >= i)
                if (lane < bw)
                    LU[(size_t)j * ldl + (j0 + lane)] = (lane >= i) ? pr : pr * rdiag_s[j0 + lane];
                if (lane == i) rdiag_s[j] = 1.0 / pr;
                if (lane == 0) {
                    dofs_s[j] = idx;
                    if (blockIdx.x == 0) a.dofs[j] = idx;
                    // residual identically zero: the next interpolation matrix is exactly singular
                    // (np.linalg.solve would raise LinAlgError at deim/deim.py:547)
                    if (k == 0ull && j + 1 < m) {
                        stop_s = 1;
                        if (blockIdx.x == 0) *a.info = j + 1;
                    }
                }
            } else {
                ++epoch;
            }
            __syncthreads();
            if (stop_s) break;
            if (i + 1 < bw) {
                const double rho = prow_s[i];
                double f[B];
#pragma unroll
                for (int l = 0; l < B; ++l) f[l] = (l > i && l < bw) ? prow_s[l] / rho : 0.0;
                for (int il = tid; il < nrows; il += kThreads) {
                    const double wv = wi[il];
#pragma unroll
                    for (int l = 1; l < B; ++l)
                        if (l > i && l < bw) W[(size_t)l * ldw + il] = fma(-f[l], wv, W[(size_t)l * ldw + il]);
                }
            }
        }
        if (stop_s) break;
        // ---------------- rows of L for the bw new points (columns 0..j0-1): warp i -> point j0 + i ----------------
        if (j0 > 0 && j0 + bw < m && warp < bw) {
            const int64_t p = dofs_s[j0 + warp];
            double v[R];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int e = r * 32 + lane;
                v[r] = (e < j0) ? U[p + (int64_t)e * ld] : 0.0;
            }
            solve_upper_T<R>(LU, rdiag_s, ldl, j0, v, lane);  // l^T = U[p, :j0] Rm_{j0}^-1
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int e = r * 32 + lane;
                if (e < j0) LU[(size_t)(j0 + warp) * ldl + e] = v[r];
            }
        }
        __syncthreads();
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

size_t greedy_smem(int m, int ldl, int B, int64_t panel_ld) {
    return ((size_t)m * ldl + (size_t)m * B + m + 32 + 8) * sizeof(double) + (size_t)(32 + m) * sizeof(int64_t) +
           (size_t)B * panel_ld * sizeof(double);
}

template <int R, int B>
int launch_greedy(rt_handle* h, GreedyArgs& a, int grid, size_t smem, cudaStream_t st) {
    void* params[] = {&a};
    const void* fn = (const void*)k_deim_greedy<R, B>;
    RT_CUDA(h, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    RT_CUDA(h, cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kThreads), params, smem, st));
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}

template <int B>
int launch_greedy_r(rt_handle* h, GreedyArgs& a, int grid, size_t smem, cudaStream_t st) {
    const int R = (a.m + 31) / 32;
    if (R <= 1) return launch_greedy<1, B>(h, a, grid, smem, st);
    if (R <= 2) return launch_greedy<2, B>(h, a, grid, smem, st);
    if (R <= 4) return launch_greedy<4, B>(h, a, grid, smem, st);
    return launch_greedy<5, B>(h, a, grid, smem, st);
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
    memset(&a, 0, sizeof(a));
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
    // contiguous row ownership: one CTA per SM, at least 256 rows per CTA
    int grid = (int)ceil_div64(Nh, 256);
    if (grid > h->sm_count) grid = h->sm_count;
    if (grid > 160) grid = 160;  // kMaxSlotsPerLane * 32
    if (grid < 1) grid = 1;
    a.rpc = ceil_div64(Nh, grid);
    grid = (int)ceil_div64(Nh, a.rpc);  // every CTA owns at least one row
    const size_t smem_max = (size_t)h->smem_optin - 1024;
    int B = 0;
    if (greedy_smem(a.m, a.ldl, 8, a.rpc) <= smem_max) B = 8;
    else if (greedy_smem(a.m, a.ldl, 4, a.rpc) <= smem_max) B = 4;
    size_t smem;
    void* slots;
    const size_t slot_bytes = 256 + (size_t)grid * 2 * kSlotWords * sizeof(unsigned long long);
    size_t wbytes = 0;
    if (B == 0) {  // panel too tall for shared memory: keep it in global memory (L2-resident), 8 columns wide
        B = 8;
        a.ldw = (Nh + 1) & ~int64_t(1);
        wbytes = (size_t)a.ldw * 8 * sizeof(double);
        smem = greedy_smem(a.m, a.ldl, 8, 0);
    } else {
        a.ldw = a.rpc;
        smem = greedy_smem(a.m, a.ldl, B, a.rpc);
    }
    int rc = rt_scratch(h, 1, slot_bytes + 256 + wbytes, &slots);
    if (rc) return rc;
    a.counter = (unsigned*)slots;
    a.slots = (unsigned long long*)((char*)slots + 256);
    RT_CUDA(h, cudaMemsetAsync(slots, 0, slot_bytes, st));  // step number 0 == "nothing published yet"
    if (wbytes) a.Wglobal = (double*)(((uintptr_t)((char*)slots + slot_bytes) + 255) & ~uintptr_t(255));
    if (B == 8) return launch_greedy_r<8>(h, a, grid, smem, st);
    return launch_greedy_r<4>(h, a, grid, smem, st);
}
