// Certification after a sweep (reference rom/hrom.py:504-626):
//   rt_rom_difference_f64 : S-ROM estimator  err[b][n] = || V_srom (uN_srom - pad(uN)) ||_2 / sqrt(Nh)
//                           for every (parameter, step)                       (utils.py:173-212)
//   rt_errors_summary_f64 : mean / median / max / min over the steps of every parameter (rom/base.py:154-163)
// The estimator is evaluated as the quadratic form sqrt(d^T G d / Nh), G = V_srom^T V_srom (Nhat x Nhat, from the
// Gram kernel): the Nh-long linear combination of the reference is never formed.  2 Nhat^2 flop against
// 8 (N + Nhat) bytes per (parameter, step) sits just above the FP64 ridge of a B200, so the rows go through
// mma.sync.m8n8k4.f64: Y^T = G D^T with warp w owning the 8 rows [8w, 8w + 8) of G as A fragments in registers for
// the whole kernel, 32 difference vectors per CTA iteration staged in shared memory as the B operand.
#include "common.cuh"

namespace {

__device__ __forceinline__ void dmma884c(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

constexpr int kRowsPerIter = 32;
constexpr int kMaxK4 = 32;  // Nhat <= 128
constexpr int kStages = 3;

__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(void* dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}

// grid: persistent; block: 32 * NT threads, NT = Nhat_pad / 8 warps.  Shared memory: kStages x {S tile, U tile}
// (32 rows x LD each, pad columns zeroed once and never written again) + the partial sums.  The rows stream in by
// cp.async, two stages ahead of the one being multiplied; d = S - U is formed when the fragments are read.
__global__ void __launch_bounds__(512) k_rom_difference(const double* __restrict__ uN, const double* __restrict__ uS, int64_t R,
                                                        int N, int Nhat, int Np, int LD, const double* __restrict__ G,
                                                        double inv_Nh, double* __restrict__ err) {
    extern __shared__ __align__(16) double sm_c[];
    const int tile = kRowsPerIter * LD;
    double* part = sm_c + 2 * kStages * tile;  // NT x kRowsPerIter partial quadratic forms
    const int t = threadIdx.x, lane = t & 31, w = t >> 5, NT = blockDim.x >> 5;
    const int g8 = lane >> 2, q = lane & 3;
    const int K4 = Np >> 2;
    for (int e = t; e < 2 * kStages * tile; e += blockDim.x) sm_c[e] = 0.0;
    // A fragments of G rows [8w, 8w + 8): a[ks] = G[8w + g8][4 ks + q], zero beyond Nhat
    double a[kMaxK4];
#pragma unroll
    for (int ks = 0; ks < kMaxK4; ++ks) {
        const int i = 8 * w + g8, j = 4 * ks + q;
        a[ks] = (ks < K4 && i < Nhat && j < Nhat) ? G[(int64_t)i * Nhat + j] : 0.0;
    }
    __syncthreads();
    const bool wide = ((Nhat | N) & 1) == 0;
    const int64_t stride = (int64_t)gridDim.x * kRowsPerIter;
    auto issue = [&](int64_t r0, int stage) {
        double* Ss = sm_c + (2 * stage) * tile;
        double* Us = Ss + tile;
        if (r0 < R) {
            if (wide) {
                const int cs = Nhat >> 1, cu = N >> 1;
                for (int e = t; e < kRowsPerIter * cs; e += blockDim.x) {
                    const int r = e / cs, c = e - r * cs;
                    if (r0 + r < R) cp_async16(Ss + r * LD + 2 * c, uS + (r0 + r) * Nhat + 2 * c);
                }
                for (int e = t; e < kRowsPerIter * cu; e += blockDim.x) {
                    const int r = e / cu, c = e - r * cu;
                    if (r0 + r < R) cp_async16(Us + r * LD + 2 * c, uN + (r0 + r) * N + 2 * c);
                }
            } else {
                for (int e = t; e < kRowsPerIter * Nhat; e += blockDim.x) {
                    const int r = e / Nhat, c = e - r * Nhat;
                    if (r0 + r < R) cp_async8(Ss + r * LD + c, uS + (r0 + r) * Nhat + c);
                }
                for (int e = t; e < kRowsPerIter * N; e += blockDim.x) {
                    const int r = e / N, c = e - r * N;
                    if (r0 + r < R) cp_async8(Us + r * LD + c, uN + (r0 + r) * N + c);
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int64_t r0 = (int64_t)blockIdx.x * kRowsPerIter;
#pragma unroll
    for (int s = 0; s < kStages - 1; ++s) issue(r0 + s * stride, s);
    int stage = 0;
    for (; r0 < R; r0 += stride) {
        asm volatile("cp.async.wait_group %0;" ::"n"(kStages - 2) : "memory");
        __syncthreads();  // this stage has landed for every thread; the stage consumed last iteration is free
        issue(r0 + (kStages - 1) * stride, (stage + kStages - 1) % kStages);
        const double* Ss = sm_c + (2 * stage) * tile;
        const double* Us = Ss + tile;
        double c[kRowsPerIter / 8][2];  // Y^T[8w + g8][rows 8 rt + 2q, + 1]: four independent accumulator chains
#pragma unroll
        for (int rt = 0; rt < kRowsPerIter / 8; ++rt) c[rt][0] = c[rt][1] = 0.0;
        const int boff = g8 * LD + q;
#pragma unroll
        for (int ks = 0; ks < kMaxK4; ++ks) {
            if (ks < K4) {
#pragma unroll
                for (int rt = 0; rt < kRowsPerIter / 8; ++rt) {
                    const int o = boff + 8 * rt * LD + 4 * ks;
                    dmma884c(c[rt][0], c[rt][1], a[ks], Ss[o] - Us[o]);
                }
            }
        }
        const int n = 8 * w + g8;
#pragma unroll
        for (int rt = 0; rt < kRowsPerIter / 8; ++rt) {
            const int o0 = (8 * rt + 2 * q) * LD + n, o1 = o0 + LD;
            double p0 = c[rt][0] * (Ss[o0] - Us[o0]), p1 = c[rt][1] * (Ss[o1] - Us[o1]);
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) {
                p0 += __shfl_xor_sync(0xffffffffu, p0, o);
                p1 += __shfl_xor_sync(0xffffffffu, p1, o);
            }
            if (g8 == 0) {
                part[w * kRowsPerIter + 8 * rt + 2 * q] = p0;
                part[w * kRowsPerIter + 8 * rt + 2 * q + 1] = p1;
            }
        }
        __syncthreads();  // `part` is complete
        if (t < kRowsPerIter && r0 + t < R) {
            double s = 0.0;
            for (int ww = 0; ww < NT; ++ww) s += part[ww * kRowsPerIter + t];  // fixed order: deterministic
            err[r0 + t] = sqrt(fmax(s, 0.0) * inv_Nh);
        }
        // `part` is rewritten only after the next iteration's first barrier, which its 32 readers reach after reading it
        stage = (stage + 1) % kStages;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// one CTA per parameter: sort the nt errors in shared memory (bitonic, padded with +inf), then
// summary[b] = {mean, median (numpy: mean of the two middle values for even nt), max, min}
__global__ void __launch_bounds__(256) k_errors_summary(const double* __restrict__ err, int nt, int np2,
                                                        double* __restrict__ summary) {
    extern __shared__ __align__(16) double sm_s[];
    double* v = sm_s;          // np2 sorted values
    double* red = sm_s + np2;  // 256 partial sums
    const int t = threadIdx.x;
    const double* e = err + (int64_t)blockIdx.x * nt;
    double s = 0.0;
    for (int i = t; i < np2; i += 256) {
        const double x = i < nt ? e[i] : INFINITY;
        v[i] = x;
        if (i < nt) s += x;
    }
    red[t] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {  // fixed tree: deterministic
        if (t < o) red[t] += red[t + o];
        __syncthreads();
    }
    for (int k = 2; k <= np2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = t; i < np2; i += 256) {
                const int l = i ^ j;
                if (l > i) {
                    const double x = v[i], y = v[l];
                    const bool up = (i & k) == 0;
                    if ((x > y) == up) { v[i] = y; v[l] = x; }
                }
            }
            __syncthreads();
        }
    }
    if (t == 0) {
        double* o = summary + (int64_t)blockIdx.x * 4;
        o[0] = red[0] / (double)nt;
        o[1] = 0.5 * (v[(nt - 1) / 2] + v[nt / 2]);
        o[2] = v[nt - 1];
        o[3] = v[0];
    }
}

}  // namespace

extern "C" int rt_rom_difference_f64(rt_handle* h, const double* uN, int64_t B, int64_t nt, int64_t N, const double* uN_srom,
                                     int64_t Nhat, const double* G, int64_t Nh, double* err, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, uN && uN_srom && G && err);
    RT_REQUIRE(h, B >= 1 && nt >= 1 && N >= 1 && Nhat >= N && Nhat <= 128 && Nh >= 1);
    RT_CUDA(h, cudaSetDevice(h->device));
    const int Np = (int)((Nhat + 7) / 8 * 8);
    const int LD = (Np + 15) / 16 * 16 + 4;  // = 4 mod 16: conflict-free B-fragment reads
    const int NT = Np / 8;
    const size_t smem = ((size_t)2 * kStages * kRowsPerIter * LD + (size_t)NT * kRowsPerIter) * sizeof(double);
    const int64_t R = B * nt;
    const int64_t iters = ceil_div64(R, kRowsPerIter);
    if (smem > 48 * 1024) RT_CUDA(h, cudaFuncSetAttribute(k_rom_difference, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int ctas_per_sm = 1;
    RT_CUDA(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, k_rom_difference, 32 * NT, smem));
    if (ctas_per_sm < 1) ctas_per_sm = 1;
    const int64_t grid = iters < (int64_t)h->sm_count * ctas_per_sm ? iters : (int64_t)h->sm_count * ctas_per_sm;
    k_rom_difference<<<(unsigned)grid, 32 * NT, smem, rt_stream(stream)>>>(uN, uN_srom, R, (int)N, (int)Nhat, Np, LD, G,
                                                                          1.0 / (double)Nh, err);
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}

extern "C" int rt_errors_summary_f64(rt_handle* h, const double* err, int64_t B, int64_t nt, double* summary, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, err && summary && B >= 1 && nt >= 1 && nt <= 16384);
    RT_CUDA(h, cudaSetDevice(h->device));
    int np2 = 1;
    while (np2 < nt) np2 <<= 1;
    const size_t smem = ((size_t)np2 + 256) * sizeof(double);
    if (smem > 48 * 1024)
        RT_CUDA(h, cudaFuncSetAttribute(k_errors_summary, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_errors_summary<<<(unsigned)B, 256, smem, rt_stream(stream)>>>(err, (int)nt, np2, summary);
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}
