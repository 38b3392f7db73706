// Certification after a sweep (reference rom/hrom.py:504-626):
//   rt_rom_difference_f64 : S-ROM estimator  err[b][n] = || V_srom (uN_srom - pad(uN)) ||_2 / sqrt(Nh)
//                           for every (parameter, step)                       (utils.py:173-212)
//   rt_errors_summary_f64 : mean / median / max / min over the steps of every parameter (rom/base.py:154-163)
// The estimator is evaluated as || R d ||_2 / sqrt(Nh) with R = chol(V_srom^T V_srom) (Nhat x Nhat, upper; the Gram
// matrix comes from the Gram kernel): the Nh-long linear combination of the reference is never formed.  About
// Nhat^2 flop against 8 (N + Nhat + 1) bytes per (parameter, step) is below the FP64 ridge of a B200: the kernel is an
// HBM stream with the small triangular products on mma.sync.m8n8k4.f64.
#include "common.cuh"

namespace {

__device__ __forceinline__ void dmma884c(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// R = chol(G) (upper, R^T R = G; a non-positive pivot zeroes its row: semi-definite G from a rank-deficient basis),
// written in A-fragment order  Rf[(nt * K4 + ks) * 32 + lane] = R[8 nt + lane / 4][kperm(ks, lane % 4)], zero padded to Np;
// kperm(ks, q) = 8 (ks / 2) + 2 q + ks % 2: the k index is permuted (in R and in d alike) so that a lane's fragments of two
// consecutive k-steps are ADJACENT columns of its row, i.e. one 16-byte load
__global__ void __launch_bounds__(128) k_chol_frag(const double* __restrict__ G, int n, int Np, double* __restrict__ Rf) {
    extern __shared__ __align__(16) double sm_a[];
    const int ld = n + 1, t = threadIdx.x;
    for (int e = t; e < n * n; e += 128) sm_a[(e / n) * ld + e % n] = G[e];
    __syncthreads();
    for (int j = 0; j < n; ++j) {
        const double piv = sm_a[j * ld + j];
        const double rd = piv > 0.0 ? 1.0 / sqrt(piv) : 0.0;
        __syncthreads();
        for (int k = j + t; k < n; k += 128) sm_a[j * ld + k] *= rd;  // row j of R (zero row if the pivot vanished)
        __syncthreads();
        for (int e = t; e < (n - j - 1) * (n - j - 1); e += 128) {
            const int i = j + 1 + e / (n - j - 1), k = j + 1 + e % (n - j - 1);
            if (k >= i) sm_a[i * ld + k] -= sm_a[j * ld + i] * sm_a[j * ld + k];
        }
        __syncthreads();
    }
    const int K4 = Np / 4;
    for (int e = t; e < (Np / 8) * K4 * 32; e += 128) {
        const int lane = e & 31, ks = (e >> 5) % K4, nt = (e >> 5) / K4;
        const int i = 8 * nt + (lane >> 2), k = 8 * (ks >> 1) + 2 * (lane & 3) + (ks & 1);  // k-permutation, see below
        Rf[e] = (i < n && k < n && k >= i) ? sm_a[i * ld + k] : 0.0;
    }
}

// err[row] = || R d_row ||_2 / sqrt(Nh),  d = uS[row] - pad(uN[row]).  Every warp works alone on RT row tiles of 8 rows
// per iteration: the difference vectors go from global memory straight into B fragments (lane (g8, q) holds
// d[8 rt + g8][kperm(ks, q)]: full sectors, RT * K4 16-byte loads in flight per lane), R comes from shared
// memory in A-fragment order (one conflict-free LDS.64 feeds RT DMMAs; only the k-tiles of the upper triangle are
// visited), the squared entries of Y^T = R D^T are summed per row in registers and reduced over the 8 lanes that
// share a row pair.  No block barrier after the set-up.
template <int K4MAX, int RT>
__global__ void __launch_bounds__(128) k_rom_difference(const double* __restrict__ uN, const double* __restrict__ uS, int64_t R,
                                                        int N, int Nhat, int Np, const double* __restrict__ Rf,
                                                        double inv_Nh, double* __restrict__ err) {
    extern __shared__ __align__(16) double sm_c[];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int g8 = lane >> 2, q = lane & 3;
    const int K4 = Np >> 2, NT8 = Np >> 3;
    for (int e = t; e < NT8 * K4 * 32; e += blockDim.x) sm_c[e] = Rf[e];
    __syncthreads();
    const bool wide = ((Nhat | N) & 1) == 0;  // 16-byte aligned rows
    const int64_t rows_per_iter = 8 * RT;
    const int64_t warp_id = (int64_t)blockIdx.x * (blockDim.x >> 5) + w, nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t r0 = warp_id * rows_per_iter; r0 < R; r0 += nwarps * rows_per_iter) {
        double fb[RT][K4MAX];
        if (wide) {
            // fragments (2j, q) and (2j + 1, q) are the adjacent columns 8 j + 2 q, + 1: one LDG.128 each for uS and uN
            double2 us[RT][K4MAX / 2], un[RT][K4MAX / 2];
#pragma unroll
            for (int rt = 0; rt < RT; ++rt) {
                const int64_t r = r0 + 8 * rt + g8;
#pragma unroll
                for (int j = 0; j < K4MAX / 2; ++j) {
                    const int k = 8 * j + 2 * q;
                    us[rt][j] = (2 * j < K4 && r < R && k < Nhat) ? __ldcs(reinterpret_cast<const double2*>(uS + r * Nhat + k))
                                                                  : make_double2(0.0, 0.0);
                    un[rt][j] = (2 * j < K4 && r < R && k < N) ? __ldcs(reinterpret_cast<const double2*>(uN + r * N + k))
                                                               : make_double2(0.0, 0.0);
                }
            }
#pragma unroll
            for (int rt = 0; rt < RT; ++rt)
#pragma unroll
                for (int j = 0; j < K4MAX / 2; ++j) {
                    fb[rt][2 * j] = us[rt][j].x - un[rt][j].x;
                    fb[rt][2 * j + 1] = us[rt][j].y - un[rt][j].y;
                }
        } else {
#pragma unroll
            for (int rt = 0; rt < RT; ++rt) {
                const int64_t r = r0 + 8 * rt + g8;
#pragma unroll
                for (int ks = 0; ks < K4MAX; ++ks) {
                    const int k = 8 * (ks >> 1) + 2 * q + (ks & 1);
                    double v = 0.0;
                    if (ks < K4 && r < R && k < Nhat) {
                        v = __ldcs(uS + r * Nhat + k);
                        if (k < N) v -= __ldcs(uN + r * N + k);
                    }
                    fb[rt][ks] = v;
                }
            }
        }
        double s0[RT], s1[RT];  // rows 8 rt + 2q and 8 rt + 2q + 1: sum over n of Y[row][n]^2 (this lane's n = 8 nt + g8)
#pragma unroll
        for (int rt = 0; rt < RT; ++rt) s0[rt] = s1[rt] = 0.0;
#pragma unroll
        for (int nt = 0; nt < K4MAX / 2; ++nt) {
            if (nt < NT8) {
                double c[RT][2];
#pragma unroll
                for (int rt = 0; rt < RT; ++rt) c[rt][0] = c[rt][1] = 0.0;
#pragma unroll
                for (int ks = 2 * nt; ks < K4MAX; ++ks) {  // R is upper triangular: k-tiles left of the diagonal are zero
                    if (ks < K4) {
                        const double a = sm_c[(nt * K4 + ks) * 32 + lane];
#pragma unroll
                        for (int rt = 0; rt < RT; ++rt) dmma884c(c[rt][0], c[rt][1], a, fb[rt][ks]);
                    }
                }
#pragma unroll
                for (int rt = 0; rt < RT; ++rt) { s0[rt] = fma(c[rt][0], c[rt][0], s0[rt]); s1[rt] = fma(c[rt][1], c[rt][1], s1[rt]); }
            }
        }
#pragma unroll
        for (int rt = 0; rt < RT; ++rt) {
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) {
                s0[rt] += __shfl_xor_sync(0xffffffffu, s0[rt], o);
                s1[rt] += __shfl_xor_sync(0xffffffffu, s1[rt], o);
            }
            if (g8 == 0) {
                const int64_t r = r0 + 8 * rt + 2 * q;
                if (r < R) err[r] = sqrt(s0[rt] * inv_Nh);
                if (r + 1 < R) err[r + 1] = sqrt(s1[rt] * inv_Nh);
            }
        }
    }
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
            for (int pr = t; pr < (np2 >> 1); pr += 256) {  // one compare-exchange pair per thread and pass
                const int i = ((pr & ~(j - 1)) << 1) | (pr & (j - 1)), l = i | j;
                const double x = v[i], y = v[l];
                const bool up = (i & k) == 0;
                if ((x > y) == up) { v[i] = y; v[l] = x; }
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

template <int K4MAX, int RT>
static int launch_rom_difference(rt_handle* h, const double* uN, const double* uS, int64_t R, int N, int Nhat, int Np,
                                 const double* Rf, double inv_Nh, double* err, cudaStream_t st) {
    auto fn = k_rom_difference<K4MAX, RT>;
    const size_t smem = (size_t)Np * Np * sizeof(double);
    if (smem > 48 * 1024) RT_CUDA(h, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    RT_CUDA(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 128, smem));
    if (per_sm < 1) per_sm = 1;
    const int64_t want = ceil_div64(R, 8 * RT * 4);
    const int64_t grid = want < (int64_t)h->sm_count * per_sm ? want : (int64_t)h->sm_count * per_sm;
    fn<<<(unsigned)grid, 128, smem, st>>>(uN, uS, R, N, Nhat, Np, Rf, inv_Nh, err);
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}

extern "C" int rt_rom_difference_f64(rt_handle* h, const double* uN, int64_t B, int64_t nt, int64_t N, const double* uN_srom,
                                     int64_t Nhat, const double* G, int64_t Nh, double* err, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, uN && uN_srom && G && err);
    RT_REQUIRE(h, B >= 1 && nt >= 1 && N >= 1 && Nhat >= N && Nhat <= 128 && Nh >= 1);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    const int Np = (int)((Nhat + 7) / 8 * 8);
    void* rf;
    int rc = rt_scratch(h, 3, (size_t)Np * Np * sizeof(double), &rf);
    if (rc) return rc;
    const size_t csmem = (size_t)Nhat * (Nhat + 1) * sizeof(double);
    if (csmem > 48 * 1024) RT_CUDA(h, cudaFuncSetAttribute(k_chol_frag, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)csmem));
    k_chol_frag<<<1, 128, csmem, st>>>(G, (int)Nhat, Np, (double*)rf);
    RT_LAUNCH_CHECK(h);
    const int64_t R = B * nt;
    if (Np <= 32) return launch_rom_difference<8, 2>(h, uN, uN_srom, R, (int)N, (int)Nhat, Np, (const double*)rf, 1.0 / (double)Nh, err, st);
    if (Np <= 64) return launch_rom_difference<16, 2>(h, uN, uN_srom, R, (int)N, (int)Nhat, Np, (const double*)rf, 1.0 / (double)Nh, err, st);
    if (Np <= 96) return launch_rom_difference<24, 1>(h, uN, uN_srom, R, (int)N, (int)Nhat, Np, (const double*)rf, 1.0 / (double)Nh, err, st);
    return launch_rom_difference<32, 1>(h, uN, uN_srom, R, (int)N, (int)Nhat, Np, (const double*)rf, 1.0 / (double)Nh, err, st);
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
