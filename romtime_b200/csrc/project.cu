// Projection of the collateral bases onto the solution reduced basis V (offline, once per ROM):
//   DEIM.project_basis   (reference deim/deim.py:495-515):   basis_rom = V^T U_f               (N x m)
//   MDEIM.project_basis  (reference deim/mdeim.py:153-192 + utils.py:96-113, :136-149):
//       A_q = csr(U[:, q]; rows, cols),  basis_rom[(i*N + j), q] = (V^T A_q V)[i, j]          (N^2 x m)
//
// The MDEIM projection is done for a whole chunk of collateral modes at once:
//   1. the COO topology (rows, cols) is sorted by row with a STABLE radix sort (duplicates keep their
//      input order and are summed, like scipy's COO -> CSR conversion) -- cub::DeviceRadixSort;
//   2. k_spmm_modes: W[r][q*N + j] = sum_{k in row r} U[k, q] V[cols[k], j]   (= [A_q V] for all q side by
//      side, one CTA per matrix row, fixed summation order -> deterministic);
//   3. one TMA + DMMA GEMM  T = V^T W  (N x mq*N, split-k over the Nh rows);
//   4. k_scatter_modes: T[i][q*N + j] -> out[(i*N + j), q] (the row-major AN.flatten() of mdeim.py:183).
#include <cub/device/device_radix_sort.cuh>

#include "gemm_util.cuh"

namespace {

__global__ void k_iota32(int32_t* v, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = (int32_t)i;
}

// first index k in [0, n) with a[k] >= key
__device__ __forceinline__ int64_t lower_bound64(const int64_t* __restrict__ a, int64_t n, int64_t key) {
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (a[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

constexpr int kSpmmThreads = 256;
constexpr int kRowBatch = 32;

__global__ void __launch_bounds__(kSpmmThreads)
k_spmm_modes(const double* __restrict__ U, int64_t usr, int64_t usc, const int64_t* __restrict__ rows_sorted,
             const int32_t* __restrict__ perm, const int64_t* __restrict__ cols, int64_t nnz,
             const double* __restrict__ V, int64_t vsr, int64_t vsc, int64_t Nh, int N, int q0, int mq,
             double* __restrict__ W, int64_t ldw) {
    __shared__ int64_t s_lo, s_hi;
    __shared__ int64_t s_k[kRowBatch], s_c[kRowBatch];
    const int64_t r = blockIdx.x;
    if (threadIdx.x == 0) s_lo = lower_bound64(rows_sorted, nnz, r);
    if (threadIdx.x == 32) s_hi = lower_bound64(rows_sorted, nnz, r + 1);
    __syncthreads();
    const int64_t lo = s_lo, hi = s_hi;
    const int width = mq * N;
    double* Wr = W + r * ldw;
    if (lo == hi) {  // empty matrix row
        for (int e = threadIdx.x; e < width; e += kSpmmThreads) Wr[e] = 0.0;
        return;
    }
    for (int64_t b0 = lo; b0 < hi; b0 += kRowBatch) {
        const int nb = (int)((hi - b0) < kRowBatch ? (hi - b0) : kRowBatch);
        __syncthreads();
        if (threadIdx.x < nb) {
            const int64_t k = perm[b0 + threadIdx.x];
            s_k[threadIdx.x] = k;
            int64_t c = cols[k];
            s_c[threadIdx.x] = (c >= 0 && c < Nh) ? c : -1;  // out-of-range column: contributes nothing
        }
        __syncthreads();
        for (int e = threadIdx.x; e < width; e += kSpmmThreads) {
            const int q = e / N, j = e - q * N;
            double acc = (b0 == lo) ? 0.0 : Wr[e];
            for (int t = 0; t < nb; ++t) {
                if (s_c[t] < 0) continue;
                acc = fma(U[s_k[t] * usr + (int64_t)(q0 + q) * usc], V[s_c[t] * vsr + (int64_t)j * vsc], acc);
            }
            Wr[e] = acc;
        }
    }
}

__global__ void k_scatter_modes(const double* __restrict__ T, int64_t ldt, int N, int q0, int mq,
                                double* __restrict__ out, int64_t osr, int64_t osc) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t total = (int64_t)N * mq * N;
    if (e >= total) return;
    const int i = (int)(e / ((int64_t)mq * N));
    const int rem = (int)(e - (int64_t)i * mq * N);
    const int q = rem / N, j = rem - q * N;
    out[((int64_t)i * N + j) * osr + (int64_t)(q0 + q) * osc] = T[(int64_t)i * ldt + rem];
}

}  // namespace

extern "C" int rt_project_vec_f64(rt_handle* h, const double* V, int64_t v_stride_row, int64_t v_stride_col,
                                  const double* U, int64_t u_stride_row, int64_t u_stride_col, int64_t Nh, int64_t N,
                                  int64_t m, double* out, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, V && U && out && Nh >= 1 && N >= 1 && m >= 1);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    dg::Operand oa{V, N, Nh, v_stride_col, v_stride_row}, ob{U, m, Nh, u_stride_col, u_stride_row}, a, b;
    int rc = dg::stage_if_needed(h, 3, oa, &a, st);
    if (rc) return rc;
    rc = dg::stage_if_needed(h, 4, ob, &b, st);
    if (rc) return rc;
    return gemm_tn(h, a, b, out, m, false, st);
}

extern "C" int rt_project_mat_f64(rt_handle* h, const double* U, int64_t u_stride_row, int64_t u_stride_col,
                                  const int64_t* rows, const int64_t* cols, int64_t nnz, const double* V,
                                  int64_t v_stride_row, int64_t v_stride_col, int64_t Nh, int64_t N, int64_t m,
                                  double* out, int64_t out_stride_row, int64_t out_stride_col, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, U && rows && cols && V && out);
    RT_REQUIRE(h, nnz >= 1 && nnz < (int64_t(1) << 31) && Nh >= 1 && N >= 1 && N <= 1024 && m >= 1);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);

    // ---- 1. stable sort of the COO entries by row ----
    int end_bit = 1;
    while (end_bit < 63 && (int64_t(1) << end_bit) < Nh) ++end_bit;
    size_t cub_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (const int64_t*)nullptr, (int64_t*)nullptr,
                                    (const int32_t*)nullptr, (int32_t*)nullptr, (int)nnz, 0, end_bit, st);
    const size_t a256 = 255;
    const size_t o_rows = 0;
    const size_t o_iota = (o_rows + (size_t)nnz * 8 + a256) & ~a256;
    const size_t o_perm = (o_iota + (size_t)nnz * 4 + a256) & ~a256;
    const size_t o_cub = (o_perm + (size_t)nnz * 4 + a256) & ~a256;
    void* sbuf;
    int rc = rt_scratch(h, 9, o_cub + cub_bytes + 256, &sbuf);
    if (rc) return rc;
    int64_t* rows_sorted = reinterpret_cast<int64_t*>((char*)sbuf + o_rows);
    int32_t* iota = reinterpret_cast<int32_t*>((char*)sbuf + o_iota);
    int32_t* perm = reinterpret_cast<int32_t*>((char*)sbuf + o_perm);
    k_iota32<<<(unsigned)ceil_div64(nnz, 256), 256, 0, st>>>(iota, nnz);
    RT_LAUNCH_CHECK(h);
    RT_CUDA(h, cub::DeviceRadixSort::SortPairs((char*)sbuf + o_cub, cub_bytes, rows, rows_sorted, iota, perm, (int)nnz, 0,
                                               end_bit, st));
    h->launches += 2;  // cub: histogram + onesweep passes (at least)

    // ---- 2-4. chunks of collateral modes ----
    const double budget = 2.0e9;  // bytes of W per chunk
    int64_t mq_max = (int64_t)(budget / ((double)Nh * N * 8.0));
    if (mq_max < 1) mq_max = 1;
    if (mq_max > m) mq_max = m;
    if ((mq_max * N) & 1) mq_max = (mq_max > 1 && ((mq_max - 1) * N) % 2 == 0) ? mq_max - 1 : mq_max;
    const int64_t ldw = (mq_max * N + 1) & ~int64_t(1);
    void *wbuf, *tbuf;
    rc = rt_scratch(h, 10, (size_t)Nh * ldw * sizeof(double), &wbuf);
    if (rc) return rc;
    rc = rt_scratch(h, 11, (size_t)N * ldw * sizeof(double), &tbuf);
    if (rc) return rc;
    dg::Operand ov{V, N, Nh, v_stride_col, v_stride_row}, a;
    rc = dg::stage_if_needed(h, 3, ov, &a, st);
    if (rc) return rc;
    for (int64_t q0 = 0; q0 < m; q0 += mq_max) {
        const int mq = (int)((m - q0) < mq_max ? (m - q0) : mq_max);
        k_spmm_modes<<<(unsigned)Nh, kSpmmThreads, 0, st>>>(U, u_stride_row, u_stride_col, rows_sorted, perm, cols, nnz, V,
                                                           v_stride_row, v_stride_col, Nh, (int)N, (int)q0, mq,
                                                           (double*)wbuf, ldw);
        RT_LAUNCH_CHECK(h);
        dg::Operand ow{(const double*)wbuf, (int64_t)mq * N, Nh, 1, ldw};
        rc = gemm_tn(h, a, ow, (double*)tbuf, ldw, false, st);
        if (rc) return rc;
        const int64_t total = N * mq * N;
        k_scatter_modes<<<(unsigned)ceil_div64(total, 256), 256, 0, st>>>((const double*)tbuf, ldw, (int)N, (int)q0, mq, out,
                                                                         out_stride_row, out_stride_col);
        RT_LAUNCH_CHECK(h);
    }
    return RT_OK;
}
