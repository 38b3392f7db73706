// POD by the method of snapshots (reference rom/pod.py:7-62): Gram matrix on TMA-fed DMMA tiles,
// column norms / normalisation fused into the Gram finalisation, basis lift as a second GEMM.
#include "dgemm_tma.cuh"
#include "stage.cuh"
#include "gemm_util.cuh"

namespace {

__global__ void k_colnorm(const double* __restrict__ G, int64_t n, double* __restrict__ cn) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) cn[j] = sqrt(G[j * n + j]);
}

// upper triangle -> (optionally D^-1 G D^-1) -> mirrored into the lower triangle
__global__ void k_sym_scale(double* __restrict__ G, int64_t n, const double* __restrict__ cn, int normalize) {
    __shared__ double tile[32][33];
    const int bi = blockIdx.y, bj = blockIdx.x;
    if (bj < bi) return;
    const int64_t i0 = (int64_t)bi * 32, j0 = (int64_t)bj * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int64_t i = i0 + r, j = j0 + threadIdx.x;
        double v = 0.0;
        if (i < n && j < n && j >= i) {
            v = G[i * n + j];
            if (normalize) v = v / (cn[i] * cn[j]);
            G[i * n + j] = v;
        }
        tile[r][threadIdx.x] = v;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int64_t j = j0 + r, i = i0 + threadIdx.x;  // write G[j][i] = tile[i-i0][j-j0]
        if (i < n && j < n && j > i) G[j * n + i] = tile[threadIdx.x][r];
    }
}

// P[i][j] <- S[i][j] / cn[j] - P[i][j]   (P row-major Nh x n; S with element strides; cn may be null)
__global__ void __launch_bounds__(256) k_deflate(const double* __restrict__ S, int64_t Nh, int64_t n, int64_t srow, int64_t scol,
                                                 const double* __restrict__ cn, double* __restrict__ P) {
    __shared__ double tile[32][33];
    const int64_t i0 = (int64_t)blockIdx.y * 32, j0 = (int64_t)blockIdx.x * 32;
    const bool col_fast = scol <= srow;  // read S along its fast axis, write P along its rows
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        if (col_fast) {
            const int64_t i = i0 + r, j = j0 + threadIdx.x;
            if (i < Nh && j < n) tile[r][threadIdx.x] = S[i * srow + j * scol];
        } else {
            const int64_t j = j0 + r, i = i0 + threadIdx.x;
            if (i < Nh && j < n) tile[threadIdx.x][r] = S[i * srow + j * scol];
        }
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int64_t i = i0 + r, j = j0 + threadIdx.x;
        if (i < Nh && j < n) {
            double v = tile[r][threadIdx.x];
            if (cn) v /= cn[j];
            P[i * n + j] = v - P[i * n + j];
        }
    }
}

int gram_finalize(rt_handle* h, double* G, int64_t n, int normalize, double* colnorm, cudaStream_t st) {
    k_colnorm<<<(unsigned)ceil_div64(n, 256), 256, 0, st>>>(G, n, colnorm);
    RT_LAUNCH_CHECK(h);
    dim3 grid((unsigned)ceil_div64(n, 32), (unsigned)ceil_div64(n, 32));
    k_sym_scale<<<grid, dim3(32, 8), 0, st>>>(G, n, colnorm, normalize);
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}

}  // namespace

extern "C" int rt_pod_gram_f64(rt_handle* h, const double* S, int64_t Nh, int64_t n, int64_t stride_row,
                               int64_t stride_col, int normalize, double* G, double* colnorm, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, S && G && colnorm && Nh >= 1 && n >= 1);
    RT_REQUIRE(h, n <= 65535);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    dg::Operand op{S, n, Nh, stride_col, stride_row}, a;
    int rc = dg::stage_if_needed(h, 3, op, &a, st);
    if (rc) return rc;
    rc = gemm_tn(h, a, a, G, n, true, st);
    if (rc) return rc;
    return gram_finalize(h, G, n, normalize, colnorm, st);
}

extern "C" int rt_gram_cross_f64(rt_handle* h, const double* A, int64_t na, int64_t a_stride_row, int64_t a_stride_col,
                                 const double* B, int64_t nb, int64_t b_stride_row, int64_t b_stride_col, int64_t Nh,
                                 double* C, int64_t ldc, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, A && B && C && Nh >= 1 && na >= 1 && nb >= 1 && ldc >= nb);
    RT_REQUIRE(h, na <= 65535);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    dg::Operand oa{A, na, Nh, a_stride_col, a_stride_row}, ob{B, nb, Nh, b_stride_col, b_stride_row}, a, b;
    int rc = dg::stage_if_needed(h, 3, oa, &a, st);
    if (rc) return rc;
    rc = dg::stage_if_needed(h, 4, ob, &b, st);
    if (rc) return rc;
    return gemm_tn(h, a, b, C, ldc, false, st);
}

extern "C" int rt_gram_finalize_f64(rt_handle* h, double* G, int64_t n, int normalize, double* colnorm, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, G && colnorm && n >= 1);
    RT_CUDA(h, cudaSetDevice(h->device));
    return gram_finalize(h, G, n, normalize, colnorm, rt_stream(stream));
}

extern "C" int rt_pod_lift_f64(rt_handle* h, const double* S, int64_t Nh, int64_t n, int64_t stride_row,
                               int64_t stride_col, const double* Wr, int64_t ldw, int64_t r, double* Q,
                               int64_t q_stride_row, int64_t q_stride_col, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, S && Wr && Q && Nh >= 1 && n >= 1 && r >= 1 && ldw >= r);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    // Q[i][c] = sum_j S[i][j] Wr[j][c] :  A = S as (mn = Nh rows, k = n), B = Wr as (mn = r, k = n)
    dg::Operand oa{S, Nh, n, stride_row, stride_col}, ob{Wr, r, n, 1, ldw}, a, b;
    int rc = dg::stage_if_needed(h, 3, oa, &a, st);
    if (rc) return rc;
    rc = dg::stage_if_needed(h, 4, ob, &b, st);
    if (rc) return rc;
    dg::Params p = dg::default_params();
    p.C = Q; p.ldc_m = q_stride_row; p.ldc_n = q_stride_col;
    return dg::launch(h, a, b, p, st);
}

// Batched forms for the tree-walk PODs: P small problems described by HOST arrays of descriptors, queued back to back on the
// stream by ONE call (the per-call cost of the Python binding, not the kernels, dominates at 100 x 50 snapshot blocks).
extern "C" int rt_pod_gram_batch_f64(rt_handle* h, const rt_gram_desc* d, int64_t P, int normalize, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, d && P >= 1);
    for (int64_t p = 0; p < P; ++p) {
        int rc = rt_pod_gram_f64(h, d[p].S, d[p].Nh, d[p].n, d[p].stride_row, d[p].stride_col, normalize, d[p].G, d[p].colnorm, stream);
        if (rc) return rc;
    }
    return RT_OK;
}

extern "C" int rt_pod_lift_batch_f64(rt_handle* h, const rt_lift_desc* d, int64_t P, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, d && P >= 1);
    for (int64_t p = 0; p < P; ++p) {
        int rc = rt_pod_lift_f64(h, d[p].S, d[p].Nh, d[p].n, d[p].stride_row, d[p].stride_col, d[p].Wr, d[p].ldw, d[p].r, d[p].Q,
                                 d[p].q_stride_row, d[p].q_stride_col, stream);
        if (rc) return rc;
    }
    return RT_OK;
}

// Deflation step of the POD's second level (romtime_b200/rom/pod.py): with P = Q (Q^T S D^-1) already in `P`
// (rt_gram_cross_f64 + rt_pod_lift_f64), P <- S D^-1 - P is what the delivered modes Q leave of the (column-normalised)
// snapshots.  colnorm == NULL: D = I.  P is Nh x n row-major.
extern "C" int rt_deflate_f64(rt_handle* h, const double* S, int64_t Nh, int64_t n, int64_t stride_row, int64_t stride_col,
                              const double* colnorm, double* P, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, S && P && Nh >= 1 && n >= 1);
    RT_CUDA(h, cudaSetDevice(h->device));
    const int64_t gy = ceil_div64(Nh, 32);
    RT_REQUIRE(h, gy <= 2147483647LL && ceil_div64(n, 32) <= 65535);
    // blockIdx.y is limited to 65535: rows go to grid.x when there are more than 2 M of them
    dim3 grid((unsigned)ceil_div64(n, 32), (unsigned)(gy <= 65535 ? gy : 65535));
    for (int64_t y0 = 0; y0 < gy; y0 += 65535) {
        const int64_t rows0 = y0 * 32, rows = (Nh - rows0) < 65535LL * 32 ? (Nh - rows0) : 65535LL * 32;
        dim3 g((unsigned)ceil_div64(n, 32), (unsigned)ceil_div64(rows, 32));
        k_deflate<<<g, dim3(32, 8), 0, rt_stream(stream)>>>(S + rows0 * stride_row, rows, n, stride_row, stride_col, colnorm,
                                                           P + rows0 * n);
        RT_LAUNCH_CHECK(h);
    }
    (void)grid;
    return RT_OK;
}
