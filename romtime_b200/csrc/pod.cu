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
