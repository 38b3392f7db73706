// Helpers shared by the translation units that drive the GEMM core: split-k partial reduction and
// the "C = A^T B" wrapper (both operands given as Nh x cols matrices with arbitrary strides).
#pragma once
#include "dgemm_tma.cuh"
#include "stage.cuh"

namespace {

// out[i][j] = sum_s P[s][i][j] in fixed order (deterministic), for tiles (ti, tj) with tj >= ti when
// `upper_tiles`, else everywhere.
__global__ void k_reduce_partials(const double* __restrict__ P, int splits, int64_t rows, int64_t cols,
                                  int64_t split_stride, double* __restrict__ out, int64_t ldo, int upper_tiles) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t i = blockIdx.y;
    if (j >= cols || i >= rows) return;
    if (upper_tiles && (j / dg::BN) < (i / dg::BM)) return;
    double s = 0.0;
    for (int k = 0; k < splits; ++k) s += P[(int64_t)k * split_stride + i * cols + j];
    out[i * ldo + j] = s;
}

// C (rows x cols, row-major ldc) = A^T B with optional split-k through partial buffers
int gemm_tn(rt_handle* h, const dg::Operand& A, const dg::Operand& B, double* C, int64_t ldc, bool symmetric,
            cudaStream_t st) {
    dg::Params p = dg::default_params();
    const int tiles_m = (int)ceil_div64(A.mn, dg::BM), tiles_n = (int)ceil_div64(B.mn, dg::BN);
    const int ntiles = symmetric ? tiles_n * (tiles_n + 1) / 2 : tiles_m * tiles_n;
    const int kt_total = (int)ceil_div64(A.k, dg::BK);
    // few output tiles (the Gram matrix of the lifted modes is ONE): as many k-splits as it takes to put two units on every SM
    const int max_splits = ntiles * 64 >= 2 * h->sm_count ? 64 : (2 * h->sm_count + ntiles - 1) / ntiles;
    const int splits = dg::choose_splits(ntiles, kt_total, h->sm_count, A.mn * B.mn, max_splits);
    p.symmetric = symmetric ? 1 : 0;
    p.splits = splits;
    if (splits == 1) {
        p.C = C; p.ldc_m = ldc; p.ldc_n = 1; p.split_stride = 0;
        return dg::launch(h, A, B, p, st);
    }
    void* part;
    int rc = rt_scratch(h, 2, (size_t)splits * A.mn * B.mn * sizeof(double), &part);
    if (rc) return rc;
    p.C = (double*)part; p.ldc_m = B.mn; p.ldc_n = 1; p.split_stride = A.mn * B.mn;
    rc = dg::launch(h, A, B, p, st);
    if (rc) return rc;
    // launch() may have reduced the split count; recompute exactly as it does
    int s2 = splits > kt_total ? kt_total : splits;
    const int ktps = (int)ceil_div64(kt_total, s2);
    s2 = (int)ceil_div64(kt_total, ktps);
    dim3 grid((unsigned)ceil_div64(B.mn, 256), (unsigned)A.mn);
    k_reduce_partials<<<grid, 256, 0, st>>>((const double*)part, s2, A.mn, B.mn, A.mn * B.mn, C, ldc, symmetric ? 1 : 0);
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}

}  // namespace
