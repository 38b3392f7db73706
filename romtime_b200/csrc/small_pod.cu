// Batched small eigensolvers for the hierarchical (tree-walk) POD: the reference consumes orth as MANY small per-parameter
// time-PODs (100-1000 dofs x 10-250 snapshots each; rom/rom.py:313-406, deim/deim.py:279-397, deim/nonlinear.py:435-466)
// followed by one stacked mu-POD.  One CTA per problem, the whole solve in shared memory, no host round trip inside:
//
//   rt_small_gram_eig_f64 : G_p = S_p^T S_p (n_p <= 256) -> diagonally pivoted Cholesky G ~= L L^T (numerical rank r <= 96, the
//                           same 8-ulp cut as rt_sym_eig_f64) -> ONE-SIDED Jacobi on the r vectors of L: their norms are the
//                           eigenvalues of G, the normalised vectors its eigenvectors (no M = L^T L, no second squaring).
//   rt_small_refine_f64   : the Hestenes refinement of romtime_b200/rom/pod.py for a batch: H_p = Y_p^T Y_p (c_p <= 96) of the
//                           lifted, nearly orthogonal graded columns -> Cholesky H = C C^T -> one-sided Jacobi on the rows of C with
//                           the rotations accumulated (Demmel-Veselic: every eigenvalue to a small RELATIVE error).
//
// Problems outside the limits are flagged in info[p] and left to the general path (rt_sym_eig_f64 / rt_sym_eig_jacobi_f64).
#include <cfloat>

#include "common.cuh"

namespace {

constexpr int kSpThreads = 256;
constexpr int kSpWarps = kSpThreads / 32;
constexpr int kSpRmax = 96;   // vectors per problem (rank of G / candidates of the refinement)
constexpr int kSpNmax = 256;  // snapshots per problem
constexpr double kSpEps = 2.220446049250313e-16;

// One-sided (Hestenes) Jacobi on nv vectors of length len, stored as rows v[c * ld + i] in shared memory; the rotations
// are also applied to the rows of x (length lenx, leading dimension ldx) when x != nullptr.  Round-robin ordering: nv - 1
// rounds of nv / 2 disjoint pairs per sweep, a warp per pair.  `flag` (shared): rotations of the running sweep.
// Returns the number of sweeps used, or -1 when 40 sweeps did not converge.  All threads of the CTA must call.
__device__ int sp_onesided_jacobi(double* v, int ld, int nv, int len, double* x, int ldx, int lenx, int* flag) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (nv < 2) return 0;
    const int m = nv + (nv & 1);  // even number of players; index nv (if any) is a bye
    for (int sweep = 0; sweep < 40; ++sweep) {
        if (threadIdx.x == 0) *flag = 0;
        __syncthreads();
        for (int rd = 0; rd < m - 1; ++rd) {
            for (int k = warp; k < m / 2; k += kSpWarps) {
                int p = k == 0 ? m - 1 : (rd + k) % (m - 1);
                int q = k == 0 ? rd : (rd - k + (m - 1)) % (m - 1);
                if (p >= nv || q >= nv) continue;
                if (p > q) { const int t = p; p = q; q = t; }
                double* vp = v + p * ld;
                double* vq = v + q * ld;
                double a = 0.0, b = 0.0, g = 0.0;
                for (int i = lane; i < len; i += 32) { const double xp = vp[i], xq = vq[i]; a = fma(xp, xp, a); b = fma(xq, xq, b); g = fma(xp, xq, g); }
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) {
                    a += __shfl_xor_sync(0xffffffffu, a, off);
                    b += __shfl_xor_sync(0xffffffffu, b, off);
                    g += __shfl_xor_sync(0xffffffffu, g, off);
                }
                if (!(a > 0.0 && b > 0.0) || !(fabs(g) > kSpEps * sqrt(a) * sqrt(b))) continue;
                const double zeta = (b - a) / (2.0 * g);
                const double t = copysign(1.0 / (fabs(zeta) + sqrt(1.0 + zeta * zeta)), zeta);
                const double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
                for (int i = lane; i < len; i += 32) {
                    const double xp = vp[i], xq = vq[i];
                    vp[i] = cs * xp - sn * xq;
                    vq[i] = sn * xp + cs * xq;
                }
                if (x) {
                    double* xp_ = x + p * ldx;
                    double* xq_ = x + q * ldx;
                    for (int i = lane; i < lenx; i += 32) {
                        const double xp = xp_[i], xq = xq_[i];
                        xp_[i] = cs * xp - sn * xq;
                        xq_[i] = sn * xp + cs * xq;
                    }
                }
                if (lane == 0) atomicAdd(flag, 1);
            }
            __syncthreads();
        }
        const int rot = *flag;
        __syncthreads();
        if (rot == 0) return sweep + 1;
    }
    return -1;
}

// squared norms of the nv vectors -> lam[c]; descending rank of every vector -> ord[rank] = c  (ties by index)
__device__ void sp_norms_and_order(const double* v, int ld, int nv, int len, double* lam, int* ord) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int c = warp; c < nv; c += kSpWarps) {
        double a = 0.0;
        for (int i = lane; i < len; i += 32) a = fma(v[c * ld + i], v[c * ld + i], a);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) a += __shfl_xor_sync(0xffffffffu, a, off);
        if (lane == 0) lam[c] = a;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < nv; c += kSpThreads) {
        int rk = 0;
        for (int j = 0; j < nv; ++j) rk += (lam[j] > lam[c] || (lam[j] == lam[c] && j < c)) ? 1 : 0;
        ord[rk] = c;
    }
    __syncthreads();
}

// block-wide arg-max of val over threads (ties: lowest index); result broadcast through red_v / red_i
__device__ void sp_block_argmax(double val, int idx, double* red_v, int* red_i, double* out_v, int* out_i) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, val, off);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, off);
        if (ov > val || (ov == val && oi < idx)) { val = ov; idx = oi; }
    }
    if (lane == 0) { red_v[warp] = val; red_i[warp] = idx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double bv = red_v[0];
        int bi = red_i[0];
        for (int w = 1; w < kSpWarps; ++w)
            if (red_v[w] > bv || (red_v[w] == bv && red_i[w] < bi)) { bv = red_v[w]; bi = red_i[w]; }
        red_v[kSpWarps] = bv;
        red_i[kSpWarps] = bi;
    }
    __syncthreads();
    *out_v = red_v[kSpWarps];
    *out_i = red_i[kSpWarps];
    __syncthreads();
}

// info: 0 ok, 1 numerical rank above kSpRmax (general path needed), 2 Jacobi did not converge
__global__ void __launch_bounds__(kSpThreads, 1) k_small_gram_eig(const double* __restrict__ G_all, int64_t stride_g,
                                                                  const int* __restrict__ n_of, int nmax, double* __restrict__ w_all,
                                                                  double* __restrict__ W_all, int64_t stride_w, int* __restrict__ rank_out,
                                                                  int* __restrict__ info) {
    extern __shared__ __align__(16) double sp_sm[];
    const int p = blockIdx.x, t = threadIdx.x;
    const int n = n_of[p];
    const double* G = G_all + (int64_t)p * stride_g;  // n x n row-major, leading dimension n
    double* w = w_all + (int64_t)p * nmax;
    double* W = W_all + (int64_t)p * stride_w;
    const int ldl = nmax | 1;
    double* L = sp_sm;                       // kSpRmax x ldl: row c = Cholesky column c
    double* d = L + (size_t)kSpRmax * ldl;   // residual diagonal (n), < 0: used as a pivot
    double* lam = d + nmax;                  // kSpRmax
    double* red_v = lam + kSpRmax;           // kSpWarps + 1
    int* red_i = reinterpret_cast<int*>(red_v + kSpWarps + 1);  // kSpWarps + 1
    int* ord = red_i + kSpWarps + 1;         // kSpRmax
    int* flag = ord + kSpRmax;
    for (int i = t; i < n; i += kSpThreads) d[i] = G[(int64_t)i * n + i];
    __syncthreads();
    double v0 = -1.0;
    int i0 = 0;
    for (int i = t; i < n; i += kSpThreads)
        if (d[i] > v0) { v0 = d[i]; i0 = i; }
    double dmax0;
    int piv;
    sp_block_argmax(v0, i0, red_v, red_i, &dmax0, &piv);
    int r = 0, status = 0;
    double dv = dmax0;
    const double cut = 8.0 * kSpEps * dmax0;  // the noise floor of an fp64 Gram matrix (same cut as k_pchol)
    while (dv > cut && dv > 0.0) {
        if (r == kSpRmax) { status = 1; break; }
        const double rs = 1.0 / sqrt(dv);
        for (int i = t; i < n; i += kSpThreads) {
            double acc = G[(int64_t)piv * n + i];  // row piv = column piv (symmetric), contiguous
            for (int c = 0; c < r; ++c) acc = fma(-L[c * ldl + i], L[c * ldl + piv], acc);
            const double l = acc * rs;
            L[r * ldl + i] = l;
            if (d[i] >= 0.0) d[i] = i == piv ? -1.0 : fmax(d[i] - l * l, 0.0);
        }
        ++r;
        __syncthreads();
        v0 = -1.0;
        i0 = 0;
        for (int i = t; i < n; i += kSpThreads)
            if (d[i] > v0) { v0 = d[i]; i0 = i; }
        sp_block_argmax(v0, i0, red_v, red_i, &dv, &piv);
    }
    if (status == 0) {
        const int sweeps = sp_onesided_jacobi(L, ldl, r, n, nullptr, 0, 0, flag);
        if (sweeps < 0) status = 2;
    }
    if (status == 0) {
        sp_norms_and_order(L, ldl, r, n, lam, ord);
        for (int k = t; k < r; k += kSpThreads) w[k] = lam[ord[k]];
        // W[i][k] = L[ord[k]][i] / sqrt(lam); eigenvector k in column k, zero beyond the rank (W was zeroed by the caller)
        for (int e = t; e < n * r; e += kSpThreads) {
            const int i = e / r, k = e - i * r;
            const int c = ord[k];
            W[(int64_t)i * n + k] = L[c * ldl + i] * (1.0 / sqrt(lam[c]));
        }
    }
    if (t == 0) { rank_out[p] = status == 0 ? r : 0; info[p] = status; }
}

// info: 0 ok, 1 H not positive definite (general path needed), 2 Jacobi did not converge
__global__ void __launch_bounds__(kSpThreads, 1) k_small_refine(const double* __restrict__ H_all, int64_t stride_h, const int* __restrict__ c_of,
                                                                int cmax, double* __restrict__ lam_all, double* __restrict__ X_all,
                                                                int64_t stride_x, double* __restrict__ defect, int* __restrict__ info) {
    extern __shared__ __align__(16) double sp_sm[];
    const int p = blockIdx.x, t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int c = c_of[p];
    const double* H = H_all + (int64_t)p * stride_h;  // c x c row-major, leading dimension c
    const int ld = cmax | 1;
    double* C = sp_sm;                  // c x ld: lower Cholesky factor, H = C C^T (rows are the Jacobi vectors)
    double* Xt = C + (size_t)cmax * ld;  // c x ld: accumulated rotations, row k = eigenvector k
    double* lam = Xt + (size_t)cmax * ld;
    double* red = lam + cmax;           // kSpWarps
    int* ord = reinterpret_cast<int*>(red + kSpWarps);
    int* flag = ord + cmax;
    int* bad = flag + 1;
    if (t == 0) *bad = 0;
    for (int e = t; e < c * c; e += kSpThreads) {
        const int i = e / c, j = e - i * c;
        C[i * ld + j] = j <= i ? H[(int64_t)i * c + j] : 0.0;
        Xt[i * ld + j] = i == j ? 1.0 : 0.0;
    }
    __syncthreads();
    // departure from orthogonality of the input columns: max |h_ij| / sqrt(h_ii h_jj)
    double dm = 0.0;
    for (int e = t; e < c * c; e += kSpThreads) {
        const int i = e / c, j = e - i * c;
        if (j < i) {
            const double pr = C[i * ld + i] * C[j * ld + j];
            if (pr > 0.0) dm = fmax(dm, fabs(C[i * ld + j]) / sqrt(pr));
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) dm = fmax(dm, __shfl_xor_sync(0xffffffffu, dm, off));
    if (lane == 0) red[warp] = dm;
    __syncthreads();
    if (t == 0) {
        double m = red[0];
        for (int w = 1; w < kSpWarps; ++w) m = fmax(m, red[w]);
        defect[p] = m;
    }
    // Cholesky in place (lower, right-looking); a non-positive pivot hands the problem to the general path
    for (int k = 0; k < c; ++k) {
        __syncthreads();
        const double akk = C[k * ld + k];
        if (!(akk > 0.0)) { if (t == 0) *bad = 1; break; }
        const double piv = sqrt(akk), rp = 1.0 / piv;
        __syncthreads();
        for (int i = k + t; i < c; i += kSpThreads) C[i * ld + k] = i == k ? piv : C[i * ld + k] * rp;
        __syncthreads();
        const int rem = c - k - 1;
        for (int e = t; e < rem * rem; e += kSpThreads) {
            const int i = k + 1 + e / rem, j = k + 1 + e % rem;
            if (j <= i) C[i * ld + j] = fma(-C[i * ld + k], C[j * ld + k], C[i * ld + j]);
        }
    }
    __syncthreads();
    int status = *bad ? 1 : 0;
    if (status == 0) {
        const int sweeps = sp_onesided_jacobi(C, ld, c, c, Xt, ld, c, flag);
        if (sweeps < 0) status = 2;
    }
    if (status == 0) {
        sp_norms_and_order(C, ld, c, c, lam, ord);
        double* lo = lam_all + (int64_t)p * cmax;
        double* X = X_all + (int64_t)p * stride_x;  // c x c row-major, eigenvector k in column k
        for (int k = t; k < c; k += kSpThreads) lo[k] = lam[ord[k]];
        for (int e = t; e < c * c; e += kSpThreads) {
            const int i = e / c, k = e - i * c;
            X[(int64_t)i * c + k] = Xt[ord[k] * ld + i];
        }
    }
    if (t == 0) info[p] = status;
}

}  // namespace

extern "C" int rt_small_gram_eig_f64(rt_handle* h, const double* G_all, int64_t stride_g, const int* n_of, int64_t P, int64_t nmax,
                                     double* w_all, double* W_all, int64_t stride_w, int* rank, int* info, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, G_all && n_of && w_all && W_all && rank && info && P >= 1 && nmax >= 1 && nmax <= kSpNmax);
    RT_REQUIRE(h, stride_g >= nmax * nmax && stride_w >= nmax * nmax);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    RT_CUDA(h, cudaMemsetAsync(w_all, 0, (size_t)P * nmax * sizeof(double), st));
    RT_CUDA(h, cudaMemsetAsync(W_all, 0, (size_t)P * stride_w * sizeof(double), st));
    const int ldl = (int)nmax | 1;
    const size_t smem = ((size_t)kSpRmax * ldl + nmax + kSpRmax + kSpWarps + 1) * sizeof(double) + (kSpWarps + 1 + kSpRmax + 4) * sizeof(int) + 16;
    int rc = rt_ensure_smem(h, 12, k_small_gram_eig, (int)smem);
    if (rc) return rc;
    k_small_gram_eig<<<(unsigned)P, kSpThreads, smem, st>>>(G_all, stride_g, n_of, (int)nmax, w_all, W_all, stride_w, rank, info);
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}

extern "C" int rt_small_refine_f64(rt_handle* h, const double* H_all, int64_t stride_h, const int* c_of, int64_t P, int64_t cmax,
                                   double* lam_all, double* X_all, int64_t stride_x, double* defect, int* info, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, H_all && c_of && lam_all && X_all && defect && info && P >= 1 && cmax >= 1 && cmax <= kSpRmax);
    RT_REQUIRE(h, stride_h >= cmax * cmax && stride_x >= cmax * cmax);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    const int ld = (int)cmax | 1;
    const size_t smem = ((size_t)2 * cmax * ld + cmax + kSpWarps) * sizeof(double) + (cmax + 4) * sizeof(int) + 16;
    int rc = rt_ensure_smem(h, 13, k_small_refine, (int)smem);
    if (rc) return rc;
    k_small_refine<<<(unsigned)P, kSpThreads, smem, st>>>(H_all, stride_h, c_of, (int)cmax, lam_all, X_all, stride_x, defect, info);
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}
