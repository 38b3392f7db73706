// Symmetric eigensolver for the POD Gram matrix: parallel-order BLOCK JACOBI in fp64.
//
// Replaces the right-singular side of scipy.linalg.svd(lapack_driver="gesvd") (rom/pod.py:38):
// with G = S^T S,  s = sqrt(eig(G)),  VT = W^T.  Jacobi is used because it delivers all n
// eigenpairs (orth returns the untruncated spectrum, rom/pod.py:59-62) with small relative errors
// on graded positive semi-definite matrices, and maps to independent 64 x 64 tile problems:
//
//   blocks of 32 indices, round-robin tournament over block pairs (nb-1 rounds per sweep); every index pair is rotated
//   ONCE per sweep: round 0 runs the full 63-step ordering of its 64 x 64 sub-problems (diagonal blocks included), the
//   other rounds only the 32 steps that pair block I with block J (10 sweeps of 0.6 ms instead of 8 of 1.05 ms at rank 260);
//   per round  (1) k_jacobi_sub : one CTA per block pair diagonalises its 64 x 64 sub-matrix in
//                                 shared memory (ONE cyclic sweep, parallel ordering, per round), keeps the
//                                 accumulated rotation W_k;
//              (2) k_jacobi_apply: A_kl <- W_k^T A_kl W_l for every tile pair k < l (DMMA, mirrored
//                                 write keeps A exactly symmetric) and V[:, l] <- V[:, l] W_l.
//   Converged when a whole sweep performs no rotation.
//   Rotation test: |a_pq| > max(eps*sqrt(|a_pp a_qq|), floor_rel*max|diag|).
//
// Plain Jacobi needs 20-30 sweeps on the rank-deficient, graded Gram matrices POD produces, so the
// solver is PRECONDITIONED (Veselic-Hari / Drmac-Veselic):
//   (a) k_pchol: diagonally pivoted Cholesky G ~= L L^T (n x r, r = numerical rank), one persistent
//       cooperative kernel (warp per row, one grid barrier per column);
//   (b) M = L^T L (r x r) on the TMA/DMMA GEMM core -- M is strongly diagonally dominant and graded;
//   (c) block Jacobi on M (5-8 sweeps);   (d) W = L X diag(lambda)^-1/2 (GEMM core).
// Eigenpairs below the pivoting tolerance (8 ulp of the largest diagonal entry -- the noise floor
// of an fp64 Gram matrix) are returned as lambda = 0 with a zero vector.
#include <cooperative_groups.h>

#include <cstdlib>

#include "gemm_util.cuh"

namespace cg = cooperative_groups;

namespace {

constexpr int BS = 32;        // block size
constexpr int TS = 2 * BS;    // tile (sub-problem) size
constexpr int LDS_ = TS + 4;  // shared-memory leading dimension (68 == 4 mod 16: conflict-free DMMA fragments)
constexpr double kEps = 2.220446049250313e-16;
// A sweep whose rotations all had |a_pq| <= kBigRotation * sqrt(a_pp a_qq) leaves off-diagonal entries of the order of the
// SQUARE of that (cyclic Jacobi converges quadratically): below the rotation threshold, so the sweep that would only verify
// "no rotation left" is not run.  The sub-problem kernel counts those rotations in the upper half of the sweep counter.
constexpr double kBigRotation = 1e-10;

__device__ __forceinline__ void rr_pair(int round, int k, int nb, int& I, int& J) {
    // circle method on nb (even) players: every unordered pair meets exactly once in nb-1 rounds
    const int m = nb - 1;
    if (k == 0) { I = round % m; J = nb - 1; }
    else { I = (round + k) % m; J = (round - k + m) % m; }
    if (I > J) { int t = I; I = J; J = t; }
}
__device__ __forceinline__ int tile_index(int I, int J, int a) { return a < BS ? I * BS + a : J * BS + (a - BS); }

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
        : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__global__ void k_maxdiag(const double* __restrict__ A, int64_t n, int64_t lda, double* out) {
    __shared__ double red[256];
    double v = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) v = fmax(v, fabs(A[i * lda + i]));
    red[threadIdx.x] = v;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) red[threadIdx.x] = fmax(red[threadIdx.x], red[threadIdx.x + s]);
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = red[0];
}

// ---- (1) 64 x 64 sub-problem: cyclic two-sided Jacobi with parallel ordering, in shared memory ----
// 1 / x to ~1 ulp: MUFU.RCP64H seed and one third-order step (x > 0, normal)
__device__ __forceinline__ double fast_rcp(double x) {
    double x0;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x0) : "d"(x));
    const double e = fma(-x, x0, 1.0);
    return fma(x0, fma(e, e, e), x0);
}

// One plane rotation: (c, s) annihilating a_pq; rot = false (c = 1, s = 0) when
// |a_pq| <= max(eps sqrt(|a_pp a_qq|), floor_rel * max|diag|); big = the entry was above kBigRotation * sqrt(|a_pp a_qq|).
__device__ __forceinline__ void jacobi_rotation(double apq_raw, double app_raw, double aqq_raw, double inv_md, double floor_rel,
                                                double& c, double& s, bool& rot, bool& big) {
    // entries relative to the largest diagonal entry: the squares below neither overflow nor underflow
    const double apq = apq_raw * inv_md, app = app_raw * inv_md, aqq = aqq_raw * inv_md;
    c = 1.0; s = 0.0;
    // |a_pq| > eps sqrt(|a_pp a_qq|), compared squared (no square root on the chain)
    const double mag2 = apq * apq, dd = fabs(app * aqq);
    rot = mag2 > (kEps * kEps) * dd && fabs(apq) > floor_rel;
    big = rot && mag2 > (kBigRotation * kBigRotation) * dd;
    if (rot) {
        // t = sign(tau) / (|tau| + sqrt(1 + tau^2)), tau = (a_qq - a_pp) / (2 a_pq), with the first division folded away:
        // t = sign(d) o / (|d| + sqrt(d^2 + o^2)),  d = a_qq - a_pp,  o = 2 a_pq
        const double d = aqq - app, o = 2.0 * apq;
        const double x = fma(d, d, o * o);
        const double hyp = x * rsqrt(x);
        const double t = copysign(1.0, d) * o * fast_rcp(fabs(d) + hyp);
        c = rsqrt(fma(t, t, 1.0));
        s = t * c;
    }
}

// The two-sided update of one 2 x 2 block (rows = pair (p, q) with rotation (c, sn), columns = pair (p2, q2) with (c2, s2)):
// column rotation, then row rotation, in registers; the diagonal block writes the annihilated entry as an exact zero.
__device__ __forceinline__ void rotate_block(double c, double sn, double c2, double s2, bool diagonal,
                                             double& xpp, double& xpq, double& xqp, double& xqq) {
    if (s2 != 0.0) {  // columns (p2, q2):  [x y] <- [c2 x - s2 y, s2 x + c2 y]
        const double t0 = c2 * xpp - s2 * xpq, t1 = s2 * xpp + c2 * xpq;
        const double t2 = c2 * xqp - s2 * xqq, t3 = s2 * xqp + c2 * xqq;
        xpp = t0; xpq = t1; xqp = t2; xqq = t3;
    }
    if (sn != 0.0) {  // rows (p, q)
        const double t0 = c * xpp - sn * xqp, t2 = sn * xpp + c * xqp;
        const double t1 = c * xpq - sn * xqq, t3 = sn * xpq + c * xqq;
        xpp = t0; xpq = t1; xqp = t2; xqq = t3;
        if (diagonal) { xpq = 0.0; xqp = 0.0; }
    }
}

// One step = 32 disjoint rotations (parallel ordering): warp 0 computes them, then 512 threads apply A <- J^T A J block
// by block while 256 threads accumulate W <- W J (the two passes side by side: the shared-memory latency of one hides
// under the other).  Measured on B200 (clock64 around the phases, 63 steps of a full sub-problem): the rotation chain
// costs 500 (nothing rotates) to 950 cycles per step, the update 300 to 1000, both through shared memory (one step moves
// both matrices: 128 KB at 128 B / clock).  Tried and not kept: the rotations of the NEXT step computed by a 25th warp
// from a second matrix buffer while the update runs (one barrier per step, but every block must then be copied even
// when nothing rotates: late sweeps slow down as much as early ones gain), and W accumulated by a second CTA from a
// global log of the rotations (the fence that publishes a step costs what the shorter update saves).
constexpr int kSubThreads = 768;

__global__ void __launch_bounds__(kSubThreads) k_jacobi_sub(double* __restrict__ A, int64_t lda, int round, int nb,
                                                            double* __restrict__ Wbuf, unsigned long long* rot_count,
                                                            const double* __restrict__ maxdiag, double floor_rel,
                                                            int offdiag_only) {
    constexpr int LD = TS + 1;
    extern __shared__ __align__(16) double sub_sm[];
    double* a = sub_sm;
    double* w = sub_sm + TS * LD;
    __shared__ double cs_c[BS], cs_s[BS];
    __shared__ int pp[BS], qq[BS];

    const int tid = threadIdx.x;
    int I, J;
    rr_pair(round, blockIdx.x, nb, I, J);
    for (int e = tid; e < TS * TS; e += kSubThreads) {
        const int r = e / TS, c = e % TS;
        a[r * LD + c] = A[(int64_t)tile_index(I, J, r) * lda + tile_index(I, J, c)];
        w[r * LD + c] = (r == c) ? 1.0 : 0.0;
    }
    __syncthreads();
    // absolute floor (relative to the largest diagonal entry) below which entries are not rotated:
    // a safeguard only -- M = L^T L from the pivoted Cholesky factor is a graded PSD matrix on which the
    // relative test is stable, so the caller passes a tiny value
    const double md = maxdiag[0];
    const double inv_md = md > 0.0 ? 1.0 / md : 1.0;

    const int ur = (tid >> 4) & 31, uc = tid & 15;   // A: block row ur, block columns uc + 16k, k < 2
    const int wr = tid & 31, wi = (tid >> 5) & 7;    // W: pair wr, rows wi + 8k, k < 8
    int rotated = 0, rotated_big = 0;                // thread 0
    // offdiag_only: the 32 x 32 pairs (p in block I, q in block J) once each, 32 steps of 32 disjoint rotations -- the pairs
    // inside the two diagonal blocks meet in the ONE round of a sweep that runs all 63 steps (every block is in exactly one
    // pair of a round), so that a sweep rotates every pair of the matrix exactly once
    const int nsteps = offdiag_only ? BS : TS - 1;
    for (int step = 0; step < nsteps; ++step) {
        if (tid < BS) {
            int p, q;
            if (offdiag_only) { p = tid; q = BS + ((tid + step) & (BS - 1)); }
            else rr_pair(step, tid, TS, p, q);
            double c, s;
            bool rot, big;
            jacobi_rotation(a[p * LD + q], a[p * LD + p], a[q * LD + q], inv_md, floor_rel, c, s, rot, big);
            rotated += __popc(__ballot_sync(0xffffffffu, rot));
            rotated_big += __popc(__ballot_sync(0xffffffffu, big));
            cs_c[tid] = c; cs_s[tid] = s; pp[tid] = p; qq[tid] = q;
        }
        __syncthreads();
        if (tid < 512) {
            const double c = cs_c[ur], sn = cs_s[ur];
            const int p = pp[ur], q = qq[ur];
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int r2 = uc + 16 * k;
                const double c2 = cs_c[r2], s2 = cs_s[r2];
                if (sn == 0.0 && s2 == 0.0) continue;
                const int p2 = pp[r2], q2 = qq[r2];
                double xpp = a[p * LD + p2], xpq = a[p * LD + q2], xqp = a[q * LD + p2], xqq = a[q * LD + q2];
                rotate_block(c, sn, c2, s2, r2 == ur, xpp, xpq, xqp, xqq);
                a[p * LD + p2] = xpp; a[p * LD + q2] = xpq; a[q * LD + p2] = xqp; a[q * LD + q2] = xqq;
            }
        } else {
            const double c2 = cs_c[wr], s2 = cs_s[wr];
            if (s2 != 0.0) {
                const int p2 = pp[wr], q2 = qq[wr];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int i = wi + 8 * k;
                    const double u = w[i * LD + p2], v = w[i * LD + q2];
                    w[i * LD + p2] = c2 * u - s2 * v;
                    w[i * LD + q2] = s2 * u + c2 * v;
                }
            }
        }
        __syncthreads();
    }
    for (int e = tid; e < TS * TS; e += kSubThreads) {
        const int r = e / TS, c = e % TS;
        A[(int64_t)tile_index(I, J, r) * lda + tile_index(I, J, c)] = a[r * LD + c];
        Wbuf[(size_t)blockIdx.x * TS * TS + e] = w[r * LD + c];
    }
    if (tid == 0 && rotated)
        atomicAdd(rot_count, (unsigned long long)rotated + ((unsigned long long)rotated_big << 32));
}

// ---- (2) tile updates with DMMA:  A_kl <- W_k^T A_kl W_l  (k < l),  V_rl <- V_rl W_l ----
// 8 warps, warp tile 16 x 32 of the 64 x 64 output.
__device__ __forceinline__ void tile_gemm(const double* __restrict__ X, const double* __restrict__ Y, bool x_transposed,
                                          int warp, int lane, double (&acc)[2][4][2]) {
    // acc = op(X) * Y, op(X)[m][k] = x_transposed ? X[k][m] : X[m][k];  Y[k][n]
    const int g = lane >> 2, tig = lane & 3;
    const int m0 = (warp >> 1) * 16, n0 = (warp & 1) * 32;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
#pragma unroll 4
    for (int k0 = 0; k0 < TS; k0 += 4) {
        double af[2], bf[4];
#pragma unroll
        for (int i = 0; i < 2; ++i)
            af[i] = x_transposed ? X[(k0 + tig) * LDS_ + m0 + 8 * i + g] : X[(m0 + 8 * i + g) * LDS_ + k0 + tig];
#pragma unroll
        for (int j = 0; j < 4; ++j) bf[j] = Y[(k0 + tig) * LDS_ + n0 + 8 * j + g];
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
    }
}

__global__ void __launch_bounds__(256) k_jacobi_apply(double* __restrict__ A, double* __restrict__ V, int64_t lda,
                                                      int round, int nb, const double* __restrict__ Wbuf, int nA) {
    extern __shared__ __align__(16) double sm[];
    double* X = sm;
    double* Wl = X + TS * LDS_;
    double* Wk = Wl + TS * LDS_;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, tig = lane & 3;
    const int m0 = (warp >> 1) * 16, n0 = (warp & 1) * 32;
    const int P = nb / 2;
    double acc[2][4][2];

    if ((int)blockIdx.x < nA) {
        // tile pair (k, l), k < l, from the linear index
        int t = blockIdx.x, k = 0;
        while (t >= P - 1 - k) { t -= P - 1 - k; ++k; }
        const int l = k + 1 + t;
        int Ik, Jk, Il, Jl;
        rr_pair(round, k, nb, Ik, Jk);
        rr_pair(round, l, nb, Il, Jl);
        for (int e = tid; e < TS * TS; e += 256) {
            const int r = e >> 6, c = e & 63;
            X[r * LDS_ + c] = A[(int64_t)tile_index(Ik, Jk, r) * lda + tile_index(Il, Jl, c)];
            Wl[r * LDS_ + c] = Wbuf[(size_t)l * TS * TS + e];
            Wk[r * LDS_ + c] = Wbuf[(size_t)k * TS * TS + e];
        }
        __syncthreads();
        tile_gemm(X, Wl, false, warp, lane, acc);  // T = X W_l
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                X[(m0 + 8 * i + g) * LDS_ + n0 + 8 * j + 2 * tig] = acc[i][j][0];
                X[(m0 + 8 * i + g) * LDS_ + n0 + 8 * j + 2 * tig + 1] = acc[i][j][1];
            }
        __syncthreads();
        tile_gemm(Wk, X, true, warp, lane, acc);  // A' = W_k^T T
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t gr = tile_index(Ik, Jk, m0 + 8 * i + g);
                const int64_t gc = tile_index(Il, Jl, n0 + 8 * j + 2 * tig);  // gc, gc+1 are in the same block
                A[gr * lda + gc] = acc[i][j][0];
                A[gr * lda + gc + 1] = acc[i][j][1];
                A[gc * lda + gr] = acc[i][j][0];
                A[(gc + 1) * lda + gr] = acc[i][j][1];
            }
    } else {
        const int t = blockIdx.x - nA;
        const int l = t % P, rb = t / P;
        int Il, Jl;
        rr_pair(round, l, nb, Il, Jl);
        for (int e = tid; e < TS * TS; e += 256) {
            const int r = e >> 6, c = e & 63;
            X[r * LDS_ + c] = V[((int64_t)rb * TS + r) * lda + tile_index(Il, Jl, c)];
            Wl[r * LDS_ + c] = Wbuf[(size_t)l * TS * TS + e];
        }
        __syncthreads();
        tile_gemm(X, Wl, false, warp, lane, acc);
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t gr = (int64_t)rb * TS + m0 + 8 * i + g;
                const int64_t gc = tile_index(Il, Jl, n0 + 8 * j + 2 * tig);
                V[gr * lda + gc] = acc[i][j][0];
                V[gr * lda + gc + 1] = acc[i][j][1];
            }
    }
}

__global__ void k_pad_copy(const double* __restrict__ G, int64_t n, int64_t ldg, double* __restrict__ A,
                           double* __restrict__ V, int64_t npad) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= npad) return;
    A[i * npad + j] = (i < n && j < n) ? G[i * ldg + j] : 0.0;
    V[i * npad + j] = (i == j) ? 1.0 : 0.0;
}

// rank of each eigenvalue in descending order (ties by index), then scatter values and vectors
__global__ void k_rank(const double* __restrict__ A, int64_t n, int64_t lda, int* __restrict__ rank, double* __restrict__ w) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double li = A[i * lda + i];
    int r = 0;
    for (int64_t j = 0; j < n; ++j) {
        const double lj = A[j * lda + j];
        r += (lj > li) || (lj == li && j < i);
    }
    rank[i] = r;
    w[r] = li;
}
__global__ void k_scatter_vectors(const double* __restrict__ V, int64_t n, int64_t lda, const int* __restrict__ rank,
                                  double* __restrict__ W, int64_t ldw) {
    __shared__ double tile[32][33];
    __shared__ int rk[32];
    const int64_t i0 = (int64_t)blockIdx.y * 32, j0 = (int64_t)blockIdx.x * 32;
    if (threadIdx.y == 0 && j0 + threadIdx.x < n) rk[threadIdx.x] = rank[j0 + threadIdx.x];
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int64_t i = i0 + r, j = j0 + threadIdx.x;
        if (i < n && j < n) tile[r][threadIdx.x] = V[i * lda + j];
    }
    __syncthreads();
    // W[i][rank[j]] = V[i][j]
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int64_t i = i0 + r, j = j0 + threadIdx.x;
        if (i < n && j < n) W[i * ldw + rk[threadIdx.x]] = tile[r][threadIdx.x];
    }
}


// ---- (a) diagonally pivoted Cholesky, one persistent cooperative kernel ----
constexpr int kPcholThreads = 1024;
struct PcholArgs {
    const double* G;  // n x n symmetric, row-major, ld = n
    int64_t n;
    double* L;        // n x ldl row-major (zero-initialised); column k = k-th Cholesky column
    int64_t ldl;
    double* d;        // n: remaining diagonal
    int* rank_out;
    double tol;       // stop when max remaining diagonal <= tol * max initial diagonal
    double* slot_val;
    int64_t* slot_idx;
};

__device__ __forceinline__ void pc_merge(double& v, int64_t& i, double ov, int64_t oi) {
    if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
}

__global__ void __launch_bounds__(kPcholThreads, 1) k_pchol(PcholArgs a) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) double rowj[];  // n
    __shared__ double red_v[32];
    __shared__ int64_t red_i[32];
    __shared__ double piv_v;
    __shared__ int64_t piv_i;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t n = a.n, ldl = a.ldl;
    const int64_t gwarp = (int64_t)blockIdx.x * (kPcholThreads / 32) + warp;
    const int64_t nwarps = (int64_t)gridDim.x * (kPcholThreads / 32);
    double d0 = 0.0;
    int k = 0;
    for (;; ++k) {
        // ---- local arg-max of the remaining diagonal (k == 0: initialise it from G) ----
        double bv = -1.0;
        int64_t bi = INT64_MAX;
        if (k == 0) {
            for (int64_t i = (int64_t)blockIdx.x * kPcholThreads + tid; i < n; i += (int64_t)gridDim.x * kPcholThreads) {
                const double v = a.G[i * n + i];
                a.d[i] = v;
                pc_merge(bv, bi, v, i);
            }
        } else {
            // column k-1:  l = (G[:, j] - L[:, :k-1] L[j, :k-1]^T) / sqrt(d_j), warp per row
            const int64_t j = piv_i;
            const double sq = sqrt(piv_v);
            const int kc = k - 1;
            for (int e = tid; e < kc; e += kPcholThreads) rowj[e] = a.L[j * ldl + e];
            __syncthreads();
            for (int64_t i = gwarp; i < n; i += nwarps) {
                const double* Li = a.L + i * ldl;
                double dot = 0.0;
                for (int e = lane; e < kc; e += 32) dot = fma(Li[e], rowj[e], dot);
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, off);
                if (lane == 0) {
                    double col, dn;
                    if (i == j) { col = sq; dn = 0.0; }
                    else { col = (a.G[j * n + i] - dot) / sq; dn = a.d[i] - col * col; }
                    a.L[i * ldl + kc] = col;
                    a.d[i] = dn;
                    pc_merge(bv, bi, dn, i);
                }
            }
        }
        // ---- CTA reduce -> slot -> grid barrier -> global reduce ----
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, bv, off);
            const int64_t oi = __shfl_xor_sync(0xffffffffu, bi, off);
            pc_merge(bv, bi, ov, oi);
        }
        if (lane == 0) { red_v[warp] = bv; red_i[warp] = bi; }
        __syncthreads();
        if (warp == 0) {
            double v = red_v[lane];
            int64_t i = red_i[lane];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                const double ov = __shfl_xor_sync(0xffffffffu, v, off);
                const int64_t oi = __shfl_xor_sync(0xffffffffu, i, off);
                pc_merge(v, i, ov, oi);
            }
            if (lane == 0) {
                a.slot_val[(size_t)(k & 1) * gridDim.x + blockIdx.x] = v;
                a.slot_idx[(size_t)(k & 1) * gridDim.x + blockIdx.x] = i;
            }
        }
        grid.sync();
        if (warp == 0) {
            double v = -1.0;
            int64_t i = INT64_MAX;
            for (unsigned s = lane; s < gridDim.x; s += 32)
                pc_merge(v, i, __ldcg(a.slot_val + (size_t)(k & 1) * gridDim.x + s),
                         __ldcg(a.slot_idx + (size_t)(k & 1) * gridDim.x + s));
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                const double ov = __shfl_xor_sync(0xffffffffu, v, off);
                const int64_t oi = __shfl_xor_sync(0xffffffffu, i, off);
                pc_merge(v, i, ov, oi);
            }
            if (lane == 0) { piv_v = v; piv_i = i; }
        }
        __syncthreads();
        if (k == 0) d0 = piv_v;
        if (k >= n || !(piv_v > a.tol * d0) || !(piv_v > 0.0)) break;
    }
    if (blockIdx.x == 0 && tid == 0) *a.rank_out = k;
}

__global__ void k_mirror_upper(double* __restrict__ M, int64_t n, int64_t ld) {
    __shared__ double tile[32][33];
    const int bi = blockIdx.y, bj = blockIdx.x;
    if (bj < bi) return;
    const int64_t i0 = (int64_t)bi * 32, j0 = (int64_t)bj * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int64_t i = i0 + r, j = j0 + threadIdx.x;
        tile[r][threadIdx.x] = (i < n && j < n) ? M[i * ld + j] : 0.0;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int64_t j = j0 + r, i = i0 + threadIdx.x;
        if (i < n && j < n && j > i) M[j * ld + i] = tile[threadIdx.x][r];
    }
}

// out[0] = max_{i != j} |a_ij| / sqrt(a_ii a_jj): the departure from orthogonality of the columns whose Gram matrix A is
// (one-sided Jacobi / Hestenes criterion).  Pairs with a non-positive diagonal product are skipped.  out is zeroed by the caller;
// non-negative doubles order like their bit patterns, so the grid-wide maximum is one 64-bit atomicMax per CTA.
__global__ void __launch_bounds__(256) k_scaled_defect(const double* __restrict__ A, int64_t n, int64_t lda, double* out) {
    __shared__ double red[256];
    const int64_t i = blockIdx.x;
    const double aii = A[i * lda + i];
    double v = 0.0;
    for (int64_t j = threadIdx.x; j < n; j += blockDim.x) {
        if (j == i) continue;
        const double d = aii * A[j * lda + j];
        if (d > 0.0) v = fmax(v, fabs(A[i * lda + j]) * rsqrt(d));
    }
    red[threadIdx.x] = v;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) red[threadIdx.x] = fmax(red[threadIdx.x], red[threadIdx.x + s]);
        __syncthreads();
    }
    if (threadIdx.x == 0) atomicMax(reinterpret_cast<unsigned long long*>(out), (unsigned long long)__double_as_longlong(red[0]));
}

// X[:, c] *= 1/sqrt(lam_c) (zero column when lam_c <= 0);  w[c] = max(lam_c, 0)
__global__ void k_scale_columns(double* __restrict__ X, int64_t r, int64_t ld, const double* __restrict__ lam,
                                double* __restrict__ w) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, row = blockIdx.y;
    if (c >= r) return;
    const double l = lam[c];
    const double sc = l > 0.0 ? rsqrt(l) : 0.0;
    X[row * ld + c] *= sc;
    if (row == 0) w[c] = l > 0.0 ? l : 0.0;
}

}  // namespace

// Block Jacobi on a dense symmetric matrix Ain (n x n, row-major, leading dimension lda_in).
// w (n, descending) and Wout (n x n row-major, leading dimension ldw, eigenvector k in column k).
static int jacobi_eig(rt_handle* h, const double* Ain, int64_t n, int64_t lda_in, double* w, double* Wout, int64_t ldw,
                      double floor_rel, int* sweeps_out, cudaStream_t st) {
    const int64_t npad = ceil_div64(n, TS) * TS;
    const int nb = (int)(npad / BS), P = nb / 2;
    void* buf;
    const size_t mat = (size_t)npad * npad * sizeof(double);
    const size_t wbytes = (size_t)P * TS * TS * sizeof(double);
    int rc = rt_scratch(h, 5, 2 * mat + wbytes + 1024, &buf);
    if (rc) return rc;
    double* A = (double*)buf;
    double* V = A + npad * npad;
    double* Wbuf = V + npad * npad;
    double* maxdiag = Wbuf + (size_t)P * TS * TS;
    unsigned long long* counter = (unsigned long long*)(maxdiag + 1);
    int* rank;
    rc = rt_scratch(h, 6, (size_t)n * sizeof(int), (void**)&rank);
    if (rc) return rc;

    k_pad_copy<<<dim3((unsigned)ceil_div64(npad, 256), (unsigned)npad), 256, 0, st>>>(Ain, n, lda_in, A, V, npad);
    RT_LAUNCH_CHECK(h);
    k_maxdiag<<<1, 256, 0, st>>>(A, n, npad, maxdiag);
    RT_LAUNCH_CHECK(h);
    const int apply_smem = 3 * TS * LDS_ * (int)sizeof(double);
    RT_CUDA(h, cudaFuncSetAttribute(k_jacobi_apply, cudaFuncAttributeMaxDynamicSharedMemorySize, apply_smem));
    const int sub_smem = 2 * TS * (TS + 1) * (int)sizeof(double);
    RT_CUDA(h, cudaFuncSetAttribute(k_jacobi_sub, cudaFuncAttributeMaxDynamicSharedMemorySize, sub_smem));
    const int nA = P * (P - 1) / 2, nV = (int)(npad / TS) * P;
    const int rounds = nb - 1;
    const int max_sweeps = 60;
    // one inner sweep per sub-problem and round: the outer iteration converges in the same number of sweeps (7-9 on the
    // POD Gram matrices) as with fully converged sub-problems, each 3-5x cheaper
    const bool pair_once = !(getenv("RT_EIG_PAIR_ONCE") && atoi(getenv("RT_EIG_PAIR_ONCE")) == 0);
    int sweeps_used = -1;
    for (int sweep = 0; sweep < max_sweeps; ++sweep) {
        RT_CUDA(h, cudaMemsetAsync(counter, 0, sizeof(unsigned long long), st));
        for (int round = 0; round < rounds; ++round) {
            k_jacobi_sub<<<P, kSubThreads, sub_smem, st>>>(A, npad, round, nb, Wbuf, counter, maxdiag, floor_rel,
                                                           pair_once && round > 0 ? 1 : 0);
            RT_LAUNCH_CHECK(h);
            k_jacobi_apply<<<nA + nV, 256, apply_smem, st>>>(A, V, npad, round, nb, Wbuf, nA);
            RT_LAUNCH_CHECK(h);
        }
        unsigned long long rot = 0;
        RT_CUDA(h, cudaMemcpyAsync(&rot, counter, sizeof(rot), cudaMemcpyDeviceToHost, st));
        RT_CUDA(h, cudaStreamSynchronize(st));
        if (getenv("RT_EIG_DEBUG"))
            fprintf(stderr, "[rt_sym_eig] n=%lld sweep %d rotations %llu (above %.0e: %llu)\n", (long long)n, sweep,
                    rot & 0xffffffffull, kBigRotation, rot >> 32);
        if ((rot >> 32) == 0) { sweeps_used = sweep + 1; break; }
    }
    k_rank<<<(unsigned)ceil_div64(n, 128), 128, 0, st>>>(A, n, npad, rank, w);
    RT_LAUNCH_CHECK(h);
    k_scatter_vectors<<<dim3((unsigned)ceil_div64(n, 32), (unsigned)ceil_div64(n, 32)), dim3(32, 8), 0, st>>>(V, n, npad, rank, Wout, ldw);
    RT_LAUNCH_CHECK(h);
    *sweeps_out = sweeps_used;
    if (sweeps_used < 0) return rt_fail(h, RT_ENOTCONV, "block Jacobi did not converge in %d sweeps (n=%lld)", max_sweeps, (long long)n);
    return RT_OK;
}

extern "C" int rt_sym_eig_f64(rt_handle* h, double* G, int64_t n, double* w, double* W, int* info, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, G && w && W && n >= 1 && n <= 16384);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);

    // ---- (a) pivoted Cholesky  G ~= L L^T ----
    const int64_t ldl = (n + 1) & ~int64_t(1);
    double* L;
    int rc = rt_scratch(h, 0, ((size_t)n * ldl + n + 8) * sizeof(double), (void**)&L);
    if (rc) return rc;
    double* d = L + (size_t)n * ldl;
    int* rank_dev = (int*)(d + n);
    PcholArgs a;
    a.G = G; a.n = n; a.L = L; a.ldl = ldl; a.d = d; a.rank_out = rank_dev;
    a.tol = 8.0 * kEps;
    int grid = (int)ceil_div64(n, kPcholThreads / 32);
    if (grid > h->sm_count) grid = h->sm_count;
    void* slots;
    rc = rt_scratch(h, 1, (size_t)grid * 2 * (sizeof(double) + sizeof(int64_t)), &slots);
    if (rc) return rc;
    a.slot_val = (double*)slots;
    a.slot_idx = (int64_t*)((double*)slots + 2 * (size_t)grid);
    const size_t smem = (size_t)n * sizeof(double) + 1024;
    RT_CUDA(h, cudaFuncSetAttribute(k_pchol, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    RT_CUDA(h, cudaMemsetAsync(L, 0, (size_t)n * ldl * sizeof(double), st));
    void* params[] = {&a};
    RT_CUDA(h, cudaLaunchCooperativeKernel((const void*)k_pchol, dim3(grid), dim3(kPcholThreads), params, smem, st));
    RT_LAUNCH_CHECK(h);
    int r = 0;
    RT_CUDA(h, cudaMemcpyAsync(&r, rank_dev, sizeof(int), cudaMemcpyDeviceToHost, st));
    RT_CUDA(h, cudaStreamSynchronize(st));
    if (getenv("RT_EIG_DEBUG")) fprintf(stderr, "[rt_sym_eig] n=%lld pivoted-Cholesky rank %d\n", (long long)n, r);
    RT_CUDA(h, cudaMemsetAsync(w, 0, (size_t)n * sizeof(double), st));
    RT_CUDA(h, cudaMemsetAsync(W, 0, (size_t)n * n * sizeof(double), st));
    int sweeps = 0;
    if (r > 0) {
        // ---- (b) M = L^T L ----
        const int64_t ldm = ((int64_t)r + 1) & ~int64_t(1);
        double* M;
        rc = rt_scratch(h, 4, ((size_t)2 * r * ldm + r) * sizeof(double), (void**)&M);
        if (rc) return rc;
        double* X = M + (size_t)r * ldm;
        double* lam = X + (size_t)r * ldm;
        dg::Operand opL{L, r, n, 1, ldl};  // (mn = column of L, k = row)
        rc = gemm_tn(h, opL, opL, M, ldm, true, st);
        if (rc) return rc;
        k_mirror_upper<<<dim3((unsigned)ceil_div64(r, 32), (unsigned)ceil_div64(r, 32)), dim3(32, 8), 0, st>>>(M, r, ldm);
        RT_LAUNCH_CHECK(h);
        // ---- (c) block Jacobi on M ----
        rc = jacobi_eig(h, M, r, ldm, lam, X, ldm, 1e-12 * kEps, &sweeps, st);
        if (rc) return rc;
        // ---- (d) W[:, :r] = L X diag(lam)^-1/2 ----
        k_scale_columns<<<dim3((unsigned)ceil_div64(r, 256), (unsigned)r), 256, 0, st>>>(X, r, ldm, lam, w);
        RT_LAUNCH_CHECK(h);
        dg::Operand opA{L, n, r, ldl, 1};   // rows of L, k = pivot column
        dg::Operand opB{X, r, r, 1, ldm};   // (mn = eigenvector index c, k): X[k][c]
        dg::Params p = dg::default_params();
        p.C = W; p.ldc_m = n; p.ldc_n = 1;
        rc = dg::launch(h, opA, opB, p, st);
        if (rc) return rc;
    }
    if (info) RT_CUDA(h, cudaMemcpyAsync(info, &sweeps, sizeof(int), cudaMemcpyHostToDevice, st));
    RT_CUDA(h, cudaStreamSynchronize(st));
    return RT_OK;
}

// Plain (un-preconditioned) block Jacobi on a symmetric matrix: ALL n eigenpairs, W a product of plane rotations
// (orthogonal to working precision), relative rotation criterion |a_pq| > eps sqrt(a_pp a_qq) -- on a scaled diagonally
// dominant matrix (the Gram matrix of nearly orthogonal, arbitrarily graded columns) every eigenvalue comes out with a
// small RELATIVE error (Demmel-Veselic).  This is the rotation step of the Hestenes refinement behind orth
// (romtime_b200/rom/pod.py): H = Y^T Y of the lifted modes Y = S W.  A (n x n, row-major, leading dimension n) is not
// modified.  defect (device double, may be NULL) receives max |a_ij| / sqrt(a_ii a_jj) of the INPUT.
extern "C" int rt_sym_eig_jacobi_f64(rt_handle* h, const double* A, int64_t n, double* w, double* W, int* info,
                                     double* defect, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, A && w && W && n >= 1 && n <= 16384);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    if (defect) {
        RT_CUDA(h, cudaMemsetAsync(defect, 0, sizeof(double), st));
        k_scaled_defect<<<(unsigned)n, 256, 0, st>>>(A, n, n, defect);
        RT_LAUNCH_CHECK(h);
    }
    int sweeps = 0;
    int rc = jacobi_eig(h, A, n, n, w, W, n, 0.0, &sweeps, st);
    if (rc) return rc;
    if (info) RT_CUDA(h, cudaMemcpyAsync(info, &sweeps, sizeof(int), cudaMemcpyHostToDevice, st));
    RT_CUDA(h, cudaStreamSynchronize(st));
    return RT_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// rt_sym_eig_jacobi_static_f64: the plain block Jacobi of rt_sym_eig_jacobi_f64 for callers that may not allocate (code under
// CUDA-graph capture set-up, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Its work space (padded copy of the matrix, rotation product, sub-problem rotations,
// counters, ordering) lives in ONE file-scope device array of fixed size, so the order is limited to kStaticEigMaxN.
// Every GPU that loads the library holds the array (about 322 KB) until exit; all calls on a device share it, so calls that
// overlap on the GPU (different streams, different host threads) race -- serialise them.  Nothing in it has to survive
// between calls: every call rewrites the padded matrix and the identity, and clears its rotation counter.
constexpr int kStaticEigMaxN = 128;
constexpr int64_t kStaticPad = (kStaticEigMaxN + TS - 1) / TS * TS;               // 128
constexpr int kStaticPairs = (int)(kStaticPad / BS / 2);                           // 2 sub-problems per round
constexpr size_t kStaticDoubles = 2 * kStaticPad * kStaticPad + (size_t)kStaticPairs * TS * TS + 16;
constexpr size_t kStaticBytes = kStaticDoubles * sizeof(double) + kStaticEigMaxN * sizeof(int);
static __device__ __align__(256) unsigned char g_static_eig_scratch[kStaticBytes];

// In host code the array's name is only a stand-in: the device address comes from cudaGetSymbolAddress, once per device,
// with that device current.
static int static_eig_scratch(rt_handle* h, unsigned char** out) {
    constexpr int kMaxDevices = 64;
    static unsigned char* addr[kMaxDevices] = {nullptr};
    if (h->device < 0 || h->device >= kMaxDevices) return rt_fail(h, RT_EINVAL, "device index %d beyond %d", h->device, kMaxDevices);
    if (!addr[h->device]) {
        void* p = nullptr;
        RT_CUDA(h, cudaGetSymbolAddress(&p, g_static_eig_scratch));
        if (!p) return rt_fail(h, RT_ECUDA, "cudaGetSymbolAddress returned a null address for the static scratch array");
        addr[h->device] = (unsigned char*)p;
    }
    *out = addr[h->device];
    return RT_OK;
}

extern "C" int rt_sym_eig_jacobi_static_f64(rt_handle* h, const double* Ain, int64_t n, double* w, double* W, int* info,
                                            double* defect, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, Ain && w && W && n >= 1);
    if (n > kStaticEigMaxN)
        return rt_fail(h, RT_EINVAL, "rt_sym_eig_jacobi_static_f64: n = %lld exceeds the static work space (n <= %d); "
                       "use rt_sym_eig_jacobi_f64", (long long)n, kStaticEigMaxN);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    unsigned char* base;
    int rc = static_eig_scratch(h, &base);
    if (rc) return rc;
    const int64_t npad = ceil_div64(n, TS) * TS;
    const int nb = (int)(npad / BS), P = nb / 2;
    double* A = (double*)base;
    double* V = A + npad * npad;
    double* Wbuf = V + npad * npad;
    double* maxdiag = Wbuf + (size_t)P * TS * TS;
    unsigned long long* counter = (unsigned long long*)(maxdiag + 1);
    int* rank = (int*)(base + kStaticDoubles * sizeof(double));

    if (defect) {
        RT_CUDA(h, cudaMemsetAsync(defect, 0, sizeof(double), st));
        k_scaled_defect<<<(unsigned)n, 256, 0, st>>>(Ain, n, n, defect);
        RT_LAUNCH_CHECK(h);
    }
    k_pad_copy<<<dim3((unsigned)ceil_div64(npad, 256), (unsigned)npad), 256, 0, st>>>(Ain, n, n, A, V, npad);
    RT_LAUNCH_CHECK(h);
    k_maxdiag<<<1, 256, 0, st>>>(A, n, npad, maxdiag);
    RT_LAUNCH_CHECK(h);
    const int apply_smem = 3 * TS * LDS_ * (int)sizeof(double);
    RT_CUDA(h, cudaFuncSetAttribute(k_jacobi_apply, cudaFuncAttributeMaxDynamicSharedMemorySize, apply_smem));
    const int sub_smem = 2 * TS * (TS + 1) * (int)sizeof(double);
    RT_CUDA(h, cudaFuncSetAttribute(k_jacobi_sub, cudaFuncAttributeMaxDynamicSharedMemorySize, sub_smem));
    const int nA = P * (P - 1) / 2, nV = (int)(npad / TS) * P;
    const int rounds = nb - 1, max_sweeps = 60;
    int sweeps = -1;
    for (int sweep = 0; sweep < max_sweeps; ++sweep) {
        RT_CUDA(h, cudaMemsetAsync(counter, 0, sizeof(unsigned long long), st));
        for (int round = 0; round < rounds; ++round) {
            k_jacobi_sub<<<P, kSubThreads, sub_smem, st>>>(A, npad, round, nb, Wbuf, counter, maxdiag, 0.0, 0);
            RT_LAUNCH_CHECK(h);
            k_jacobi_apply<<<nA + nV, 256, apply_smem, st>>>(A, V, npad, round, nb, Wbuf, nA);
            RT_LAUNCH_CHECK(h);
        }
        unsigned long long rot = 0;
        RT_CUDA(h, cudaMemcpyAsync(&rot, counter, sizeof(rot), cudaMemcpyDeviceToHost, st));
        RT_CUDA(h, cudaStreamSynchronize(st));
        if ((rot >> 32) == 0) { sweeps = sweep + 1; break; }
    }
    k_rank<<<(unsigned)ceil_div64(n, 128), 128, 0, st>>>(A, n, npad, rank, w);
    RT_LAUNCH_CHECK(h);
    k_scatter_vectors<<<dim3((unsigned)ceil_div64(n, 32), (unsigned)ceil_div64(n, 32)), dim3(32, 8), 0, st>>>(V, n, npad, rank, W, n);
    RT_LAUNCH_CHECK(h);
    if (sweeps < 0) return rt_fail(h, RT_ENOTCONV, "block Jacobi did not converge in %d sweeps (n=%lld)", max_sweeps, (long long)n);
    if (info) RT_CUDA(h, cudaMemcpyAsync(info, &sweeps, sizeof(int), cudaMemcpyHostToDevice, st));
    RT_CUDA(h, cudaStreamSynchronize(st));
    return RT_OK;
}
