// Online stage of the hyper-reduced ROM (reference rom/rom.py:430-555, :877-929; deim/deim.py:416-493):
// theta solve, reduced-operator assembly and the batched implicit BDF sweep.
//
// Per time step and per chunk of parameters (Bc = 37*128 so that the assembly GEMM fills 148 SMs
// in whole waves) the sweep launches
//   k_state      : features phi(mu,t) (piston closed forms), u* and the BDF history vector w
//   dgemm (TU)   : state-dependent part of the trilinear sampled entries  T u*
//   k_entries    : sampled entries  f = scale * (C phi + b0 T u*)              (deim/deim.py:437-438)
//   (setup)      : Atil = basis_rom PT_U^-1 for all six reductors in ONE block-diagonal GEMM, once per model,
//                  so that sum_q theta_q A_q == sum_k f_k Atil_k and no theta solve is left per step  (:491-492)
//   dgemm (K)    : [K | M] = F^T Atil -- sum_q theta_q A_q^r for 5 operators, K = bdf M + dt(...)
//                  accumulated in one pass; M is emitted at the k-boundary (hook)   (:445, rom.py:906)
//   dgemm (f)    : reduced forcing  dt * basis_rhs theta_rhs
//   solve        : b = M w + f, u = K^-1 b with partial pivoting (rom.py:492): k_gj_smem (blocked Gauss-Jordan on DMMA
//                  tiles, 32 < N <= 64), k_lu_shift (register LU, N <= 32), k_lu_solve (shared-memory LU, N > 64)
// All GEMMs run on the TMA + DMMA core (dgemm_tma.cuh).  No collective: parameters are independent.
#include <utility>
#include <vector>

#include "gemm_util.cuh"

namespace {

// ---- optional per-kernel-class timing (rt_profile_enable / rt_profile_read) ----
struct Profile {
    bool on = false;
    std::vector<cudaEvent_t> pool;
    std::vector<int> cls;   // class of event pair i (events 2i, 2i+1)
    size_t used = 0;        // pairs in flight
    double ms[RT_PROFILE_CLASSES] = {0};
    int64_t cnt[RT_PROFILE_CLASSES] = {0};
};
struct ProfScope {
    Profile* p; cudaStream_t st; size_t idx;
    ProfScope(Profile* prof, bool active, int c, cudaStream_t s) : p(active ? prof : nullptr), st(s), idx(0) {
        if (!p) return;
        if (p->pool.size() < 2 * (p->used + 1)) {
            cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
            p->pool.push_back(a); p->pool.push_back(b); p->cls.push_back(c);
        }
        idx = p->used++;
        p->cls[idx] = c;
        cudaEventRecord(p->pool[2 * idx], st);
    }
    ~ProfScope() { if (p) cudaEventRecord(p->pool[2 * idx + 1], st); }
};

constexpr double kGamma = 1.4;  // fom/nonlinear.py:31
constexpr int kNOps = 6;
constexpr int kTril = 4;        // ops: mass, stiffness, convection, nonlinear_lifting, trilinear, rhs
constexpr int kRhs = 5;

// ------------------------------------------------------------------ small dense inverse
// In-place Gauss-Jordan with partial pivoting in shared memory (one CTA), m <= 160.
__global__ void __launch_bounds__(1024) k_invert(const double* __restrict__ A, int m, double* __restrict__ Ainv, int* info) {
    extern __shared__ __align__(16) double sm[];
    const int ld = m | 1;
    double* a = sm;                                   // m * ld
    int* piv = reinterpret_cast<int*>(a + (size_t)m * ld);  // m
    __shared__ int prow;
    __shared__ double pval;
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int e = tid; e < m * m; e += nt) a[(e / m) * ld + (e % m)] = A[e];
    __syncthreads();
    for (int k = 0; k < m; ++k) {
        if (tid < 32) {
            double best = -1.0;
            int bi = k;
            for (int i = k + tid; i < m; i += 32) {
                const double v = fabs(a[i * ld + k]);
                if (v > best) { best = v; bi = i; }
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                const double ov = __shfl_xor_sync(0xffffffffu, best, off);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
                if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
            }
            if (tid == 0) {
                prow = bi;
                piv[k] = bi;
                pval = a[bi * ld + k];
                if (best == 0.0 && *info == 0) *info = k + 1;
            }
        }
        __syncthreads();
        const int p = prow;
        if (p != k)
            for (int j = tid; j < m; j += nt) { const double t = a[k * ld + j]; a[k * ld + j] = a[p * ld + j]; a[p * ld + j] = t; }
        __syncthreads();
        const double rinv = 1.0 / pval;
        for (int j = tid; j < m; j += nt) a[k * ld + j] = (j == k) ? rinv : a[k * ld + j] * rinv;
        __syncthreads();
        // eliminate column k from every other row; column k receives -f * rinv
        for (int e = tid; e < m * m; e += nt) {
            const int i = e / m, j = e % m;
            if (i == k || j == k) continue;
            a[i * ld + j] = fma(-a[i * ld + k], a[k * ld + j], a[i * ld + j]);
        }
        __syncthreads();
        for (int i = tid; i < m; i += nt)
            if (i != k) a[i * ld + k] = -a[i * ld + k] * rinv;
        __syncthreads();
    }
    // undo the row interchanges as column interchanges, in reverse order
    for (int k = m - 1; k >= 0; --k) {
        const int p = piv[k];
        if (p != k)
            for (int i = tid; i < m; i += nt) { const double t = a[i * ld + k]; a[i * ld + k] = a[i * ld + p]; a[i * ld + p] = t; }
        __syncthreads();
    }
    for (int e = tid; e < m * m; e += nt) Ainv[e] = a[(e / m) * ld + (e % m)];
}

// ------------------------------------------------------------------ sweep plan
struct SweepPlan {
    int N, Npad, N2, ldk;       // ldk: row stride of K / M / basis_cat rows (even)
    int m[kNOps], mpad[kNOps], qoff[kNOps + 1];
    int Qmat, Qall;             // rows of the 5 matrix operators / all six (multiples of 16)
    int evaluator, n_mu;
    int Bc, Bcpad;              // parameters in this chunk / padded row stride of q-major buffers
    int ldu;                    // row stride of the state vectors Un / Unm1 (Npad internally, N for caller-owned state)
    int qsum;                   // sum of the unpadded m[i] (row length of caller-supplied sampled entries)
    int msum[kNOps + 1];        // prefix sums of m[i]
};

// phi(mu, t): problems/piston.py:18-60 ("rest" / "sudden"), fom/nonlinear.py:159-166 (b0),
// :203-233 (lifting g = bL x / L), :277-299 (mesh velocity w = x L'/L)
__device__ __forceinline__ void piston_features(int which, double a0, double om, double de, double t, double* phi) {
    double L, dL, bL, dbL, s, c;
    sincos(om * t, &s, &c);
    if (which == 0) {
        L = 1.0 - de * (1.0 - c);
        dL = -om * de * s;
        bL = -de * (om / a0) * s;
        dbL = -de * om * (om / a0) * c;
    } else {
        L = 1.0 - de * s;
        dL = -om * de * c;
        bL = -de * (om / a0) * c;
        dbL = de * om * (om / a0) * s;
    }
    const double b0 = (kGamma + 1.0) / 2.0 * a0;
    phi[0] = 1.0; phi[1] = L; phi[2] = 1.0 / L; phi[3] = dL; phi[4] = a0; phi[5] = b0 * bL;
    phi[6] = dbL * L; phi[7] = b0 * bL * bL; phi[8] = a0 * bL; phi[9] = dL * bL; phi[10] = bL / L; phi[11] = b0;
}

// MFP1 heat problem (problems/mfp1.py:27-75): L = 1 - sin(omega t) (omega = 0: fixed domain), boundary data b0 = 1 - E,
// bL = (1 - E)(1 + delta^2 L^2), forcing beta E (1 + delta^2 x^2) - 2 delta^2 alpha_0 (1 - E); the features the P1 weak forms
// of fom/heat.py:80-297 and the lifting of fom/base.py:377-495 reduce to (romtime_b200/evaluators.py Mfp1P1)
__device__ __forceinline__ void mfp1_features(double de, double be, double al, double om, double t, double* phi) {
    double s, c;
    sincos(om * t, &s, &c);
    const double L = 1.0 - s, dL = -om * c, E = exp(-be * t), d2 = de * de;
    phi[0] = 1.0; phi[1] = L; phi[2] = al / L; phi[3] = dL; phi[4] = be * E * L; phi[5] = be * E * d2 * L * L * L;
    phi[6] = al * (1.0 - E) * d2 * L; phi[7] = (1.0 - E) * d2 * dL * L * L;
    phi[8] = 0.0; phi[9] = 0.0; phi[10] = 0.0; phi[11] = 0.0;
}

// Phi[f][p], Ustar[p][j] = u* (rom.py:897-901), Wv[p][j] = BDF history vector (rom.py:919-925)
__global__ void k_state(SweepPlan pl, const double* __restrict__ mu, double t, int have_prev, int bdf2,
                        const double* __restrict__ Un, const double* __restrict__ Unm1, double* __restrict__ Phi,
                        double* __restrict__ Ustar, double* __restrict__ Wv) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (mu && e < pl.Bc) {
        double phi[RT_NF];
        const double* m = mu + e * pl.n_mu;
        if (pl.evaluator == 2) mfp1_features(m[0], m[1], m[2], m[3], t, phi);
        else piston_features(pl.evaluator, m[0], m[1], m[2], t, phi);
#pragma unroll
        for (int f = 0; f < RT_NF; ++f) Phi[(int64_t)f * pl.Bcpad + e] = phi[f];
    }
    if (e < (int64_t)pl.Bc * pl.Npad) {
        const int64_t p = e / pl.Npad;
        const int j = (int)(e - p * pl.Npad);
        const double un = j < pl.N ? Un[p * pl.ldu + j] : 0.0, um = j < pl.N ? Unm1[p * pl.ldu + j] : 0.0;
        Ustar[e] = have_prev ? 2.0 * un - um : un;
        Wv[e] = bdf2 ? 2.0 * un - 0.5 * um : un;
    }
}

// Caller-supplied sampled entries (rt_online_step_f64): Fext[p][msum[op] + k] -> F[qoff[op] + k][p], scaled
// like k_entries (mass unscaled, everything else times dt); padding rows are zero.
__global__ void __launch_bounds__(256) k_entries_ext(SweepPlan pl, const double* __restrict__ Fext, double dt,
                                                     double* __restrict__ F) {
    const int p = blockIdx.x * 256 + threadIdx.x;
    if (p >= pl.Bc) return;
    const int q0 = blockIdx.y * 32;
    for (int q = q0; q < q0 + 32 && q < pl.Qall; ++q) {
        int op = 0;
        while (q >= pl.qoff[op + 1]) ++op;
        const int k = q - pl.qoff[op];
        double v = 0.0;
        if (k < pl.m[op]) {
            v = Fext[(int64_t)p * pl.qsum + pl.msum[op] + k];
            if (op != 0) v *= dt;
        }
        F[(int64_t)q * pl.Bcpad + p] = v;
    }
}

// F[q][p] = scale_op * ( sum_f Ccat[q][f] Phi[f][p] + [trilinear] b0_p * TU[q - qoff][p] )
// blockIdx.y = a group of 16 rows q (operator blocks start at multiples of 16, so a group has ONE operator);
// the group's 16 x RT_NF coefficients are staged in shared memory, one thread per parameter.
__global__ void __launch_bounds__(128) k_entries(SweepPlan pl, const double* __restrict__ Ccat, const double* __restrict__ Phi,
                                                 const double* __restrict__ TU, double dt, double* __restrict__ F) {
    // two parameters per thread (Bcpad is a multiple of 128: every row of Phi / TU / F is 16-byte aligned at an even p),
    // 16-byte loads and stores: half the instructions and twice the bytes in flight per thread of the one-parameter form
    __shared__ double cs[16 * RT_NF];
    const int q0 = blockIdx.y * 16;
    for (int e = threadIdx.x; e < 16 * RT_NF; e += 128) cs[e] = Ccat[(int64_t)q0 * RT_NF + e];
    __syncthreads();
    const int p = 2 * (blockIdx.x * 128 + threadIdx.x);
    if (p >= pl.Bc) return;
    int op = 0;
    while (q0 >= pl.qoff[op + 1]) ++op;
    const int k0 = q0 - pl.qoff[op];
    const int nvalid = pl.m[op] - k0 < 16 ? pl.m[op] - k0 : 16;  // rows beyond m[op] are zero padding
    double2 phi[RT_NF];
#pragma unroll
    for (int f = 0; f < RT_NF; ++f) phi[f] = *reinterpret_cast<const double2*>(Phi + (int64_t)f * pl.Bcpad + p);
    const bool tril = op == kTril && TU != nullptr;
    const double scale = op == 0 ? 1.0 : dt;
    double2 tu[16];
    if (tril) {
#pragma unroll
        for (int r = 0; r < 16; ++r)
            tu[r] = r < nvalid ? *reinterpret_cast<const double2*>(TU + (int64_t)(k0 + r) * pl.Bcpad + p) : make_double2(0.0, 0.0);
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) {
        double2 v = make_double2(0.0, 0.0);
        if (r < nvalid) {
#pragma unroll
            for (int f = 0; f < RT_NF; ++f) { v.x = fma(cs[r * RT_NF + f], phi[f].x, v.x); v.y = fma(cs[r * RT_NF + f], phi[f].y, v.y); }
            if (tril) { v.x = fma(phi[11].x, tu[r].x, v.x); v.y = fma(phi[11].y, tu[r].y, v.y); }
            v.x *= scale; v.y *= scale;
        }
        *reinterpret_cast<double2*>(F + (int64_t)(q0 + r) * pl.Bcpad + p) = v;
    }
}

// Skinny products of the step (N <= 64, m <= 128): 32 parameters per CTA on the FP64 FMA pipe.  On the 128 x 128
// tiles of the GEMM core they fill 37 of 148 SMs and pay a pipeline fill each; here every SM takes 32 parameters.
//   TU[q][p] = sum_j T[q][j] u*[p][j]        (trilinear sampled entries, deim/nonlinear.py:247-283)
// thread (pp = t % 32, qg = t / 32): 16 rows q for one parameter; T and the 32 u* rows are staged in shared memory
__global__ void __launch_bounds__(256) k_tu(SweepPlan pl, const double* __restrict__ Tpad, const double* __restrict__ Ustar,
                                            int mT, double* __restrict__ TU) {
    extern __shared__ __align__(16) double sm_tu[];
    const int Np = pl.Npad, ldu = Np + 2;
    double* Ts = sm_tu;            // 128 x Np (rows >= mT zero)
    double* us = Ts + 128 * Np;    // 32 x ldu
    const int t = threadIdx.x, pp = t & 31, qg = t >> 5;
    const int64_t p0 = (int64_t)blockIdx.x * 32;
#pragma unroll 8
    for (int e = t; e < 128 * Np; e += 256) Ts[e] = e < mT * Np ? Tpad[e] : 0.0;
#pragma unroll 8
    for (int e = t; e < 32 * Np; e += 256) {
        const int r = e / Np, j = e - r * Np;
        us[r * ldu + j] = p0 + r < pl.Bc ? Ustar[(p0 + r) * Np + j] : 0.0;
    }
    __syncthreads();
    double acc[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) acc[r] = 0.0;
    const double* trow = Ts + qg * 16 * Np;
    for (int j = 0; j < Np; j += 2) {
        const double2 u = *reinterpret_cast<const double2*>(us + pp * ldu + j);
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            const double2 tv = *reinterpret_cast<const double2*>(trow + r * Np + j);
            acc[r] = fma(tv.x, u.x, acc[r]);
            acc[r] = fma(tv.y, u.y, acc[r]);
        }
    }
    if (p0 + pp < pl.Bc) {
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            const int q = qg * 16 + r;
            if (q < mT) TU[(int64_t)q * pl.Bcpad + p0 + pp] = acc[r];
        }
    }
}

//   From[p][i] = sum_k F_rhs[k][p] rhs_til[k][i]     (reduced forcing, rom.py:911-929)
// thread (pp = t % 32, ig = t / 32): 8 components i for one parameter
__global__ void __launch_bounds__(256) k_forcing(SweepPlan pl, const double* __restrict__ Frhs, const double* __restrict__ rhs_til,
                                                 int mp, double* __restrict__ From) {
    // both operands are staged with every load of the CTA in flight at once (an L2 round trip per k otherwise)
    extern __shared__ __align__(16) double sm_f[];
    double* fs = sm_f;             // F_rhs[k][32 parameters of this CTA]
    double* rs = sm_f + 128 * 32;  // rhs_til (mp x Npad)
    const int t = threadIdx.x, pp = t & 31, i0 = (t >> 5) * 8;
    const int Np = pl.Npad;
    const int64_t p0 = (int64_t)blockIdx.x * 32;
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        const int k = (t >> 5) + 8 * e;
        fs[k * 32 + pp] = (k < mp && p0 + pp < pl.Bc) ? Frhs[(int64_t)k * pl.Bcpad + p0 + pp] : 0.0;
    }
    {
        const double2* src = reinterpret_cast<const double2*>(rhs_til);
        double2* dst = reinterpret_cast<double2*>(rs);
        const int n2 = mp * Np / 2;
#pragma unroll 16
        for (int e = t; e < n2; e += 256) dst[e] = __ldg(src + e);
    }
    __syncthreads();
    if (i0 >= Np) return;
    double acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.0;
#pragma unroll 4
    for (int k = 0; k < mp; ++k) {
        const double f = fs[k * 32 + pp];
        const double* r = rs + k * Np + i0;
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
            if (i0 + e < Np) {  // Npad is even: the pair is inside the row
                const double2 rv = *reinterpret_cast<const double2*>(r + e);
                acc[e] = fma(f, rv.x, acc[e]);
                acc[e + 1] = fma(f, rv.y, acc[e + 1]);
            }
        }
    }
    if (p0 + pp < pl.Bc) {
#pragma unroll
        for (int e = 0; e < 8; ++e)
            if (i0 + e < Np) From[(p0 + pp) * Np + i0 + e] = acc[e];
    }
}

// ------------------------------------------------------------------ fused pre-step (N <= 64, m_tril <= 128, m_rhs <= 128)
// k_state + k_tu + k_entries + k_forcing of one time step in ONE persistent kernel on the FP64 tensor pipe.  A CTA keeps T
// (trilinear table) and rhs_til (folded forcing basis) in shared memory and walks over blocks of 32 parameters:
//   (a) phi(mu_p, t), u*_p, w_p (w goes to global memory for the assembly GEMM's M w hook);
//   (b) F[q][p] = scale_op ( C[q][:] phi_p  +  [trilinear rows] T[q][:] (b0_p u*_p) )  as DMMA tiles of 8 rows x 8 parameters
//       (3 k-steps for the 12 features, Npad / 4 more for the trilinear rows); the rows of the forcing operator stay in
//       shared memory, the others are the A operand of the assembly GEMM;
//   (c) From[p][:] = rhs_til^T F_rhs[:, p]  (reduced forcing).
// The four separate kernels took 116 us per 14,208-parameter step (each a launch with its own staging, 1.5 waves of CTAs,
// scalar FMAs fed by broadcast shared-memory loads); this one is bound by 3,200 DMMAs per block.
constexpr int kPreLd = 68;    // rows of T / rhs_til / u*: 68 == 4 (mod 16), conflict-free DMMA fragments in both orientations
constexpr int kPreLdp = 36;   // rows of the [k][32 parameters] operands (phi, F_rhs): 36 == 4 (mod 16)
constexpr int kPreBlock = 32;
constexpr size_t kPreSmemBytes = ((size_t)2 * 128 * kPreLd + (size_t)kPreBlock * kPreLd + (size_t)RT_NF * kPreLdp +
                                  (size_t)128 * kPreLdp) * sizeof(double);

__device__ __forceinline__ void pre_dmma(double (&c)[2], double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

constexpr int kPreThreads = 512;   // 16 warps: four per scheduler keep the DMMA pipe fed through the short chains of a q-tile

__global__ void __launch_bounds__(kPreThreads, 1) k_prestep(SweepPlan pl, const double* __restrict__ mu, double t, int have_prev, int bdf2,
                                                    const double* __restrict__ Un, const double* __restrict__ Unm1,
                                                    const double* __restrict__ Ccat, const double* __restrict__ Tpad,
                                                    const double* __restrict__ rhs_til, double dt, double* __restrict__ Wv,
                                                    double* __restrict__ F, double* __restrict__ From) {
    extern __shared__ __align__(16) double sm_pre[];
    double* Ts = sm_pre;                          // 128 x kPreLd   T[q][j], zero beyond (m_tril, Npad)
    double* Rs = Ts + 128 * kPreLd;               // 128 x kPreLd   rhs_til[k][i]
    double* us = Rs + 128 * kPreLd;               // 32 x kPreLd    b0_p u*_p[j]
    double* phis = us + kPreBlock * kPreLd;       // RT_NF x kPreLdp  phi[f][p]
    double* fs = phis + RT_NF * kPreLdp;          // 128 x kPreLdp  F_rhs[k][p]
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, g = lane >> 2, tig = lane & 3;
    const int Np = pl.Npad, N = pl.N;
    const int mT = Tpad ? pl.m[kTril] : 0, mR = rhs_til ? pl.mpad[kRhs] : 0;
    {
        // both tables with every load of a thread in flight at once (16 + 16 elements per thread)
        constexpr int kPer = 128 * 64 / kPreThreads;
        double tv[kPer], rv[kPer];
#pragma unroll
        for (int i = 0; i < kPer; ++i) {
            const int e = tid + i * kPreThreads, r = e >> 6, j = e & 63;
            tv[i] = (r < mT && j < Np) ? __ldg(Tpad + r * Np + j) : 0.0;
            rv[i] = (r < mR && j < Np) ? __ldg(rhs_til + r * Np + j) : 0.0;
        }
#pragma unroll
        for (int i = 0; i < kPer; ++i) {
            const int e = tid + i * kPreThreads, r = e >> 6, j = e & 63;
            Ts[r * kPreLd + j] = tv[i];
            Rs[r * kPreLd + j] = rv[i];
        }
    }
    const int ksteps_u = (Np + 3) / 4;            // k-steps over the reduced dimension
    const int qtiles = pl.Qall / 8;
    const int nblk = (pl.Bc + kPreBlock - 1) / kPreBlock;
    for (int blk = blockIdx.x; blk < nblk; blk += gridDim.x) {
        const int64_t p0 = (int64_t)blk * kPreBlock;
        // ---- (a) features, then u* (scaled by b0 = phi[11], the trilinear coefficient) and w; the state loads are issued
        // first, their latency hides under the feature evaluation
        double un_reg[kPreBlock * 64 / kPreThreads], um_reg[kPreBlock * 64 / kPreThreads];
#pragma unroll
        for (int i = 0; i < kPreBlock * 64 / kPreThreads; ++i) {
            const int e = tid + i * kPreThreads, r = e >> 6, j = e & 63;
            const bool in = p0 + r < pl.Bc && j < N;
            un_reg[i] = in ? Un[(p0 + r) * pl.ldu + j] : 0.0;
            um_reg[i] = in ? Unm1[(p0 + r) * pl.ldu + j] : 0.0;
        }
        if (tid < kPreBlock) {
            double phi[RT_NF];
#pragma unroll
            for (int f = 0; f < RT_NF; ++f) phi[f] = 0.0;
            if (p0 + tid < pl.Bc) {
                const double* m = mu + (p0 + tid) * pl.n_mu;
                if (pl.evaluator == 2) mfp1_features(m[0], m[1], m[2], m[3], t, phi);
                else piston_features(pl.evaluator, m[0], m[1], m[2], t, phi);
            }
#pragma unroll
            for (int f = 0; f < RT_NF; ++f) phis[f * kPreLdp + tid] = phi[f];
        }
        __syncthreads();
        {
            constexpr int kPer = kPreBlock * 64 / kPreThreads;
#pragma unroll
            for (int i = 0; i < kPer; ++i) {
                const int e = tid + i * kPreThreads, r = e >> 6, j = e & 63;
                const double un = un_reg[i], um = um_reg[i];
                const double ustar = have_prev ? 2.0 * un - um : un;
                us[r * kPreLd + j] = phis[11 * kPreLdp + r] * ustar;
                if (p0 + r < pl.Bc && j < Np) Wv[(p0 + r) * Np + j] = bdf2 ? 2.0 * un - 0.5 * um : un;
            }
        }
        __syncthreads();
        // ---- (b) sampled entries: q-tile qt = 8 rows, all 32 parameters (4 tiles of 8)
        double cfrag[RT_NF / 4];   // C fragment of the warp's next q-tile, loaded one tile ahead (L2 latency off the chain)
        if (w < qtiles) {
#pragma unroll
            for (int ks = 0; ks < RT_NF / 4; ++ks) cfrag[ks] = __ldg(Ccat + (int64_t)(8 * w + g) * RT_NF + tig + 4 * ks);
        }
        for (int qt = w; qt < qtiles; qt += kPreThreads / 32) {
            const int q0 = 8 * qt;
            // operator of the tile, from constant-index reads of the plan (a dynamic index would put the plan on the stack)
            const int op = (q0 >= pl.qoff[1]) + (q0 >= pl.qoff[2]) + (q0 >= pl.qoff[3]) + (q0 >= pl.qoff[4]) + (q0 >= pl.qoff[5]);
            double acc[4][2];
#pragma unroll
            for (int jn = 0; jn < 4; ++jn) acc[jn][0] = acc[jn][1] = 0.0;
            double ca[RT_NF / 4];
#pragma unroll
            for (int ks = 0; ks < RT_NF / 4; ++ks) ca[ks] = cfrag[ks];
            if (qt + kPreThreads / 32 < qtiles) {
#pragma unroll
                for (int ks = 0; ks < RT_NF / 4; ++ks)
                    cfrag[ks] = __ldg(Ccat + (int64_t)(q0 + kPreThreads / 4 + g) * RT_NF + tig + 4 * ks);
            }
#pragma unroll
            for (int ks = 0; ks < RT_NF / 4; ++ks) {
#pragma unroll
                for (int jn = 0; jn < 4; ++jn) pre_dmma(acc[jn], ca[ks], phis[(4 * ks + tig) * kPreLdp + 8 * jn + g]);
            }
            if (op == kTril && mT > 0) {
                const double* trow = Ts + (q0 - pl.qoff[kTril] + g) * kPreLd + tig;
#pragma unroll 4
                for (int ks = 0; ks < ksteps_u; ++ks) {
                    const double a = trow[4 * ks];
#pragma unroll
                    for (int jn = 0; jn < 4; ++jn) pre_dmma(acc[jn], a, us[(8 * jn + g) * kPreLd + 4 * ks + tig]);
                }
            }
            const double scale = op == 0 ? 1.0 : dt;
            if (op == kRhs) {
                double* frow = fs + (q0 - pl.qoff[kRhs] + g) * kPreLdp + 2 * tig;
#pragma unroll
                for (int jn = 0; jn < 4; ++jn)
                    *reinterpret_cast<double2*>(frow + 8 * jn) = make_double2(acc[jn][0] * scale, acc[jn][1] * scale);
            } else {
                double* frow = F + (int64_t)(q0 + g) * pl.Bcpad + p0 + 2 * tig;
#pragma unroll
                for (int jn = 0; jn < 4; ++jn)
                    if (p0 + 8 * jn + 2 * tig < pl.Bc)
                        *reinterpret_cast<double2*>(frow + 8 * jn) = make_double2(acc[jn][0] * scale, acc[jn][1] * scale);
            }
        }
        __syncthreads();
        // ---- (c) reduced forcing: warp w takes rows i = 8 (w & 7) .. + 7 of From^T for 16 of the 32 parameters
        if (mR > 0 && 8 * (w & 7) < Np) {
            const int it = w & 7, jn0 = 2 * (w >> 3);
            double acc[2][2];
#pragma unroll
            for (int jn = 0; jn < 2; ++jn) acc[jn][0] = acc[jn][1] = 0.0;
            const double* rcol = Rs + tig * kPreLd + 8 * it + g;
            const double* fcol = fs + tig * kPreLdp + 8 * jn0 + g;
#pragma unroll 8
            for (int ks = 0; ks < mR / 4; ++ks) {
                const double a = rcol[4 * ks * kPreLd];
#pragma unroll
                for (int jn = 0; jn < 2; ++jn) pre_dmma(acc[jn], a, fcol[4 * ks * kPreLdp + 8 * jn]);
            }
            const int i = 8 * it + g;
            if (i < Np) {
#pragma unroll
                for (int jn = 0; jn < 2; ++jn) {
                    const int64_t p = p0 + 8 * (jn0 + jn) + 2 * tig;
                    if (p < pl.Bc) From[p * Np + i] = acc[jn][0];
                    if (p + 1 < pl.Bc) From[(p + 1) * Np + i] = acc[jn][1];
                }
            }
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------ dense LU solve per parameter
// One CTA per parameter: K (N x N) and the right-hand side b = M w + f as an augmented column in
// shared memory, LU with partial pivoting (dgesv-like, replaces GMRES rom.py:36/:492), back
// substitution by warp 0 with the solution in registers.  Also performs the BDF history shuffle
// (rom.py:499-503) and writes the reduced solution (rom.py:537-549 "rom" storage).
constexpr int kLuThreads = 256;

// Optional dense additive terms of rt_online_step_f64 (operators without a hyper-reductor, projected by
// the caller): K += Kadd[p], M += Madd[p] (both N x N row-major per parameter), b += fadd[p] (N).
struct LuExtras {
    const double* Kadd;
    const double* Madd;
    const double* fadd;
};

template <int R>
__global__ void __launch_bounds__(kLuThreads) k_lu_solve(SweepPlan pl, const double* __restrict__ Kmat,
                                                        const double* __restrict__ Mmat, const double* __restrict__ From,
                                                        const double* __restrict__ Wv, LuExtras ex, double* __restrict__ Un,
                                                        double* __restrict__ Unm1, double* __restrict__ uN_out,
                                                        int64_t out_stride_p, int64_t out_off, int* __restrict__ info,
                                                        int step) {
    extern __shared__ __align__(16) double sm[];
    const int N = pl.N, ld = (N + 1) | 1;  // odd leading dimension, column N holds b
    double* a = sm;               // N * ld
    double* rdiag = a + (size_t)N * ld;  // N
    double* wv = rdiag + N;       // N
    __shared__ int prow;
    __shared__ int singular;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t p = blockIdx.x;
    const double* Kp = Kmat + p * pl.ldk;
    const double* Mp = Mmat + p * pl.ldk;
    const double* Ka = ex.Kadd ? ex.Kadd + p * N * N : nullptr;
    const double* Ma = ex.Madd ? ex.Madd + p * N * N : nullptr;
    for (int e = tid; e < N * N; e += kLuThreads) a[(e / N) * ld + (e % N)] = Kp[e] + (Ka ? Ka[e] : 0.0);
    if (tid < N) wv[tid] = Wv[p * pl.Npad + tid];
    if (tid == 0) singular = 0;
    __syncthreads();
    for (int i = warp; i < N; i += kLuThreads / 32) {
        double dot = 0.0;
        for (int j = lane; j < N; j += 32) dot = fma(Mp[i * N + j] + (Ma ? Ma[i * N + j] : 0.0), wv[j], dot);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, off);
        if (lane == 0) a[i * ld + N] = dot + From[p * pl.Npad + i] + (ex.fadd ? ex.fadd[p * N + i] : 0.0);
    }
    __syncthreads();
    const int ty = tid >> 5, tx = lane;
    for (int k = 0; k < N; ++k) {
        if (warp == 0) {
            double best = -1.0;
            int bi = k;
            for (int i = k + lane; i < N; i += 32) {
                const double v = fabs(a[i * ld + k]);
                if (v > best) { best = v; bi = i; }
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                const double ov = __shfl_xor_sync(0xffffffffu, best, off);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
                if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
            }
            if (lane == 0) { prow = bi; if (best == 0.0) singular = 1; }
        }
        __syncthreads();
        const int pr = prow;
        if (pr != k && tid <= N - k) {
            const int j = k + tid;  // columns k..N (N = rhs)
            const double t = a[k * ld + j]; a[k * ld + j] = a[pr * ld + j]; a[pr * ld + j] = t;
        }
        __syncthreads();
        const double rinv = 1.0 / a[k * ld + k];
        if (tid == 0) rdiag[k] = rinv;
        for (int i = k + 1 + ty; i < N; i += kLuThreads / 32) {
            const double l = a[i * ld + k] * rinv;
            for (int j = k + 1 + tx; j <= N; j += 32) a[i * ld + j] = fma(-l, a[k * ld + j], a[i * ld + j]);
        }
        __syncthreads();
    }
    if (warp == 0) {
        double bv[R];
#pragma unroll
        for (int r = 0; r < R; ++r) { const int e = r * 32 + lane; bv[r] = e < N ? a[e * ld + N] : 0.0; }
#pragma unroll
        for (int kr = R - 1; kr >= 0; --kr) {
            for (int kl = 31; kl >= 0; --kl) {
                const int k = kr * 32 + kl;
                if (k >= N) continue;
                const double xk = __shfl_sync(0xffffffffu, bv[kr], kl) * rdiag[k];
                if (lane == kl) bv[kr] = xk;
#pragma unroll
                for (int er = 0; er <= kr; ++er) {
                    const int e = er * 32 + lane;
                    if (e < k) bv[er] = fma(-a[e * ld + k], xk, bv[er]);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int e = r * 32 + lane;
            if (e < N) {
                const int64_t s = p * pl.ldu + e;
                Unm1[s] = Un[s];
                Un[s] = bv[r];
                if (uN_out) uN_out[p * out_stride_p + out_off + e] = bv[r];
            }
        }
        if (lane == 0 && singular && info[p] == 0) info[p] = step + 1;
    }
}

// ------------------------------------------------------------------ staging helper of the register-resident LU
// coalesced staging of an N x N row-major matrix into padded shared memory: warp w takes rows w, w+NW, ...;
// 16 rows (32 loads per lane) are issued before the first store so that one round trip covers them
template <int NW, int LD>
__device__ __forceinline__ void stage_matrix(double* __restrict__ stage, const double* __restrict__ src,
                                             const double* __restrict__ add, int N, int warp, int lane) {
    for (int i0 = warp; i0 < N; i0 += 16 * NW) {
        double v0[16], v1[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const int i = i0 + u * NW;
            v0[u] = (i < N && lane < N) ? __ldg(src + i * N + lane) : 0.0;
            v1[u] = (i < N && lane + 32 < N) ? __ldg(src + i * N + lane + 32) : 0.0;
        }

#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const int i = i0 + u * NW;
            if (i < N && lane < N) stage[i * LD + lane] = v0[u];
            if (i < N && lane + 32 < N) stage[i * LD + lane + 32] = v1[u];
        }
    }
    if (add) {  // rt_online_step_f64 only: dense additive term, each warp revisits the rows it staged
        __syncwarp();
        for (int i = warp; i < N; i += NW)
            for (int j = lane; j < N; j += 32) stage[i * LD + j] += __ldg(add + i * N + j);
    }
}

// ------------------------------------------------------------------ register-resident LU, shifting rows (used for N <= 32)
// Thread t keeps row t of [K | b] in registers (rows / columns beyond N are identity padding); RUNTIME loop over the
// elimination steps: after step k every row is shifted left by one while it is updated
// (r[j-1] = r[j] - l * u[j]), so the pivot column is always r[0] and every register index stays static.  The
// loop body is ~100 instructions instead of 13,000 unrolled ones (no instruction-cache misses), and per step
//   redux.sync arg-max -> cand (8 bytes per warp) -> barrier -> the GLOBAL winner alone stores its row, once,
//   straight into the U storage used by the back substitution -> barrier -> broadcast LDS.128 + DFMA.
// Row k of U is stored shifted: U[k][c - k] is the entry of column c >= k, U[k][NP - k] the right-hand side.
// First half of the steps works on NP + 1 entries per row, second half on NP / 2 + 1.
template <int NP, int W>
__device__ __forceinline__ void lu_shift_step(double (&r)[NP + 1], bool& done, bool& sing, int k, double* U, double* rdiag,
                                              unsigned long long (*cand)[NP / 32], int t, int lane, int warp) {
    constexpr int NW = NP / 32;
    constexpr int LDU = NP + 2;
    const int par = k & 1;
    const double rinv = 1.0 / r[0];  // speculative: off the critical path of the arg-max
    const unsigned long long key =
        done ? 0ull : ((unsigned long long)__double_as_longlong(fabs(r[0])) & ~127ull) | 64ull | (unsigned long long)(63 - t);
    const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
    const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
    const unsigned ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
    if (lane == 0) cand[par][warp] = ((unsigned long long)mh << 32) | ml;
    __syncthreads();
    unsigned long long gk = cand[par][0];
#pragma unroll
    for (int w = 1; w < NW; ++w) gk = cand[par][w] > gk ? cand[par][w] : gk;
    const int gt = 63 - (int)(gk & 63ull);
    double* urow = U + k * LDU;
    if (t == gt && !done) {
#pragma unroll
        for (int j2 = 0; j2 < W; j2 += 2) *reinterpret_cast<double2*>(urow + j2) = make_double2(r[j2], r[j2 + 1]);
        urow[W] = r[W];
        rdiag[k] = rinv;
        done = true;
        if (r[0] == 0.0) sing = true;
    }
    __syncthreads();
    if (!done) {
        const double l = r[0] * rdiag[k];
#pragma unroll
        for (int j2 = 0; j2 < W; j2 += 2) {
            const double2 q = *reinterpret_cast<const double2*>(urow + j2);
            if (j2 > 0) r[j2 - 1] = fma(-l, q.x, r[j2]);
            r[j2] = fma(-l, q.y, r[j2 + 1]);
        }
        r[W - 1] = fma(-l, urow[W], r[W]);
        r[W] = 0.0;
    }
}

template <int NP>
__global__ void __launch_bounds__(NP, (NP == 64 ? 5 : 8)) k_lu_shift(SweepPlan pl, const double* __restrict__ Kmat,
                                                                   const double* __restrict__ Mmat,
                                                                   const double* __restrict__ From, const double* __restrict__ Wv,
                                                                   LuExtras ex, double* __restrict__ Un, double* __restrict__ Unm1,
                                                                   double* __restrict__ uN_out, int64_t out_stride_p,
                                                                   int64_t out_off, int* __restrict__ info, int step) {
    constexpr int LD = NP + 1;   // staging stride (odd: conflict-free row reads)
    constexpr int LDU = NP + 2;  // U storage stride (even: 16-byte aligned rows)
    constexpr int NW = NP / 32;
    extern __shared__ __align__(16) double sm[];
    double* U = sm;                  // NP * LDU  (also the staging area for M and K)
    double* wv = U + NP * LDU;       // NP
    double* rdiag = wv + NP;         // NP
    __shared__ unsigned long long cand[2][NW];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int N = pl.N;
    const int64_t p = blockIdx.x;
    const double* Kp = Kmat + p * pl.ldk;
    const double* Mp = Mmat + p * pl.ldk;

    stage_matrix<NW, LD>(U, Mp, ex.Madd ? ex.Madd + p * N * N : nullptr, N, warp, lane);
    wv[t] = t < N ? Wv[p * pl.Npad + t] : 0.0;
    __syncthreads();
    double bt = 0.0;
    if (t < N) {
        double b0 = 0.0, b1 = 0.0;
        int j = 0;
        for (; j + 1 < N; j += 2) { b0 = fma(U[t * LD + j], wv[j], b0); b1 = fma(U[t * LD + j + 1], wv[j + 1], b1); }
        if (j < N) b0 = fma(U[t * LD + j], wv[j], b0);
        bt = (b0 + b1) + From[p * pl.Npad + t];
        if (ex.fadd) bt += ex.fadd[p * N + t];
    }
    __syncthreads();
    stage_matrix<NW, LD>(U, Kp, ex.Kadd ? ex.Kadd + p * N * N : nullptr, N, warp, lane);
    __syncthreads();
    double r[NP + 1];
#pragma unroll
    for (int j = 0; j < NP; ++j) r[j] = (t < N && j < N) ? U[t * LD + j] : (j == t ? 1.0 : 0.0);
    r[NP] = bt;
    __syncthreads();

    bool done = t >= N, sing = false;  // identity padding rows take no part
    const int half = N < NP / 2 ? N : NP / 2;
    for (int k = 0; k < half; ++k) lu_shift_step<NP, NP>(r, done, sing, k, U, rdiag, cand, t, lane, warp);
    for (int k = half; k < N; ++k) lu_shift_step<NP, NP / 2>(r, done, sing, k, U, rdiag, cand, t, lane, warp);
    __syncthreads();
    if (warp == 0) {
        // back substitution: x_k = (rhs_k - sum_{c > k} U[k][c - k] x_c) * rdiag[k];  rhs_k = U[k][NP - k]
        double xv[NW];
#pragma unroll
        for (int q = 0; q < NW; ++q) { const int e = q * 32 + lane; xv[q] = e < N ? U[e * LDU + (NP - e)] : 0.0; }
#pragma unroll
        for (int kr = NW - 1; kr >= 0; --kr) {
#pragma unroll 4
            for (int kl = 31; kl >= 0; --kl) {
                const int k = kr * 32 + kl;
                if (k >= N) continue;
                const double xk = __shfl_sync(0xffffffffu, xv[kr], kl) * rdiag[k];
                if (lane == kl) xv[kr] = xk;
#pragma unroll
                for (int er = 0; er <= kr; ++er) {
                    const int e = er * 32 + lane;
                    if (e < k) xv[er] = fma(-U[e * LDU + (k - e)], xk, xv[er]);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < NW; ++q) {
            const int e = q * 32 + lane;
            if (e < N) {
                const int64_t s = p * pl.ldu + e;
                Unm1[s] = Un[s];
                Un[s] = xv[q];
                if (uN_out) uN_out[p * out_stride_p + out_off + e] = xv[q];
            }
        }
    }
    if (sing) atomicCAS(&info[p], 0, step + 1);
}

__device__ __forceinline__ void dmma884(double2& c, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c.x), "+d"(c.y)
                 : "d"(a), "d"(b));
}

// element pair (i, j), (i, j + 1) of the identity-padded N x N row-major matrix A (+ optional dense term)
__device__ __forceinline__ double2 gj_load2(const double* __restrict__ A, const double* __restrict__ add, int N, bool vec,
                                            int i, int j) {
    double2 v;
    if (vec && i < N && j < N) {
        v = __ldg(reinterpret_cast<const double2*>(A + i * N + j));
    } else {
        v.x = (i < N && j < N) ? __ldg(A + i * N + j) : (i == j ? 1.0 : 0.0);
        v.y = (i < N && j + 1 < N) ? __ldg(A + i * N + j + 1) : (i == j + 1 ? 1.0 : 0.0);
    }
    if (add) {
        if (i < N && j < N) v.x += __ldg(add + i * N + j);
        if (i < N && j + 1 < N) v.y += __ldg(add + i * N + j + 1);
    }
    return v;
}

// ------------------------------------------------------------------ blocked Gauss-Jordan on DMMA tiles (32 < N <= 64)
// u = K^-1 b by Gauss-Jordan elimination with implicit partial pivoting.  The padded 64 x 64 matrix lives in shared
// memory as 8 x 8 tiles in DMMA accumulator order; it is processed in 8 PANELS of 8 columns (one tile column each),
// every panel as two sub-panels of 4 columns.  The kernel is bound by the LATENCY of the one warp that factors a panel
// (six parameters per SM is all the shared memory holds), so everything that is not a pivot step is kept off that
// warp's dependent chain:
//   pivot step: the 64 x 4 slab is in registers (a lane owns two rows).  Arg-max over the rows that were not pivots yet
//              by ONE redux.sync on a 32-bit key; the pivot lane drops its record (4 doubles) into shared memory and all
//              lanes read it back (56 cycles; eight or-reductions cost 124, eight shuffles 73); 1 / pivot by MUFU.RCP64H +
//              a third-order correction beside the reduction; the pivot column is eliminated in EVERY other row.  The
//              4 steps of a sub-panel compose to  X <- X - W Xraw[pivot rows]  with a 64 x 4 matrix W accumulated next
//              to the slab (3 DFMA per row and step).
//   panel:     sub-panel 1 gets the rank-4 update of sub-panel 0 in registers (the panel warp holds W1 and the slab);
//              both (W1, g1), (W2, g2) are published as they are.
//   trailing:  X <- X - W1 Xraw[g1] - W2 X'[g2],  X'[g2] = Xraw[g2] - W1[g2, :] Xraw[g1]  (the rows g2 after the first
//              update, 4 DFMA per lane): two mma.sync.m8n8k4.f64 per 8 x 8 tile and panel, ONE LDS.128 + STS.128 per tile
//              and panel -- half the shared-memory traffic of rank-4 updates.
//   rhs:       b = M w + f rides along as a vector, updated by a trailing warp one panel behind.
// After 8 panels the matrix is a row-permuted diagonal: x_c = rhs[row_c] / pivot_c -- no triangular solves, no stored
// factors.  1 / pivot_c is parked in the (dead) pivot entry itself, the column a row was the pivot of in a byte array:
// the solver state lives in shared memory, so the panel role can ROTATE over the four warps from panel to panel (a
// warp's scheduler is warp id mod 4: a fixed panel warp puts the chains of all six resident parameters on one of the
// four schedulers of the SM).  In iteration T one warp applies the pending update of panel T to tile column T + 1 and
// factors it (look-ahead) while the other three update the remaining live tile columns; one block barrier per panel
// plus a named barrier that protects the single W buffer.
// Shared-memory layouts:
//   tiles: element (i, j) in tile (i / 8, j / 8), slot ((i % 8) * 4 + (j % 8) / 2) ^ ((i / 8) & 3), component j & 1 -- a
//          lane's accumulator slot XOR-ed by the tile row, so that the panel warp's row-wise LDS.128 of a quarter warp
//          (rows of four different tile rows) hit eight distinct 16-byte bank groups; tile accesses stay conflict-free;
//   nw[h]: -W of sub-panel h, element (row, k) at k * 68 + row: conflict-free for the panel warp's row-wise STS.64 and
//          for the A-fragment LDS.64 of the trailing warps (lane (g8, q) reads q * 68 + 8 rt + g8).
constexpr int kGjLdw = 68;
struct GjsShared {
    double2 tiles[64 * 32];
    double nw[2][4 * kGjLdw];
    double rhs[64];
    double2 rec[4][2];        // pivot record of step J of the running sub-panel
    int gr[8];                // pivot rows of the published panel (g1, g2)
    unsigned char colof[64];  // column that row i was the pivot of (0xff: none yet)
};

// offset (in doubles, tile column 0) of element (i, 8 ct + n) -- add 64 ct
__device__ __forceinline__ int gjs_rowoff(int i, int n) {
    const int rt = i >> 3;
    return ((rt * 256 + ((((i & 7) << 2) | (n >> 1)) ^ (rt & 3))) << 1) | (n & 1);
}

// 1 / x to ~1 ulp: MUFU.RCP64H seed (20 bits) and one third-order step x0 (1 + e + e^2), e = 1 - x x0 -- three dependent
// DFMA (40 cycles) instead of the rcp.rn.f64 sequence (76).  x = 0 (or subnormal) gives NaN / inf.
__device__ __forceinline__ double gjs_rcp(double x) {
    double x0;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x0) : "d"(x));
    const double e = fma(-x, x0, 1.0);
    const double p = fma(e, e, e);
    return fma(x0, p, x0);
}

// one pivot step of a sub-panel; the lane owns rows row0 and row0 + 32 (a0 / a1: slab, w0 / w1: multipliers).
// rb0 / rb1 = 64 | (63 - row): the 7 low key bits (candidate flag, row); cand0 / cand1 = all ones while the row has not been a pivot.
template <int J>
__device__ __forceinline__ void gjs_step(GjsShared& sh, double (&a0)[4], double (&a1)[4], double (&w0)[4], double (&w1)[4],
                                         unsigned& cand0, unsigned& cand1, int (&gr)[4], double& my_rinv, unsigned rb0, unsigned rb1,
                                         int lane) {
    // key: high word of |a_J| over the row bits: the pivot is within 2^-13 of the column maximum, the lowest row wins ties
    const unsigned k0 = (((unsigned)__double2hiint(a0[J]) & 0x7fffff80u) | rb0) & cand0;
    const unsigned k1 = (((unsigned)__double2hiint(a1[J]) & 0x7fffff80u) | rb1) & cand1;
    const bool sel = k1 > k0;
    const unsigned mx = __reduce_max_sync(0xffffffffu, sel ? k1 : k0);
    // reciprocal of the lane's own candidate: independent of the reduction, its chain runs beside it
    const double rinv_loc = gjs_rcp(sel ? a1[J] : a0[J]);
    double rec[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) rec[j] = j < J ? (sel ? w1[j] : w0[j]) : (j == J ? rinv_loc : (sel ? a1[j] : a0[j]));
    const bool me0 = ((mx ^ rb0) & 127u) == 0u, me1 = ((mx ^ rb1) & 127u) == 0u;
    if (me0 | me1) {
        sh.rec[J][0] = make_double2(rec[0], rec[1]);
        sh.rec[J][1] = make_double2(rec[2], rec[3]);
    }
    __syncwarp();
    const double2 ra = sh.rec[J][0], rb = sh.rec[J][1];
    const double r[4] = {ra.x, ra.y, rb.x, rb.y};
    gr[J] = 63 - (int)(mx & 63u);
    my_rinv = lane == J ? r[J] : my_rinv;  // lane J keeps 1 / pivot of step J for the bookkeeping after the sub-panel
    const double l0 = me0 ? 0.0 : a0[J] * r[J];
    const double l1 = me1 ? 0.0 : a1[J] * r[J];
#pragma unroll
    for (int j = J + 1; j < 4; ++j) { a0[j] = fma(-l0, r[j], a0[j]); a1[j] = fma(-l1, r[j], a1[j]); }
#pragma unroll
    for (int j = 0; j < J; ++j) { w0[j] = fma(-l0, r[j], w0[j]); w1[j] = fma(-l1, r[j], w1[j]); }
    w0[J] = l0;
    w1[J] = l1;
    cand0 = me0 ? 0u : cand0;
    cand1 = me1 ? 0u : cand1;
}

// 4 pivot steps on the slab (a0, a1) = sub-panel h of tile column T: W in (w0, w1), pivot rows gr; publishes -W, the pivot
// rows, 1 / pivot (parked in the pivot entry) and the rows' columns
__device__ __forceinline__ void gjs_subpanel(GjsShared& sh, int T, int h, double (&a0)[4], double (&a1)[4], double (&w0)[4],
                                             double (&w1)[4], int (&gr)[4], unsigned& cand0, unsigned& cand1, int row0, int lane,
                                             bool guard, int* __restrict__ info_p, int step) {
#pragma unroll
    for (int j = 0; j < 4; ++j) { w0[j] = 0.0; w1[j] = 0.0; }
    const unsigned rb0 = 64u | (unsigned)(63 - row0), rb1 = 64u | (unsigned)(31 - row0);
    double rv = 0.0;
    gjs_step<0>(sh, a0, a1, w0, w1, cand0, cand1, gr, rv, rb0, rb1, lane);
    gjs_step<1>(sh, a0, a1, w0, w1, cand0, cand1, gr, rv, rb0, rb1, lane);
    gjs_step<2>(sh, a0, a1, w0, w1, cand0, cand1, gr, rv, rb0, rb1, lane);
    gjs_step<3>(sh, a0, a1, w0, w1, cand0, cand1, gr, rv, rb0, rb1, lane);
    // the other warps have read what they need of the previous panel: sh.nw / sh.gr may be overwritten
    if (guard) asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll
    for (int k = 0; k < 4; ++k) { sh.nw[h][k * kGjLdw + row0] = -w0[k]; sh.nw[h][k * kGjLdw + row0 + 32] = -w1[k]; }
    if (lane < 4) {
        // lane J records pivot J: 1 / pivot is parked in the pivot entry (g, col) -- the column is finished, nobody reads it
        // again as a matrix entry --, the column in colof[g]; a vanished pivot (1 / pivot not finite) flags the parameter
        const int g = lane == 0 ? gr[0] : (lane == 1 ? gr[1] : (lane == 2 ? gr[2] : gr[3]));
        sh.gr[4 * h + lane] = g;
        reinterpret_cast<double*>(sh.tiles)[64 * T + gjs_rowoff(g, 4 * h + lane)] = rv;
        sh.colof[g] = (unsigned char)(8 * T + 4 * h + lane);
        if (!(fabs(rv) < 1.7976931348623157e308)) atomicCAS(info_p, 0, step + 1);
    }
}

// Factor panel T (tile column T, every earlier update applied) and publish it: -W1, -W2 in sh.nw, pivot rows in sh.gr.
// One warp; `guard`: wait (named barrier) until the other warps are done with the previous panel before sh.nw / sh.gr are overwritten.
__device__ __forceinline__ void gjs_factor(GjsShared& sh, int T, int lane, bool guard, int* __restrict__ info_p, int step) {
    // rows of this lane (row0, row0 + 32): a quarter warp covers two rows of four different tile rows (conflict-free slab
    // LDS.128, see the layout note above) and a half warp sixteen different rows mod 16 (conflict-free row-wise STS.64)
    const int tra = (lane >> 1) & 3;
    const int row0 = 8 * tra + (lane & 1) + 2 * ((tra >> 1) ^ (lane >> 4)) + 4 * ((lane >> 3) & 1);
    const int rt = row0 >> 3, ri = row0 & 7, sw = rt & 3;
    const double2* lo = sh.tiles + (rt * 8 + T) * 32;
    const double2* hi = lo + 4 * 8 * 32;  // row + 32: four tile rows further
    double a0[4], a1[4];
    {
        const int s0 = (ri * 4) ^ sw, s1 = (ri * 4 + 1) ^ sw;
        const double2 x0 = lo[s0], x1 = lo[s1], y0 = hi[s0], y1 = hi[s1];
        a0[0] = x0.x; a0[1] = x0.y; a0[2] = x1.x; a0[3] = x1.y;
        a1[0] = y0.x; a1[1] = y0.y; a1[2] = y1.x; a1[3] = y1.y;
    }
    unsigned cand0 = sh.colof[row0] == 0xff ? ~0u : 0u, cand1 = sh.colof[row0 + 32] == 0xff ? ~0u : 0u;
    double w0[4], w1[4];
    int g1[4], g2[4];
    gjs_subpanel(sh, T, 0, a0, a1, w0, w1, g1, cand0, cand1, row0, lane, guard, info_p, step);
    {
        // sub-panel 1 <- itself - W1 Xraw[g1, its 4 columns]: the raw rows come from the tiles (nobody writes column T now)
        const int s2 = (ri * 4 + 2) ^ sw, s3 = (ri * 4 + 3) ^ sw;
        const double2 x2 = lo[s2], x3 = lo[s3], y2 = hi[s2], y3 = hi[s3];
        a0[0] = x2.x; a0[1] = x2.y; a0[2] = x3.x; a0[3] = x3.y;
        a1[0] = y2.x; a1[1] = y2.y; a1[2] = y3.x; a1[3] = y3.y;
        const double2* tc = sh.tiles + 32 * T;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int rtk = g1[k] >> 3, rik = (g1[k] & 7) << 2, swk = rtk & 3;
            const double2 xa = tc[rtk * 256 + ((rik | 2) ^ swk)], xb = tc[rtk * 256 + ((rik | 3) ^ swk)];  // columns 4, 5 and 6, 7
            a0[0] = fma(-w0[k], xa.x, a0[0]); a0[1] = fma(-w0[k], xa.y, a0[1]); a0[2] = fma(-w0[k], xb.x, a0[2]); a0[3] = fma(-w0[k], xb.y, a0[3]);
            a1[0] = fma(-w1[k], xa.x, a1[0]); a1[1] = fma(-w1[k], xa.y, a1[1]); a1[2] = fma(-w1[k], xb.x, a1[2]); a1[3] = fma(-w1[k], xb.y, a1[3]);
        }
    }
    gjs_subpanel(sh, T, 1, a0, a1, w0, w1, g2, cand0, cand1, row0, lane, false, info_p, step);
}

// B fragments of one tile column for the published panel: Xraw[g1] and X'[g2] = Xraw[g2] - W1[g2, :] Xraw[g1] (4 DFMA).
// Read before the warp stores any tile of that column.
struct GjsB {
    double b1, b2;
};
struct GjsRows {  // per lane: offsets of rows g1[0..3], g1[q], g2[q] at its column g8, and the coupling -W1[g2[q], :]
    double c21[4];
    int o1[4], o1q, o2;
};
__device__ __forceinline__ void gjs_load_rows(const GjsShared& sh, GjsRows& r, int lane) {
    const int g8 = lane >> 2, q = lane & 3;
    const int g2q = sh.gr[4 + q];
#pragma unroll
    for (int k = 0; k < 4; ++k) { r.o1[k] = gjs_rowoff(sh.gr[k], g8); r.c21[k] = sh.nw[0][k * kGjLdw + g2q]; }
    r.o1q = gjs_rowoff(sh.gr[q], g8);
    r.o2 = gjs_rowoff(g2q, g8);
}
__device__ __forceinline__ GjsB gjs_load_b(const GjsShared& sh, const GjsRows& r, int ct) {
    const double* tl = reinterpret_cast<const double*>(sh.tiles) + 64 * ct;
    double x1[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) x1[k] = tl[r.o1[k]];
    GjsB b;
    b.b1 = tl[r.o1q];
    b.b2 = tl[r.o2];
#pragma unroll
    for (int k = 0; k < 4; ++k) b.b2 = fma(r.c21[k], x1[k], b.b2);
    return b;
}
// one warp: tile columns ca (and cb) <- themselves - W1 Xraw[g1] - W2 X'[g2] (rank 8: two DMMA per tile).  Four tile rows
// at a time: their A fragments (-W1, -W2: 16 registers) are loaded once and used for both columns.
__device__ __forceinline__ void gjs_update_cols(GjsShared& sh, const GjsB& ba, const GjsB& bb, int ca, int cb, bool has_b, int lane) {
    const int g8 = lane >> 2, q = lane & 3;
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
        double a1[4], a2[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            a1[i] = sh.nw[0][q * kGjLdw + 8 * (4 * hf + i) + g8];
            a2[i] = sh.nw[1][q * kGjLdw + 8 * (4 * hf + i) + g8];
        }
        double2 c[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { const int rt = 4 * hf + i; c[i] = sh.tiles[(rt * 8 + ca) * 32 + (lane ^ (rt & 3))]; }
#pragma unroll
        for (int i = 0; i < 4; ++i) dmma884(c[i], a1[i], ba.b1);
#pragma unroll
        for (int i = 0; i < 4; ++i) dmma884(c[i], a2[i], ba.b2);
#pragma unroll
        for (int i = 0; i < 4; ++i) { const int rt = 4 * hf + i; sh.tiles[(rt * 8 + ca) * 32 + (lane ^ (rt & 3))] = c[i]; }
        if (has_b) {
#pragma unroll
            for (int i = 0; i < 4; ++i) { const int rt = 4 * hf + i; c[i] = sh.tiles[(rt * 8 + cb) * 32 + (lane ^ (rt & 3))]; }
#pragma unroll
            for (int i = 0; i < 4; ++i) dmma884(c[i], a1[i], bb.b1);
#pragma unroll
            for (int i = 0; i < 4; ++i) dmma884(c[i], a2[i], bb.b2);
#pragma unroll
            for (int i = 0; i < 4; ++i) { const int rt = 4 * hf + i; sh.tiles[(rt * 8 + cb) * 32 + (lane ^ (rt & 3))] = c[i]; }
        }
    }
}

// right-hand side of row i after the published panel: rhs - W1 rhs[g1] - W2 rhs'[g2], rhs'[g2] = rhs[g2] - W1[g2, :] rhs[g1]
__device__ __forceinline__ double gjs_rhs_row(const GjsShared& sh, int i) {
    double r1[4], r2[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { r1[k] = sh.rhs[sh.gr[k]]; r2[k] = sh.rhs[sh.gr[4 + k]]; }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int g = sh.gr[4 + k];
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) r2[k] = fma(sh.nw[0][kk * kGjLdw + g], r1[kk], r2[k]);
    }
    double s = sh.rhs[i];
#pragma unroll
    for (int k = 0; k < 4; ++k) s = fma(sh.nw[0][k * kGjLdw + i], r1[k], s);
#pragma unroll
    for (int k = 0; k < 4; ++k) s = fma(sh.nw[1][k * kGjLdw + i], r2[k], s);
    return s;
}

// PERSISTENT: a grid smaller than the chunk strides over it (the two-lane experiment of DESIGN.md 8, RT_GJ_GRID); the shipped
// launch is one system per CTA with no loop and no trailing barrier
template <bool PERSISTENT>
__global__ void __launch_bounds__(128, 6) k_gj_smem(SweepPlan pl, const double* __restrict__ Kmat, const double* __restrict__ Mmat,
                                                    const double* __restrict__ Mw,  // M w from the GEMM's matvec hook, or null
                                                    const double* __restrict__ From, const double* __restrict__ Wv, LuExtras ex,
                                                    double* __restrict__ Un, double* __restrict__ Unm1,
                                                    double* __restrict__ uN_out, int64_t out_stride_p, int64_t out_off,
                                                    int* __restrict__ info_all, int step) {
    __shared__ __align__(16) GjsShared sh;
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int g8 = lane >> 2, q = lane & 3;
    const int N = pl.N;
    const bool vec = (N & 1) == 0;
#pragma unroll 1
    for (int64_t p = blockIdx.x; p < (PERSISTENT ? (int64_t)pl.Bc : (int64_t)blockIdx.x + 1); p += gridDim.x) {
    int* info = info_all + p;
    const double* Kp = Kmat + p * pl.ldk;
    const double* Mp = Mmat + p * pl.ldk;
    const double* Ka = ex.Kadd ? ex.Kadd + p * N * N : nullptr;
    const double* Ma = ex.Madd ? ex.Madd + p * N * N : nullptr;

    // ---- b = M w + f (thread t: row t / 2, every other column pair, partner summed by shuffle) and K into the tiles
    if (t < 64) { sh.nw[0][t] = t < N ? __ldg(Wv + p * pl.Npad + t) : 0.0; sh.colof[t] = 0xff; }  // w staged in the (still unused) W buffer
    const int row = t >> 1, hp = t & 1;
    double s0 = 0.0, s1 = 0.0;
    // tile idx = w * 16 + i (rt = 2 w + i / 8): slot lane ^ (rt & 3)
    const int la = lane ^ ((2 * w) & 3), lb = lane ^ ((2 * w + 1) & 3);
    if (Mw) {
        // M w already formed inside the assembly GEMM (N == 64, no dense extras): only K is loaded, two batches in flight
        const double2* ktile = reinterpret_cast<const double2*>(Kp + g8 * 64 + 2 * q);  // + (8 rt) * 64 + 8 ct doubles
        double2 b0[8], b1[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { const int idx = w * 16 + i; b0[i] = __ldg(ktile + (idx >> 3) * 256 + (idx & 7) * 4); }
#pragma unroll
        for (int i = 0; i < 8; ++i) { const int idx = w * 16 + 8 + i; b1[i] = __ldg(ktile + (idx >> 3) * 256 + (idx & 7) * 4); }
        if (hp == 0) s0 = __ldg(Mw + p * pl.Npad + row);
#pragma unroll
        for (int i = 0; i < 8; ++i) sh.tiles[(w * 16 + i) * 32 + la] = b0[i];
#pragma unroll
        for (int i = 0; i < 8; ++i) sh.tiles[(w * 16 + 8 + i) * 32 + lb] = b1[i];
    } else if (N == 64 && !Ma && !Ka) {
        // fast path: four batches of eight 16-byte loads, two batches in flight
        const double2* mrow = reinterpret_cast<const double2*>(Mp + row * 64) + hp;
        const double2* ktile = reinterpret_cast<const double2*>(Kp + g8 * 64 + 2 * q);  // + (8 rt) * 64 + 8 ct doubles
        double2 b0[8], b1[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) b0[i] = __ldg(mrow + 2 * i);
#pragma unroll
        for (int i = 0; i < 8; ++i) b1[i] = __ldg(mrow + 2 * (i + 8));
        __syncthreads();  // w visible
        const double2* w2 = reinterpret_cast<const double2*>(sh.nw[0]) + hp;
#pragma unroll
        for (int i = 0; i < 8; ++i) { const double2 wv = w2[2 * i]; s0 = fma(b0[i].x, wv.x, s0); s1 = fma(b0[i].y, wv.y, s1); }
#pragma unroll
        for (int i = 0; i < 8; ++i) { const int idx = w * 16 + i; b0[i] = __ldg(ktile + (idx >> 3) * 256 + (idx & 7) * 4); }
#pragma unroll
        for (int i = 0; i < 8; ++i) { const double2 wv = w2[2 * (i + 8)]; s0 = fma(b1[i].x, wv.x, s0); s1 = fma(b1[i].y, wv.y, s1); }
#pragma unroll
        for (int i = 0; i < 8; ++i) { const int idx = w * 16 + 8 + i; b1[i] = __ldg(ktile + (idx >> 3) * 256 + (idx & 7) * 4); }
#pragma unroll
        for (int i = 0; i < 8; ++i) sh.tiles[(w * 16 + i) * 32 + la] = b0[i];
#pragma unroll
        for (int i = 0; i < 8; ++i) sh.tiles[(w * 16 + 8 + i) * 32 + lb] = b1[i];
    } else {
        __syncthreads();
        if (row < N) {
            for (int j = 2 * hp; j < N; j += 4) {
                const double2 m = gj_load2(Mp, Ma, N, vec, row, j);
                s0 = fma(m.x, sh.nw[0][j], s0);
                if (j + 1 < N) s1 = fma(m.y, sh.nw[0][j + 1], s1);
            }
        }
#pragma unroll 4
        for (int i = 0; i < 16; ++i) {
            const int idx = w * 16 + i, rt = idx >> 3, ct = idx & 7;
            sh.tiles[idx * 32 + (lane ^ (rt & 3))] = gj_load2(Kp, Ka, N, vec, 8 * rt + g8, 8 * ct + 2 * q);
        }
    }
    {
        double s = s0 + s1;
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        if (hp == 0) {
            if (row < N) {
                s += From[p * pl.Npad + row];
                if (ex.fadd) s += ex.fadd[p * N + row];
            }
            sh.rhs[row] = s;
        }
    }
    __syncthreads();

    const int rot = (int)(p & 3);
    if (w == rot) gjs_factor(sh, 0, lane, false, info, step);
    __syncthreads();
#pragma unroll 1
    for (int T = 0; T < 7; ++T) {
        // panel T is published: its update is pending on tile columns T + 1 .. 7 and on the right-hand side.
        // role 0: column T + 1, then factors it; roles 1-3: columns T + 1 + role, T + 4 + role (if any); role 1: the right-hand side
        const int role = (w - (T + 1 + rot)) & 3;
        const int ca = T + 1 + role, cb = ca + 3;
        const bool has_a = ca < 8, has_b = role != 0 && cb < 8;
        GjsB ba, bb;
        if (has_a) {
            GjsRows rows;
            gjs_load_rows(sh, rows, lane);
            ba = gjs_load_b(sh, rows, ca);
            bb = ba;
            if (has_b) bb = gjs_load_b(sh, rows, cb);
        }
        if (role == 0) {
            // the fragments of column T + 1 are read before the named barrier inside gjs_factor; the A fragments too
            gjs_update_cols(sh, ba, ba, ca, ca, false, lane);
            __syncwarp();
            gjs_factor(sh, T + 1, lane, true, info, step);
        } else {
            double s0 = 0.0, s1 = 0.0;
            if (role == 1) { s0 = gjs_rhs_row(sh, lane); s1 = gjs_rhs_row(sh, lane + 32); }
            if (has_a) gjs_update_cols(sh, ba, bb, ca, cb, has_b, lane);
            // every read of the published panel (sh.nw, sh.gr) is done: the panel warp may overwrite it
            asm volatile("bar.arrive 1, 128;" ::: "memory");
            if (role == 1) { __syncwarp(); sh.rhs[lane] = s0; sh.rhs[lane + 32] = s1; }
        }
        __syncthreads();
    }
    // ---- x_c = rhs[row] / pivot (rhs after the last panel), c = the column row was the pivot of; history shuffle, output
    if (t < 64) {
        const int c = sh.colof[t];
        if (c < N) {
            const double x = gjs_rhs_row(sh, t) * reinterpret_cast<const double*>(sh.tiles)[64 * (c >> 3) + gjs_rowoff(t, c & 7)];
            const int64_t s = p * pl.ldu + c;
            Unm1[s] = Un[s];
            Un[s] = x;
            if (uN_out) uN_out[p * out_stride_p + out_off + c] = x;
        }
    }
    if (PERSISTENT) __syncthreads();  // the next system of a persistent CTA re-uses sh
    else break;
    }
}

}  // namespace

// ------------------------------------------------------------------ stand-alone entry points

extern "C" int rt_invert_f64(rt_handle* h, const double* A, int64_t m, double* Ainv, int* info, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, A && Ainv && info && m >= 1 && m <= 160);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    RT_CUDA(h, cudaMemsetAsync(info, 0, sizeof(int), st));
    const int ld = (int)m | 1;
    const size_t smem = (size_t)m * ld * sizeof(double) + (size_t)m * sizeof(int) + 16;
    RT_CUDA(h, cudaFuncSetAttribute(k_invert, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_invert<<<1, 1024, smem, st>>>(A, (int)m, Ainv, info);
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}

extern "C" int rt_theta_f64(rt_handle* h, const double* PT_Uinv, int64_t m, const double* F, int64_t B, double* theta,
                            void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, PT_Uinv && F && theta && m >= 1 && B >= 1);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    // theta[b][q'] = sum_q F[b][q] Inv[q'][q]
    dg::Operand oa{F, B, m, m, 1}, ob{PT_Uinv, m, m, m, 1}, a, b;
    int rc = dg::stage_if_needed(h, 3, oa, &a, st);
    if (rc) return rc;
    rc = dg::stage_if_needed(h, 4, ob, &b, st);
    if (rc) return rc;
    dg::Params p = dg::default_params();
    p.C = theta; p.ldc_m = m; p.ldc_n = 1;
    return dg::launch(h, a, b, p, st);
}

extern "C" int rt_assemble_batch_f64(rt_handle* h, const double* basis, int64_t R, int64_t m, int64_t b_stride_row,
                                     int64_t b_stride_col, const double* theta, int64_t B, double* out, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, basis && theta && out && R >= 1 && m >= 1 && B >= 1);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    // out[b][r] = sum_q theta[b][q] basis[r][q]
    dg::Operand oa{theta, B, m, m, 1}, ob{basis, R, m, b_stride_row, b_stride_col}, a, b;
    int rc = dg::stage_if_needed(h, 3, oa, &a, st);
    if (rc) return rc;
    rc = dg::stage_if_needed(h, 4, ob, &b, st);
    if (rc) return rc;
    dg::Params p = dg::default_params();
    p.C = out; p.ldc_m = R; p.ldc_n = 1;
    return dg::launch(h, a, b, p, st);
}

// ------------------------------------------------------------------ the batched BDF sweep
namespace {

inline int64_t even64(int64_t x) { return (x + 1) & ~int64_t(1); }
inline int64_t pad16(int64_t x) { return (x + 15) & ~int64_t(15); }

// Parameters per chunk.  The assembly GEMM has (chunk/128) x tiles_n output tiles on `sms` persistent CTAs;
// the chunk (a multiple of 128, at most 40*128 so that the K/M buffers stay ~330 MB) is chosen to
// minimise the total number of tile waves over the whole batch, i.e. the wave-quantisation loss of the
// dominant kernel: 37*128 parameters x 32 tiles = 8 full waves of 148 for N = 64 on a B200.
int64_t plan_chunk(int64_t B, int N, int sms) {
    const int tiles_n = (int)ceil_div64((int64_t)N * N, dg::BN);
    const int64_t rows_total = ceil_div64(B, 128);
    // the solve kernel has its own quantisation: k_gj_smem keeps 6 parameters per SM resident (one "solve wave" =
    // 6 * sms parameters, 0.4 of a GEMM tile wave at N = 64); 4736 parameters are 5.33 such waves, 14208 exactly 16
    const bool gj = N > 32 && N <= 64;
    const int64_t solve_wave = 6 * (int64_t)sms;
    const double solve_weight = gj ? 0.4 : 0.0;
    int64_t best_k = 1;
    double best_cost = 1e300;
    for (int64_t k = 1; k <= 120; ++k) {
        const int64_t nfull = rows_total / k, rem = rows_total - nfull * k;
        const int64_t rem_params = B - nfull * k * 128 > 0 ? B - nfull * k * 128 : 0;
        const double waves = (double)nfull * (double)ceil_div64(k * tiles_n, sms) + (double)ceil_div64(rem * tiles_n, sms);
        const double swaves = (double)nfull * (double)ceil_div64(k * 128, solve_wave) + (double)ceil_div64(rem_params, solve_wave);
        const double chunks = (double)(nfull + (rem ? 1 : 0));
        const double cost = waves + solve_weight * swaves + 0.25 * chunks;  // + per-chunk-step cost of the small kernels
        if (cost < best_cost - 1e-9) { best_cost = cost; best_k = k; }
    }
    return best_k * 128;
}

struct Carver {
    char* base;
    size_t off = 0;
    template <typename T>
    T* take(size_t count) {
        off = (off + 255) & ~size_t(255);
        T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
        off += count * sizeof(T);
        return p;
    }
};

struct SweepBuffers {
    double *basis_cat, *basis_rhs, *basis_til, *rhs_til, *Ccat, *Invcat, *Tpad;
    int2* kt_ranges;
    double *Phi, *Ustar, *Wv, *Un, *Unm1, *TU, *F, *Kmat, *Mmat, *From, *Mw;
};

size_t carve(Carver& c, const SweepPlan& pl, SweepBuffers& b) {
    const size_t Bcp = pl.Bcpad;
    b.basis_cat = c.take<double>((size_t)pl.Qmat * pl.ldk);
    b.basis_rhs = c.take<double>((size_t)pl.mpad[kRhs] * pl.Npad);
    b.basis_til = c.take<double>((size_t)pl.Qmat * pl.ldk);
    b.rhs_til = c.take<double>((size_t)pl.mpad[kRhs] * pl.Npad);
    b.Ccat = c.take<double>((size_t)pl.Qall * RT_NF);
    b.Invcat = c.take<double>((size_t)pl.Qall * pl.Qall);
    b.Tpad = c.take<double>((size_t)pl.mpad[kTril] * pl.Npad);
    b.kt_ranges = c.take<int2>((size_t)(pl.Qall + 127) / 128);
    const size_t setup_end = c.off;
    b.Phi = c.take<double>(RT_NF * Bcp);
    b.Ustar = c.take<double>(Bcp * pl.Npad);
    b.Wv = c.take<double>(Bcp * pl.Npad);
    b.Un = c.take<double>(Bcp * pl.Npad);
    b.Unm1 = c.take<double>(Bcp * pl.Npad);
    b.TU = c.take<double>((size_t)pl.mpad[kTril] * Bcp);
    b.F = c.take<double>((size_t)pl.Qall * Bcp);
    b.Kmat = c.take<double>(Bcp * pl.ldk);
    b.Mmat = c.take<double>(Bcp * pl.ldk);
    b.From = c.take<double>(Bcp * pl.Npad);
    b.Mw = c.take<double>(Bcp * pl.Npad);
    return setup_end;
}

struct StepOut {
    double* uN_out;        // nullable
    int64_t stride_p, off;
    int* info;
    int step;
};

template <int R>
int launch_lu(rt_handle* h, const SweepPlan& pl, const SweepBuffers& b, const LuExtras& ex, const StepOut& o,
              cudaStream_t st) {
    const int ld = (pl.N + 1) | 1;
    const size_t smem = ((size_t)pl.N * ld + 2 * pl.N) * sizeof(double);
    // per handle (= per device) and grown on demand: N = 70 then N = 96 share R = 3 but need 40 KB and 76 KB
    int rc = rt_ensure_smem(h, R, k_lu_solve<R>, (int)smem);
    if (rc) return rc;
    k_lu_solve<R><<<pl.Bc, kLuThreads, smem, st>>>(pl, b.Kmat, b.Mmat, b.From, b.Wv, ex, b.Un, b.Unm1, o.uN_out, o.stride_p, o.off,
                                                   o.info, o.step);
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}

template <int NP>
int launch_lu_reg(rt_handle* h, const SweepPlan& pl, const SweepBuffers& b, const LuExtras& ex, const StepOut& o,
                  cudaStream_t st) {
    const size_t smem = ((size_t)NP * (NP + 2) + 2 * NP) * sizeof(double) + 16;
    int rc = rt_ensure_smem(h, 8, k_lu_shift<NP>, 64 * 1024);
    if (rc) return rc;
    k_lu_shift<NP><<<pl.Bc, NP, smem, st>>>(pl, b.Kmat, b.Mmat, b.From, b.Wv, ex, b.Un, b.Unm1, o.uN_out, o.stride_p, o.off,
                                            o.info, o.step);
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}

// N == 64 without dense extras: the assembly GEMM contracts M(mu, t) with the history vector at its k-boundary instead of
// storing M (466 MB per 14,208-parameter chunk never written nor read back)
inline bool mw_in_gemm(const SweepPlan& pl, const LuExtras& ex) {
    return pl.N == 64 && pl.ldk == 4096 && !ex.Kadd && !ex.Madd && pl.mpad[0] > 0 && pl.mpad[0] < pl.Qmat;
}

int launch_gj(rt_handle* h, const SweepPlan& pl, const SweepBuffers& b, const LuExtras& ex, const StepOut& o, cudaStream_t st) {
#ifdef RT_GEMM_LAUNCH_REGS
    RT_CUDA(h, cudaFuncSetAttribute(k_gj_smem<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
#endif
    int gj_grid = pl.Bc;
    if (getenv("RT_GJ_GRID")) { gj_grid = atoi(getenv("RT_GJ_GRID")); if (gj_grid > pl.Bc) gj_grid = pl.Bc; }
    const double* mw = mw_in_gemm(pl, ex) ? b.Mw : nullptr;
    if (gj_grid < pl.Bc)
        k_gj_smem<true><<<gj_grid, 128, 0, st>>>(pl, b.Kmat, b.Mmat, mw, b.From, b.Wv, ex, b.Un, b.Unm1, o.uN_out, o.stride_p, o.off,
                                                 o.info, o.step);
    else
        k_gj_smem<false><<<pl.Bc, 128, 0, st>>>(pl, b.Kmat, b.Mmat, mw, b.From, b.Wv, ex, b.Un, b.Unm1, o.uN_out, o.stride_p, o.off,
                                                o.info, o.step);
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}

// Everything that is fixed for one model and one maximum chunk size: the plan, the carved scratch and the
// concatenated zero-padded operator tables.
struct SweepCtx {
    SweepPlan pl;
    SweepBuffers b;
    bool has_T;
    int64_t chunk;
};

// The staged model of the last sweep / step, kept in the handle: a caller that marks its rt_rom_model with a non-zero
// `generation` (unique per set of device tables) pays the table copies and the fold GEMM once, not once per call --
// RomConstructor.solve drives rt_online_step_f64 once per time step with the same model.
struct SweepCache {
    bool valid = false;
    rt_rom_model model;
    bool need_evaluator = false;
    int64_t B = 0;
    void* ws = nullptr;
    size_t ws_bytes = 0;
    cudaEvent_t ready = nullptr;
    SweepCtx cx;
};

int sweep_setup(rt_handle* h, const rt_rom_model* model, int64_t B, bool need_evaluator, cudaStream_t st, SweepCtx& cx) {
    RT_REQUIRE(h, model->n_ops == kNOps);
    if (model->N < 1 || model->N > 128)
        return rt_fail(h, RT_EINVAL, "reduced dimension N = %d is outside the supported range 1..128", model->N);
    if (!h->sweep_cache) h->sweep_cache = new SweepCache();
    SweepCache* sc = (SweepCache*)h->sweep_cache;
    if (sc->valid && model->generation != 0 && memcmp(&sc->model, model, sizeof(rt_rom_model)) == 0 &&
        sc->need_evaluator == need_evaluator && sc->B == B && sc->ws == h->ws[8] && sc->ws_bytes == h->ws_bytes[8]) {
        cx = sc->cx;
        RT_CUDA(h, cudaStreamWaitEvent(st, sc->ready, 0));
        return RT_OK;
    }
    sc->valid = false;
    RT_REQUIRE(h, model->evaluator >= 0 && model->evaluator <= 2);
    RT_REQUIRE(h, model->n_mu == 0 || model->n_mu == (model->evaluator == 2 ? 4 : 3));
    SweepPlan& pl = cx.pl;
    SweepBuffers& b = cx.b;
    memset(&pl, 0, sizeof(pl));
    pl.N = model->N;
    pl.Npad = (int)even64(pl.N);
    pl.ldu = pl.Npad;
    pl.N2 = pl.N * pl.N;
    pl.ldk = (int)even64(pl.N2);
    pl.evaluator = model->evaluator;
    pl.n_mu = model->n_mu ? model->n_mu : 3;
    int off = 0, msum = 0;
    for (int i = 0; i < kNOps; ++i) {
        const rt_rom_operator& op = model->ops[i];
        RT_REQUIRE(h, op.m >= 0);  // m == 0: operator absent (e.g. no convection for a fixed-domain ROM)
        if (op.m > 0) {
            RT_REQUIRE(h, op.basis_rom && op.PT_Uinv);
            RT_REQUIRE(h, !need_evaluator || op.C);
            RT_REQUIRE(h, (i == kRhs) == (op.is_vector != 0));
            RT_REQUIRE(h, op.ld >= (i == kRhs ? pl.N : pl.N2));
        }
        pl.m[i] = op.m;
        pl.mpad[i] = (int)pad16(op.m);
        pl.qoff[i] = off;
        pl.msum[i] = msum;
        off += pl.mpad[i];
        msum += op.m;
    }
    pl.qoff[kNOps] = off;
    pl.msum[kNOps] = msum;
    pl.qsum = msum;
    pl.Qall = off;
    pl.Qmat = pl.qoff[kRhs];
    RT_REQUIRE(h, pl.Qall >= 1);
    cx.chunk = plan_chunk(B, pl.N, h->sm_count);
    const int64_t Bc_max = B < cx.chunk ? B : cx.chunk;
    pl.Bcpad = (int)(ceil_div64(Bc_max, 128) * 128);
    pl.Bc = (int)Bc_max;

    Carver sizing{nullptr};
    carve(sizing, pl, b);
    void* ws;
    int rc = rt_scratch(h, 8, sizing.off + 1024, &ws);
    if (rc) return rc;
    Carver real{(char*)ws};
    const size_t setup_bytes = carve(real, pl, b);

    // ---- one-off model staging: concatenated, zero-padded operator tables ----
    RT_CUDA(h, cudaMemsetAsync(ws, 0, setup_bytes, st));
    for (int i = 0; i < kNOps; ++i) {
        const rt_rom_operator& op = model->ops[i];
        if (op.m == 0) continue;
        const int q0 = pl.qoff[i];
        if (i != kRhs)
            RT_CUDA(h, cudaMemcpy2DAsync(b.basis_cat + (size_t)q0 * pl.ldk, (size_t)pl.ldk * 8, op.basis_rom, (size_t)op.ld * 8,
                                         (size_t)pl.N2 * 8, op.m, cudaMemcpyDeviceToDevice, st));
        else
            RT_CUDA(h, cudaMemcpy2DAsync(b.basis_rhs, (size_t)pl.Npad * 8, op.basis_rom, (size_t)op.ld * 8, (size_t)pl.N * 8,
                                         op.m, cudaMemcpyDeviceToDevice, st));
        if (op.C)
            RT_CUDA(h, cudaMemcpyAsync(b.Ccat + (size_t)q0 * RT_NF, op.C, (size_t)op.m * RT_NF * 8, cudaMemcpyDeviceToDevice, st));
        RT_CUDA(h, cudaMemcpy2DAsync(b.Invcat + (size_t)q0 * pl.Qall + q0, (size_t)pl.Qall * 8, op.PT_Uinv, (size_t)op.m * 8,
                                     (size_t)op.m * 8, op.m, cudaMemcpyDeviceToDevice, st));
    }
    cx.has_T = need_evaluator && model->ops[kTril].m > 0 && model->ops[kTril].T != nullptr;
    if (cx.has_T)
        RT_CUDA(h, cudaMemcpy2DAsync(b.Tpad, (size_t)pl.Npad * 8, model->ops[kTril].T, (size_t)pl.N * 8, (size_t)pl.N * 8,
                                     model->ops[kTril].m, cudaMemcpyDeviceToDevice, st));
    {
        // k-tile range of every 128-row tile of the block-diagonal inverse
        const int tiles = (pl.Qall + 127) / 128;
        int2 host_ranges[64];
        RT_REQUIRE(h, tiles <= 64);
        for (int t = 0; t < tiles; ++t) {
            const int r0 = t * 128, r1 = (r0 + 128 < pl.Qall) ? r0 + 128 : pl.Qall;
            int lo = pl.Qall, hi = 0;
            for (int i = 0; i < kNOps; ++i) {
                if (pl.qoff[i] < r1 && pl.qoff[i + 1] > r0) {
                    if (pl.qoff[i] < lo) lo = pl.qoff[i];
                    if (pl.qoff[i + 1] > hi) hi = pl.qoff[i + 1];
                }
            }
            host_ranges[t] = make_int2(lo / 16, hi / 16);
        }
        RT_CUDA(h, cudaMemcpyAsync(b.kt_ranges, host_ranges, tiles * sizeof(int2), cudaMemcpyHostToDevice, st));
        RT_CUDA(h, cudaStreamSynchronize(st));  // host_ranges is on the stack
    }
    // ---- fold the interpolation solves into the bases, once per model:
    //   sum_q theta_q A_q with theta = PT_U^-1 f   ==   sum_k f_k Atil_k,   Atil_k = sum_q (PT_U^-1)[q][k] A_q
    // so a time step goes from the sampled entries straight to the assembly GEMM (no per-step theta GEMM).
    if (pl.Qmat > 0) {
        dg::Operand oa{b.Invcat, pl.Qmat, pl.Qmat, 1, pl.Qall}, ob{b.basis_cat, pl.N2, pl.Qmat, 1, pl.ldk};
        dg::Params p = dg::default_params();
        p.C = b.basis_til; p.ldc_m = pl.ldk; p.ldc_n = 1;
        p.kt_ranges = b.kt_ranges;
        int rc2 = dg::launch(h, oa, ob, p, st);
        if (rc2) return rc2;
    }
    if (pl.m[kRhs] > 0) {
        const double* inv_rhs = b.Invcat + (size_t)pl.qoff[kRhs] * pl.Qall + pl.qoff[kRhs];
        dg::Operand oa{inv_rhs, pl.mpad[kRhs], pl.mpad[kRhs], 1, pl.Qall}, ob{b.basis_rhs, pl.N, pl.mpad[kRhs], 1, pl.Npad};
        dg::Params p = dg::default_params();
        p.C = b.rhs_til; p.ldc_m = pl.Npad; p.ldc_n = 1;
        int rc2 = dg::launch(h, oa, ob, p, st);
        if (rc2) return rc2;
    }
    if (model->generation != 0) {
        if (!sc->ready) RT_CUDA(h, cudaEventCreateWithFlags(&sc->ready, cudaEventDisableTiming));
        RT_CUDA(h, cudaEventRecord(sc->ready, st));
        sc->model = *model;
        sc->need_evaluator = need_evaluator;
        sc->B = B;
        sc->ws = h->ws[8];
        sc->ws_bytes = h->ws_bytes[8];
        sc->cx = cx;
        sc->valid = true;
    }
    return RT_OK;
}

// One implicit step for the current chunk (pl.Bc parameters; state in b.Un / b.Unm1 with stride pl.ldu).
//   Fext == nullptr: sampled entries from the device evaluator (mu_c, t);
//   Fext != nullptr: caller-supplied sampled entries (Bc x qsum row-major).
int sweep_step(rt_handle* h, SweepCtx& cx, const double* mu_c, const double* Fext, double t, double dt, int bdf, int64_t n,
               const LuExtras& ex, const StepOut& o, bool profiled, cudaStream_t st) {
    const SweepPlan& pl = cx.pl;
    const SweepBuffers& b = cx.b;
    int rc;
    const int have_prev = (bdf == 2 && n > 0) ? 1 : 0;
    const double bdfc = have_prev ? 1.5 : 1.0;  // first step uses bdf = 1.0 (rom.py:481-483)
    const int64_t nstate = (int64_t)pl.Bc * pl.Npad;
    Profile* prof = (Profile*)h->profile;
    const bool pa = prof && prof->on && profiled;
    // state, trilinear product, sampled entries and reduced forcing in one kernel when the tables fit its shared memory
    const bool fused = !Fext && mu_c && pl.Npad <= 64 && (!cx.has_T || pl.m[kTril] <= 128) && pl.mpad[kRhs] <= 128 &&
                       pl.Qmat > 0 && !(getenv("RT_FUSED_PRESTEP") && atoi(getenv("RT_FUSED_PRESTEP")) == 0);
    if (fused) {
        ProfScope ps(prof, pa, 7, st);
        rc = rt_ensure_smem(h, 11, k_prestep, (int)kPreSmemBytes);
        if (rc) return rc;
        int grid = (pl.Bc + kPreBlock - 1) / kPreBlock;
        if (grid > h->sm_count) grid = h->sm_count;
        k_prestep<<<grid, kPreThreads, kPreSmemBytes, st>>>(pl, mu_c, t, have_prev, bdf == 2, b.Un, b.Unm1, b.Ccat,
                                                   cx.has_T ? b.Tpad : nullptr, pl.m[kRhs] > 0 ? b.rhs_til : nullptr, dt, b.Wv, b.F,
                                                   b.From);
        RT_LAUNCH_CHECK(h);
    } else {
    {
        ProfScope ps(prof, pa, 0, st);
        k_state<<<(unsigned)ceil_div64(nstate, 256), 256, 0, st>>>(pl, mu_c, t, have_prev, bdf == 2, b.Un, b.Unm1, b.Phi, b.Ustar,
                                                                   b.Wv);
    }
    RT_LAUNCH_CHECK(h);
    const dim3 egrid((unsigned)ceil_div64(pl.Bc, 256), (unsigned)ceil_div64(pl.Qall, 32));
    if (Fext) {
        ProfScope ps(prof, pa, 2, st);
        k_entries_ext<<<egrid, 256, 0, st>>>(pl, Fext, dt, b.F);
        RT_LAUNCH_CHECK(h);
    } else {
        if (cx.has_T) {
            ProfScope ps(prof, pa, 1, st);
            if (pl.N <= 64 && pl.m[kTril] <= 128) {
                const size_t smem = ((size_t)128 * pl.Npad + 32 * (pl.Npad + 2)) * sizeof(double);
                rc = rt_ensure_smem(h, 9, k_tu, (128 * 64 + 32 * 66) * 8);
                if (rc) return rc;
                k_tu<<<(unsigned)ceil_div64(pl.Bc, 32), 256, smem, st>>>(pl, b.Tpad, b.Ustar, pl.m[kTril], b.TU);
                RT_LAUNCH_CHECK(h);
            } else {
                dg::Operand oa{b.Tpad, pl.m[kTril], pl.Npad, pl.Npad, 1}, ob{b.Ustar, pl.Bc, pl.Npad, pl.Npad, 1};
                dg::Params p = dg::default_params();
                p.C = b.TU; p.ldc_m = pl.Bcpad; p.ldc_n = 1;
                rc = dg::launch(h, oa, ob, p, st);
                if (rc) return rc;
            }
        }
        ProfScope ps(prof, pa, 2, st);
        const dim3 egrid2((unsigned)ceil_div64(pl.Bc, 256), (unsigned)(pl.Qall / 16));
        k_entries<<<egrid2, 128, 0, st>>>(pl, b.Ccat, b.Phi, cx.has_T ? b.TU : nullptr, dt, b.F);
        RT_LAUNCH_CHECK(h);
    }
    }  // !fused
    if (pl.Qmat > 0) {
        ProfScope ps(prof, pa, 4, st);  // K = bdf M + dt (A + C + N^ + N(u*)),  M emitted at the k-boundary
        dg::Operand oa{b.F, pl.Bc, pl.Qmat, 1, pl.Bcpad}, ob{b.basis_til, pl.N2, pl.Qmat, 1, pl.ldk};
        dg::Params p = dg::default_params();
        p.C = b.Kmat; p.ldc_m = pl.ldk; p.ldc_n = 1;
        p.C2 = b.Mmat; p.hook_kt = pl.mpad[0] / 16; p.hook_scale = bdfc;
        if (mw_in_gemm(pl, ex)) { p.C2 = b.Mw; p.hook_w = b.Wv; p.hook_ldw = pl.Npad; p.hook_ldo = pl.Npad; }
        if (p.hook_kt >= pl.Qmat / 16) {  // only a mass reductor: K = bdf M, no k-tile left to trigger the hook
            p.hook_kt = -1;
            p.alpha = 1.0;
            p.C = b.Mmat;
            rc = dg::launch(h, oa, ob, p, st);
            if (rc) return rc;
            p.C = b.Kmat;
            p.alpha = bdfc;
        }
        rc = dg::launch(h, oa, ob, p, st);
        if (rc) return rc;
    } else {
        RT_CUDA(h, cudaMemsetAsync(b.Kmat, 0, (size_t)pl.Bc * pl.ldk * sizeof(double), st));
        RT_CUDA(h, cudaMemsetAsync(b.Mmat, 0, (size_t)pl.Bc * pl.ldk * sizeof(double), st));
    }
    if (fused) {
        if (pl.m[kRhs] == 0) RT_CUDA(h, cudaMemsetAsync(b.From, 0, (size_t)pl.Bc * pl.Npad * sizeof(double), st));
    } else if (pl.m[kRhs] > 0) {
        ProfScope ps(prof, pa, 5, st);  // reduced forcing dt * basis_rhs theta_rhs
        if (pl.N <= 64 && pl.mpad[kRhs] <= 128) {
            rc = rt_ensure_smem(h, 10, k_forcing, (128 * 32 + 128 * 64) * 8);
            if (rc) return rc;
            const size_t smem = ((size_t)128 * 32 + (size_t)pl.mpad[kRhs] * pl.Npad) * sizeof(double);
            k_forcing<<<(unsigned)ceil_div64(pl.Bc, 32), 256, smem, st>>>(pl, b.F + (size_t)pl.qoff[kRhs] * pl.Bcpad, b.rhs_til,
                                                                      pl.mpad[kRhs], b.From);
            RT_LAUNCH_CHECK(h);
        } else {
            dg::Operand oa{b.F + (size_t)pl.qoff[kRhs] * pl.Bcpad, pl.Bc, pl.mpad[kRhs], 1, pl.Bcpad};
            dg::Operand ob{b.rhs_til, pl.N, pl.mpad[kRhs], 1, pl.Npad};
            dg::Params p = dg::default_params();
            p.C = b.From; p.ldc_m = pl.Npad; p.ldc_n = 1;
            rc = dg::launch(h, oa, ob, p, st);
            if (rc) return rc;
        }
    } else {
        RT_CUDA(h, cudaMemsetAsync(b.From, 0, (size_t)pl.Bc * pl.Npad * sizeof(double), st));
    }
    ProfScope ps_lu(prof, pa, 6, st);
    const int R = (pl.N + 31) / 32;
    if (pl.N <= 32) rc = launch_lu_reg<32>(h, pl, b, ex, o, st);
    else if (pl.N <= 64) rc = launch_gj(h, pl, b, ex, o, st);
    else switch (R) {
        case 1: rc = launch_lu<1>(h, pl, b, ex, o, st); break;
        case 2: rc = launch_lu<2>(h, pl, b, ex, o, st); break;
        case 3: rc = launch_lu<3>(h, pl, b, ex, o, st); break;
        default: rc = launch_lu<4>(h, pl, b, ex, o, st); break;
    }
    return rc;
}

// Steps n0 .. n1 - 1 of the nt-step sweep.  Un_state / Unm1_state == nullptr: the whole sweep from the zero initial condition
// with the state in the handle's scratch (n0 must be 0); otherwise the caller-owned state (B x N row-major, updated in
// place) carries the sweep from one segment to the next.
int sweep_impl(rt_handle* h, const rt_rom_model* model, const double* mu, int64_t B, int64_t nt, int64_t n0, int64_t n1, double dt,
               int bdf, double* Un_state, double* Unm1_state, double* uN_out, double* uN_last, int* info, cudaStream_t st) {
    RT_REQUIRE(h, model && mu && info && B >= 1 && nt >= 1 && n0 >= 0 && n0 < n1 && n1 <= nt);
    RT_REQUIRE(h, bdf == 1 || bdf == 2);
    RT_REQUIRE(h, (Un_state != nullptr) == (Unm1_state != nullptr));
    RT_REQUIRE(h, Un_state || n0 == 0);
    SweepCtx cx;
    int rc = sweep_setup(h, model, B, true, st, cx);
    if (rc) return rc;
    SweepPlan& pl = cx.pl;
    SweepBuffers& b = cx.b;
    if (n0 == 0) RT_CUDA(h, cudaMemsetAsync(info, 0, (size_t)B * sizeof(int), st));
    if (Un_state) pl.ldu = pl.N;
    const LuExtras none{nullptr, nullptr, nullptr};
    double t0 = 0.0;
    for (int64_t n = 0; n < n0; ++n) t0 += dt;  // the same accumulation as the uninterrupted loop (rom.py:476-479)
    for (int64_t p0 = 0; p0 < B; p0 += cx.chunk) {
        pl.Bc = (int)((B - p0) < cx.chunk ? (B - p0) : cx.chunk);
        if (Un_state) {
            b.Un = Un_state + p0 * pl.N;
            b.Unm1 = Unm1_state + p0 * pl.N;
            if (n0 == 0) {
                RT_CUDA(h, cudaMemsetAsync(b.Un, 0, (size_t)pl.Bc * pl.N * sizeof(double), st));
                RT_CUDA(h, cudaMemsetAsync(b.Unm1, 0, (size_t)pl.Bc * pl.N * sizeof(double), st));
            }
        } else {
            const size_t state_bytes = (size_t)pl.Bcpad * pl.Npad * sizeof(double);
            RT_CUDA(h, cudaMemsetAsync(b.Un, 0, state_bytes, st));    // zero initial condition (rom.py:451-452)
            RT_CUDA(h, cudaMemsetAsync(b.Unm1, 0, state_bytes, st));  // uN_n1 = zeros, not None (rom.py:456-460)
        }
        const double* mu_c = mu + p0 * pl.n_mu;
        double t = t0;
        for (int64_t n = n0; n < n1; ++n) {
            t += dt;  // same accumulation as the reference loop (rom.py:476-479)
            StepOut o{nullptr, 0, 0, info + p0, (int)n};
            if (uN_out) { o.uN_out = uN_out + p0 * nt * pl.N; o.stride_p = nt * pl.N; o.off = n * pl.N; }
            else if (uN_last && n == nt - 1) { o.uN_out = uN_last + p0 * pl.N; o.stride_p = pl.N; o.off = 0; }
            rc = sweep_step(h, cx, mu_c, nullptr, t, dt, bdf, n, none, o, p0 == 0, st);
            if (rc) return rc;
        }
    }
    return RT_OK;
}

}  // namespace

extern "C" int rt_online_sweep_f64(rt_handle* h, const rt_rom_model* model, const double* mu, int64_t B, int64_t nt,
                                   double dt, int bdf, double* uN_out, int* info, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, uN_out);
    RT_CUDA(h, cudaSetDevice(h->device));
    return sweep_impl(h, model, mu, B, nt, 0, nt, dt, bdf, nullptr, nullptr, uN_out, nullptr, info, rt_stream(stream));
}

extern "C" int rt_online_sweep_segment_f64(rt_handle* h, const rt_rom_model* model, const double* mu, int64_t B, int64_t nt,
                                           int64_t n0, int64_t n1, double dt, int bdf, double* Un, double* Unm1, double* uN_out,
                                           int* info, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, uN_out && Un && Unm1);
    RT_CUDA(h, cudaSetDevice(h->device));
    return sweep_impl(h, model, mu, B, nt, n0, n1, dt, bdf, Un, Unm1, uN_out, nullptr, info, rt_stream(stream));
}

extern "C" int rt_online_sweep_last_f64(rt_handle* h, const rt_rom_model* model, const double* mu, int64_t B, int64_t nt,
                                        double dt, int bdf, double* uN_last, int* info, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, uN_last);
    RT_CUDA(h, cudaSetDevice(h->device));
    return sweep_impl(h, model, mu, B, nt, 0, nt, dt, bdf, nullptr, nullptr, nullptr, uN_last, info, rt_stream(stream));
}

extern "C" int rt_online_step_f64(rt_handle* h, const rt_rom_model* model, const double* F, int64_t B, double dt, int bdf,
                                  int64_t step, const double* Kadd, const double* Madd, const double* fadd, double* Un,
                                  double* Unm1, int* info, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, model && F && Un && Unm1 && info && B >= 1 && step >= 0);
    RT_REQUIRE(h, bdf == 1 || bdf == 2);
    RT_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = rt_stream(stream);
    SweepCtx cx;
    int rc = sweep_setup(h, model, B, false, st, cx);
    if (rc) return rc;
    SweepPlan& pl = cx.pl;
    pl.ldu = pl.N;  // caller-owned state, B x N row-major
    const int64_t N2 = pl.N2;
    for (int64_t p0 = 0; p0 < B; p0 += cx.chunk) {
        pl.Bc = (int)((B - p0) < cx.chunk ? (B - p0) : cx.chunk);
        cx.b.Un = Un + p0 * pl.N;
        cx.b.Unm1 = Unm1 + p0 * pl.N;
        const LuExtras ex{Kadd ? Kadd + p0 * N2 : nullptr, Madd ? Madd + p0 * N2 : nullptr, fadd ? fadd + p0 * pl.N : nullptr};
        StepOut o{nullptr, 0, 0, info + p0, (int)step};
        rc = sweep_step(h, cx, nullptr, F + p0 * pl.qsum, 0.0, dt, bdf, step, ex, o, false, st);
        if (rc) return rc;
    }
    return RT_OK;
}

// called by rt_destroy (capi.cu): the handle-owned state of this translation unit
void rt_online_release(rt_handle* h) {
    if (h->sweep_cache) {
        SweepCache* sc = (SweepCache*)h->sweep_cache;
        if (sc->ready) cudaEventDestroy(sc->ready);
        delete sc;
        h->sweep_cache = nullptr;
    }
    if (h->profile) {
        Profile* p = (Profile*)h->profile;
        for (cudaEvent_t e : p->pool) cudaEventDestroy(e);
        delete p;
        h->profile = nullptr;
    }
}

extern "C" int64_t rt_sweep_chunk(int64_t B, int N, int sm_count) {
    if (B < 1 || N < 1 || sm_count < 1) return 0;
    return plan_chunk(B, N, sm_count);
}

extern "C" int rt_profile_enable(rt_handle* h, int on) {
    if (!h) return RT_EINVAL;
    if (!h->profile) h->profile = new Profile();
    ((Profile*)h->profile)->on = on != 0;
    return RT_OK;
}

extern "C" int rt_profile_read(rt_handle* h, double* ms, int64_t* launches) {
    if (!h || !ms || !launches) return RT_EINVAL;
    RT_CUDA(h, cudaSetDevice(h->device));
    RT_CUDA(h, cudaDeviceSynchronize());
    Profile* p = (Profile*)h->profile;
    for (int c = 0; c < RT_PROFILE_CLASSES; ++c) { ms[c] = 0; launches[c] = 0; }
    if (!p) return RT_OK;
    for (size_t i = 0; i < p->used; ++i) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, p->pool[2 * i], p->pool[2 * i + 1]) == cudaSuccess) {
            p->ms[p->cls[i]] += t;
            p->cnt[p->cls[i]] += 1;
        }
    }
    p->used = 0;
    for (int c = 0; c < RT_PROFILE_CLASSES; ++c) { ms[c] = p->ms[c]; launches[c] = p->cnt[c]; p->ms[c] = 0; p->cnt[c] = 0; }
    return RT_OK;
}
