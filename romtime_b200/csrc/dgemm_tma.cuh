// FP64 GEMM core for sm_100a: TMA (cp.async.bulk.tensor, 128B swizzle) -> shared-memory ring ->
// DMMA (mma.sync.m8n8k4.f64, the only native FP64 tensor shape on sm_100a: SASS DMMA.8x8x4).
//
//   C[M x N] = alpha * A * B,   A given as X_A[m][k], B as X_B[n][k]  (both "mn x k" operands)
//
// Each operand is described by a 2-D tensor map and a "major":
//   MAJOR_K : k contiguous in global memory  (box = 16 k x 128 mn -> smem [mn][16], one TMA op)
//   MAJOR_MN: mn contiguous                 (box = 16 mn x 16 k  -> smem [k][16], 8 TMA ops per tile)
// Both land in 128-byte shared-memory rows with the 128B XOR swizzle; the DMMA fragment reads use
// the k permutation  k(s,t) = 8*(t>>1) + 2*((t+s)&3) + (t&1)  (t = lane&3, s = k-step 0..3), which
// makes every 8-byte fragment load bank-conflict free for both majors (DESIGN.md "GEMM core").
//
// CTA = 8 DMMA warps (warp tile 32 x 64, 64 fp64 accumulators per thread, setmaxnreg 240) + one
// producer warpgroup (setmaxnreg 24) whose elected lane issues TMA;
// 128 x 128 x 16 tiles, 6-stage full/empty mbarrier ring; persistent over (tile, k-split) units.
// Out-of-range rows/columns/k are zero-filled by TMA, so any M, N, K works without tail code.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace dg {

enum { MAJOR_K = 0, MAJOR_MN = 1 };

constexpr int BM = 128, BN = 128, BK = 16;
#ifndef RT_GEMM_STAGES
#define RT_GEMM_STAGES 6
#endif
#ifndef RT_GEMM_CONSUMER_REGS
#define RT_GEMM_CONSUMER_REGS 240
#endif
#define RT_STR2(x) #x
#define RT_STR(x) RT_STR2(x)
#ifdef RT_GEMM_LAUNCH_REGS
#define RT_GEMM_BOUNDS __maxnreg__(RT_GEMM_LAUNCH_REGS)
#else
#define RT_GEMM_BOUNDS __launch_bounds__(THREADS, 1)
#endif
constexpr int STAGES = RT_GEMM_STAGES;
constexpr int CONSUMER_WARPS = 8;
constexpr int PRODUCER_WARPS = 4;                // one warpgroup; lane 0 of its first warp issues TMA
constexpr int THREADS = (CONSUMER_WARPS + PRODUCER_WARPS) * 32;
constexpr int TILE_BYTES = BM * BK * 8;           // 16 KB per operand per stage
constexpr int STAGE_BYTES = 2 * TILE_BYTES;       // 32 KB
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr int MI = 4, NI = 8;                     // warp tile 32 (m) x 64 (n)
constexpr int WARPS_M = 4, WARPS_N = 2;

struct Params {
    int64_t M, N, K;
    double* C;
    int64_t ldc_m, ldc_n;      // element (m, n) at C[m*ldc_m + n*ldc_n]
    double alpha;
    int tiles_m, tiles_n;
    int symmetric;             // 1: only tiles with tn >= tm (A == B, square)
    int splits;                // k-splits; split s covers k-tiles [s*kt_per_split, ...)
    int kt_per_split;
    int64_t split_stride;      // element offset between the partial outputs of consecutive splits
    // optional second output: when k-tile index == hook_kt (within split 0, splits must be 1)
    // the accumulators are stored to C2 (same layout as C) and then multiplied by hook_scale
    int hook_kt;               // -1: disabled
    double* C2;
    double hook_scale;
    // optional block-sparse K: k-tile range [x, y) per M-tile (device pointer, tiles_m entries);
    // used for block-diagonal left operands (theta solve of all operators in one launch)
    const int2* kt_ranges;
    // matvec form of the second output (when hook_w != nullptr): instead of storing the accumulators, every group of 64
    // columns [64 r, 64 r + 64) is contracted with the row's vector  C2[m * hook_ldo + r] = sum_j acc(m, 64 r + j) *
    // hook_w[m * hook_ldw + j]  (the online step's M(mu, t) w: M is never materialised)
    const double* hook_w;
    int64_t hook_ldw, hook_ldo;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    const uint32_t addr = smem_u32(bar);
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ double lds_f64(uint32_t addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
        : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// byte offset of element (mn, k) inside one 16 KB operand tile
template <int MAJOR>
__device__ __forceinline__ uint32_t frag_off(int mn, int k) {
    if (MAJOR == MAJOR_K)
        return (uint32_t)(mn * 128 + ((((k >> 1) ^ (mn & 7))) << 4) + ((k & 1) << 3));
    else
        return (uint32_t)((mn >> 4) * 2048 + k * 128 + (((((mn & 15) >> 1) ^ (k & 7))) << 4) + ((mn & 1) << 3));
}

__device__ __forceinline__ void tile_coords(const Params& p, int tile, int& tm, int& tn) {
    if (p.symmetric) {
        const int T = p.tiles_n;
        int row = 0, off = 0;
        while (off + (T - row) <= tile) { off += T - row; ++row; }
        tm = row;
        tn = row + (tile - off);
    } else {
        tm = tile / p.tiles_n;
        tn = tile % p.tiles_n;
    }
}

template <int MI_, int NI_>
__device__ __forceinline__ void store_acc(const double (&acc)[MI_][NI_][2], double* C, int64_t ldc_m,
                                          int64_t ldc_n, int64_t M, int64_t N, int64_t m_base,
                                          int64_t n_base, int g, int tig, double alpha) {
    const bool vec = (ldc_n == 1) && ((ldc_m & 1) == 0) && ((reinterpret_cast<uintptr_t>(C) & 15) == 0);
#pragma unroll
    for (int i = 0; i < MI_; ++i) {
        const int64_t m = m_base + 8 * i + g;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < NI_; ++j) {
            const int64_t n = n_base + 8 * j + 2 * tig;
            if (n >= N) continue;
            double* dst = C + m * ldc_m + n * ldc_n;
            const double v0 = alpha * acc[i][j][0], v1 = alpha * acc[i][j][1];
            if (vec && n + 1 < N) {
                *reinterpret_cast<double2*>(dst) = make_double2(v0, v1);
            } else {
                dst[0] = v0;
                if (n + 1 < N) dst[ldc_n] = v1;
            }
        }
    }
}

// NT > 0 (N <= 8 NT <= 96, one column of tiles): the right part of the 128 x 128 tile would be all padding, so the eight
// consumer warps split the rows instead (16 rows x 8 NT columns each) and the DMMAs on the padding disappear (the POD lift
// S W_c with c = retained + guard modes, the Gram matrix of the lifted modes).  A warp whose 16 rows lie beyond M (the
// Gram matrix of c <= 96 columns) only releases the stages.
template <int AMAJOR, int BMAJOR, int NT>
__global__ void RT_GEMM_BOUNDS
k_dgemm(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const Params p) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
    uint64_t* empty = full + STAGES;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], CONSUMER_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int ntiles = p.symmetric ? p.tiles_n * (p.tiles_n + 1) / 2 : p.tiles_m * p.tiles_n;
    const int nunits = ntiles * p.splits;
    const int kt_total = (int)((p.K + BK - 1) / BK);

    if (warp < PRODUCER_WARPS) {
        // ===================== TMA producer warpgroup (one elected lane) =====================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
        if (warp == 0 && lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
                const int split = u / ntiles, tile = u % ntiles;
                int tm, tn;
                tile_coords(p, tile, tm, tn);
                int kt0 = split * p.kt_per_split;
                int kt1 = min(kt_total, kt0 + p.kt_per_split);
                if (p.kt_ranges) { const int2 r = p.kt_ranges[tm]; kt0 = r.x; kt1 = min(kt_total, r.y); }
                for (int kt = kt0; kt < kt1; ++kt) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    mbar_expect_tx(&full[stage], STAGE_BYTES);
                    unsigned char* sa = smem + stage * STAGE_BYTES;
                    unsigned char* sb = sa + TILE_BYTES;
                    if (AMAJOR == MAJOR_K) {
                        tma_load_2d(sa, &mapA, &full[stage], kt * BK, tm * BM);
                    } else {
#pragma unroll
                        for (int b = 0; b < 8; ++b) tma_load_2d(sa + b * 2048, &mapA, &full[stage], tm * BM + 16 * b, kt * BK);
                    }
                    if (BMAJOR == MAJOR_K) {
                        tma_load_2d(sb, &mapB, &full[stage], kt * BK, tn * BN);
                    } else {
#pragma unroll
                        for (int b = 0; b < 8; ++b) tma_load_2d(sb + b * 2048, &mapB, &full[stage], tn * BN + 16 * b, kt * BK);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        return;
    }

    // ===================== DMMA consumers (two warpgroups) =====================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 " RT_STR(RT_GEMM_CONSUMER_REGS) ";");
    const int cw = warp - PRODUCER_WARPS;
    const int g = lane >> 2, tig = lane & 3;
    constexpr bool NARROW = NT > 0;
    constexpr int MI_ = NARROW ? 2 : MI, NI_ = NARROW ? NT : NI;
    const int wm0 = NARROW ? cw * 16 : (cw / WARPS_N) * (MI * 8), wn0 = NARROW ? 0 : (cw % WARPS_N) * (NI * 8);
    // Fragment addresses as 32-bit shared-space offsets (LDS with immediate offsets instead of generic 64-bit
    // loads).  For k-step s the thread reads k index kq = 8*(tig>>1) + 2*((tig+s)&3) + (tig&1); in both majors rows
    // mn and mn + 16 are exactly 2048 bytes apart and the swizzle term only depends on (mn & 15, k), so two base
    // offsets per k-step (rows w0 + g and w0 + 8 + g) plus immediates cover the whole warp tile.
    const uint32_t smem_base = smem_u32(smem);
    uint32_t a_off[2][4], b_off[2][4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const int kq = 8 * (tig >> 1) + 2 * ((tig + s) & 3) + (tig & 1);
        a_off[0][s] = frag_off<AMAJOR>(wm0 + g, kq);
        a_off[1][s] = frag_off<AMAJOR>(wm0 + 8 + g, kq);
        b_off[0][s] = TILE_BYTES + frag_off<BMAJOR>(wn0 + g, kq);
        b_off[1][s] = TILE_BYTES + frag_off<BMAJOR>(wn0 + 8 + g, kq);
    }

    double acc[MI_][NI_][2];
    int stage = 0;
    uint32_t phase = 0;
    for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
        const int split = u / ntiles, tile = u % ntiles;
        int tm, tn;
        tile_coords(p, tile, tm, tn);
        int kt0 = split * p.kt_per_split;
        int kt1 = min(kt_total, kt0 + p.kt_per_split);
        if (p.kt_ranges) { const int2 r = p.kt_ranges[tm]; kt0 = r.x; kt1 = min(kt_total, r.y); }
#pragma unroll
        for (int i = 0; i < MI_; ++i)
#pragma unroll
            for (int j = 0; j < NI_; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
        // narrow tiles: rows of this warp beyond M are padding (TMA zero fill) -- nothing to multiply, nothing to store
        const bool idle = NARROW && (int64_t)tm * BM + wm0 >= p.M;

        for (int kt = kt0; kt < kt1; ++kt) {
            if constexpr (NT == 0 || NT == NI) if (kt == p.hook_kt) {
                if (p.hook_w) {
                    const int r = 2 * tn + (wn0 >> 6);
                    // one row of accumulator tiles at a time: its 8 loads are issued back to back (volatile asm keeps them
                    // together; left to itself the compiler re-uses one register pair and serialises the L2 round trips)
#pragma unroll
                    for (int i = 0; i < MI_; ++i) {
                        const int64_t m = (int64_t)tm * BM + wm0 + 8 * i + g;
                        const double* wrow = p.hook_w + (m < p.M ? m : 0) * p.hook_ldw + 2 * tig;
                        double wx[NI], wy[NI];
#pragma unroll
                        for (int j = 0; j < NI; ++j)
                            asm volatile("ld.global.nc.v2.f64 {%0, %1}, [%2];" : "=d"(wx[j]), "=d"(wy[j]) : "l"(wrow + 8 * j));
                        // all 16 values live at once: the FMAs may not start before the last load was issued
                        asm volatile("" : "+d"(wx[0]), "+d"(wx[1]), "+d"(wx[2]), "+d"(wx[3]), "+d"(wx[4]), "+d"(wx[5]), "+d"(wx[6]), "+d"(wx[7]));
                        asm volatile("" : "+d"(wy[0]), "+d"(wy[1]), "+d"(wy[2]), "+d"(wy[3]), "+d"(wy[4]), "+d"(wy[5]), "+d"(wy[6]), "+d"(wy[7]));
                        double s0 = 0.0, s1 = 0.0;
#pragma unroll
                        for (int j = 0; j < NI; ++j) {
                            s0 = fma(acc[i][j][0], wx[j], s0);
                            s1 = fma(acc[i][j][1], wy[j], s1);
                        }
                        double sum = s0 + s1;
                        sum += __shfl_xor_sync(0xffffffffu, sum, 1);
                        sum += __shfl_xor_sync(0xffffffffu, sum, 2);
                        if (tig == 0 && m < p.M) p.C2[m * p.hook_ldo + r] = sum;
                    }
                } else {
                    store_acc(acc, p.C2, p.ldc_m, p.ldc_n, p.M, p.N, (int64_t)tm * BM + wm0, (int64_t)tn * BN + wn0, g, tig, 1.0);
                }
#pragma unroll
                for (int i = 0; i < MI_; ++i)
#pragma unroll
                    for (int j = 0; j < NI; ++j) { acc[i][j][0] *= p.hook_scale; acc[i][j][1] *= p.hook_scale; }
            }
            mbar_wait(&full[stage], phase);
            const uint32_t st_addr = smem_base + stage * STAGE_BYTES;
            if (!idle) {
#pragma unroll
                for (int s = 0; s < 4; ++s) {
                    double af[MI_], bf[NI_];
#pragma unroll
                    for (int i = 0; i < MI_; ++i) af[i] = lds_f64(st_addr + a_off[i & 1][s] + (uint32_t)(i >> 1) * 2048u);
#pragma unroll
                    for (int j = 0; j < NI_; ++j) bf[j] = lds_f64(st_addr + b_off[j & 1][s] + (uint32_t)(j >> 1) * 2048u);
#pragma unroll
                    for (int i = 0; i < MI_; ++i)
#pragma unroll
                        for (int j = 0; j < NI_; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        double* Cout = p.C + (int64_t)split * p.split_stride;
        if (!idle) store_acc(acc, Cout, p.ldc_m, p.ldc_n, p.M, p.N, (int64_t)tm * BM + wm0, (int64_t)tn * BN + wn0, g, tig, p.alpha);
    }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline int get_encoder(rt_handle* h, EncodeTiledFn* fn) {
    if (!h->tma_encode) {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        RT_CUDA(h, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q));
        if (!f || q != cudaDriverEntryPointSuccess)
            return rt_fail(h, RT_ECUDA, "cuTensorMapEncodeTiled not available from the driver");
        h->tma_encode = f;
    }
    *fn = reinterpret_cast<EncodeTiledFn>(h->tma_encode);
    return RT_OK;
}

// An "mn x k" operand view: element (mn, k) at ptr[mn*stride_mn + k*stride_k]; one stride must be 1.
struct Operand {
    const double* ptr;
    int64_t mn, k;
    int64_t stride_mn, stride_k;
    int major() const { return stride_k == 1 ? MAJOR_K : MAJOR_MN; }
    bool tma_ok() const {
        if (reinterpret_cast<uintptr_t>(ptr) & 15) return false;
        if (stride_k == 1) return (stride_mn % 2 == 0 || mn == 1) && stride_mn >= k;
        if (stride_mn == 1) return (stride_k % 2 == 0 || k == 1) && stride_k >= mn;
        return false;
    }
};

inline int make_map(rt_handle* h, const Operand& op, CUtensorMap* map) {
    EncodeTiledFn enc;
    int rc = get_encoder(h, &enc);
    if (rc) return rc;
    cuuint64_t dims[2], strides[1];
    cuuint32_t box[2], estr[2] = {1, 1};
    if (op.major() == MAJOR_K) {
        dims[0] = (cuuint64_t)op.k; dims[1] = (cuuint64_t)op.mn;
        strides[0] = (cuuint64_t)op.stride_mn * 8;
        box[0] = 16; box[1] = 128;
    } else {
        dims[0] = (cuuint64_t)op.mn; dims[1] = (cuuint64_t)op.k;
        strides[0] = (cuuint64_t)op.stride_k * 8;
        box[0] = 16; box[1] = 16;
    }
    if (dims[1] == 1) strides[0] = ((dims[0] * 8 + 15) / 16) * 16;  // unused, must be a multiple of 16
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(op.ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return rt_fail(h, RT_ECUDA, "cuTensorMapEncodeTiled failed with %d (mn=%lld k=%lld smn=%lld sk=%lld)", (int)r,
                       (long long)op.mn, (long long)op.k, (long long)op.stride_mn, (long long)op.stride_k);
    return RT_OK;
}

template <int AMAJOR, int BMAJOR, int NT>
inline int launch_n(rt_handle* h, const CUtensorMap& ma, const CUtensorMap& mb, const Params& p, int grid, cudaStream_t st) {
    auto fn = k_dgemm<AMAJOR, BMAJOR, NT>;
    RT_CUDA(h, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
#ifdef RT_GEMM_LAUNCH_REGS
    RT_CUDA(h, cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
#endif
    fn<<<grid, THREADS, SMEM_BYTES, st>>>(ma, mb, p);
    RT_LAUNCH_CHECK(h);
    return RT_OK;
}

template <int AMAJOR, int BMAJOR>
inline int launch_t(rt_handle* h, const CUtensorMap& ma, const CUtensorMap& mb, const Params& p, int grid, cudaStream_t st) {
    // one column of tiles: the narrow instantiations (a symmetric product with a single tile is one too; the accumulator
    // hook exists in the 64-column one only, as before)
    if (p.tiles_n == 1 && (!p.symmetric || p.tiles_m == 1)) {
        if (p.N <= 64) return launch_n<AMAJOR, BMAJOR, 8>(h, ma, mb, p, grid, st);
        if (p.hook_kt < 0 && p.N <= 80) return launch_n<AMAJOR, BMAJOR, 10>(h, ma, mb, p, grid, st);
        if (p.hook_kt < 0 && p.N <= 96) return launch_n<AMAJOR, BMAJOR, 12>(h, ma, mb, p, grid, st);
    }
    return launch_n<AMAJOR, BMAJOR, 0>(h, ma, mb, p, grid, st);
}

// C = alpha * A B^T-style product of two "mn x k" operands.  `p` carries C, strides, split and hook
// settings; M, N, K, tiles are filled here.
inline int launch(rt_handle* h, const Operand& A, const Operand& B, Params p, cudaStream_t st) {
    if (A.k != B.k) return rt_fail(h, RT_EINVAL, "dgemm: inner dimensions differ");
    if (!A.tma_ok() || !B.tma_ok()) return rt_fail(h, RT_EINVAL, "dgemm: operand not TMA-compatible (caller must stage it)");
    p.M = A.mn; p.N = B.mn; p.K = A.k;
    p.tiles_m = (int)ceil_div64(p.M, BM);
    p.tiles_n = (int)ceil_div64(p.N, BN);
    const int kt_total = (int)ceil_div64(p.K, BK);
    if (p.splits < 1) p.splits = 1;
    if (p.splits > kt_total) p.splits = kt_total > 0 ? kt_total : 1;
    p.kt_per_split = (int)ceil_div64(kt_total, p.splits);
    p.splits = (int)ceil_div64(kt_total, p.kt_per_split);
    if (p.splits < 1) p.splits = 1;
    CUtensorMap ma, mb;
    int rc = make_map(h, A, &ma);
    if (rc) return rc;
    rc = make_map(h, B, &mb);
    if (rc) return rc;
    const int ntiles = p.symmetric ? p.tiles_n * (p.tiles_n + 1) / 2 : p.tiles_m * p.tiles_n;
    int grid = ntiles * p.splits;
    if (grid > h->sm_count) grid = h->sm_count;
    if (grid < 1) grid = 1;
    const int am = A.major(), bm = B.major();
    if (am == MAJOR_K && bm == MAJOR_K) return launch_t<MAJOR_K, MAJOR_K>(h, ma, mb, p, grid, st);
    if (am == MAJOR_K && bm == MAJOR_MN) return launch_t<MAJOR_K, MAJOR_MN>(h, ma, mb, p, grid, st);
    if (am == MAJOR_MN && bm == MAJOR_K) return launch_t<MAJOR_MN, MAJOR_K>(h, ma, mb, p, grid, st);
    return launch_t<MAJOR_MN, MAJOR_MN>(h, ma, mb, p, grid, st);
}

inline Params default_params() {
    Params p;
    memset(&p, 0, sizeof(p));
    p.alpha = 1.0;
    p.splits = 1;
    p.hook_kt = -1;
    p.hook_scale = 1.0;
    return p;
}

}  // namespace dg
