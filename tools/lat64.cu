// Dependent-issue latencies on B200 of the instructions on the pivot chain: DFMA, DMUL, MUFU.RCP64H, DMMA, SHFL, redux.
#include <cstdio>
#include <cuda_runtime.h>
template <int OP>
__global__ void k(double* out, long long* cyc, int iters, double a, double b) {
    double x = threadIdx.x * 1e-3 + 1.0, y = 2.0;
    unsigned u = threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        if (OP == 0) { x = fma(x, a, b); }
        if (OP == 1) { x = x * a; }
        if (OP == 2) { x = 1.0 / x; }
        if (OP == 3) { asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(x), "+d"(y) : "d"(a), "d"(b)); }
        if (OP == 4) { x = __shfl_sync(0xffffffffu, x, (threadIdx.x + 1) & 31); }
        if (OP == 5) { u = __reduce_max_sync(0xffffffffu, u) + threadIdx.x; }
        if (OP == 6) {  // 8 independent or-reductions (broadcast of 4 doubles without the MIO pipe)
            unsigned r0 = __reduce_or_sync(0xffffffffu, u), r1 = __reduce_or_sync(0xffffffffu, u + 1), r2 = __reduce_or_sync(0xffffffffu, u + 2),
                     r3 = __reduce_or_sync(0xffffffffu, u + 3), r4 = __reduce_or_sync(0xffffffffu, u + 4), r5 = __reduce_or_sync(0xffffffffu, u + 5),
                     r6 = __reduce_or_sync(0xffffffffu, u + 6), r7 = __reduce_or_sync(0xffffffffu, u + 7);
            u = (r0 ^ r1 ^ r2 ^ r3 ^ r4 ^ r5 ^ r6 ^ r7) + threadIdx.x;
        }
        if (OP == 7) {  // 8 independent 32-bit shuffles (= 4 doubles)
            const int src = (u + i) & 31;
            unsigned r0 = __shfl_sync(0xffffffffu, u, src), r1 = __shfl_sync(0xffffffffu, u + 1, src), r2 = __shfl_sync(0xffffffffu, u + 2, src),
                     r3 = __shfl_sync(0xffffffffu, u + 3, src), r4 = __shfl_sync(0xffffffffu, u + 4, src), r5 = __shfl_sync(0xffffffffu, u + 5, src),
                     r6 = __shfl_sync(0xffffffffu, u + 6, src), r7 = __shfl_sync(0xffffffffu, u + 7, src);
            u = (r0 ^ r1 ^ r2 ^ r3 ^ r4 ^ r5 ^ r6 ^ r7) + threadIdx.x;
        }
        if (OP == 8) {  // MUFU.RCP64H seed + third-order step (3 dependent DFMA)
            double x0; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x0) : "d"(x));
            const double e = fma(-x, x0, 1.0); const double p = fma(e, e, e); x = fma(x0, p, x0) + a;
        }
        if (OP == 9) {  // one lane stores 4 doubles, __syncwarp, all lanes read them back (broadcast through shared memory)
            __shared__ double2 buf[2];
            if (threadIdx.x == (u & 31)) { buf[0] = make_double2(x, y); buf[1] = make_double2(y, x); }
            __syncwarp();
            const double2 p0 = buf[0], p1 = buf[1];
            __syncwarp();
            x = p0.x + p1.y * 1e-30; u = u * 5 + 1;
        }
        if (OP == 10) {  // dependent LDS.64 chain
            __shared__ double tab[32];
            if (i == 0) { tab[threadIdx.x] = (double)((threadIdx.x * 7 + 3) & 31); __syncwarp(); }
            x = tab[(int)x & 31];
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[OP] = t1 - t0;
    if (x + y + u == 123.456) out[0] = x;
}
int main() {
    double* out; long long* cyc; cudaMalloc(&out, 8); cudaMallocManaged(&cyc, 128);
    const int it = 4096;
    k<0><<<1, 32>>>(out, cyc, it, 1.0000001, 1e-9); k<1><<<1, 32>>>(out, cyc, it, 1.0000001, 1e-9);
    k<2><<<1, 32>>>(out, cyc, it, 1.0000001, 1e-9); k<3><<<1, 32>>>(out, cyc, it, 1.0000001, 1e-9);
    k<4><<<1, 32>>>(out, cyc, it, 1.0000001, 1e-9); k<5><<<1, 32>>>(out, cyc, it, 1.0000001, 1e-9);
    k<6><<<1, 32>>>(out, cyc, it, 1.0000001, 1e-9); k<7><<<1, 32>>>(out, cyc, it, 1.0000001, 1e-9);
    k<8><<<1, 32>>>(out, cyc, it, 1.0000001, 1e-9); k<9><<<1, 32>>>(out, cyc, it, 1.0000001, 1e-9); k<10><<<1, 32>>>(out, cyc, it, 1.0000001, 1e-9);
    cudaDeviceSynchronize();
    const char* nm[] = {"dfma", "dmul", "rcp(1/x)", "dmma884", "shfl64", "redux_max+add", "8x redux_or group", "8x shfl32 group", "fast rcp (+dadd)", "sts->syncwarp->2 lds.128 broadcast", "lds.64 (+cvt)"};
    for (int i = 0; i < 11; ++i) printf("%s: %.1f cycles per dependent op\n", nm[i], (double)cyc[i] / it);
    return 0;
}
