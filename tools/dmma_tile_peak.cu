// DMMA ceiling with the GEMM core's register tile: 4 x 8 accumulator tiles per warp (64 accumulator registers),
// distinct A / B fragment registers, 8 warps per CTA, 1 CTA per SM -- no shared-memory traffic at all.
// Compare with rt_fp64_peak_tflops (same operands for every DMMA): the gap is what operand fetch costs.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256) k(double* out, int iters, double s) {
    double acc[4][8][2], af[4], bf[8];
    for (int i = 0; i < 4; ++i) af[i] = s + i + threadIdx.x;
    for (int j = 0; j < 8; ++j) bf[j] = s * j - threadIdx.x;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 8; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j)
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(acc[i][j][0]), "+d"(acc[i][j][1]) : "d"(af[i]), "d"(bf[j]));
    }
    double t = 0;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 8; ++j) t += acc[i][j][0] + acc[i][j][1];
    if (t == 123.456) out[0] = t;
}
// the same tile with its fragments re-loaded from shared memory every k-step (12 LDS.64 per 32 DMMA, conflict-free
// addresses, no barriers, no TMA): what the GEMM main loop would reach without its pipeline synchronisation
__global__ void __launch_bounds__(256) k_lds(double* out, int iters, double s) {
    __shared__ double sm[4096];
    for (int e = threadIdx.x; e < 4096; e += 256) sm[e] = s + e;
    __syncthreads();
    double acc[4][8][2];
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 8; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    const int lane = threadIdx.x & 31;
    const unsigned base = (unsigned)__cvta_generic_to_shared(sm) + lane * 8;
    for (int it = 0; it < iters; ++it) {
        const unsigned off = base + (it & 7) * 3072;
        double af[4], bf[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) asm volatile("ld.shared.f64 %0, [%1];" : "=d"(af[i]) : "r"(off + i * 256));
#pragma unroll
        for (int j = 0; j < 8; ++j) asm volatile("ld.shared.f64 %0, [%1];" : "=d"(bf[j]) : "r"(off + 1024 + j * 256));
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j)
                asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                    : "+d"(acc[i][j][0]), "+d"(acc[i][j][1]) : "d"(af[i]), "d"(bf[j]));
    }
    double t = 0;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 8; ++j) t += acc[i][j][0] + acc[i][j][1];
    if (t == 123.456) out[0] = t;
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    double* out; cudaMalloc(&out, 8);
    const int iters = 20000;
    for (int bps = 1; bps <= 2; ++bps) {
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        k<<<p.multiProcessorCount * bps, 256>>>(out, iters, 1.0000001);
        cudaDeviceSynchronize();
        float best = 1e30f;
        for (int r = 0; r < 3; ++r) {
            cudaEventRecord(e0); k<<<p.multiProcessorCount * bps, 256>>>(out, iters, 1.0000001); cudaEventRecord(e1);
            cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        printf("register tile 4x8, %d CTA(s) of 8 warps per SM: %.2f TFLOP/s\n", bps,
               (double)p.multiProcessorCount * bps * 8 * 32 * (double)iters * 512.0 / best * 1e-9);
    }
    {
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        k_lds<<<p.multiProcessorCount, 256>>>(out, iters, 1.0000001);
        cudaDeviceSynchronize();
        float best = 1e30f;
        for (int r = 0; r < 3; ++r) {
            cudaEventRecord(e0); k_lds<<<p.multiProcessorCount, 256>>>(out, iters, 1.0000001); cudaEventRecord(e1);
            cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        printf("register tile 4x8 with LDS fragments every k-step: %.2f TFLOP/s\n",
               (double)p.multiProcessorCount * 8 * 32 * (double)iters * 512.0 / best * 1e-9);
    }
    return 0;
}
