// Do DMMA (mma.sync f64) and DFMA share one execution resource on B200?  Warps [0, dw) of every CTA run a DMMA
// loop, the rest a DFMA loop; reports each class's TFLOP/s when run alone and together.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_mix fp64_mix.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) k_mix(double* out, int iters_mma, int iters_fma, int dmma_warps, double a, double b) {
    const int warp = threadIdx.x >> 5;
    double s = 0;
    if (warp < dmma_warps) {
        double c0[8], c1[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { c0[i] = i; c1[i] = threadIdx.x; }
        for (int it = 0; it < iters_mma; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(c0[i]), "+d"(c1[i]) : "d"(a), "d"(b));
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) s += c0[i] + c1[i];
    } else {
        double acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = threadIdx.x * 1e-9 + i;
        for (int it = 0; it < iters_fma; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[i] = fma(acc[i], a, b);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) s += acc[i];
    }
    if (s == 123.456) out[0] = s;
}

static float run(double* out, int grid, int im, int ifm, int dw) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_mix<<<grid, 256>>>(out, im, ifm, dw, 1.0000001, 1e-9);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        cudaEventRecord(e0);
        k_mix<<<grid, 256>>>(out, im, ifm, dw, 1.0000001, 1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int grid = p.multiProcessorCount * 4;
    double* out; cudaMalloc(&out, 8);
    const int im = 20000;
    // DMMA: 8 ilp * 512 flop per warp-iteration; DFMA: 8 ilp * 64 flop per warp-iteration (32 lanes * 2)
    const int ifm = im * 4;  // a DFMA warp instruction is 64 flop vs 512: run 4x the iterations so both take similar time
    float t_mma = run(out, grid, im, 0, 4);
    float t_fma = run(out, grid, 0, ifm, 4);
    float t_mix = run(out, grid, im, ifm, 4);
    double f_mma = (double)grid * 4 * 8 * im * 512.0, f_fma = (double)grid * 4 * 8 * (double)ifm * 64.0;
    printf("{\"dmma_alone_4warps_tflops\": %.2f, \"dfma_alone_4warps_tflops\": %.2f, \"mixed_ms\": %.3f, \"dmma_alone_ms\": %.3f, "
           "\"dfma_alone_ms\": %.3f, \"mixed_total_tflops\": %.2f}\n",
           f_mma / t_mma * 1e-9, f_fma / t_fma * 1e-9, t_mix, t_mma, t_fma, (f_mma + f_fma) / t_mix * 1e-9);
    float t8 = run(out, grid, im, 0, 8);
    printf("{\"dmma_8warps_tflops\": %.2f}\n", (double)grid * 8 * 8 * im * 512.0 / t8 * 1e-9);
    return 0;
}
