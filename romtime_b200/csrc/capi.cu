// Life-cycle entry points of the C ABI + the FP64 DMMA peak probe.
#include "common.cuh"

extern "C" int rt_version(void) { return 100; }

extern "C" int rt_create(int device, rt_handle** out) {
    if (!out) return RT_EINVAL;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        cudaGetLastError();
        return RT_ENODEVICE;  // no CPU fallback: the caller must fail loudly
    }
    if (device < 0 || device >= count) return RT_EINVAL;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return RT_ECUDA;
    if (prop.major != 10) return RT_ENODEVICE;  // sm_100a cubins only
    if (cudaSetDevice(device) != cudaSuccess) return RT_ECUDA;
    rt_handle* h = new rt_handle();
    h->device = device;
    h->sm_count = prop.multiProcessorCount;
    h->smem_optin = (int)prop.sharedMemPerBlockOptin;
    *out = h;
    return RT_OK;
}

void rt_online_release(rt_handle* h);  // online.cu

extern "C" int rt_destroy(rt_handle* h) {
    if (!h) return RT_EINVAL;
    cudaSetDevice(h->device);
    rt_online_release(h);
    for (int i = 0; i < rt_handle::kSlots; ++i)
        if (h->ws[i]) cudaFree(h->ws[i]);
    delete h;
    return RT_OK;
}

extern "C" const char* rt_last_error(const rt_handle* h) { return h ? h->err : "null handle"; }

extern "C" int64_t rt_launch_count(const rt_handle* h) { return h ? h->launches : 0; }

// Let kernels of this handle's device read buffers that live on `peer` (NVLink / NVSwitch peer memory, e.g. a neighbour
// rank's snapshot columns opened through CUDA IPC): the column-sharded Gram blocks then pull their B operand tile by tile
// with TMA over NVLink while the DMMAs run -- no staging copy, no send / receive.
extern "C" int rt_enable_peer_access(rt_handle* h, int peer) {
    if (!h) return RT_EINVAL;
    RT_CUDA(h, cudaSetDevice(h->device));
    if (peer == h->device) return RT_OK;
    int can = 0;
    RT_CUDA(h, cudaDeviceCanAccessPeer(&can, h->device, peer));
    if (!can) return rt_fail(h, RT_EINVAL, "device %d cannot access device %d as a peer", h->device, peer);
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return RT_OK; }
    RT_CUDA(h, e);
    return RT_OK;
}

// Strided device-to-host / host-to-device copy on a stream (cudaMemcpy2DAsync): a time slab [:, n0:n1, :] of the sweep
// history (B x nt x N) goes to pinned host memory as ONE copy while later steps are computed.
extern "C" int rt_memcpy2d_async(rt_handle* h, void* dst, int64_t dst_pitch, const void* src, int64_t src_pitch, int64_t width_bytes,
                                 int64_t rows, int to_host, void* stream) {
    if (!h) return RT_EINVAL;
    RT_REQUIRE(h, dst && src && width_bytes >= 1 && rows >= 1 && dst_pitch >= width_bytes && src_pitch >= width_bytes);
    RT_CUDA(h, cudaSetDevice(h->device));
    RT_CUDA(h, cudaMemcpy2DAsync(dst, (size_t)dst_pitch, src, (size_t)src_pitch, (size_t)width_bytes, (size_t)rows,
                                 to_host ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice, rt_stream(stream)));
    return RT_OK;
}

// Register-resident DMMA loop: 8 independent m8n8k4 accumulators per warp (see tools/fp64_peak.cu).
__global__ void __launch_bounds__(256) k_fp64_peak(double* out, int iters, double a, double b) {
    double c0[8], c1[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { c0[i] = i; c1[i] = threadIdx.x; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c0[i]), "+d"(c1[i]) : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += c0[i] + c1[i];
    if (s == 123.456) out[0] = s;
}

extern "C" int rt_fp64_peak_tflops(rt_handle* h, double* tflops) {
    if (!h || !tflops) return RT_EINVAL;
    if (h->fp64_peak > 0) { *tflops = h->fp64_peak; return RT_OK; }
    RT_CUDA(h, cudaSetDevice(h->device));
    void* out;
    int rc = rt_scratch(h, 7, 256, &out);
    if (rc) return rc;
    const int iters = 20000, grid = h->sm_count * 4;
    cudaEvent_t e0, e1;
    RT_CUDA(h, cudaEventCreate(&e0));
    RT_CUDA(h, cudaEventCreate(&e1));
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        RT_CUDA(h, cudaEventRecord(e0, 0));
        k_fp64_peak<<<grid, 256>>>((double*)out, iters, 1.0000001, 1e-9);
        RT_LAUNCH_CHECK(h);
        RT_CUDA(h, cudaEventRecord(e1, 0));
        RT_CUDA(h, cudaEventSynchronize(e1));
        float ms;
        RT_CUDA(h, cudaEventElapsedTime(&ms, e0, e1));
        if (r > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    h->fp64_peak = (double)grid * 8 /*warps*/ * 8 /*ilp*/ * iters * 512.0 / best * 1e-9;
    *tflops = h->fp64_peak;
    return RT_OK;
}
