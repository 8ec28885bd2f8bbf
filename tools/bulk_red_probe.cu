// Probe: scatter-add throughput of 128-byte (and larger) segments into a large fp32 buffer, two ways:
//   mode 0  red.global.add.v4.f32, 8 lanes x 16 B per segment (what the fused backward's scatter does)
//   mode 1  cp.reduce.async.bulk.global.shared::cta.add.f32, one bulk op per segment issued by the owning thread
// Segments land on pseudo-random SEG-byte aligned offsets of a 192 MB buffer (the d_feats size at B=8).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o bulk_red_probe tools/bulk_red_probe.cu && ./bulk_red_probe
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

template <int SEG>   // bytes per segment
__global__ void __launch_bounds__(128) red_kernel(float* buf, uint32_t nseg, int iters) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int LPS = SEG / 16;             // lanes per segment
    constexpr int SPW = 32 / LPS;             // segments per warp instruction
    const float4 v = make_float4(1.f, 1.f, 1.f, 1.f);
    for (int it = 0; it < iters; ++it) {
        const uint32_t id = ((blockIdx.x * iters + it) * 4 + warp) * SPW + lane / LPS;
        float* p = buf + (size_t)(hash32(id) % nseg) * (SEG / 4) + (lane % LPS) * 4;
        asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    }
}

template <int SEG, int SLOTS>
__global__ void __launch_bounds__(128) bulk_kernel(float* buf, uint32_t nseg, int iters) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int PITCH = SEG + 16;
    unsigned char* mine = smem + (size_t)threadIdx.x * SLOTS * PITCH;
    for (int it = 0; it < iters; ++it) {
        unsigned char* slot = mine + (it % SLOTS) * PITCH;
        // the slot's previous bulk op must have read its source
        asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(SLOTS - 1) : "memory");
        for (int q = 0; q < SEG / 16; ++q) *reinterpret_cast<float4*>(slot + q * 16) = make_float4(1.f, 1.f, 1.f, 1.f);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        const uint32_t id = (blockIdx.x * iters + it) * 128 + threadIdx.x;
        float* p = buf + (size_t)(hash32(id) % nseg) * (SEG / 4);
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(p),
                     "r"(smem_u32(slot)), "n"(SEG)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <int SEG>
static void run(float* buf, size_t bytes) {
    const uint32_t nseg = (uint32_t)(bytes / SEG);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 2; ++mode) {
        const int grid = 148 * 4;
        const int segs_per_cta_iter = 128;                       // both modes: 128 segments per CTA iteration
        const int iters = 4096;
        const int iters_red = iters * segs_per_cta_iter / (4 * (32 / (SEG / 16)));
        float ms = 0.f;
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(e0);
            if (mode == 0) red_kernel<SEG><<<grid, 128>>>(buf, nseg, iters_red);
            else {
                constexpr int SLOTS = 4;
                const size_t smem = (size_t)128 * SLOTS * (SEG + 16);
                cudaFuncSetAttribute(bulk_kernel<SEG, SLOTS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                bulk_kernel<SEG, SLOTS><<<grid, 128, smem>>>(buf, nseg, iters);
            }
            cudaEventRecord(e1);
            cudaError_t e = cudaEventSynchronize(e1);
            if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return; }
            cudaEventElapsedTime(&ms, e0, e1);
        }
        const double total = (double)grid * iters * segs_per_cta_iter * SEG;
        printf("SEG=%4d  %s  %.3f ms  %.1f GB/s\n", SEG, mode ? "bulk reduce" : "red.v4     ", ms, total / ms * 1e-6);
    }
}

int main() {
    const size_t bytes = (size_t)192 << 20;
    float* buf;
    cudaMalloc(&buf, bytes);
    cudaMemset(buf, 0, bytes);
    run<128>(buf, bytes);
    run<256>(buf, bytes);
    // sanity: total added must equal the number of floats written
    double sum = 0;
    {
        float* h = (float*)malloc(bytes);
        cudaMemcpy(h, buf, bytes, cudaMemcpyDeviceToHost);
        for (size_t i = 0; i < bytes / 4; ++i) sum += h[i];
        free(h);
    }
    printf("sum of buffer = %.0f (expect %.0f)\n", sum, 3.0 * 2 * 148 * 4 * 4096.0 * 128 * (128 + 256) / 4);
    return 0;
}
