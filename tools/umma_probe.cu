// Probe: one tcgen05.mma kind::tf32 tile (M=128, N=64, K=32 = 4 instructions) with hand-built descriptors, operands
// written by ordinary threads in the K-major SWIZZLE_128B shared-memory layout, result read back with tcgen05.ld and
// compared with a CPU product.  Used to validate the descriptor encoding before it goes into a product kernel.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_probe tools/umma_probe.cu && timeout 30 ./umma_probe
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <math.h>

constexpr int M = 128, N = 64, K = 32;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major, 128-byte swizzle: row r (M or N index), 32 tf32 = 128 bytes per row, 8-row atoms of 1024 bytes;
// 16-byte chunk c of row r is stored at chunk (c ^ (r & 7))
__host__ __device__ inline int sw128_index(int r, int k) {   // float index inside the tile
    const int chunk = (k >> 2) ^ (r & 7);
    return (r >> 3) * 256 + (r & 7) * 32 + chunk * 4 + (k & 3);
}

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);            // start address, 16-byte units, bits [0,14)
    d |= (uint64_t)1 << 16;                             // leading byte offset (unused for swizzled K-major), bits [16,30)
    d |= (uint64_t)(1024 >> 4) << 32;                   // stride byte offset between 8-row atoms, bits [32,46)
    d |= (uint64_t)1 << 46;                             // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                             // SWIZZLE_128B
    return d;
}

__global__ void __launch_bounds__(128) probe(const float* A, const float* B, float* D) {
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* base = smem + ((1024u - (smem_u32(smem) & 1023u)) & 1023u);   // SWIZZLE_128B atoms: 1024-byte aligned
    float* sA = reinterpret_cast<float*>(base);                  // 128 x 32 floats = 16 KB
    float* sB = reinterpret_cast<float*>(base + 16384);          // 64 x 32 floats  =  8 KB
    __shared__ uint32_t tmem_base;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < M * K; i += 128) sA[sw128_index(i / K, i % K)] = A[i];
    for (int i = tid; i < N * K; i += 128) sB[sw128_index(i / K, i % K)] = B[i];
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic smem writes -> visible to the MMA's async proxy
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base;
    if (tid == 0) {
        // instruction descriptor: D=F32 (bits 4-5 = 1), A=B=TF32 (2 at bits 7-9 and 10-12), K-major both, N>>3 at 17, M>>4 at 24
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint64_t da = make_desc(smem_u32(sA)), db = make_desc(smem_u32(sB));
        for (int k = 0; k < K / 8; ++k) {
            const uint64_t a_k = da + (uint64_t)((k * 32) >> 4), b_k = db + (uint64_t)((k * 32) >> 4);
            const uint32_t acc = k > 0 ? 1u : 0u;
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem),
                "l"(a_k), "l"(b_k), "r"(idesc), "r"(acc)
                : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar))
                     : "memory");
    }
    // everyone waits for the MMAs
    {
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                : "=r"(done)
                : "r"(smem_u32(&bar))
                : "memory");
        }
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // warp w reads TMEM lanes 32w..32w+31 (= rows), 64 columns, 16 at a time
    const int row = tid;
    for (int c0 = 0; c0 < N; c0 += 16) {
        uint32_t v[16];
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
              "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
            : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int i = 0; i < 16; ++i) D[row * N + c0 + i] = __uint_as_float(v[i]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem));
}

int main() {
    float *hA = (float*)malloc(M * K * 4), *hB = (float*)malloc(N * K * 4), *hD = (float*)malloc(M * N * 4);
    srand(1);
    for (int i = 0; i < M * K; ++i) hA[i] = (float)((rand() % 17) - 8) * 0.25f;     // exactly representable in TF32
    for (int i = 0; i < N * K; ++i) hB[i] = (float)((rand() % 13) - 6) * 0.5f;
    float *dA, *dB, *dD;
    cudaMalloc(&dA, M * K * 4); cudaMalloc(&dB, N * K * 4); cudaMalloc(&dD, M * N * 4);
    cudaMemcpy(dA, hA, M * K * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB, N * K * 4, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0, M * N * 4);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
    probe<<<1, 128, 16384 + 8192 + 1024>>>(dA, dB, dD);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    cudaMemcpy(hD, dD, M * N * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0;
    int bad = 0;
    for (int m = 0; m < M; ++m)
        for (int n = 0; n < N; ++n) {
            double ref = 0;
            for (int k = 0; k < K; ++k) ref += (double)hA[m * K + k] * hB[n * K + k];
            const double err = fabs(ref - hD[m * N + n]);
            if (err > maxerr) maxerr = err;
            if (err > 1e-3 && bad++ < 8) printf("  D[%d][%d] = %f, want %f\n", m, n, hD[m * N + n], ref);
        }
    printf("max |err| = %g, mismatches = %d of %d\n", maxerr, bad, M * N);
    return bad ? 2 : 0;
}
