// Probe: tcgen05.mma kind::tf32 with BOTH operands MN-major (the layout a TMA box {32 channels, KP rows} with
// SWIZZLE_128B leaves in shared memory when the global tensor is [rows = K][channels = M or N], channels contiguous):
//     D[m][n] = sum_k A[k][m] * B[k][n]        M = 128, N = 64, K = 32 (4 instructions of K = 8)
// Canonical MN-major SWIZZLE_128B layout (cute/atom/mma_traits_sm100.hpp, in 16-byte units):
//     ((8, n_atoms), (8, k_groups)) : ((1, LBO), (8, SBO))
// i.e. one 128-byte row per k holding 32 consecutive MN elements, 8 rows = one 1024-byte swizzle atom, atoms of later
// k-groups SBO bytes further, the next 32 MN elements LBO bytes further.  The probe tries the descriptor field
// assignments (LBO, SBO) and (SBO, LBO) and reports which one reproduces the CPU product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_mn_probe tools/umma_mn_probe.cu && timeout 30 ./umma_mn_probe
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

constexpr int M = 128, N = 64, K = 32;
constexpr int kAtomBytes = 1024;               // 8 k-rows x 128 B = one K = 8 instruction's slice of a 32-element MN block
constexpr int kColBlockBytes = (K / 8) * kAtomBytes;   // one 32-element MN block over all of K

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// float index of element (mn, k) in the MN-major tile: one 128-byte row per k holding 32 consecutive MN elements.
//   swz32 = 0: SWIZZLE_128B          16-byte chunks XOR (k & 7), 8-row atoms          (layout type 2)
//   swz32 = 1: SWIZZLE_128B_BASE32B  32-byte chunks XOR (k & 3), 4-row atoms of 512 B (layout type 1) -- what CUTLASS
//              uses for MN-major TF32 ("SW128_32B is the only available smem layout", sm100_common.inl:92)
__host__ __device__ inline int mn_index(int mn, int k, int swz32) {
    const int blk = mn >> 5, e = mn & 31;
    if (swz32) {
        const int chunk = (e >> 3) ^ (k & 3);
        return (blk * kColBlockBytes + k * 128 + chunk * 32) / 4 + (e & 7);
    }
    const int chunk = (e >> 2) ^ (k & 7);
    return (blk * kColBlockBytes + (k >> 3) * kAtomBytes + (k & 7) * 128 + chunk * 16) / 4 + (e & 3);
}

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lead_bytes, uint32_t stride_bytes, int swz32) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lead_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((stride_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(swz32 ? 1 : 2) << 61;   // SWIZZLE_128B_BASE32B : SWIZZLE_128B
    return d;
}

__global__ void __launch_bounds__(128) probe(const float* A, const float* B, float* D, int variant) {
    const int swz32 = variant >> 1;
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* base = smem + ((1024u - (smem_u32(smem) & 1023u)) & 1023u);
    float* sA = reinterpret_cast<float*>(base);                           // 4 blocks x 4 KB
    float* sB = reinterpret_cast<float*>(base + (M / 32) * kColBlockBytes);   // 2 blocks x 4 KB
    __shared__ uint32_t tmem_base;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < K * M; i += 128) sA[mn_index(i % M, i / M, swz32)] = A[i];      // A is [K][M]
    for (int i = tid; i < K * N; i += 128) sB[mn_index(i % N, i / N, swz32)] = B[i];      // B is [K][N]
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base;
    if (tid == 0) {
        // D=F32, A=B=TF32, A and B MN-major (bits 15, 16), N>>3 at 17, M>>4 at 24
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(N >> 3) << 17) |
                               ((uint32_t)(M >> 4) << 24);
        const uint32_t lbo = kColBlockBytes, sbo = swz32 ? 512 : kAtomBytes;   // MN-block stride, k-atom stride
        for (int k = 0; k < K / 8; ++k) {
            const uint32_t oa = smem_u32(sA) + k * kAtomBytes, ob = smem_u32(sB) + k * kAtomBytes;
            const uint64_t a_k = (variant & 1) == 0 ? make_desc(oa, lbo, sbo, swz32) : make_desc(oa, sbo, lbo, swz32);
            const uint64_t b_k = (variant & 1) == 0 ? make_desc(ob, lbo, sbo, swz32) : make_desc(ob, sbo, lbo, swz32);
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
    float *hA = (float*)malloc(K * M * 4), *hB = (float*)malloc(K * N * 4), *hD = (float*)malloc(M * N * 4);
    srand(1);
    for (int i = 0; i < K * M; ++i) hA[i] = (float)((rand() % 17) - 8) * 0.25f;     // exactly representable in TF32
    for (int i = 0; i < K * N; ++i) hB[i] = (float)((rand() % 13) - 6) * 0.5f;
    float *dA, *dB, *dD;
    cudaMalloc(&dA, K * M * 4); cudaMalloc(&dB, K * N * 4); cudaMalloc(&dD, M * N * 4);
    cudaMemcpy(dA, hA, K * M * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB, K * N * 4, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    int ok = 0;
    for (int variant = 0; variant < 4; ++variant) {
        cudaMemset(dD, 0, M * N * 4);
        probe<<<1, 128, (M / 32 + N / 32) * kColBlockBytes + 1024>>>(dA, dB, dD, variant);
        cudaError_t e = cudaDeviceSynchronize();
        printf("variant %d (%s, %s): kernel %s\n", variant, (variant >> 1) ? "SWIZZLE_128B_BASE32B" : "SWIZZLE_128B",
               (variant & 1) == 0 ? "lead=MN-block stride, stride=k-atom stride" : "swapped", cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
        cudaMemcpy(hD, dD, M * N * 4, cudaMemcpyDeviceToHost);
        double maxerr = 0;
        int bad = 0;
        for (int m = 0; m < M; ++m)
            for (int n = 0; n < N; ++n) {
                double ref = 0;
                for (int k = 0; k < K; ++k) ref += (double)hA[k * M + m] * hB[k * N + n];
                const double err = fabs(ref - hD[m * N + n]);
                if (err > maxerr) maxerr = err;
                if (err > 1e-3 && bad++ < 4) printf("  D[%d][%d] = %f, want %f\n", m, n, hD[m * N + n], ref);
            }
        printf("  max |err| = %g, mismatches = %d of %d\n", maxerr, bad, M * N);
        if (!bad) ok |= 1 << variant;
    }
    printf("working variants mask: %d\n", ok);
    return ok ? 0 : 2;
}
