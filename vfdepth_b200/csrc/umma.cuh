// tcgen05 / TMEM / TMA helpers for the dense products of the fused gather (dense_proj.cu).
//   - shared-memory matrix descriptors for K-major SWIZZLE_128B tiles (8-row x 128-byte atoms, 1024 bytes each; the
//     encoding was validated against a CPU product by tools/umma_probe.cu in round 1),
//   - tcgen05.mma kind::tf32 issue / commit / fences,
//   - cp.async.bulk.tensor (TMA) tile loads that complete on an mbarrier; the tensor maps are encoded on the host with
//     cuTensorMapEncodeTiled, reached through cudaGetDriverEntryPoint so the library does not link libcuda.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace vfd {

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// arrives on `bar` when all tcgen05.mma issued so far by this thread have finished
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// shared-memory matrix descriptor: K-major, SWIZZLE_128B, 8-row atoms of 1024 bytes
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
           ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// instruction descriptor of kind::tf32, fp32 accumulate, both operands K-major, M x N tile
__host__ __device__ inline uint32_t umma_idesc_tf32(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// MN-major operand (rows = K, 32 consecutive M/N elements per 128-byte row, 4-row atoms of 512 bytes with 32-byte chunks
// XOR row & 3 = what a TMA box {32 floats, rows} with SWIZZLE_128B_ATOM_32B leaves in shared memory): layout type
// SWIZZLE_128B_BASE32B, leading byte offset = distance between 32-element M/N blocks, stride byte offset = 512 (between
// the two 4-row atoms one K = 8 instruction reads).  Validated by tools/umma_mn_probe.cu (profiles/r02_umma_mn_probe.txt).
__device__ __forceinline__ uint64_t umma_desc_mn(uint32_t saddr, uint32_t block_stride_bytes) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)((block_stride_bytes >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)(512 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)1 << 61);
}
// instruction descriptor of kind::tf32, fp32 accumulate, both operands MN-major
__host__ __device__ inline uint32_t umma_idesc_tf32_mn(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(N >> 3) << 17) |
           ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
        "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
        : "memory");
}

// 3-D tiled TMA load: box of the tensor map at element coordinates (c0 innermost, c1, c2) -> shared memory, completes
// the box's bytes on `bar`.  Rows / columns of the box that fall outside the tensor are zero-filled.
__device__ __forceinline__ void tma_load_3d(void* sdst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
            smem_u32(sdst)),
        "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

// 16 consecutive fp32 accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// ---- host: tensor-map encoding ----
typedef CUresult (*vfd_encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline vfd_encode_tiled_fn encode_tiled_entry() {
    static vfd_encode_tiled_fn fn = []() -> vfd_encode_tiled_fn {   // resolved once; the entry point never changes
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return nullptr;
        return reinterpret_cast<vfd_encode_tiled_fn>(p);
    }();
    return fn;
}

// fp32 tensor [d2][d1][d0] (d0 contiguous), box [1][box1][32 floats = 128 bytes], zero fill.  swizzle32 = false:
// SWIZZLE_128B (16-byte chunks XOR row & 7: the K-major operand layout); true: SWIZZLE_128B_ATOM_32B (32-byte chunks XOR
// row & 3: the only MN-major layout kind::tf32 accepts -- UMMA layout type SWIZZLE_128B_BASE32B, tools/umma_mn_probe.cu)
inline bool make_map_3d(CUtensorMap* map, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint32_t box1,
                        bool swizzle32 = false) {
    vfd_encode_tiled_fn enc = encode_tiled_entry();
    if (!enc) return false;
    const cuuint64_t dims[3] = {d0, d1, d2};
    const cuuint64_t strides[2] = {d0 * sizeof(float), d0 * d1 * sizeof(float)};
    const cuuint32_t box[3] = {32, box1, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
               CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace vfd
