// Pieces shared by the voxel-gather kernels (voxel_gather.cu) and the fused gather + MLP kernels (voxel_fuse.cu):
// the per-(voxel, camera) projection record, the validity decision of volumetric_fusionnet.py:183-195 and the
// bilinear corner walk.  One source for both so the decisions stay bit-identical.
#pragma once
#include "common.cuh"

namespace vfd {

struct CamSlot {   // one (voxel, camera) projection
    int idx;       // iy0 * w + ix0
    float wx, wy;  // fractional parts
    float zrel;    // v.z / depth_div
};

// Project the voxel centre (x,y,z) into one camera (Einv: 16 floats, K: 16 floats, mask_cam: [h,w] resized
// self-occlusion mask).  Returns the validity of :183-195 and fills `s` (zeros when invalid) and g[2] (the
// normalised sampling coordinate of :166-181, bit-exact).
__device__ __forceinline__ bool project_cam(const vfd_fusion_desc& d, const float* mask_cam, const float* Einv,
                                            const float* K, float x, float y, float z, CamSlot& s, float* g) {
    float v[3];
    vfd_voxel_to_pixel(Einv, K, x, y, z, d.w, d.h, v, g);
    const bool inb = vfd_in_bounds2(g[0], g[1]);
    const float ix = vfd_unnormalize(g[0], d.w), iy = vfd_unnormalize(g[1], d.h);
    bool ok = false;
    s.idx = 0; s.wx = 0.f; s.wy = 0.f; s.zrel = 0.f;
    if (inb && v[2] > 0.f) {
        // nearest-neighbour self-occlusion lookup: ATen uses nearbyint (half-to-even) -> rintf
        const float fx = rintf(ix), fy = rintf(iy);
        if (fx >= 0.f && fx <= (float)(d.w - 1) && fy >= 0.f && fy <= (float)(d.h - 1))
            ok = __ldg(mask_cam + (int)fy * d.w + (int)fx) > 0.5f;
        const float x0 = floorf(ix), y0 = floorf(iy);
        s.idx = (int)y0 * d.w + (int)x0;
        s.wx = ix - x0;
        s.wy = iy - y0;
        s.zrel = VFD_DIV(v[2], d.depth_div);
    }
    return ok;
}

// bilinear corner walk shared by forward (gather) and backward (scatter)
struct Corners {
    int off[4];
    float wgt[4];
};
__device__ __forceinline__ Corners make_corners(const CamSlot& s, int w, int h, int C) {
    Corners q;
    int x0 = s.idx % w, y0 = s.idx / w;
    bool x1 = (x0 + 1) < w, y1 = (y0 + 1) < h;
    float ex = 1.f - s.wx, ey = 1.f - s.wy;
    q.off[0] = s.idx * C;            q.wgt[0] = ex * ey;
    q.off[1] = (s.idx + 1) * C;      q.wgt[1] = x1 ? s.wx * ey : 0.f;
    q.off[2] = (s.idx + w) * C;      q.wgt[2] = y1 ? ex * s.wy : 0.f;
    q.off[3] = (s.idx + w + 1) * C;  q.wgt[3] = (x1 && y1) ? s.wx * s.wy : 0.f;
    if (!x1) { q.off[1] = q.off[0]; q.off[3] = q.off[0]; }
    if (!y1) { q.off[2] = q.off[0]; q.off[3] = q.off[0]; }
    return q;
}

// Per-voxel corner table written by vfd_voxel_classify for the voxels that reach an MLP (camera count 1 or 2): up to two
// entries per voxel row (b*N+n), in camera order, each = the four bilinear corner offsets (floats inside the camera's
// feature map), the four corner weights and {b*num_cams + c, bits of v.z / depth_div}; meta.x = -1 marks an unused
// entry.  The fused gather + MLP kernels read it instead of re-projecting in every pass.
struct CornerTable {
    int4* off;     // [2][rows]
    float4* wgt;   // [2][rows]
    int2* meta;    // [2][rows]
};
__host__ __device__ inline size_t corner_table_bytes(size_t rows) { return rows * 2 * (16 + 16 + 8); }
__host__ __device__ inline CornerTable carve_table(void* p, size_t rows) {
    CornerTable t;
    t.off = reinterpret_cast<int4*>(p);
    t.wgt = reinterpret_cast<float4*>(t.off + 2 * rows);
    t.meta = reinterpret_cast<int2*>(t.wgt + 2 * rows);
    return t;
}

// ---- legacy-path tensor-core helpers (mma.sync m16n8k8 tf32, fp32 accumulate) for the fused 1x1 MLPs ----
// x = hi + lo (+ a remainder below 2^-20 |x|) with hi, lo both exactly representable in TF32 (13 low mantissa bits
// zero): three products hi*hi + hi*lo + lo*hi recover fp32-level accuracy ("3xTF32").  Truncation instead of
// cvt.rna.tf32 (which sm_100a emulates with ~6 instructions): hi carries 10 explicit mantissa bits of x, lo the next
// 10, and x - hi is exact in fp32.
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
    hi = __float_as_uint(x) & 0xffffe000u;
    lo = __float_as_uint(x - __uint_as_float(hi)) & 0xffffe000u;
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

}  // namespace vfd
