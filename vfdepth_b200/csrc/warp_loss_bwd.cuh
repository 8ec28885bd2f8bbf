// Arguments and constants shared by the loss-backward kernels (warp_loss_bwd.cu: all three loss sets in one kernel;
// warp_loss_sets.cu: the training mode, one lean kernel per loss set).
#pragma once
#include "warp_tile.cuh"

namespace vfd {

struct LossBwdArgs {
    vfd_warp_desc d;
    const float* depth;
    const float* T;
    const float* params;
    const float4* raw;
    const uint8_t* vbits;
    const float* noise;
    const float* d_num;   // [n_call,3]
    float* d_depth;
    float* d_T;
    float* partial;       // [B*n_call][tiles][kMaxW*12]
    // fused forward + backward (training): the loss sums leave this kernel too, gradients are written per loss set and
    // scaled by the upstream once the sums are known
    double* sums;         // [n_call][8]
    float* g3;            // [B,n_call,3,H,W] d loss_set / d depth with unit upstream
    // training mode, between the two set kernels (workspace behind the partials)
    uint8_t* tile_flag;   // [B*n_call][tiles]: the tile's region holds a spatial / spatio-temporal warp
    uint8_t* auto_bit;    // [B*n_call][H*W]: the auto-mask decision (reprojection <= identity) of every pixel
};

#ifndef VFD_LOSS_STRIP
#define VFD_LOSS_STRIP 8
#endif
constexpr int kStrip = VFD_LOSS_STRIP;   // tiles a CTA of the backward walks, top to bottom
// Width of a tile of the per-set training kernels (warp_loss_sets.cu).  28, not 32: the region a strip step fills is
// then 32 x 8 = one pixel per thread (at 30 and 32 a second pass of one warp followed, with seven at the barrier), the
// ring on which the SSIM runs is 30 x 8 = one ring row per warp, and the float2 tap loads of a warp stay inside one region
// row (no bank conflicts: 17 % of the shared-memory wavefronts at 32).  Phase C runs 28 of 32 lanes.  Measured at B = 8:
// 1.59 / 1.56 / 1.51 ms at 32 / 30 / 28.
#ifndef VFD_SET_TILE_W
#define VFD_SET_TILE_W 28
#endif
constexpr int kSetW = VFD_SET_TILE_W;

// tiles per image of the training-mode kernels (kSetW wide; never fewer than the 32-wide tiling of warp_loss_bwd_kernel)
__host__ __device__ inline int warp_set_tiles(const vfd_warp_desc& d) {
    return ((d.W + kSetW - 1) / kSetW) * ((d.H + kTileH - 1) / kTileH);
}
// bytes of the d_T partial sums at the head of the backward workspace (16-byte multiple)
__host__ __device__ inline size_t warp_partial_bytes(const vfd_warp_desc& d) {
    return (size_t)d.B * d.n_call * warp_set_tiles(d) * kMaxW * 12 * sizeof(float);
}
__host__ __device__ inline size_t warp_tile_flag_bytes(const vfd_warp_desc& d) {
    return ((size_t)d.B * d.n_call * warp_set_tiles(d) + 15) & ~(size_t)15;
}

// how many of pixel p's 3 horizontal (or vertical) window taps land on q, with reflection padding of 1
__device__ __forceinline__ int tap_mult(int p, int q, int n) {
    int m = 0;
#pragma unroll
    for (int dlt = -1; dlt <= 1; ++dlt) m += (reflect1(p + dlt, n) == q) ? 1 : 0;
    return m;
}

// warp sum of 12 values as a reduce-scatter butterfly (16 shuffles instead of 60, fixed order): after the steps over lane
// bits 4..1 lane l holds element l >> 1 summed over half the warp; the even lanes of the first 24 store to dst[0..12)
__device__ __forceinline__ void warp_sum12_store(const float* dP, int lane, float* dst) {
    float v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = i < 12 ? dP[i] : 0.f;
#pragma unroll
    for (int n = 8; n >= 1; n >>= 1) {
        const bool up = (lane & (2 * n)) != 0;
#pragma unroll
        for (int i = 0; i < n; ++i) {
            const float keep = up ? v[i + n] : v[i], send = up ? v[i] : v[i + n];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2 * n);
        }
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
    if (!(lane & 1) && (lane >> 1) < 12) dst[lane >> 1] = v[0];
}

// the training-mode launches (warp_loss_sets.cu); returns a VFD status
int launch_loss_sets(const LossBwdArgs& a, cudaStream_t st);

}  // namespace vfd
