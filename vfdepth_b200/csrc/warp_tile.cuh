// Shared-memory tile of the fused photometric-loss kernels.  Pass 1 (warp_stats.cu) left the raw warp of every
// source image as one float4 (r, g, b, warp mask) per pixel; here a (32+2h) x (8+2h) region of one target camera is
// normalised (intensity alignment, view_rendering.py:58-59), the two neighbours of each frame are summed
// (view_rendering.py:193-198) and the eight images the loss compares are laid out as float2 pairs for the packed SSIM
// of warp_ssim.cuh:  img[channel][pair][pixel], pairs as listed there.
#pragma once
#include "warp_ssim.cuh"

namespace vfd {

// 32 x 8 tiles, one thread per pixel.  (30-pixel-wide tiles, which make the backward's 32-wide ring exactly one pass of the
// CTA through the SSIM evaluation per strip step, were measured at 2.51 ms against 2.47 ms: with two CTAs per SM the
// barrier wait of one is filled by the other.)
constexpr int kTileW = 32, kTileH = 8, kTileThreads = 256, kTilePixels = kTileW * kTileH;

template <int HALO>
struct Region {
    static constexpr int RW = kTileW + 2 * HALO, RH = kTileH + 2 * HALO, R = RW * RH;
    static constexpr int kImg = 0;                                   // float2 [3][kPairs][R]
    static constexpr int kOmask = kImg + 2 * 3 * kPairs * R;         // float  [1+VFD_WARP_SRC][R] summed neighbour masks
    static constexpr int kCmask = kOmask + (1 + VFD_WARP_SRC) * R;   // float  [VFD_WARP_SRC][R]   temporal warp masks
    static constexpr int kFloats = kCmask + VFD_WARP_SRC * R;
};

// normalised prediction of warp j from its raw sample (view_rendering.py:58-59), in the reference's operation order
//   ((w - w_mean) / (w_std + 1e-8)) * s_std + s_mean, then * mask
// prm = {w_mean, w_std+1e-8, s_std, s_mean, apply, RN(1 / (w_std+1e-8))}.  The quotient is formed as q = t*r;
// q += fma(-den, q, t) * r (Markstein), which is the correctly rounded t / den without a division instruction.
// apply == 0 (alignment off, or the batch-wide early exit of :50-53): the raw, un-masked warp.
__device__ __forceinline__ float warp_normalise(const float* prm, float w, float m) {
    if (prm[4] == 0.f) return w;
    const float t = VFD_SUB(w, prm[0]);
    float q = VFD_MUL(t, prm[5]);
    q = VFD_FMA(VFD_FMA(-prm[1], q, t), prm[5], q);
    return VFD_MUL(VFD_ADD(VFD_MUL(q, prm[2]), prm[3]), m);
}
// d pred / d w
__device__ __forceinline__ float warp_normalise_grad(const float* prm, float m) {
    return prm[4] == 0.f ? 1.f : m * prm[2] * prm[5];
}

// where warp j's normalised image and mask go in the tile: float offset of its (pair, component) inside one
// channel's block of pairs, and its mask plane
__device__ __forceinline__ void warp_slot(int j, int n_src, int R, int omask0, int cmask0, int* img_off, int* msk_off) {
    if (j < n_src) {                 // temporal warp of frame 1+j -> pair 0, component j
        *img_off = j;
        *msk_off = cmask0 + j * R;
    } else {
        const int fi = (j - n_src) / VFD_WARP_NBR;          // 0: spatial -> pair 2.x; 1, 2: spatio-temporal -> pair 1.x/.y
        *img_off = fi == 0 ? 2 * R * 2 : 1 * R * 2 + (fi - 1);
        *msk_off = omask0 + fi * R;
    }
}

// Phase A.  Caller synchronises afterwards.  Pixels of the region outside the image are never read.
// NSRC (= n_src, 1 or 2) is a template parameter so that the warp -> plane mapping is static: all raw warps and the
// un-warped images of a pixel are loaded in one batch (8 x 16 B + 9 x 4 B in flight per thread), combined in registers
// in the reference's order (neighbour k = 0, then 1, view_rendering.py:193-198) and written once.
// Returns this thread's spatial-presence flags: bit 0 = some pixel it filled has a non-zero spatial / spatio-temporal
// warp (or such a warp's parameters are not finite), bit 1 = the same restricted to region rows >= kTileH (the rows a
// strip-walking caller carries over to the next tile).  A tile whose whole region has none skips the overlap images.
template <int HALO, int NSRC, int THREADS>
__device__ __forceinline__ unsigned fill_region_n(const WarpCtx& ctx, const vfd_warp_desc& d, int b, int ci, int ox, int oy,
                                                  const float4* __restrict__ raw, const uint8_t* __restrict__ vbits,
                                                  float* sm, int row0) {
    using Rg = Region<HALO>;
    constexpr int NWc = NSRC + (1 + NSRC) * VFD_WARP_NBR;
    constexpr unsigned kSpatialBits = ((1u << NWc) - 1u) & ~((1u << NSRC) - 1u);
    const int W = d.W, H = d.H, HW = H * W, cam = d.cam0 + ci;
    const float* ref = d.color[0] + (size_t)(b * d.num_cams + cam) * 3 * HW;
    const float* src0 = d.color[1] + (size_t)(b * d.num_cams + cam) * 3 * HW;
    const float* src1 = d.color[NSRC] + (size_t)(b * d.num_cams + cam) * 3 * HW;   // == src0 when NSRC == 1 (unused)
    float2* img = reinterpret_cast<float2*>(sm + Rg::kImg);
    // bit j of `live`: warp j exists; bit j of `always`: its parameters are not finite, so even an all-zero sample goes
    // through the arithmetic (NaN * 0 must stay NaN as in the reference)
    unsigned live = 0, always = 0;
#pragma unroll
    for (int j = 0; j < NWc; ++j) {
        if (ctx.active[j] != 0) {
            live |= 1u << j;
            if (ctx.prm[j][6] == 0.f) always |= 1u << j;
        }
    }
    unsigned flags = (always & kSpatialBits) ? 3u : 0u;
    for (int r = row0 * Rg::RW + threadIdx.x; r < Rg::R; r += THREADS) {      // rows below row0 are already there
        const int x = ox + r % Rg::RW, y = oy + r / Rg::RW;
        if (x < 0 || x >= W || y < 0 || y >= H) continue;
        const int p = y * W + x;
        const unsigned vb = __ldg(vbits + p);
        const unsigned use = live & (vb | always);
        if (vb & kSpatialBits) flags |= (r >= kTileH * Rg::RW) ? 3u : 1u;
        // The temporal warps are stored at nearly every pixel: their loads do not wait for the validity byte (an entry
        // that was never written is read and dropped -- the buffer is allocated in full), so one L2 round trip covers
        // them, the byte and the un-warped images; only the sparse spatial / spatio-temporal entries depend on the byte.
        float4 w[NWc];
#pragma unroll
        for (int j = 0; j < NSRC; ++j)
            w[j] = (live & (1u << j)) ? __ldg(raw + (size_t)j * HW + p) : make_float4(0.f, 0.f, 0.f, 0.f);
        float vr[3], v0[3], v1[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            vr[c] = __ldg(ref + c * HW + p);
            v0[c] = __ldg(src0 + c * HW + p);
            v1[c] = __ldg(src1 + c * HW + p);
        }
#pragma unroll
        for (int j = 0; j < NSRC; ++j)
            if (!(vb & (1u << j))) w[j] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int j = NSRC; j < NWc; ++j)
            w[j] = (live & vb & (1u << j)) ? __ldg(raw + (size_t)j * HW + p) : make_float4(0.f, 0.f, 0.f, 0.f);
        // temporal warps -> pair 0 (.x = frame 1, .y = frame 2); their masks -> cmask planes
        float tw[2][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}}, cm[2] = {0.f, 0.f};
#pragma unroll
        for (int j = 0; j < NSRC; ++j) {
            if (use & (1u << j)) {
                tw[j][0] = VFD_ADD(0.f, warp_normalise(ctx.prm[j], w[j].x, w[j].w));
                tw[j][1] = VFD_ADD(0.f, warp_normalise(ctx.prm[j], w[j].y, w[j].w));
                tw[j][2] = VFD_ADD(0.f, warp_normalise(ctx.prm[j], w[j].z, w[j].w));
                cm[j] = VFD_ADD(0.f, w[j].w);
            }
        }
        // neighbour warps of frame fi, summed over the two neighbours: fi = 0 -> pair 2.x, fi = 1, 2 -> pair 1.x / .y
        float ov[3][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}}, om[3] = {0.f, 0.f, 0.f};
        if (use & kSpatialBits) {
#pragma unroll
            for (int fi = 0; fi <= NSRC; ++fi) {
#pragma unroll
                for (int k = 0; k < VFD_WARP_NBR; ++k) {
                    const int j = NSRC + fi * VFD_WARP_NBR + k;
                    if (use & (1u << j)) {
                        ov[fi][0] = VFD_ADD(ov[fi][0], warp_normalise(ctx.prm[j], w[j].x, w[j].w));
                        ov[fi][1] = VFD_ADD(ov[fi][1], warp_normalise(ctx.prm[j], w[j].y, w[j].w));
                        ov[fi][2] = VFD_ADD(ov[fi][2], warp_normalise(ctx.prm[j], w[j].z, w[j].w));
                        om[fi] = VFD_ADD(om[fi], w[j].w);
                    }
                }
            }
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float2* pc = img + c * kPairs * Rg::R + r;
            pc[0 * Rg::R] = make_float2(tw[0][c], tw[1][c]);
            pc[1 * Rg::R] = make_float2(ov[1][c], ov[2][c]);
            pc[2 * Rg::R] = make_float2(ov[0][c], v0[c]);
            pc[3 * Rg::R] = make_float2(v1[c], vr[c]);
        }
#pragma unroll
        for (int f = 0; f < 1 + VFD_WARP_SRC; ++f) sm[Rg::kOmask + f * Rg::R + r] = om[f];
#pragma unroll
        for (int f = 0; f < VFD_WARP_SRC; ++f) sm[Rg::kCmask + f * Rg::R + r] = cm[f];
    }
    return flags;
}

template <int HALO, int THREADS = kTileThreads>
__device__ __forceinline__ unsigned fill_region(const WarpCtx& ctx, const vfd_warp_desc& d, int b, int ci, int ox, int oy,
                                                const float4* __restrict__ raw, const uint8_t* __restrict__ vbits,
                                                float* sm, int row0 = 0) {
    if (d.n_src == 2) return fill_region_n<HALO, 2, THREADS>(ctx, d, b, ci, ox, oy, raw, vbits, sm, row0);
    return fill_region_n<HALO, 1, THREADS>(ctx, d, b, ci, ox, oy, raw, vbits, sm, row0);
}

// CTA-wide bitwise OR of every thread's `flags` (one shared word, zeroed by the caller before the previous barrier)
__device__ __forceinline__ unsigned cta_or(unsigned flags, unsigned* s_word) {
    const unsigned w = __reduce_or_sync(0xffffffffu, flags);
    if ((threadIdx.x & 31) == 0 && w) atomicOr(s_word, w);
    __syncthreads();
    return *s_word;
}

// Everything the loss decides at one pixel (single_cam_loss.py:17-55, multi_cam_loss.py:16-59).
struct PixelLoss {
    float reproj, automask;          // min_f L(color_f), (reproj <= identity) * ref_mask
    float spatial, sp_mask;          // L(overlap_0), ref_mask * overlap_mask_0
    float st, st_mask;               // min_f L(overlap_f), max_f(ref_mask * overlap_mask_f * automask)
    int sel_reproj, sel_st;          // which frame won each min (gradient routing)
};

// g (GRAD): SSIM derivative coefficients of pairs 0..2 (see warp_ssim.cuh).  SP = false (no spatial / spatio-temporal
// warp anywhere in the tile's region): the overlap images and masks are all zero, their terms are exactly zero.
template <int HALO, bool GRAD, bool SP = true>
__device__ __forceinline__ PixelLoss pixel_loss(const float* sm, const vfd_warp_desc& d, int x, int y, int ox, int oy,
                                                float rmask, const float* noise /*[VFD_WARP_SRC]*/, SsimGrad2* g) {
    using Rg = Region<HALO>;
    const bool two = d.n_src > 1;
    int idx[9];
    tap_indices(x, y, d.W, d.H, ox, oy, Rg::RW, idx);
    const int ctr = (y - oy) * Rg::RW + (x - ox);
    float2 L[kPairs];
    photometric8<GRAD, SP>(reinterpret_cast<const float2*>(sm + Rg::kImg), Rg::R, idx, L, g);
    PixelLoss o;
    // temporal reprojection: torch.min over frames keeps the first minimum on a tie
    o.sel_reproj = (two && L[0].y < L[0].x) ? 1 : 0;
    o.reproj = o.sel_reproj ? L[0].y : L[0].x;
    if (two && L[0].y != L[0].y) o.reproj = L[0].y;          // torch.min returns NaN if either frame's loss is NaN
    // identity reprojection + tie-break noise (single_cam_loss.py:44-47)
    float ident = VFD_ADD(L[2].y, VFD_MUL(0.00001f, noise[0]));
    if (two) ident = nan_min(ident, VFD_ADD(L[3].x, VFD_MUL(0.00001f, noise[1])));
    // argmin over cat([reproj, identity]) == 0: the first minimum wins a tie (loss_util.py:15-17)
    o.automask = VFD_MUL((o.reproj <= ident) ? 1.f : 0.f, rmask);
    if (!SP) {
        o.sp_mask = o.st_mask = 0.f;
        o.spatial = o.st = 0.f;
        o.sel_st = 0;
        return o;
    }
    o.sp_mask = VFD_MUL(rmask, sm[Rg::kOmask + ctr]);
    o.spatial = L[2].x;
    o.sel_st = (two && L[1].y < L[1].x) ? 1 : 0;
    o.st = o.sel_st ? L[1].y : L[1].x;
    if (two && L[1].y != L[1].y) o.st = L[1].y;
    o.st_mask = VFD_MUL(VFD_MUL(rmask, sm[Rg::kOmask + Rg::R + ctr]), o.automask);
    if (two) o.st_mask = nan_max(o.st_mask, VFD_MUL(VFD_MUL(rmask, sm[Rg::kOmask + 2 * Rg::R + ctr]), o.automask));
    return o;
}

}  // namespace vfd
