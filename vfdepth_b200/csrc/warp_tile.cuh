// Shared-memory tile of the fused warp + photometric-loss kernels: every pixel of a (32+2h) x (8+2h) region of one
// target camera is warped NW times ONCE, normalised, and the per-frame predictions the loss consumes are kept as
// channel planes:
//   pred[k]   k < VFD_WARP_SRC: temporal warp of frame 1+k (('color', f, 0));  k = VFD_WARP_SRC + fi: sum over the
//             neighbours of the warps of frame fi (('overlap', f, 0), view_rendering.py:193-198)
//   tgt       the target image, idn[k] the un-warped source frames of the same camera (identity reprojection,
//             single_cam_loss.py:33-43), omask[fi] the summed neighbour masks, cmask[k] the temporal warp masks.
#pragma once
#include "warp_ssim.cuh"

namespace vfd {

constexpr int kTileW = 32, kTileH = 8, kTileThreads = kTileW * kTileH;
constexpr int kNPred = VFD_WARP_SRC + 1 + VFD_WARP_SRC;   // 5

template <int HALO>
struct Region {
    static constexpr int RW = kTileW + 2 * HALO, RH = kTileH + 2 * HALO, R = RW * RH;
    static constexpr int kPred = 0;                          // [kNPred][3][R]
    static constexpr int kTgt = kPred + kNPred * 3 * R;      // [3][R]
    static constexpr int kIdn = kTgt + 3 * R;                // [VFD_WARP_SRC][3][R]
    static constexpr int kOmask = kIdn + VFD_WARP_SRC * 3 * R;   // [1+VFD_WARP_SRC][R]
    static constexpr int kCmask = kOmask + (1 + VFD_WARP_SRC) * R;   // [VFD_WARP_SRC][R]
    static constexpr int kFloats = kCmask + VFD_WARP_SRC * R;
};

// Phase A: warp every in-image pixel of the region and fill the planes.  Caller synchronises afterwards.
template <int HALO>
__device__ __forceinline__ void fill_region(const WarpCtx& ctx, const vfd_warp_desc& d, int b, int ci, int ox, int oy,
                                            const float* __restrict__ depth, float* sm) {
    using Rg = Region<HALO>;
    const int W = d.W, H = d.H, HW = H * W, cam = d.cam0 + ci, NW = ctx.NW, n_src = d.n_src;
    const float* ref = d.color[0] + (size_t)(b * d.num_cams + cam) * 3 * HW;
    for (int r = threadIdx.x; r < Rg::R; r += kTileThreads) {
        const int x = ox + r % Rg::RW, y = oy + r / Rg::RW;
        if (x < 0 || x >= W || y < 0 || y >= H) continue;
        const int p = y * W + x;
        float X[3];
        vfd_backproject(ctx.invK, (float)x, (float)y, __ldg(depth + p), X);
        float ov[1 + VFD_WARP_SRC][3], om[1 + VFD_WARP_SRC];
#pragma unroll
        for (int f = 0; f <= VFD_WARP_SRC; ++f) { ov[f][0] = ov[f][1] = ov[f][2] = 0.f; om[f] = 0.f; }
        for (int j = 0; j < NW; ++j) {
            if (!ctx.active[j]) continue;
            const WarpSample s = warp_sample<false>(ctx.P[j], ctx.img[j], ctx.msk[j], X, W, H);
            float n[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) n[c] = warp_normalise(ctx.prm[j], s.v[c], s.m);
            if (j < n_src) {
#pragma unroll
                for (int c = 0; c < 3; ++c) sm[Rg::kPred + (j * 3 + c) * Rg::R + r] = n[c];
                sm[Rg::kCmask + j * Rg::R + r] = s.m;
            } else {
                const int fi = (j - n_src) / VFD_WARP_NBR;
#pragma unroll
                for (int f = 0; f <= VFD_WARP_SRC; ++f)
                    if (f == fi) {
#pragma unroll
                        for (int c = 0; c < 3; ++c) ov[f][c] = VFD_ADD(ov[f][c], n[c]);
                        om[f] = VFD_ADD(om[f], s.m);
                    }
            }
        }
        for (int f = 0; f <= n_src; ++f) {
#pragma unroll
            for (int c = 0; c < 3; ++c) sm[Rg::kPred + ((VFD_WARP_SRC + f) * 3 + c) * Rg::R + r] = ov[f][c];
            sm[Rg::kOmask + f * Rg::R + r] = om[f];
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) sm[Rg::kTgt + c * Rg::R + r] = __ldg(ref + c * HW + p);
        for (int k = 0; k < n_src; ++k) {
            const float* src = d.color[1 + k] + (size_t)(b * d.num_cams + cam) * 3 * HW;
#pragma unroll
            for (int c = 0; c < 3; ++c) sm[Rg::kIdn + (k * 3 + c) * Rg::R + r] = __ldg(src + c * HW + p);
        }
    }
}

// Everything the loss decides at one pixel (single_cam_loss.py:17-55, multi_cam_loss.py:16-59).
struct PixelLoss {
    float reproj, automask;          // min_f L(color_f), (reproj <= identity) * ref_mask
    float spatial, sp_mask;          // L(overlap_0), ref_mask * overlap_mask_0
    float st, st_mask;               // min_f L(overlap_f), max_f(ref_mask * overlap_mask_f * automask)
    int sel_reproj, sel_st;          // which frame won each min (gradient routing)
};

template <int HALO, bool GRAD>
__device__ __forceinline__ PixelLoss pixel_loss(const float* sm, const vfd_warp_desc& d, int x, int y, int ox, int oy,
                                                float rmask, const float* noise /*[n_src]*/, SsimGrad* g /*[kNPred]*/) {
    using Rg = Region<HALO>;
    const int n_src = d.n_src;
    int idx[9];
    tap_indices(x, y, d.W, d.H, ox, oy, Rg::RW, idx);
    const int ctr = (y - oy) * Rg::RW + (x - ox);
    const float* tgt = sm + Rg::kTgt;
    const TargetStats ts = target_stats(tgt, Rg::R, idx);
    PixelLoss o;
    float ident = 0.f;
    o.reproj = 0.f; o.sel_reproj = 0; o.st = 0.f; o.sel_st = 0; o.st_mask = 0.f;
#pragma unroll
    for (int k = 0; k < VFD_WARP_SRC; ++k) {
        if (k >= n_src) break;
        const float l = photometric<GRAD>(sm + Rg::kPred + k * 3 * Rg::R, tgt, Rg::R, idx, ctr, ts, GRAD ? g + k : nullptr);
        if (k == 0 || l < o.reproj) { o.reproj = l; o.sel_reproj = k; }
        float li = photometric<false>(sm + Rg::kIdn + k * 3 * Rg::R, tgt, Rg::R, idx, ctr, ts, nullptr);
        li = VFD_ADD(li, VFD_MUL(0.00001f, noise[k]));
        if (k == 0 || li < ident) ident = li;
    }
    // argmin over cat([reproj, identity]) == 0: the first minimum wins a tie (loss_util.py:15-17)
    o.automask = VFD_MUL((o.reproj <= ident) ? 1.f : 0.f, rmask);
    o.sp_mask = VFD_MUL(rmask, sm[Rg::kOmask + ctr]);
    o.spatial = photometric<GRAD>(sm + Rg::kPred + VFD_WARP_SRC * 3 * Rg::R, tgt, Rg::R, idx, ctr, ts,
                                  GRAD ? g + VFD_WARP_SRC : nullptr);
#pragma unroll
    for (int k = 0; k < VFD_WARP_SRC; ++k) {
        if (k >= n_src) break;
        const float l = photometric<GRAD>(sm + Rg::kPred + (VFD_WARP_SRC + 1 + k) * 3 * Rg::R, tgt, Rg::R, idx, ctr, ts,
                                          GRAD ? g + VFD_WARP_SRC + 1 + k : nullptr);
        if (k == 0 || l < o.st) { o.st = l; o.sel_st = k; }
        const float m = VFD_MUL(VFD_MUL(rmask, sm[Rg::kOmask + (1 + k) * Rg::R + ctr]), o.automask);
        o.st_mask = (k == 0) ? m : fmaxf(o.st_mask, m);
    }
    return o;
}

}  // namespace vfd
