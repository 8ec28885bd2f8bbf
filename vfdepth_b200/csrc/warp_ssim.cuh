// SSIM + L1 photometric loss at one pixel from shared-memory tiles (models/losses/loss_util.py:43-78), forward value
// and the derivative coefficients the backward pass gathers.
//
// The 3x3 means are summed tap by tap in row-major order and divided by 9 -- the order ATen's avg_pool2d uses -- with
// unfused multiplies and adds, so E[x^2] - mu^2 (a difference of two nearly equal numbers) carries the same rounding
// as the reference instead of an independent 3e-8 error on a 0.25-sized quantity.
#pragma once
#include "warp_common.cuh"

namespace vfd {

constexpr float kC1 = (float)(0.01 * 0.01);   // python: 0.01**2 in double, cast to fp32 when it meets the tensor
constexpr float kC2 = (float)(0.03 * 0.03);

// reflection padding of 1 (ReflectionPad2d(1), loss_util.py:47-49)
__device__ __forceinline__ int reflect1(int i, int n) { return i < 0 ? -i : (i >= n ? 2 * (n - 1) - i : i); }

// the 9 tile-local tap indices of pixel (x, y); tile origin (ox, oy) in image coordinates, row pitch `pitch`
__device__ __forceinline__ void tap_indices(int x, int y, int W, int H, int ox, int oy, int pitch, int* idx) {
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy) {
        const int yy = reflect1(y + dy, H) - oy;
#pragma unroll
        for (int dx = -1; dx <= 1; ++dx) idx[(dy + 1) * 3 + dx + 1] = yy * pitch + reflect1(x + dx, W) - ox;
    }
}

struct TargetStats {   // per channel: mean and E[t^2] of the target window
    float mu[3], sq[3];
};

__device__ __forceinline__ TargetStats target_stats(const float* tgt, int plane, const int* idx) {
    TargetStats t;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float s = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const float v = tgt[c * plane + idx[k]];
            s = VFD_ADD(s, v);
            s2 = VFD_ADD(s2, VFD_MUL(v, v));
        }
        t.mu[c] = VFD_DIV(s, 9.f);
        t.sq[c] = VFD_DIV(s2, 9.f);
    }
    return t;
}

struct SsimGrad {   // d loss / d (window mean of x, of x^2, of x*y) per channel, before the upstream weight
    float a[3], b[3], c[3];
};

// 0.85 * mean_c clamp((1 - ssim_c) / 2, 0, 1) + 0.15 * mean_c |t_c - p_c|     (loss_util.py:70-78)
// pred / tgt: 3 channel planes of `plane` floats; idx: taps; ctr: tile-local index of the pixel itself
template <bool GRAD>
__device__ __forceinline__ float photometric(const float* pred, const float* tgt, int plane, const int* idx, int ctr,
                                             const TargetStats& ts, SsimGrad* g) {
    float ssum = 0.f, lsum = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float* pp = pred + c * plane;
        const float* tp = tgt + c * plane;
        float sx = 0.f, sxx = 0.f, sxy = 0.f;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const float xv = pp[idx[k]], yv = tp[idx[k]];
            sx = VFD_ADD(sx, xv);
            sxx = VFD_ADD(sxx, VFD_MUL(xv, xv));
            sxy = VFD_ADD(sxy, VFD_MUL(xv, yv));
        }
        const float mu_x = VFD_DIV(sx, 9.f), mu_y = ts.mu[c];
        const float musq_x = VFD_MUL(mu_x, mu_x), musq_y = VFD_MUL(mu_y, mu_y), mu_xy = VFD_MUL(mu_x, mu_y);
        const float sig_x = VFD_SUB(VFD_DIV(sxx, 9.f), musq_x);
        const float sig_y = VFD_SUB(ts.sq[c], musq_y);
        const float sig_xy = VFD_SUB(VFD_DIV(sxy, 9.f), mu_xy);
        const float n1 = VFD_ADD(VFD_MUL(2.f, mu_xy), kC1), n2 = VFD_ADD(VFD_MUL(2.f, sig_xy), kC2);
        const float d1 = VFD_ADD(VFD_ADD(musq_x, musq_y), kC1), d2 = VFD_ADD(VFD_ADD(sig_x, sig_y), kC2);
        const float den = VFD_ADD(VFD_MUL(d1, d2), 1e-8f);
        const float ssim = VFD_DIV(VFD_MUL(n1, n2), den);
        const float raw = VFD_DIV(VFD_SUB(1.f, ssim), 2.f);
        const float l = fminf(fmaxf(raw, 0.f), 1.f);
        ssum = VFD_ADD(ssum, l);
        lsum = VFD_ADD(lsum, fabsf(VFD_SUB(tp[ctr], pp[ctr])));
        if (GRAD) {
            // d l / d ssim = -1/2 inside the clamp (torch: gradient passes where 0 <= raw <= 1), then * 0.85 / 3
            const float gs = (raw >= 0.f && raw <= 1.f) ? (-0.5f * 0.85f / 3.f) : 0.f;
            const float inv = 1.f / den;
            // n2 = 2 (sxy - mu_x mu_y) + C2 ; d2 = (sxx - mu_x^2) + sig_y + C2
            const float dn_dmu = 2.f * mu_y * n2 - 2.f * mu_y * n1;
            const float dd_dmu = 2.f * mu_x * d2 - 2.f * mu_x * d1;
            g->a[c] = gs * (dn_dmu - ssim * dd_dmu) * inv;
            g->b[c] = gs * (-ssim * d1) * inv;
            g->c[c] = gs * (2.f * n1) * inv;
        }
    }
    const float s_mean = VFD_DIV(ssum, 3.f), l_mean = VFD_DIV(lsum, 3.f);
    return VFD_ADD(VFD_MUL(0.85f, s_mean), VFD_MUL(0.15f, l_mean));
}

}  // namespace vfd
