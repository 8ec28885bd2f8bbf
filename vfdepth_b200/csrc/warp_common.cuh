// View-synthesis warp shared by the statistics, loss-forward and loss-backward kernels
// (models/geometry/geometry_util.py:33-82 + models/geometry/view_rendering.py:61-82).
//
// One CTA works on one target camera of one rig.  Its NW warps (temporal + spatial + spatio-temporal) share the
// back-projected 3-D point of a pixel; each warp j then has its own pre-multiplied projection P_j = (K_src @ T_j)[:3]
// (built with the reference's no-FMA 4x4 product, vfd_coords.h), source image and source self-occlusion mask.
#pragma once
#include "common.cuh"

namespace vfd {

constexpr int kMaxW = VFD_WARP_MAXW;

struct WarpCtx {
    float P[kMaxW][12];       // (K_src @ T_j)[:3,:]
    float prm[kMaxW][8];      // intensity-alignment parameters of vfd_warp_stats (may be unused)
    const float* img[kMaxW];  // source image planes [3,H,W] of (b, src cam, frame)
    const float* msk[kMaxW];  // source mask [H,W]
    int active[kMaxW];
    float invK[12];
    int NW;
};

__device__ __forceinline__ int warp_count(const vfd_warp_desc& d) { return d.n_src + (1 + d.n_src) * VFD_WARP_NBR; }

// frame index into desc.color[] and source camera of warp j of target camera `cam` (-1: warp does not exist)
__device__ __forceinline__ void warp_source(const vfd_warp_desc& d, int cam, int j, int* fi, int* src) {
    if (j < d.n_src) {
        *fi = 1 + j;
        *src = cam;
    } else {
        const int jj = j - d.n_src;
        *fi = jj / VFD_WARP_NBR;
        const int s = d.nbr[cam][jj % VFD_WARP_NBR];
        *src = (s >= 0 && s < d.num_cams) ? s : -1;
    }
}

// filled by the first NW threads of the CTA; caller synchronises
__device__ __forceinline__ void warp_ctx_init(WarpCtx& c, const vfd_warp_desc& d, const float* T, const float* params,
                                              int b, int ci) {
    const int cam = d.cam0 + ci;
    const int NW = warp_count(d);
    const size_t HW = (size_t)d.H * d.W;
    if (threadIdx.x == 0) c.NW = NW;
    if (threadIdx.x < 12) c.invK[threadIdx.x] = __ldg(d.invK + (size_t)(b * d.num_cams + cam) * 16 + threadIdx.x);
    if ((int)threadIdx.x < NW) {
        const int j = threadIdx.x;
        int fi, src;
        warp_source(d, cam, j, &fi, &src);
        c.active[j] = src >= 0;
        const int s = src >= 0 ? src : cam;
        c.img[j] = d.color[fi] + (size_t)(b * d.num_cams + s) * 3 * HW;
        c.msk[j] = d.mask + (size_t)(b * d.num_cams + s) * HW;
        float K[16], Tm[16];
        const float* Kp = d.K + (size_t)(b * d.num_cams + s) * 16;
        const float* Tp = T + ((size_t)(b * d.n_call + ci) * NW + j) * 16;
#pragma unroll
        for (int i = 0; i < 16; ++i) { K[i] = __ldg(Kp + i); Tm[i] = __ldg(Tp + i); }
        float P[12];
        vfd_premul_KT(K, Tm, P);
#pragma unroll
        for (int i = 0; i < 12; ++i) c.P[j][i] = P[i];
        if (params) {
            const float* pp = params + ((size_t)(b * d.n_call + ci) * NW + j) * 8;
#pragma unroll
            for (int i = 0; i < 8; ++i) c.prm[j][i] = __ldg(pp + i);
        }
    }
}

// Pass 1 -> pass 2 scratch ("raw"): one float4 (r, g, b, warp mask) per pixel and warp, [B,n_call,NW,H*W], followed by
// one validity byte per pixel, [B,n_call,H*W]: bit j set <=> warp j's float4 at this pixel is not all-zero.  Only those
// float4 are written and read back -- a spatial / spatio-temporal warp leaves the neighbour camera's image for most
// pixels of the target (zeros padding, mask 0), so ~2/3 of the buffer is never touched.
__host__ __device__ inline size_t warp_raw_f4(const vfd_warp_desc& d) {
    return (size_t)d.B * d.n_call * (d.n_src + (1 + d.n_src) * VFD_WARP_NBR) * d.H * d.W;
}
__host__ __device__ inline size_t warp_raw_bytes(const vfd_warp_desc& d) {
    return warp_raw_f4(d) * sizeof(float4) + (size_t)d.B * d.n_call * d.H * d.W;
}

struct WarpSample {
    float v[3];   // bilinear sample of the source image (NaN -> 2.0, view_rendering.py:73-74)
    float m;      // warped mask: in-bounds(strict) * nearest(src mask)   (:70, :77-82)
    float gx, gy; // normalised sampling coordinate (bit-exact with the reference)
    // derivative information (GRAD only)
    float dvx[3], dvy[3];  // d v / d ix, d v / d iy   (ATen grid_sampler_2d_backward)
    float cam[3];          // P @ [X;1]
    float den;             // cam.z + 1e-7
    bool nan;              // the sample was replaced by 2.0 or lies outside the image: no gradient flows
};

template <bool GRAD>
__device__ __forceinline__ WarpSample warp_sample(const float* P, const float* __restrict__ img,
                                                  const float* __restrict__ msk, const float* X, int W, int H) {
    WarpSample s;
    float cam[3], g[2];
    vfd_reproject(P, X, W, H, cam, g);
    s.gx = g[0];
    s.gy = g[1];
    const float ix = vfd_unnormalize(g[0], W), iy = vfd_unnormalize(g[1], H);
    if (GRAD) {
        s.cam[0] = cam[0]; s.cam[1] = cam[1]; s.cam[2] = cam[2];
        s.den = VFD_ADD(cam[2], 1e-7f);
        s.nan = false;
#pragma unroll
        for (int c = 0; c < 3; ++c) s.dvx[c] = s.dvy[c] = 0.f;
    }
    const bool inb = vfd_in_bounds2(g[0], g[1]);
    if (isnan(ix) || isnan(iy)) {
        s.v[0] = s.v[1] = s.v[2] = 2.0f;
        s.m = 0.f;   // nearest sample of a NaN coordinate is 0 (:75-76)
        if (GRAD) s.nan = true;
        return s;
    }
    const size_t HW = (size_t)H * W;
    // nearest (half-to-even) self-occlusion mask of the source camera
    float mval = 0.f;
    {
        const float rx = rintf(ix), ry = rintf(iy);
        if (rx >= 0.f && rx <= (float)(W - 1) && ry >= 0.f && ry <= (float)(H - 1))
            mval = __ldg(msk + (size_t)(int)ry * W + (int)rx);
    }
    s.m = inb ? mval : 0.f;
    // bilinear, zeros padding, align_corners=True; weights as ATen forms them
    const float x0f = floorf(ix), y0f = floorf(iy);
    s.v[0] = s.v[1] = s.v[2] = 0.f;
    if (!(x0f >= -1.f && x0f <= (float)(W - 1) && y0f >= -1.f && y0f <= (float)(H - 1))) {
        if (GRAD) s.nan = true;    // all four corners outside the image: zeros, and no gradient either
        return s;
    }
    const int x0 = (int)x0f, y0 = (int)y0f;
    const float tx1 = (x0f + 1.f) - ix, tx0 = ix - x0f, ty1 = (y0f + 1.f) - iy, ty0 = iy - y0f;
    const bool vx0 = x0 >= 0, vx1 = x0 + 1 <= W - 1, vy0 = y0 >= 0, vy1 = y0 + 1 <= H - 1;
    const float wnw = tx1 * ty1, wne = tx0 * ty1, wsw = tx1 * ty0, wse = tx0 * ty0;
    const float* p = img + (ptrdiff_t)y0 * W + x0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float* pc = p + c * HW;
        const float nw = (vx0 && vy0) ? __ldg(pc) : 0.f;
        const float ne = (vx1 && vy0) ? __ldg(pc + 1) : 0.f;
        const float sw = (vx0 && vy1) ? __ldg(pc + W) : 0.f;
        const float se = (vx1 && vy1) ? __ldg(pc + W + 1) : 0.f;
        float val = nw * wnw;
        val = fmaf(ne, wne, val);
        val = fmaf(sw, wsw, val);
        val = fmaf(se, wse, val);
        if (isnan(val)) {
            val = 2.0f;
            if (GRAD) s.nan = true;
        }
        s.v[c] = val;
        if (GRAD) {
            s.dvx[c] = (ne - nw) * ty1 + (se - sw) * ty0;
            s.dvy[c] = (sw - nw) * tx1 + (se - ne) * tx0;
        }
    }
    return s;
}

}  // namespace vfd
