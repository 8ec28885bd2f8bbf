// Pass 2 of the view-synthesis path, backward: d loss / d depth (per pixel) and d loss / d T (4x4 per warp).
// The forward saved nothing but the statistics and pass 1's raw warps: the CTA rebuilds the normalised images of a
// (32+4) x (8+4) region from them, re-evaluates the loss decisions (min over frames, auto-mask, masks) on the (32+2) x (8+2) ring around its tile, turns them into
// SSIM derivative coefficients in shared memory, gathers those over each interior pixel's 3x3 neighbourhood
// (reflection padding handled by tap multiplicity), and pushes the result through normalisation, the bilinear
// sampler and the projection.  Every d_depth pixel is written exactly once; d_T goes through per-CTA partial sums
// that a second kernel adds in a fixed order -- the whole backward is deterministic.
//   autograd of: view_rendering.py:58-82, geometry_util.py:52-77, loss_util.py:21-25,43-78,
//                single_cam_loss.py:17-55, multi_cam_loss.py:16-59
#include "warp_loss_bwd.cuh"

namespace vfd {

extern __shared__ __align__(16) float vfd_warp_smem[];

// coefficient planes over the 1-halo ring (R1 pixels): 3 sets (reproj, spatial, spatio-temporal) x 3 channels x (a,b,c),
// the set's upstream weight (for the L1 term) and the frame each min selected
// Threads per CTA.  The tile is 32 x 8 pixels (phase C: 8 warps), but a strip step fills 36 x 8 = 288 region pixels and
// evaluates 34 x 8 = 272 ring pixels: with 8 warps each of those phases is one full pass plus a second pass of ONE warp
// while seven wait at the barrier (a quarter of the kernel's stall samples).  -DVFD_LOSS_THREADS=320 takes every phase in
// one pass at 96 registers (two CTAs per SM): measured 2.382 ms against 2.389 ms at 256 threads / 128 registers -- the
// other CTA of the SM already fills the barrier wait, so the default stays 256.  What did help is occupancy: see
// warp_loss_sets.cu, which serves the training mode.
#ifndef VFD_LOSS_THREADS
#define VFD_LOSS_THREADS 256
#endif
constexpr int kBwdThreads = VFD_LOSS_THREADS;
// training mode (vfd_warp_loss_train_fwd): 1 = one lean kernel per loss set (warp_loss_sets.cu), 0 = this file's kernel
#ifndef VFD_LOSS_SPLIT
#define VFD_LOSS_SPLIT 1
#endif

// Stored as float4 planes -- (a, b, c, -) per (set, channel) and (w_reproj, w_spatial, w_st, selection bits) per pixel --
// so that one 128-bit shared access moves what three 32-bit ones did (the scalar stores / loads of these planes were a
// quarter of the kernel's stall samples).
struct Coef {
    static constexpr int RW = kTileW + 2, RH = kTileH + 2, R = RW * RH;
    static constexpr int kAbc = 0;                 // float4 [3 sets][3 ch][R]
    static constexpr int kWgt = 9;                 // float4 [R]: (wr, ws, wt, int bits: sel_reproj | sel_st << 1)
    static constexpr int kPlanes = 10;
    static constexpr int kFloats = kPlanes * 4 * R;
};

constexpr int kNPred = VFD_WARP_SRC + 1 + VFD_WARP_SRC;   // differentiable images: 2 temporal, overlap_0, 2 spatio-temporal

__device__ __forceinline__ float pick5(const float (*v)[3], int k, int c) {
    float r = v[0][c];
#pragma unroll
    for (int i = 1; i < kNPred; ++i) r = (k == i) ? v[i][c] : r;
    return r;
}

// Phase B at ring position r: loss decisions + SSIM derivative coefficients -> coefficient planes.  SP = false: the tile
// holds no spatial / spatio-temporal warp, the planes of those two sets are zero.
template <bool SP>
__device__ __forceinline__ PixelLoss ring_pixel(const float* sm, float* cf, const vfd_warp_desc& d, int x, int y, int ox,
                                                int oy, int r, float rmask, const float* nz, const float* w_set) {
    SsimGrad2 g[3];
    const PixelLoss o = pixel_loss<2, true, SP>(sm, d, x, y, ox, oy, rmask, nz, g);
    const float wr = w_set[0] * o.automask, ws = SP ? w_set[1] * o.sp_mask : 0.f, wt = SP ? w_set[2] * o.st_mask : 0.f;
    const bool s1 = o.sel_reproj == 1, t1 = o.sel_st == 1;
    float4* cf4 = reinterpret_cast<float4*>(cf);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        cf4[(Coef::kAbc + c) * Coef::R + r] = make_float4(wr * (s1 ? g[0].a[c].y : g[0].a[c].x), wr * (s1 ? g[0].b[c].y : g[0].b[c].x),
                                                          wr * (s1 ? g[0].c[c].y : g[0].c[c].x), 0.f);
        if (SP) {
            cf4[(Coef::kAbc + 3 + c) * Coef::R + r] = make_float4(ws * g[2].a[c].x, ws * g[2].b[c].x, ws * g[2].c[c].x, 0.f);
            cf4[(Coef::kAbc + 6 + c) * Coef::R + r] = make_float4(wt * (t1 ? g[1].a[c].y : g[1].a[c].x), wt * (t1 ? g[1].b[c].y : g[1].b[c].x),
                                                                  wt * (t1 ? g[1].c[c].y : g[1].c[c].x), 0.f);
        } else {      // a later tile of the strip may carry these ring rows over and read them
            cf4[(Coef::kAbc + 3 + c) * Coef::R + r] = f4_zero();
            cf4[(Coef::kAbc + 6 + c) * Coef::R + r] = f4_zero();
        }
    }
    cf4[Coef::kWgt * Coef::R + r] = make_float4(wr, ws, wt, __int_as_float((s1 ? 1 : 0) | (t1 ? 2 : 0)));
    return o;
}

// FUSED: training mode.  One launch evaluates the loss sums (what warp_loss_fwd_kernel computes) AND the gradients with
// unit upstream, per loss set (every differentiable image belongs to exactly one set: temporal warps -> reprojection,
// overlap_0 -> spatial, overlap_+-1 -> spatio-temporal); vfd_warp_loss_train_bwd scales them by d total / d num_set once
// autograd knows it.  The backward kernel recomputed the whole forward anyway, so the separate forward launch goes away.
template <bool FUSED>
__global__ void __launch_bounds__(kBwdThreads, 2) warp_loss_bwd_kernel(const LossBwdArgs a) {
    using Rg = Region<2>;
    const vfd_warp_desc& d = a.d;
    __shared__ WarpCtx ctx;
    __shared__ float s_dP[kTilePixels / 32][kMaxW * 12];     // phase C runs on the tile's 8 warps
    __shared__ double sacc[8];
    __shared__ unsigned s_flags[2];
    float* sm = vfd_warp_smem;
    float* cf = sm + Rg::kFloats;
    const int b = blockIdx.z / d.n_call, ci = blockIdx.z % d.n_call, cam = d.cam0 + ci;
    const int W = d.W, H = d.H, HW = H * W, n_src = d.n_src;
    // a CTA walks a vertical strip of kStrip tiles: the bottom 4 region rows / 2 ring rows of a tile are the top rows of
    // the next one, so they are moved up inside shared memory instead of being refilled / recomputed, and the per-CTA
    // prologue (warp context) is paid once per strip
    const int tiles_y = (H + kTileH - 1) / kTileH;
    const int ty_first = blockIdx.y * kStrip, ty_last = min(tiles_y, ty_first + kStrip);
    const int strip_y0 = ty_first * kTileH, strip_y1 = ty_last * kTileH;
    const int tx0 = blockIdx.x * kTileW, ox = tx0 - 2;
    warp_ctx_init(ctx, d, a.T, a.params, b, ci);
    if (FUSED && threadIdx.x < 8) sacc[threadIdx.x] = 0.0;
    if (threadIdx.x < 2) s_flags[threadIdx.x] = 0u;
    __syncthreads();
    unsigned carried = 0u;     // spatial presence in the region rows carried over from the previous tile of the strip
    const float* depth = a.depth + (size_t)(b * d.n_call + ci) * HW;
    const float* rmaskp = d.mask + (size_t)(b * d.num_cams + cam) * HW;
    float w_set[3] = {1.f, 1.f, 1.f};
    if (!FUSED) {
        w_set[0] = __ldg(a.d_num + ci * 3 + 0); w_set[1] = __ldg(a.d_num + ci * 3 + 1); w_set[2] = __ldg(a.d_num + ci * 3 + 2);
    }
    float facc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int tyi = ty_first; tyi < ty_last; ++tyi) {
    const int ty0 = tyi * kTileH, oy = ty0 - 2;
    const bool cont = tyi > ty_first;
    if (cont) {
        // rows 8..11 of the region -> rows 0..3, ring rows 8, 9 -> 0, 1 (sources and destinations are disjoint)
        float2* img2 = reinterpret_cast<float2*>(sm + Rg::kImg);
        for (int i = threadIdx.x; i < 3 * kPairs * 4 * Rg::RW; i += kBwdThreads) {
            const int pl = i / (4 * Rg::RW), o = i - pl * (4 * Rg::RW);
            img2[pl * Rg::R + o] = img2[pl * Rg::R + kTileH * Rg::RW + o];
        }
        float* mk = sm + Rg::kOmask;     // the 1 + VFD_WARP_SRC overlap-mask planes and VFD_WARP_SRC colour-mask planes
        for (int i = threadIdx.x; i < (1 + 2 * VFD_WARP_SRC) * 4 * Rg::RW; i += kBwdThreads) {
            const int pl = i / (4 * Rg::RW), o = i - pl * (4 * Rg::RW);
            mk[pl * Rg::R + o] = mk[pl * Rg::R + kTileH * Rg::RW + o];
        }
        float4* cfm = reinterpret_cast<float4*>(cf);
        for (int i = threadIdx.x; i < Coef::kPlanes * 2 * Coef::RW; i += kBwdThreads) {
            const int pl = i / (2 * Coef::RW), o = i - pl * (2 * Coef::RW);
            cfm[pl * Coef::R + o] = cfm[pl * Coef::R + kTileH * Coef::RW + o];
        }
        __syncthreads();      // the fill below overwrites the rows that were just read
    }
    for (int i = threadIdx.x; i < (kTilePixels / 32) * kMaxW * 12; i += kBwdThreads) (&s_dP[0][0])[i] = 0.f;
    const unsigned fl = fill_region<2, kBwdThreads>(ctx, d, b, ci, ox, oy, a.raw + (size_t)(b * d.n_call + ci) * ctx.NW * HW,
                                       a.vbits + (size_t)(b * d.n_call + ci) * HW, sm, cont ? 4 : 0);
    // any spatial / spatio-temporal warp in this tile's region?  (flag words alternate between tiles; the other one is
    // cleared for the next tile now that every thread is past the previous tile's read of it)
    const unsigned any = cta_or(fl, &s_flags[tyi & 1]);
    if (threadIdx.x == 0) s_flags[(tyi + 1) & 1] = 0u;
    const bool sp = ((any & 1u) | (cont ? carried : 0u)) != 0u;
    carried = (any >> 1) & 1u;

    // ---- phase B: loss decisions + SSIM derivative coefficients on the ring ----
    for (int r = (cont ? 2 * Coef::RW : 0) + threadIdx.x; r < Coef::R; r += kBwdThreads) {
        const int x = tx0 - 1 + r % Coef::RW, y = ty0 - 1 + r / Coef::RW;
        if (x < 0 || x >= W || y < 0 || y >= H) continue;
        const int p = y * W + x;
        float nz[VFD_WARP_SRC] = {0.f, 0.f};
        {   // static indices: a run-time index would put nz[] in local memory
            const float* np = a.noise + (size_t)(b * d.n_call + ci) * n_src * HW + p;
            nz[0] = __ldg(np);
            if (n_src > 1) nz[1] = __ldg(np + HW);
        }
        const float rm = __ldg(rmaskp + p);
        const PixelLoss o = sp ? ring_pixel<true>(sm, cf, d, x, y, ox, oy, r, rm, nz, w_set)
                               : ring_pixel<false>(sm, cf, d, x, y, ox, oy, r, rm, nz, w_set);
        // every ring row is evaluated once per strip: its loss counts here if the row belongs to the strip
        if (FUSED && x >= tx0 && x < tx0 + kTileW && y >= strip_y0 && y < strip_y1) {
            facc[0] += VFD_MUL(o.reproj, o.automask); facc[1] += o.automask;
            facc[2] += VFD_MUL(o.spatial, o.sp_mask); facc[3] += o.sp_mask;
            facc[4] += VFD_MUL(o.st, o.st_mask);      facc[5] += o.st_mask;
        }
    }
    __syncthreads();

    // ---- phase C: gather the coefficients at the interior pixel, then back through normalise / sampler / projection ----
    const int x = tx0 + (threadIdx.x % kTileW), y = ty0 + (threadIdx.x / kTileW);
    const bool live = threadIdx.x < kTilePixels && x < W && y < H;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    float dpred[kNPred][3];
#pragma unroll
    for (int k = 0; k < kNPred; ++k) dpred[k][0] = dpred[k][1] = dpred[k][2] = 0.f;
    float X[3] = {0.f, 0.f, 0.f}, ray[3] = {0.f, 0.f, 0.f};
    if (live) {
        float SA[kNPred][3], SB[kNPred][3], SC[kNPred][3];
#pragma unroll
        for (int k = 0; k < kNPred; ++k)
#pragma unroll
            for (int c = 0; c < 3; ++c) SA[k][c] = SB[k][c] = SC[k][c] = 0.f;
        // how often the window of neighbour (x + dx, y + dy) contains this pixel: 1 away from the image border (reflection
        // padding only folds taps within two pixels of it), 0 for a neighbour outside the image
        float mx[3] = {1.f, 1.f, 1.f}, my[3] = {1.f, 1.f, 1.f};
        if (x < 2 || x >= W - 2) {
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx) mx[dx + 1] = (x + dx < 0 || x + dx >= W) ? 0.f : (float)tap_mult(x + dx, x, W);
        }
        if (y < 2 || y >= H - 2) {
#pragma unroll
            for (int dy = -1; dy <= 1; ++dy) my[dy + 1] = (y + dy < 0 || y + dy >= H) ? 0.f : (float)tap_mult(y + dy, y, H);
        }
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy) {
            const int py = y + dy;
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx) {
                const int px = x + dx;
                const float mult = my[dy + 1] * mx[dx + 1];
                if (mult == 0.f) continue;
                const int r = (py - (ty0 - 1)) * Coef::RW + (px - (tx0 - 1));
                const float4* cf4 = reinterpret_cast<const float4*>(cf);
                const int sel = __float_as_int(cf4[Coef::kWgt * Coef::R + r].w);
                // the min's selection as tap weights: w[k] = mult for the selected frame, 0 for the other (mult is 0, 1, 2
                // or 4, so w * q is exact and the fused form below rounds like the separate multiply and add it replaces;
                // a non-finite coefficient now also reaches the unselected frame's sum, as NaN)
                const float wr0 = (sel & 1) ? 0.f : mult, wr1 = mult - wr0;
                const float wt0 = (sel >> 1) ? 0.f : mult, wt1 = mult - wt0;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float4 q0 = cf4[(Coef::kAbc + c) * Coef::R + r];
                    SA[0][c] = fmaf(q0.x, wr0, SA[0][c]); SB[0][c] = fmaf(q0.y, wr0, SB[0][c]); SC[0][c] = fmaf(q0.z, wr0, SC[0][c]);
                    if (VFD_WARP_SRC > 1) {
                        SA[1][c] = fmaf(q0.x, wr1, SA[1][c]); SB[1][c] = fmaf(q0.y, wr1, SB[1][c]); SC[1][c] = fmaf(q0.z, wr1, SC[1][c]);
                    }
                    if (sp) {      // CTA-uniform: the spatial / spatio-temporal planes are all zero otherwise
                        const float4 q1 = cf4[(Coef::kAbc + 3 + c) * Coef::R + r], q2 = cf4[(Coef::kAbc + 6 + c) * Coef::R + r];
                        constexpr int S = VFD_WARP_SRC;
                        SA[S][c] = fmaf(q1.x, mult, SA[S][c]); SB[S][c] = fmaf(q1.y, mult, SB[S][c]); SC[S][c] = fmaf(q1.z, mult, SC[S][c]);
                        SA[S + 1][c] = fmaf(q2.x, wt0, SA[S + 1][c]); SB[S + 1][c] = fmaf(q2.y, wt0, SB[S + 1][c]);
                        SC[S + 1][c] = fmaf(q2.z, wt0, SC[S + 1][c]);
                        if (VFD_WARP_SRC > 1) {
                            SA[S + 2][c] = fmaf(q2.x, wt1, SA[S + 2][c]); SB[S + 2][c] = fmaf(q2.y, wt1, SB[S + 2][c]);
                            SC[S + 2][c] = fmaf(q2.z, wt1, SC[S + 2][c]);
                        }
                    }
                }
            }
        }
        const int ctr2 = (y - oy) * Rg::RW + (x - ox);
        const int r0 = (y - (ty0 - 1)) * Coef::RW + (x - (tx0 - 1));
        const float4 w4 = reinterpret_cast<const float4*>(cf)[Coef::kWgt * Coef::R + r0];
        const int sel0 = __float_as_int(w4.w);
        const float wl[3] = {w4.x, w4.y, w4.z};
#pragma unroll
        for (int k = 0; k < kNPred; ++k) {
            if (!sp && k >= VFD_WARP_SRC) continue;     // dpred stays 0
            // which set's L1 term reaches pred k at this pixel
            float w1;
            if (k < VFD_WARP_SRC) w1 = ((sel0 & 1) == k) ? wl[0] : 0.f;
            else if (k == VFD_WARP_SRC) w1 = wl[1];
            else w1 = ((sel0 >> 1) == k - VFD_WARP_SRC - 1) ? wl[2] : 0.f;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                // image k of the tile: 0,1 = pair 0 (.x,.y); 2 = pair 2.x; 3,4 = pair 1 (.x,.y); target = pair 3.y
                const float2* ic = reinterpret_cast<const float2*>(sm + Rg::kImg) + c * kPairs * Rg::R + ctr2;
                const float2 pr = ic[(k < VFD_WARP_SRC ? 0 : (k == VFD_WARP_SRC ? 2 : 1)) * Rg::R];
                const float xv = (k == 1 || k == VFD_WARP_SRC + 2) ? pr.y : pr.x, yv = ic[3 * Rg::R].y;
                const float df = xv - yv;
                const float sgn = df > 0.f ? 1.f : (df < 0.f ? -1.f : 0.f);
                dpred[k][c] = (SA[k][c] + 2.f * xv * SB[k][c] + yv * SC[k][c]) * (1.f / 9.f) + (0.15f / 3.f) * w1 * sgn;
            }
        }
        vfd_pixel_ray(ctx.invK, (float)x, (float)y, ray);
        const float dep = __ldg(depth + y * W + x);
        X[0] = VFD_MUL(dep, ray[0]); X[1] = VFD_MUL(dep, ray[1]); X[2] = VFD_MUL(dep, ray[2]);
    }
    float gdepth = 0.f, gset[3] = {0.f, 0.f, 0.f};
    for (int j = 0; j < ctx.NW; ++j) {
        if (!ctx.active[j] || (!sp && j >= n_src)) continue;      // CTA-uniform
        float dP[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) dP[i] = 0.f;
        // Most (pixel, warp) pairs carry no gradient: the min picked the other frame in the whole 3x3 neighbourhood, or
        // the pixel lies outside the spatial overlap.  Those skip the sampler; a warp whose 32 pixels all skip also
        // skips the reduction (its slots of s_dP stay zero, so the sums are unchanged).
        bool contrib = false;
        if (live) {
            const int k = j < n_src ? j : VFD_WARP_SRC + (j - n_src) / VFD_WARP_NBR;
            const float up[3] = {pick5(dpred, k, 0), pick5(dpred, k, 1), pick5(dpred, k, 2)};
            if (up[0] != 0.f || up[1] != 0.f || up[2] != 0.f) {
                const WarpSample s = warp_sample<true>(ctx.P[j], ctx.img[j], ctx.msk[j], X, W, H);
                if (!s.nan) {
                    contrib = true;
                    const float ng = warp_normalise_grad(ctx.prm[j], s.m);
                    float gix = 0.f, giy = 0.f;
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const float dw = up[c] * ng;
                        gix = fmaf(dw, s.dvx[c], gix);
                        giy = fmaf(dw, s.dvy[c], giy);
                    }
                    // ix == cam.x / den, iy == cam.y / den  (normalise then un-normalise cancel exactly in the derivative)
                    const float inv = 1.f / s.den;
                    const float dc0 = gix * inv, dc1 = giy * inv;
                    const float dc2 = -(gix * s.cam[0] + giy * s.cam[1]) * inv * inv;
                    const float* P = ctx.P[j];
                    // d(ix)/d(depth) = (a0 * den - cam.x * a2) / den^2 with a = P[:, :3] @ ray, cam = depth * a + P[:, 3].
                    // The depth-proportional parts cancel EXACTLY in that numerator (a pure rotation moves no pixel when
                    // depth changes); written as below the fp32 result keeps ~7 digits even when depth*|a| >> |t|, where
                    // the chain-rule form loses 2-3 of them.
                    const float a0 = P[0] * ray[0] + P[1] * ray[1] + P[2] * ray[2];
                    const float a1 = P[4] * ray[0] + P[5] * ray[1] + P[6] * ray[2];
                    const float a2 = P[8] * ray[0] + P[9] * ray[1] + P[10] * ray[2];
                    const float t2 = P[11] + 1e-7f;
                    const float gj = (gix * (a0 * t2 - P[3] * a2) + giy * (a1 * t2 - P[7] * a2)) * inv * inv;
                    if (FUSED) {
                        if (k < VFD_WARP_SRC) gset[0] += gj;
                        else if (k == VFD_WARP_SRC) gset[1] += gj;
                        else gset[2] += gj;
                    } else {
                        gdepth += gj;
                    }
                    const float dc[3] = {dc0, dc1, dc2};
#pragma unroll
                    for (int rr = 0; rr < 3; ++rr) {
                        dP[rr * 4 + 0] = dc[rr] * X[0];
                        dP[rr * 4 + 1] = dc[rr] * X[1];
                        dP[rr * 4 + 2] = dc[rr] * X[2];
                        dP[rr * 4 + 3] = dc[rr];
                    }
                }
            }
        }
        if (!__any_sync(0xffffffffu, contrib)) continue;
        warp_sum12_store(dP, lane, &s_dP[wid][j * 12]);
    }
    if (live) {
        if (FUSED) {
            float* g3 = a.g3 + (size_t)(b * d.n_call + ci) * 3 * HW + y * W + x;
            g3[0] = gset[0]; g3[HW] = gset[1]; g3[2 * HW] = gset[2];
        } else {
            a.d_depth[(size_t)(b * d.n_call + ci) * HW + y * W + x] = gdepth;
        }
    }
    __syncthreads();
    const int ntile = gridDim.x * tiles_y, tile = tyi * gridDim.x + blockIdx.x;
    float* part = a.partial + ((size_t)blockIdx.z * ntile + tile) * (kMaxW * 12);
    for (int i = threadIdx.x; i < ctx.NW * 12; i += kBwdThreads) {
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < kTilePixels / 32; ++w) v += s_dP[w][i];
        part[i] = v;
    }
    __syncthreads();      // s_dP, the region and the ring are rewritten by the next tile
    }   // tiles of the strip
    if (FUSED) {   // loss sums: fp32 inside a warp, fp64 above (as warp_loss_fwd_kernel)
#pragma unroll
        for (int q = 0; q < 6; ++q) {
            float v = facc[q];
#pragma unroll
            for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if ((threadIdx.x & 31) == 0 && v != 0.f) atomicAdd(&sacc[q], (double)v);
        }
    }
    __syncthreads();
    if (FUSED && threadIdx.x < 6 && sacc[threadIdx.x] != 0.0) atomicAdd(a.sums + ci * 8 + threadIdx.x, sacc[threadIdx.x]);
}

// d_T[b,ci,j] = K_src^T[:, :3] @ (sum over tiles of dP): kDtParts threads per element each add a contiguous range of
// tiles in index order (fp64), the ranges are then added in order -- a fixed summation tree, so d_T is deterministic.
constexpr int kDtParts = 8, kDtThreads = kMaxW * 12 * kDtParts;
__global__ void __launch_bounds__(kDtThreads) warp_dT_reduce_kernel(const LossBwdArgs a, int ntile, bool scale_sets) {
    const vfd_warp_desc& d = a.d;
    __shared__ double s_part[kDtParts][kMaxW * 12];
    __shared__ double s_sum[kMaxW * 12];
    const int bc = blockIdx.x, b = bc / d.n_call, ci = bc % d.n_call, cam = d.cam0 + ci;
    const int NW = warp_count(d);
    const float* part = a.partial + (size_t)bc * ntile * (kMaxW * 12);
    {
        const int i = threadIdx.x % (kMaxW * 12), q = threadIdx.x / (kMaxW * 12);
        const int per = (ntile + kDtParts - 1) / kDtParts, t0 = q * per, t1 = min(ntile, t0 + per);
        // four independent chains per thread (combined in a fixed order): the loop was one dependent load + add per tile,
        // i.e. pure L2 latency (0.035 ms for 48 CTAs at 8 % issue)
        double s4[4] = {0.0, 0.0, 0.0, 0.0};
        if (i < NW * 12) {
            int t = t0;
            for (; t + 4 <= t1; t += 4) {
#pragma unroll
                for (int k = 0; k < 4; ++k) s4[k] += (double)part[(size_t)(t + k) * (kMaxW * 12) + i];
            }
            for (; t < t1; ++t) s4[0] += (double)part[(size_t)t * (kMaxW * 12) + i];
        }
        s_part[q][i] = (s4[0] + s4[1]) + (s4[2] + s4[3]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < NW * 12; i += blockDim.x) {
        double s = 0.0;
#pragma unroll
        for (int q = 0; q < kDtParts; ++q) s += s_part[q][i];
        if (scale_sets) {   // partials of the fused kernel carry unit upstream: apply d total / d num of the warp's set
            const int j = i / 12;
            const int set = j < d.n_src ? 0 : ((j - d.n_src) / VFD_WARP_NBR == 0 ? 1 : 2);
            s *= (double)__ldg(a.d_num + ci * 3 + set);
        }
        s_sum[i] = s;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < NW * 16; i += blockDim.x) {
        const int j = i / 16, k = (i % 16) / 4, c = i % 4;
        int fi, src;
        warp_source(d, cam, j, &fi, &src);
        float v = 0.f;
        if (src >= 0) {
            const float* K = d.K + (size_t)(b * d.num_cams + src) * 16;
            double s = 0.0;
            for (int r = 0; r < 3; ++r) s += (double)__ldg(K + r * 4 + k) * s_sum[j * 12 + r * 4 + c];
            v = (float)s;
        }
        a.d_T[((size_t)bc * NW + j) * 16 + i % 16] = v;
    }
}

// d_depth = sum over the three loss sets of (d total / d num_set) * g3_set
__global__ void __launch_bounds__(256) warp_grad_combine_kernel(const LossBwdArgs a) {
    const vfd_warp_desc& d = a.d;
    const int bc = blockIdx.y, ci = bc % d.n_call, HW = d.H * d.W;
    const float w0 = __ldg(a.d_num + ci * 3 + 0), w1 = __ldg(a.d_num + ci * 3 + 1), w2 = __ldg(a.d_num + ci * 3 + 2);
    const float* g = a.g3 + (size_t)bc * 3 * HW;
    float* o = a.d_depth + (size_t)bc * HW;
    for (int p = blockIdx.x * 256 + threadIdx.x; p < HW; p += gridDim.x * 256)
        o[p] = w0 * __ldg(g + p) + w1 * __ldg(g + HW + p) + w2 * __ldg(g + 2 * HW + p);
}

int check_warp_desc(const vfd_warp_desc* d);   // warp_stats.cu

static size_t bwd_smem_bytes() { return (Region<2>::kFloats + Coef::kFloats) * sizeof(float); }

}  // namespace vfd

using namespace vfd;

// d_T partial sums, then (training mode) one flag byte per tile and one auto-mask decision byte per pixel
extern "C" size_t vfd_warp_bwd_workspace_bytes(const vfd_warp_desc* d) {
    if (check_warp_desc(d)) return 0;
    return warp_partial_bytes(*d) + warp_tile_flag_bytes(*d) + (size_t)d->B * d->n_call * d->H * d->W;
}

extern "C" int vfd_warp_loss_bwd(const vfd_warp_desc* d, const float* depth, const float* T, const float* params,
                                 const float* raw, const float* noise, const float* d_num, float* d_depth, float* d_T, void* ws,
                                 size_t ws_bytes, void* stream) {
    int rc = check_warp_desc(d);
    if (rc) return rc;
    if (!depth || !T || !params || !raw || !noise || !d_num || !d_depth || !d_T || !ws) return VFD_ERR_NULL;
    if (reinterpret_cast<uintptr_t>(raw) & 15) return VFD_ERR_ALIGN;
    if (ws_bytes < vfd_warp_bwd_workspace_bytes(d)) return VFD_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    LossBwdArgs a = {};
    a.d = *d; a.depth = depth; a.T = T; a.params = params; a.raw = reinterpret_cast<const float4*>(raw); a.noise = noise; a.d_num = d_num;
    a.vbits = reinterpret_cast<const uint8_t*>(a.raw + warp_raw_f4(*d));
    a.d_depth = d_depth; a.d_T = d_T; a.partial = (float*)ws;
    const size_t smem = bwd_smem_bytes();
    // > 48 KB of dynamic shared memory needs the opt-in; setting it is idempotent and per device, so it is simply
    // repeated on every call (no global state to keep in step with the current device)
    if (cudaFuncSetAttribute(warp_loss_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return VFD_ERR_CUDA;
    const int tiles_x = (d->W + kTileW - 1) / kTileW, tiles_y = (d->H + kTileH - 1) / kTileH;
    dim3 grid(tiles_x, (tiles_y + kStrip - 1) / kStrip, d->B * d->n_call);
    warp_loss_bwd_kernel<false><<<grid, kBwdThreads, smem, st>>>(a);
    VFD_CHECK_LAUNCH();
    warp_dT_reduce_kernel<<<d->B * d->n_call, kDtThreads, 0, st>>>(a, tiles_x * tiles_y, false);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_warp_loss_train_fwd(const vfd_warp_desc* d, const float* depth, const float* T, const float* params,
                                       const float* raw, const float* noise, double* sums, float* g3, void* ws,
                                       size_t ws_bytes, void* stream) {
    int rc = check_warp_desc(d);
    if (rc) return rc;
    if (!depth || !T || !params || !raw || !noise || !sums || !g3 || !ws) return VFD_ERR_NULL;
    if (reinterpret_cast<uintptr_t>(raw) & 15) return VFD_ERR_ALIGN;
    if (ws_bytes < vfd_warp_bwd_workspace_bytes(d)) return VFD_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(sums, 0, (size_t)d->n_call * 8 * sizeof(double), st) != cudaSuccess) return VFD_ERR_CUDA;
    LossBwdArgs a = {};
    a.d = *d; a.depth = depth; a.T = T; a.params = params; a.raw = reinterpret_cast<const float4*>(raw); a.noise = noise;
    a.vbits = reinterpret_cast<const uint8_t*>(a.raw + warp_raw_f4(*d));
    a.sums = sums; a.g3 = g3; a.partial = (float*)ws;
#if VFD_LOSS_SPLIT
    a.tile_flag = reinterpret_cast<uint8_t*>(ws) + warp_partial_bytes(*d);
    a.auto_bit = a.tile_flag + warp_tile_flag_bytes(*d);
    return launch_loss_sets(a, st);
#else
    const size_t smem = bwd_smem_bytes();
    if (cudaFuncSetAttribute(warp_loss_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return VFD_ERR_CUDA;
    const int tiles_x = (d->W + kTileW - 1) / kTileW, tiles_y = (d->H + kTileH - 1) / kTileH;
    dim3 grid(tiles_x, (tiles_y + kStrip - 1) / kStrip, d->B * d->n_call);
    warp_loss_bwd_kernel<true><<<grid, kBwdThreads, smem, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
#endif
}

extern "C" int vfd_warp_loss_train_bwd(const vfd_warp_desc* d, const float* d_num, const float* g3, const void* ws,
                                       size_t ws_bytes, float* d_depth, float* d_T, void* stream) {
    int rc = check_warp_desc(d);
    if (rc) return rc;
    if (!d_num || !g3 || !ws || !d_depth || !d_T) return VFD_ERR_NULL;
    if (ws_bytes < vfd_warp_bwd_workspace_bytes(d)) return VFD_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    LossBwdArgs a = {};
    a.d = *d; a.d_num = d_num; a.g3 = const_cast<float*>(g3); a.partial = (float*)const_cast<void*>(ws);
    a.d_depth = d_depth; a.d_T = d_T;
    int bx = (d->H * d->W + 1023) / 1024;
    if (bx > 256) bx = 256;
    warp_grad_combine_kernel<<<dim3(bx, d->B * d->n_call), 256, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
#if VFD_LOSS_SPLIT
    const int tiles = warp_set_tiles(*d);
#else
    const int tiles = ((d->W + kTileW - 1) / kTileW) * ((d->H + kTileH - 1) / kTileH);
#endif
    warp_dT_reduce_kernel<<<d->B * d->n_call, kDtThreads, 0, st>>>(a, tiles, true);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
