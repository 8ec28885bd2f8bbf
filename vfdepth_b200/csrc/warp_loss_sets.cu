// Training mode of pass 2 (vfd_warp_loss_train_fwd): loss sums + unit-upstream gradients, ONE LEAN KERNEL PER LOSS SET.
//
// warp_loss_bwd_kernel evaluates the three loss sets of a camera (temporal reprojection, spatial, spatio-temporal) in one
// CTA: 8 images x 3 channels per region pixel and 10 float4 coefficient planes = 104 KB of shared memory and 128
// registers, i.e. 16 warps per SM -- ncu showed it latency-bound at 42 % issue with every pipe below 30 %.  The sets only
// meet in two places: the spatio-temporal mask multiplies the auto-mask (multi_cam_loss.py:49-56) and d_depth adds them.
// So the work is split by set:
//   SET 0 (every tile)   temporal warps + identity images: reprojection min, auto-mask (single_cam_loss.py:17-55).
//                        5 images per pixel, 3 coefficient planes: 37 KB, 3 CTAs per SM at 80 registers.  Writes the
//                        auto-mask decision of every pixel (one byte) and, per tile, whether its region holds a spatial /
//                        spatio-temporal warp at all (the validity bytes of pass 1 say so).
//   SET 1 (flagged tiles only, ~40 %)  overlap images: spatial + spatio-temporal losses (multi_cam_loss.py:16-59), reading
//                        the auto-mask byte.  4 images per pixel, 6 coefficient planes: 52 KB, 3 CTAs per SM.
// Per pixel the arithmetic (SSIM tap order, decisions, coefficient products, sampler) is the one of warp_loss_bwd_kernel
// statement for statement, so g3 / partials / sums carry the same bits; what changes is occupancy and that tiles without
// spatial warps never touch the overlap code.  A CTA walks a strip of kStrip tiles (kSetW = 28 wide, warp_loss_bwd.cuh)
// top to bottom; region rows and ring rows live in CIRCULAR shared-memory buffers (12 and 10 rows), so the 4 region rows /
// 2 ring rows the next tile shares are neither moved nor recomputed.  Work a zero mask makes zero is skipped: the SSIM of
// ring pixels under a zero self-occlusion mask, and the 3x3 coefficient gather of pixels whose nine taps all carry a zero
// upstream weight (a bit per ring pixel) -- both switched off once a non-finite value enters the strip's region.
// Measured and dropped: requesting the next tile's inputs with cp.async into a per-thread staging area one tile ahead
// (the region fill and the barrier behind it are a third of the temporal kernel's stall samples): 1.539 vs 1.550 ms --
// the wait is for the slowest warp of the phase, not for the loads.
//   autograd of: view_rendering.py:58-82, geometry_util.py:52-77, loss_util.py:21-25,43-78,
//                single_cam_loss.py:17-55, multi_cam_loss.py:16-59
#include "warp_loss_bwd.cuh"

namespace vfd {

extern __shared__ __align__(16) float vfd_warp_smem[];

// Region of a set kernel (halo 2): float2 planes [3 channels][2 pairs][R], then three float planes [R].
//   SET 0: pair 0 = (color_-1, color_+1) temporal warps, pair 1 = (frame -1, frame +1) identity images, aux = target
//   SET 1: pair 0 = (overlap_-1, overlap_+1),            pair 1 = (overlap_0, target),       aux = overlap masks 0, -1, +1
struct SetRegion {
    static constexpr int RW = kSetW + 4, RH = kTileH + 4, R = RW * RH;
    static constexpr int kAux = 2 * 3 * 2 * R;
    static constexpr int kFloats = kAux + 3 * R;
};
// Coefficient planes over the 1-halo ring, float4 each:
//   SET 0: [c] = wr * (a, b, c) of the selected temporal frame; .w of plane 0 = wr, of plane 1 = selection bit
//   SET 1: [c] = ws * (a, b, c) of overlap_0, .w of plane 0 / 1 / 2 = ws / wt / selection bit; [3 + c] = wt * (a, b, c)
//          of the selected spatio-temporal frame
template <int SET>
struct SetCoef {
    static constexpr int RW = kSetW + 2, RH = kTileH + 2, R = RW * RH;
    static constexpr int kPlanes = SET == 0 ? 3 : 6;
    static constexpr int kFloats = kPlanes * 4 * R;
};

__device__ __forceinline__ int wrap_row(int v, int n) { return v >= n ? v - n : v; }

// ---- phase A ------------------------------------------------------------------------------------------------------
// Region rows [row0, RH) of the tile at (ox, oy) -> circular slots (row + base) mod RH.  live / always: bit j = warp j
// exists / its alignment parameters are not finite, so even an all-zero sample goes through the arithmetic (NaN * 0 must
// stay NaN as in the reference).  Returns the spatial-presence
// flags of fill_region_n (warp_tile.cuh): bit 0 = some pixel filled now has a spatial / spatio-temporal warp (or such a
// warp's parameters are not finite), bit 1 = the same for rows >= kTileH (the rows the next tile of the strip shares),
// bit 2 = some value written to the region is not finite.
template <int SET, int NSRC>
__device__ __forceinline__ unsigned fill_set(const WarpCtx& ctx, const vfd_warp_desc& d, int b, int ci, int ox, int oy,
                                             const float4* __restrict__ raw, const uint8_t* __restrict__ vbits, float* sm,
                                             int row0, int base, unsigned live, unsigned always) {
    using Rg = SetRegion;
    constexpr int NWc = NSRC + (1 + NSRC) * VFD_WARP_NBR;
    constexpr unsigned kSpatialBits = ((1u << NWc) - 1u) & ~((1u << NSRC) - 1u);
    const int W = d.W, H = d.H, HW = H * W, cam = d.cam0 + ci;
    const float* ref = d.color[0] + (size_t)(b * d.num_cams + cam) * 3 * HW;
    const float* src0 = d.color[1] + (size_t)(b * d.num_cams + cam) * 3 * HW;
    const float* src1 = d.color[NSRC] + (size_t)(b * d.num_cams + cam) * 3 * HW;   // == src0 when NSRC == 1 (unused)
    float2* img = reinterpret_cast<float2*>(sm);
    unsigned flags = (always & kSpatialBits) ? 3u : 0u;
    for (int r = row0 * Rg::RW + threadIdx.x; r < Rg::R; r += kTileThreads) {
        const int rr = r / Rg::RW, cx = r - rr * Rg::RW;
        const int x = ox + cx, y = oy + rr;
        if (x < 0 || x >= W || y < 0 || y >= H) continue;
        const int p = y * W + x;
        const int s = wrap_row(rr + base, Rg::RH) * Rg::RW + cx;
        const unsigned vb = __ldg(vbits + p);
        const unsigned use = live & (vb | always);
        if (vb & kSpatialBits) flags |= (rr >= kTileH) ? 3u : 1u;
        if (SET == 0) {
            // the temporal warps are stored at nearly every pixel: their loads do not wait for the validity byte (an entry
            // pass 1 never wrote is read and dropped; the buffer is allocated in full)
            float4 w[NSRC];
#pragma unroll
            for (int j = 0; j < NSRC; ++j)
                w[j] = (live & (1u << j)) ? __ldg(raw + (size_t)j * HW + p) : f4_zero();
            float vr[3], v0[3], v1[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                vr[c] = __ldg(ref + c * HW + p);
                v0[c] = __ldg(src0 + c * HW + p);
                v1[c] = __ldg(src1 + c * HW + p);
            }
            float tw[2][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
#pragma unroll
            for (int j = 0; j < NSRC; ++j) {
                if (!(vb & (1u << j))) w[j] = f4_zero();
                if (use & (1u << j)) {
                    tw[j][0] = VFD_ADD(0.f, warp_normalise(ctx.prm[j], w[j].x, w[j].w));
                    tw[j][1] = VFD_ADD(0.f, warp_normalise(ctx.prm[j], w[j].y, w[j].w));
                    tw[j][2] = VFD_ADD(0.f, warp_normalise(ctx.prm[j], w[j].z, w[j].w));
                }
            }
            float chk = 0.f;      // a non-finite value anywhere in the region switches the masked-pixel skip of phase B off
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                img[(c * 2 + 0) * Rg::R + s] = make_float2(tw[0][c], tw[1][c]);
                img[(c * 2 + 1) * Rg::R + s] = make_float2(v0[c], v1[c]);
                sm[Rg::kAux + c * Rg::R + s] = vr[c];
                chk += (tw[0][c] + tw[1][c]) + (v0[c] + v1[c]) + vr[c];
            }
            if (!isfinite(chk)) flags |= 4u;
        } else {
            float4 w[NWc];
#pragma unroll
            for (int j = NSRC; j < NWc; ++j)
                w[j] = (live & vb & (1u << j)) ? __ldg(raw + (size_t)j * HW + p) : f4_zero();
            float vr[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) vr[c] = __ldg(ref + c * HW + p);
            // neighbour warps of frame fi, summed over the two neighbours in the reference's order (k = 0, then 1,
            // view_rendering.py:193-198): fi = 0 -> pair 1.x, fi = 1, 2 -> pair 0.x / .y
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
            float chk = 0.f;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                img[(c * 2 + 0) * Rg::R + s] = make_float2(ov[1][c], ov[2][c]);
                img[(c * 2 + 1) * Rg::R + s] = make_float2(ov[0][c], vr[c]);
                chk += (ov[1][c] + ov[2][c]) + (ov[0][c] + vr[c]);
            }
#pragma unroll
            for (int f = 0; f < 3; ++f) sm[Rg::kAux + f * Rg::R + s] = om[f];
            if (!isfinite(chk + (om[0] + om[1] + om[2]))) flags |= 4u;
        }
    }
    return flags;
}

// ---- phase B ------------------------------------------------------------------------------------------------------
// Photometric loss of the set's two image pairs against the target (loss_util.py:43-78) at the pixel whose nine tap
// slots are idx[]: the packed SSIM of warp_ssim.cuh (photometric8) restricted to two pairs, the same operations in the
// same order per lane.  SET 0: the target comes from its own plane (scalar sums; add / mul .rn like the packed lanes), the
// derivative coefficients of pair 0 are returned in g0.  SET 1: the target is pair 1's .y lane (that lane compares the
// target with itself and is ignored); g0 = pair 0 (both lanes), g1 = pair 1 (.x lane).
template <int SET>
__device__ __forceinline__ void photometric_set(const float* sm, const int* idx, float2* L, SsimGrad2* g0, SsimGrad2* g1) {
    using Rg = SetRegion;
    const float2* img = reinterpret_cast<const float2*>(sm);
    float2 ssum[2], lsum[2];
#pragma unroll
    for (int q = 0; q < 2; ++q) ssum[q] = lsum[q] = splat2(0.f);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float2* pc = img + c * 2 * Rg::R;
        const float* py = sm + Rg::kAux + c * Rg::R;
        float2 sx[2], sxx[2], sxy[2], xc[2];
        float sy = 0.f, syy = 0.f, ycv = 0.f;
#pragma unroll
        for (int q = 0; q < 2; ++q) sx[q] = sxx[q] = sxy[q] = xc[q] = splat2(0.f);
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            float2 x[2];
            x[0] = pc[idx[k]];
            x[1] = pc[Rg::R + idx[k]];
            const float yv = SET == 0 ? py[idx[k]] : x[1].y;
            const float2 yy = splat2(yv);
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                sx[q] = add2(sx[q], x[q]);
                sxx[q] = fma2(x[q], x[q], sxx[q]);      // fused, as in photometric8 (see the note there)
                sxy[q] = fma2(x[q], yy, sxy[q]);
                if (k == 4) xc[q] = x[q];
            }
            if (SET == 0) {     // the target's sums, as the scalar twin of a packed lane
                sy = __fadd_rn(sy, yv);
                syy = __fmaf_rn(yv, yv, syy);
                if (k == 4) ycv = yv;
            }
        }
        if (SET == 1) { sy = sx[1].y; syy = sxx[1].y; ycv = xc[1].y; }
        const float mu_y = div_const<9>(sy), e_yy = div_const<9>(syy);
        const float musq_y = __fmul_rn(mu_y, mu_y), sig_y = __fsub_rn(e_yy, musq_y);
        const float2 yc = splat2(ycv);
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const float2 mu = div_const2<9>(sx[q]);
            const float2 musq = mul2(mu, mu), mu_xy = mul2(mu, splat2(mu_y));
            const float2 sig_x = sub2(div_const2<9>(sxx[q]), musq);
            const float2 sig_xy = sub2(div_const2<9>(sxy[q]), mu_xy);
            const float2 n1 = fma2(splat2(2.f), mu_xy, splat2(kC1)), n2 = fma2(splat2(2.f), sig_xy, splat2(kC2));
            const float2 d1 = add2(add2(musq, splat2(musq_y)), splat2(kC1));
            const float2 d2 = add2(add2(sig_x, splat2(sig_y)), splat2(kC2));
            const float2 den = add2(mul2(d1, d2), splat2(1e-8f));
            const float2 num = mul2(n1, n2);
            const float2 inv = make_float2(fast_rcp(den.x), fast_rcp(den.y));
            const float2 ssim = mul2(num, inv);
            const float2 raw = mul2(sub2(splat2(1.f), ssim), splat2(0.5f));
            const float2 l = make_float2(fminf(fmaxf(raw.x, 0.f), 1.f), fminf(fmaxf(raw.y, 0.f), 1.f));
            ssum[q] = add2(ssum[q], fma2(ssim, splat2(0.f), l));      // keeps a NaN ssim (torch.clamp), see photometric8
            const float2 df = sub2(yc, xc[q]);
            lsum[q] = add2(lsum[q], make_float2(fabsf(df.x), fabsf(df.y)));
            if (q == 0 || SET == 1) {
                const float gk = -0.5f * 0.85f / 3.f;
                const float2 gs = make_float2((raw.x >= 0.f && raw.x <= 1.f) ? gk : 0.f,
                                              (raw.y >= 0.f && raw.y <= 1.f) ? gk : 0.f);
                const float2 gi = mul2(gs, inv);
                const float2 dn = mul2(splat2(2.f * mu_y), sub2(n2, n1));
                const float2 dd = mul2(mul2(splat2(2.f), mu), sub2(d2, d1));
                SsimGrad2* g = q == 0 ? g0 : g1;
                g->a[c] = mul2(gi, sub2(dn, mul2(ssim, dd)));
                g->b[c] = mul2(gi, mul2(ssim, mul2(splat2(-1.f), d1)));
                g->c[c] = mul2(gi, mul2(splat2(2.f), n1));
            }
        }
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        const float2 s_mean = div_const2<3>(ssum[q]), l_mean = div_const2<3>(lsum[q]);
        L[q] = add2(mul2(splat2(0.85f), s_mean), mul2(splat2(0.15f), l_mean));
    }
}

// nine tap slots of pixel (x, y) in the circular region (reflection padding of 1, loss_util.py:47-49)
__device__ __forceinline__ void tap_slots(int x, int y, int W, int H, int ox, int oy, int base, int* idx) {
    using Rg = SetRegion;
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy) {
        const int row = wrap_row(reflect1(y + dy, H) - oy + base, Rg::RH) * Rg::RW;
#pragma unroll
        for (int dx = -1; dx <= 1; ++dx) idx[(dy + 1) * 3 + dx + 1] = row + reflect1(x + dx, W) - ox;
    }
}

#ifndef VFD_SET0_CTAS
#define VFD_SET0_CTAS 3
#endif
#ifndef VFD_SET1_CTAS
#define VFD_SET1_CTAS 3
#endif
template <int SET, bool FUSED>
__global__ void __launch_bounds__(kTileThreads, SET == 0 ? VFD_SET0_CTAS : VFD_SET1_CTAS) warp_loss_set_kernel(const LossBwdArgs a) {
    using Rg = SetRegion;
    using Cf = SetCoef<SET>;
    constexpr int kSlots = SET == 0 ? VFD_WARP_SRC : (1 + VFD_WARP_SRC) * VFD_WARP_NBR;     // warps this kernel differentiates
    const vfd_warp_desc& d = a.d;
    __shared__ WarpCtx ctx;
    __shared__ float s_dP[2][kTilePixels / 32][kSlots * 12];
    __shared__ double sacc[4];
    __shared__ unsigned s_flags[2];
    __shared__ unsigned s_nz[SetCoef<SET>::RH];      // per ring row (circular like the coefficient planes): bit = column
    float* sm = vfd_warp_smem;
    float4* cf4 = reinterpret_cast<float4*>(sm + Rg::kFloats);
    const int b = blockIdx.z / d.n_call, ci = blockIdx.z % d.n_call, cam = d.cam0 + ci;
    const int W = d.W, H = d.H, HW = H * W, n_src = d.n_src;
    const bool two = n_src > 1;
    const int tiles_y = (H + kTileH - 1) / kTileH, ntile = gridDim.x * tiles_y;
    const int ty_first = blockIdx.y * kStrip, ty_last = min(tiles_y, ty_first + kStrip);
    const int strip_y0 = ty_first * kTileH, strip_y1 = ty_last * kTileH;
    const int tx0 = blockIdx.x * kSetW, ox = tx0 - 2;
    const size_t bc = (size_t)(b * d.n_call + ci);
    if (SET == 1) {     // nothing to do in this strip?  (uniform over the CTA; no barrier has been reached yet)
        bool any_tile = false;
        for (int tyi = ty_first; tyi < ty_last; ++tyi) any_tile |= a.tile_flag[bc * ntile + tyi * gridDim.x + blockIdx.x] != 0;
        if (!any_tile) return;
    }
    warp_ctx_init(ctx, d, a.T, a.params, b, ci);
    if (threadIdx.x < 4) sacc[threadIdx.x] = 0.0;
    if (threadIdx.x < 2) s_flags[threadIdx.x] = 0u;
    __syncthreads();
    unsigned live_w = 0, always_w = 0;      // per warp j: exists / parameters not finite (once per CTA, not per tile)
    for (int j = 0; j < ctx.NW; ++j) {
        if (ctx.active[j] != 0) {
            live_w |= 1u << j;
            if (ctx.prm[j][6] == 0.f) always_w |= 1u << j;
        }
    }
    const float* depth = a.depth + bc * HW;
    const float* rmaskp = d.mask + (size_t)(b * d.num_cams + cam) * HW;
    const float4* raw = a.raw + bc * ctx.NW * HW;
    const uint8_t* vbits = a.vbits + bc * HW;
    uint8_t* abit = a.auto_bit + bc * HW;
    float w_set[3] = {1.f, 1.f, 1.f};
    if (!FUSED) {
        w_set[0] = __ldg(a.d_num + ci * 3 + 0); w_set[1] = __ldg(a.d_num + ci * 3 + 1); w_set[2] = __ldg(a.d_num + ci * 3 + 2);
    }
    float facc[4] = {0.f, 0.f, 0.f, 0.f};
    unsigned carried = 0u;     // SET 0: spatial presence in the region rows shared with the previous tile
    bool cont = false;         // the previous tile of the strip was processed: its last 4 region / 2 ring rows are in place
    int base = 0, cbase = 0;   // circular offsets of the region (mod RH) and of the ring (mod Cf::RH)
    int nproc = 0;             // tiles processed so far: parity picks the s_dP buffer and the flag word
    bool nonfinite = false;    // a non-finite value has entered the region since the last fresh fill
    for (int tyi = ty_first; tyi < ty_last; ++tyi) {
        const int tile = tyi * gridDim.x + blockIdx.x;
        if (SET == 1 && a.tile_flag[bc * ntile + tile] == 0) {      // CTA-uniform
            cont = false;
            continue;
        }
        const int ty0 = tyi * kTileH, oy = ty0 - 2;
        if (cont) {
            base = wrap_row(base + kTileH, Rg::RH);
            cbase = wrap_row(cbase + kTileH, Cf::RH);
        } else {
            base = cbase = 0;
        }
        float (*dPs)[kSlots * 12] = s_dP[nproc & 1];
        for (int i = threadIdx.x; i < (kTilePixels / 32) * kSlots * 12; i += kTileThreads) (&dPs[0][0])[i] = 0.f;
        const unsigned fl = n_src == 2 ? fill_set<SET, 2>(ctx, d, b, ci, ox, oy, raw, vbits, sm, cont ? 4 : 0, base, live_w, always_w)
                                       : fill_set<SET, 1>(ctx, d, b, ci, ox, oy, raw, vbits, sm, cont ? 4 : 0, base, live_w, always_w);
        bool sp = true;
        // CTA-wide OR of the fill's flags (flag words alternate between tiles; the other one is cleared for the next tile now
        // that every thread is past the previous tile's read of it)
        const unsigned any = cta_or(fl, &s_flags[nproc & 1]);
        if (threadIdx.x == 0) s_flags[(nproc + 1) & 1] = 0u;
        // a pixel whose self-occlusion mask is 0 has every loss mask 0: its SSIM is skipped (coefficients and sums are
        // exactly the zeros the masked products give) -- unless something non-finite has entered the strip's region, in
        // which case the arithmetic runs so that NaN * 0 stays NaN as in the reference
        if (!cont) nonfinite = false;
        nonfinite = nonfinite || (any & 4u) != 0u || (always_w != 0u);
        if (SET == 0) {      // any spatial / spatio-temporal warp in this tile's region?
            sp = ((any & 1u) | (cont ? carried : 0u)) != 0u;
            carried = (any >> 1) & 1u;
            if (threadIdx.x == 0) a.tile_flag[bc * ntile + tile] = sp ? 1 : 0;
        }

        // ---- phase B: loss decisions + SSIM derivative coefficients on the ring ----
        // a warp per ring row (lane = column; lanes past the ring's width idle when kSetW < 30)
        static_assert(Cf::RW <= 32, "ring row = one warp");
        for (int rr = (cont ? 2 : 0) + (threadIdx.x >> 5); rr < Cf::RH; rr += kTileThreads / 32) {
            const int cx = threadIdx.x & 31;
            bool nz = false;      // this ring pixel carries a non-zero upstream weight (its coefficients can be non-zero)
            do {
            if (cx >= Cf::RW) break;
            const int x = tx0 - 1 + cx, y = ty0 - 1 + rr;
            if (x < 0 || x >= W || y < 0 || y >= H) break;
            const int p = y * W + x;
            const int rs = wrap_row(rr + cbase, Cf::RH) * Cf::RW + cx;
            const float rm = __ldg(rmaskp + p);
            if (rm == 0.f && !nonfinite) {      // masked out (see above): zero coefficients, no contribution to any sum
#pragma unroll
                for (int c = 0; c < Cf::kPlanes; ++c) cf4[c * Cf::R + rs] = f4_zero();
                if (SET == 0 && x >= tx0 && x < tx0 + kSetW && y >= strip_y0 && y < strip_y1) abit[p] = 0;
                break;
            }
            int idx[9];
            tap_slots(x, y, W, H, ox, oy, base, idx);
            float2 L[2];
            SsimGrad2 g0, g1;
            const bool mine = x >= tx0 && x < tx0 + kSetW && y >= strip_y0 && y < strip_y1;   // counted / owned by this CTA
            if (SET == 0) {
                float nz0 = 0.f, nz1 = 0.f;
                {
                    const float* np = a.noise + bc * n_src * HW + p;
                    nz0 = __ldg(np);
                    if (two) nz1 = __ldg(np + HW);
                }
                photometric_set<0>(sm, idx, L, &g0, &g1);
                // temporal reprojection: torch.min over frames keeps the first minimum on a tie, NaN if either is NaN
                const bool s1 = two && L[0].y < L[0].x;
                float reproj = s1 ? L[0].y : L[0].x;
                if (two && L[0].y != L[0].y) reproj = L[0].y;
                // identity reprojection + tie-break noise (single_cam_loss.py:44-47)
                float ident = VFD_ADD(L[1].x, VFD_MUL(0.00001f, nz0));
                if (two) ident = nan_min(ident, VFD_ADD(L[1].y, VFD_MUL(0.00001f, nz1)));
                const bool keep = reproj <= ident;      // argmin over cat([reproj, identity]) == 0 (loss_util.py:15-17)
                const float automask = VFD_MUL(keep ? 1.f : 0.f, rm);
                const float wr = w_set[0] * automask;
                nz = wr != 0.f;      // (a NaN weight counts)
#pragma unroll
                for (int c = 0; c < 3; ++c)
                    cf4[c * Cf::R + rs] = make_float4(wr * (s1 ? g0.a[c].y : g0.a[c].x), wr * (s1 ? g0.b[c].y : g0.b[c].x),
                                                      wr * (s1 ? g0.c[c].y : g0.c[c].x),
                                                      c == 0 ? wr : (c == 1 ? __int_as_float(s1 ? 1 : 0) : 0.f));
                if (mine) {
                    abit[p] = keep ? 1 : 0;
                    if (FUSED) { facc[0] += VFD_MUL(reproj, automask); facc[1] += automask; }
                }
            } else {
                const float automask = VFD_MUL(abit[p] ? 1.f : 0.f, rm);
                const int ctr = idx[4];
                photometric_set<1>(sm, idx, L, &g0, &g1);
                const float sp_mask = VFD_MUL(rm, sm[Rg::kAux + ctr]);
                const float spatial = L[1].x;
                const bool t1 = two && L[0].y < L[0].x;
                float stl = t1 ? L[0].y : L[0].x;
                if (two && L[0].y != L[0].y) stl = L[0].y;
                float st_mask = VFD_MUL(VFD_MUL(rm, sm[Rg::kAux + Rg::R + ctr]), automask);
                if (two) st_mask = nan_max(st_mask, VFD_MUL(VFD_MUL(rm, sm[Rg::kAux + 2 * Rg::R + ctr]), automask));
                const float ws = w_set[1] * sp_mask, wt = w_set[2] * st_mask;
                nz = ws != 0.f || wt != 0.f;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    cf4[c * Cf::R + rs] = make_float4(ws * g1.a[c].x, ws * g1.b[c].x, ws * g1.c[c].x,
                                                      c == 0 ? ws : (c == 1 ? wt : __int_as_float(t1 ? 1 : 0)));
                    cf4[(3 + c) * Cf::R + rs] = make_float4(wt * (t1 ? g0.a[c].y : g0.a[c].x), wt * (t1 ? g0.b[c].y : g0.b[c].x),
                                                            wt * (t1 ? g0.c[c].y : g0.c[c].x), 0.f);
                }
                if (FUSED && mine) {
                    facc[0] += VFD_MUL(spatial, sp_mask); facc[1] += sp_mask;
                    facc[2] += VFD_MUL(stl, st_mask);     facc[3] += st_mask;
                }
            }
            } while (0);
            // which ring pixels of this row can contribute: phase C skips the 3x3 gather of a pixel whose nine taps are all
            // zero (with the auto-mask off -- the usual case in a static scene -- or outside the overlap, most are)
            const unsigned nzm = __ballot_sync(0xffffffffu, nz);
            if (cx == 0) s_nz[wrap_row(rr + cbase, Cf::RH)] = nzm;
        }
        __syncthreads();

        // ---- phase C: gather the coefficients at the interior pixel, then back through normalise / sampler / projection ----
        const int x = tx0 + (threadIdx.x & 31), y = ty0 + (threadIdx.x >> 5);      // a warp per tile row
        const bool live = (threadIdx.x & 31) < kSetW && x < W && y < H;
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
        constexpr int NP = SET == 0 ? VFD_WARP_SRC : 1 + VFD_WARP_SRC;      // differentiable images of this set
        float dpred[NP][3];
#pragma unroll
        for (int k = 0; k < NP; ++k) dpred[k][0] = dpred[k][1] = dpred[k][2] = 0.f;
        float X[3] = {0.f, 0.f, 0.f}, ray[3] = {0.f, 0.f, 0.f};
        bool taps = false;
        if (live) {
            const int cxm = x - tx0;      // ring column of the left tap
            const unsigned m3 = s_nz[wrap_row(y - ty0 + cbase, Cf::RH)] | s_nz[wrap_row(y - ty0 + 1 + cbase, Cf::RH)] |
                                s_nz[wrap_row(y - ty0 + 2 + cbase, Cf::RH)];
            taps = ((m3 >> cxm) & 7u) != 0u || nonfinite;
        }
        if (taps) {
            const float dep = __ldg(depth + y * W + x);
            // window sums of the coefficients as packed pairs: AB = (sum a, sum b), CJ = (sum c, unused lane -- the .w of a
            // coefficient plane rides along).  One fma.rn.f32x2 where two three-register FFMA stood: the same bits, 2/3 of
            // the instructions and of the fma-pipe cycles of this gather (math-pipe throttle was 11 % of the stall samples).
            float2 AB[NP][3], CJ[NP][3];
#pragma unroll
            for (int k = 0; k < NP; ++k)
#pragma unroll
                for (int c = 0; c < 3; ++c) AB[k][c] = CJ[k][c] = splat2(0.f);
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
                const int rrow = wrap_row(y + dy - (ty0 - 1) + cbase, Cf::RH) * Cf::RW;
#pragma unroll
                for (int dx = -1; dx <= 1; ++dx) {
                    const float mult = my[dy + 1] * mx[dx + 1];
                    if (mult == 0.f) continue;
                    const int r = rrow + (x + dx - (tx0 - 1));
                    // the min's selection as tap weights: w[k] = mult for the selected frame, 0 for the other (mult is 0, 1, 2
                    // or 4, so w * q is exact and the fused form rounds like a separate multiply and add)
                    float4 q[3];
#pragma unroll
                    for (int c = 0; c < 3; ++c) q[c] = cf4[c * Cf::R + r];
                    if (SET == 0) {
                        const int sel = __float_as_int(q[1].w);
                        const float w0 = sel ? 0.f : mult, w1 = mult - w0;
                        const float2 W0 = splat2(w0), W1 = splat2(w1);
#pragma unroll
                        for (int c = 0; c < 3; ++c) {
                            const float2 ab = make_float2(q[c].x, q[c].y), cj = make_float2(q[c].z, q[c].w);
                            AB[0][c] = fma2(ab, W0, AB[0][c]); CJ[0][c] = fma2(cj, W0, CJ[0][c]);
                            AB[1][c] = fma2(ab, W1, AB[1][c]); CJ[1][c] = fma2(cj, W1, CJ[1][c]);
                        }
                    } else {
                        const int sel = __float_as_int(q[2].w);
                        const float w0 = sel ? 0.f : mult, w1 = mult - w0;
                        const float2 WM = splat2(mult), W0 = splat2(w0), W1 = splat2(w1);
#pragma unroll
                        for (int c = 0; c < 3; ++c) {
                            const float4 q2 = cf4[(3 + c) * Cf::R + r];
                            const float2 ab = make_float2(q[c].x, q[c].y), cj = make_float2(q[c].z, q[c].w);
                            const float2 ab2 = make_float2(q2.x, q2.y), cj2 = make_float2(q2.z, q2.w);
                            AB[0][c] = fma2(ab, WM, AB[0][c]); CJ[0][c] = fma2(cj, WM, CJ[0][c]);
                            AB[1][c] = fma2(ab2, W0, AB[1][c]); CJ[1][c] = fma2(cj2, W0, CJ[1][c]);
                            AB[2][c] = fma2(ab2, W1, AB[2][c]); CJ[2][c] = fma2(cj2, W1, CJ[2][c]);
                        }
                    }
                }
            }
            const int ctr = wrap_row(y - oy + base, Rg::RH) * Rg::RW + (x - ox);
            const int r0 = wrap_row(y - (ty0 - 1) + cbase, Cf::RH) * Cf::RW + (x - (tx0 - 1));
            const float2* img = reinterpret_cast<const float2*>(sm);
            float wl[NP];
            if (SET == 0) {
                const float wr = cf4[r0].w;
                const int sel0 = __float_as_int(cf4[Cf::R + r0].w);
                wl[0] = sel0 ? 0.f : wr;
                wl[1] = sel0 ? wr : 0.f;
            } else {
                const float ws = cf4[r0].w, wt = cf4[Cf::R + r0].w;
                const int sel0 = __float_as_int(cf4[2 * Cf::R + r0].w);
                wl[0] = ws;
                wl[1] = sel0 ? 0.f : wt;
                wl[2] = sel0 ? wt : 0.f;
            }
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float2 p0 = img[(c * 2 + 0) * Rg::R + ctr], p1 = img[(c * 2 + 1) * Rg::R + ctr];
                const float yv = SET == 0 ? sm[Rg::kAux + c * Rg::R + ctr] : p1.y;
#pragma unroll
                for (int k = 0; k < NP; ++k) {
                    // SET 0: pred 0, 1 = pair 0 .x, .y;  SET 1: pred 0 = overlap_0 = pair 1.x, pred 1, 2 = pair 0 .x, .y
                    const float xv = SET == 0 ? (k == 0 ? p0.x : p0.y) : (k == 0 ? p1.x : (k == 1 ? p0.x : p0.y));
                    const float df = xv - yv;
                    const float sgn = df > 0.f ? 1.f : (df < 0.f ? -1.f : 0.f);
                    dpred[k][c] = (AB[k][c].x + 2.f * xv * AB[k][c].y + yv * CJ[k][c].x) * (1.f / 9.f) + (0.15f / 3.f) * wl[k] * sgn;
                }
            }
            vfd_pixel_ray(ctx.invK, (float)x, (float)y, ray);
            X[0] = VFD_MUL(dep, ray[0]); X[1] = VFD_MUL(dep, ray[1]); X[2] = VFD_MUL(dep, ray[2]);
        }
        float gset[SET == 0 ? 1 : 2] = {};
        // One copy of the sampler / back-projection code, run once per warp of this set (a rolled loop: unrolled over the
        // six overlap warps the kernel was 95 KB of code and 13 % of its stall samples were instruction-cache misses).
        //   SET 0: warp j = slot (temporal frame k = slot);  SET 1: warp j = n_src + slot, k = slot / NBR (frame k of
        //   neighbour slot % NBR)
        const int n_slots = SET == 0 ? n_src : (1 + n_src) * VFD_WARP_NBR;
#pragma unroll 1
        for (int slot = 0; slot < n_slots; ++slot) {
            {
                const int k = SET == 0 ? slot : slot / VFD_WARP_NBR;
                const int j = SET == 0 ? slot : n_src + slot;
                if (!ctx.active[j]) continue;      // CTA-uniform
                float up[3];      // dpred[k][:] (static indices: a run-time index would put dpred in local memory)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    float v = dpred[0][c];
#pragma unroll
                    for (int kk = 1; kk < NP; ++kk) v = (k == kk) ? dpred[kk][c] : v;
                    up[c] = v;
                }
                float dP[12];
#pragma unroll
                for (int i = 0; i < 12; ++i) dP[i] = 0.f;
                // Most (pixel, warp) pairs carry no gradient: the min picked the other frame in the whole 3x3 neighbourhood,
                // or the pixel lies outside the spatial overlap.  Those skip the sampler; a warp whose 32 pixels all skip
                // also skips the reduction (its slots of s_dP stay zero, so the sums are unchanged).
                bool contrib = false;
                if (live && (up[0] != 0.f || up[1] != 0.f || up[2] != 0.f)) {
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
                        // d(ix)/d(depth) = (a0 * den - cam.x * a2) / den^2 with a = P[:, :3] @ ray, cam = depth * a + P[:, 3]:
                        // the depth-proportional parts cancel exactly in that numerator (see warp_loss_bwd.cu)
                        const float a0 = P[0] * ray[0] + P[1] * ray[1] + P[2] * ray[2];
                        const float a1 = P[4] * ray[0] + P[5] * ray[1] + P[6] * ray[2];
                        const float a2 = P[8] * ray[0] + P[9] * ray[1] + P[10] * ray[2];
                        const float t2 = P[11] + 1e-7f;
                        const float gj = (gix * (a0 * t2 - P[3] * a2) + giy * (a1 * t2 - P[7] * a2)) * inv * inv;
                        if (SET == 0) gset[0] += gj;
                        else gset[k == 0 ? 0 : 1] += gj;
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
                if (!__any_sync(0xffffffffu, contrib)) continue;
                warp_sum12_store(dP, lane, &dPs[wid][slot * 12]);
            }
        }
        if (live) {
            float* g3 = a.g3 + bc * 3 * HW + y * W + x;
            if (SET == 0) {
                g3[0] = gset[0];
                if (!sp) { g3[HW] = 0.f; g3[2 * HW] = 0.f; }     // no SET 1 pass over this tile
            } else {
                g3[HW] = gset[0]; g3[2 * HW] = gset[1];
            }
        }
        __syncthreads();
        {   // d_T partial sums of the tile: this set's warps (and zeros for the other set's where no SET 1 pass follows)
            float* part = a.partial + (bc * ntile + tile) * (kMaxW * 12);
            const int i0 = SET == 0 ? 0 : n_src * 12, i1 = SET == 0 ? n_src * 12 : ctx.NW * 12;
            for (int i = i0 + threadIdx.x; i < i1; i += kTileThreads) {
                float v = 0.f;
#pragma unroll
                for (int w = 0; w < kTilePixels / 32; ++w) v += dPs[w][i - i0];
                part[i] = v;
            }
            if (SET == 0 && !sp)
                for (int i = i1 + threadIdx.x; i < ctx.NW * 12; i += kTileThreads) part[i] = 0.f;
        }
        cont = true;
        ++nproc;
        // no barrier here: the next tile zeroes the OTHER s_dP buffer, fills region slots and writes ring slots that this
        // tile's phase C (finished at the barrier above) was the last to read
    }   // tiles of the strip
    if (FUSED) {   // loss sums: fp32 inside a warp, fp64 above (as warp_loss_fwd_kernel)
#pragma unroll
        for (int q = 0; q < (SET == 0 ? 2 : 4); ++q) {
            float v = facc[q];
#pragma unroll
            for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if ((threadIdx.x & 31) == 0 && v != 0.f) atomicAdd(&sacc[q], (double)v);
        }
        __syncthreads();
        if (threadIdx.x < (SET == 0 ? 2 : 4) && sacc[threadIdx.x] != 0.0)
            atomicAdd(a.sums + ci * 8 + (SET == 0 ? 0 : 2) + threadIdx.x, sacc[threadIdx.x]);
    }
}

template <int SET>
static size_t set_smem_bytes() { return (SetRegion::kFloats + SetCoef<SET>::kFloats) * sizeof(float); }

int launch_loss_sets(const LossBwdArgs& a, cudaStream_t st) {
    const vfd_warp_desc& d = a.d;
    const int tiles_x = (d.W + kSetW - 1) / kSetW, tiles_y = (d.H + kTileH - 1) / kTileH;
    dim3 grid(tiles_x, (tiles_y + kStrip - 1) / kStrip, d.B * d.n_call);
    // > 48 KB of dynamic shared memory needs the opt-in; setting it is idempotent and per device, so it is simply
    // repeated on every call (no global state to keep in step with the current device)
    if (cudaFuncSetAttribute(warp_loss_set_kernel<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)set_smem_bytes<0>()) != cudaSuccess ||
        cudaFuncSetAttribute(warp_loss_set_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)set_smem_bytes<1>()) != cudaSuccess)
        return VFD_ERR_CUDA;
    warp_loss_set_kernel<0, true><<<grid, kTileThreads, set_smem_bytes<0>(), st>>>(a);
    VFD_CHECK_LAUNCH();
    warp_loss_set_kernel<1, true><<<grid, kTileThreads, set_smem_bytes<1>(), st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

}  // namespace vfd
