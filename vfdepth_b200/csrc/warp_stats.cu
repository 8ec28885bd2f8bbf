// Pass 1 of the view-synthesis path: intensity-alignment statistics of every warp
// (models/geometry/view_rendering.py:30-59: get_mean_std + the early exit of get_norm_image_single).
//
// Per (rig, target camera, warp j):  count of pixels with ref_mask*warp_mask != 0, the masked sums of the target
// and of the warped image, and the UNMASKED sums / sums of squares the reference's variance runs over (:36).
// The raw warps (r, g, b, mask) are also written out, one float4 per pixel and warp, so that the loss kernels
// normalise and compare them without projecting and sampling again -- but only where they are not all-zero (one
// validity byte per pixel says which): most pixels of a spatial warp fall outside the neighbour camera's image.
// Per-thread partial sums are fp32 over a handful of pixels; everything above that (warp, CTA, grid) is fp64, so the
// one-pass variance  (S2 - 2 m S1 + n m^2) / n  is as accurate as the reference's two-pass form.
#include "warp_common.cuh"

namespace vfd {

constexpr int kStatThreads = 256;
constexpr int kStatPx = 8;   // pixels per thread
// A CTA owns a kStatTW x kStatTH pixel tile (thread = one column of a 64-wide tile, every 4th row): a compact tile lets
// the CTA prove, from the tile's depth range, that a spatial / spatio-temporal warp leaves the neighbour camera's image
// for ALL of its pixels and skip that warp -- the common case, the overlap with a neighbour is a band at one side.
constexpr int kStatTW = 64, kStatTH = (kStatThreads / kStatTW) * kStatPx;

struct StatsArgs {
    vfd_warp_desc d;
    const float* depth;
    const float* T;
    double* stats;
    float* params;
    float4* raw;   // [B,n_call,NW,H*W] (r, g, b, warp mask) of every warp: read back by the loss kernels
    uint8_t* vbits;  // [B,n_call,H*W] which of a pixel's warps are not all-zero (only those are stored in raw)
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Can warp j sample anything but zeros somewhere in the tile?  cam = P @ [d * ray(u, v); 1] is linear in (d*u, d*v, d), so
// over the box [u0,u1] x [v0,v1] x [dmin,dmax] it stays inside the hull of the 8 corner values; when the 8 denominators
// share a sign and are well away from 0 (relative to the terms they are summed from), the sampling coordinate
// cam.xy / den stays inside the hull of the 8 corner coordinates.  The warp is skipped when that hull is clear of
// [-1, W] x [-1, H] (outside of which bilinear zeros-padding sample, nearest mask and in-bounds flag are all 0,
// warp_sample) by a margin that dwarfs the fp32 error of the per-pixel evaluation.  Anything non-finite or close
// to the camera plane fails a comparison and keeps the warp.  Called by 8 consecutive lanes (lane & 7 = corner).
__device__ __forceinline__ bool warp_leaves_image(const float* invK, const float* P, float u0, float u1, float v0, float v1,
                                                  float dmin, float dmax, int W, int H, int corner) {
    float X[3], cam[3], g[2];
    vfd_backproject(invK, (corner & 1) ? u1 : u0, (corner & 2) ? v1 : v0, (corner & 4) ? dmax : dmin, X);
    vfd_reproject(P, X, W, H, cam, g);
    const float ix = vfd_unnormalize(g[0], W), iy = vfd_unnormalize(g[1], H);
    const float den = cam[2] + 1e-7f;
    const float mag = fabsf(P[8] * X[0]) + fabsf(P[9] * X[1]) + fabsf(P[10] * X[2]) + fabsf(P[11]);
    // sign of den with a relative floor; 0 = too close to the camera plane (or NaN)
    int sgn = (den > 1e-3f * mag) ? 1 : ((den < -1e-3f * mag) ? -1 : 0);
    float xlo = ix - (4.f + 2e-3f * fabsf(ix)), xhi = ix + (4.f + 2e-3f * fabsf(ix));
    float ylo = iy - (4.f + 2e-3f * fabsf(iy)), yhi = iy + (4.f + 2e-3f * fabsf(iy));
    bool same = true;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        // (shuffles outside the && chains: every lane must execute them)
        const int other_sgn = __shfl_xor_sync(0xffffffffu, sgn, o);
        same = same && (other_sgn == sgn);
        const int other_same = __shfl_xor_sync(0xffffffffu, (int)same, o);
        same = same && (other_same != 0);
        xlo = fminf(xlo, __shfl_xor_sync(0xffffffffu, xlo, o));
        xhi = fmaxf(xhi, __shfl_xor_sync(0xffffffffu, xhi, o));
        ylo = fminf(ylo, __shfl_xor_sync(0xffffffffu, ylo, o));
        yhi = fmaxf(yhi, __shfl_xor_sync(0xffffffffu, yhi, o));
    }
    // fminf / fmaxf drop a NaN operand: a NaN coordinate must keep the warp
    const bool finite = !(isnan(ix) || isnan(iy) || isinf(ix) || isinf(iy));
    bool all_finite = finite;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        const int other = __shfl_xor_sync(0xffffffffu, (int)all_finite, o);
        all_finite = all_finite && (other != 0);
    }
    if (!same || sgn == 0 || !all_finite) return false;
    return xhi < -1.f || xlo > (float)W || yhi < -1.f || ylo > (float)H;
}

__global__ void __launch_bounds__(kStatThreads, 5) warp_stats_kernel(const StatsArgs a) {
    const vfd_warp_desc& d = a.d;
    __shared__ WarpCtx ctx;
    __shared__ double sacc[kMaxW][8];
    __shared__ float s_dmin[kStatThreads / 32], s_dmax[kStatThreads / 32];
    __shared__ int s_bad[kStatThreads / 32], s_skip[kMaxW];
    // per pixel of the tile: the target's self-occlusion mask and channel sum -- the same for all NW warps, and read by
    // the warp loop right after the sampler's own (dependent) loads, where one more global round trip hurt
    __shared__ float s_rmask[kStatPx][kStatThreads], s_rsum[kStatPx][kStatThreads];
    const int b = blockIdx.z, ci = blockIdx.y, cam = d.cam0 + ci;
    const int W = d.W, H = d.H, HW = H * W;
    warp_ctx_init(ctx, d, a.T, nullptr, b, ci);
    for (int i = threadIdx.x; i < kMaxW * 8; i += kStatThreads) (&sacc[0][0])[i] = 0.0;
    const int tiles_x = (W + kStatTW - 1) / kStatTW;
    const int tx0 = (blockIdx.x % tiles_x) * kStatTW, ty0 = (blockIdx.x / tiles_x) * kStatTH;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int u = tx0 + (threadIdx.x % kStatTW), v0 = ty0 + threadIdx.x / kStatTW;
    constexpr int kRowStep = kStatThreads / kStatTW;
    const float* depth = a.depth + (size_t)(b * d.n_call + ci) * HW;
    const float* ref = d.color[0] + (size_t)(b * d.num_cams + cam) * 3 * HW;
    const float* rmask = d.mask + (size_t)(b * d.num_cams + cam) * HW;

    // the tile's depth range (the depths are read again, from L1 / L2, by the warp loop)
    {
        float lo = INFINITY, hi = -INFINITY;
        bool bad = false;
#pragma unroll
        for (int k = 0; k < kStatPx; ++k) {
            const int v = v0 + k * kRowStep;
            if (u < W && v < H) {
                const float dep = __ldg(depth + v * W + u);
                lo = fminf(lo, dep);
                hi = fmaxf(hi, dep);
                bad = bad || !isfinite(dep);
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        const bool any_bad = __any_sync(0xffffffffu, bad);
        if (lane == 0) { s_dmin[wid] = lo; s_dmax[wid] = hi; s_bad[wid] = any_bad; }
    }
    __syncthreads();
    const int NW = ctx.NW;
    if (threadIdx.x < kMaxW * 8) {     // 8 lanes per warp j, one per corner of the (u, v, depth) box
        const int j = threadIdx.x >> 3;
        float lo = s_dmin[0], hi = s_dmax[0];
        int bad = s_bad[0];
#pragma unroll
        for (int w = 1; w < kStatThreads / 32; ++w) { lo = fminf(lo, s_dmin[w]); hi = fmaxf(hi, s_dmax[w]); bad |= s_bad[w]; }
        const int jj = j < NW ? j : 0;
        const bool out = warp_leaves_image(ctx.invK, ctx.P[jj], (float)tx0, (float)min(tx0 + kStatTW - 1, W - 1), (float)ty0,
                                           (float)min(ty0 + kStatTH - 1, H - 1), lo, hi, W, H, threadIdx.x & 7);
        if ((threadIdx.x & 7) == 0) s_skip[j] = (j < NW && out && !bad && lo <= hi) ? 1 : 0;
    }

    // target image: unmasked sum and sum of squares (shared by all warps of this camera)
    {
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < kStatPx; ++k) {
            const int v = v0 + k * kRowStep;
            if (u < W && v < H) {
                const int p = v * W + u;
                const float r0 = __ldg(ref + p), r1 = __ldg(ref + HW + p), r2 = __ldg(ref + 2 * HW + p);
                s1 += r0; s2 = fmaf(r0, r0, s2);
                s1 += r1; s2 = fmaf(r1, r1, s2);
                s1 += r2; s2 = fmaf(r2, r2, s2);
                s_rsum[k][threadIdx.x] = r0 + r1 + r2;
                s_rmask[k][threadIdx.x] = __ldg(rmask + p);
            }
        }
        s1 = warp_sum(s1);
        s2 = warp_sum(s2);
        if (lane == 0) {
            atomicAdd(&sacc[0][5], (double)s1);
            atomicAdd(&sacc[0][6], (double)s2);
        }
    }
    __syncthreads();
    unsigned long long vb = 0ull;      // validity bits of this thread's kStatPx pixels, one byte each
    for (int j = 0; j < NW; ++j) {
        if (!ctx.active[j] || s_skip[j]) continue;
        float4* raw = a.raw + ((size_t)(b * d.n_call + ci) * NW + j) * HW;
        float cnt = 0.f, srm = 0.f, swm = 0.f, sw1 = 0.f, sw2 = 0.f;
#pragma unroll 2
        for (int k = 0; k < kStatPx; ++k) {
            const int v = v0 + k * kRowStep;
            if (u >= W || v >= H) continue;
            const int p = v * W + u;
            float X[3];
            vfd_backproject(ctx.invK, (float)u, (float)v, __ldg(depth + p), X);
            if (j >= d.n_src) {
                // Inside a kept tile most pixels of a spatial / spatio-temporal warp still land outside the neighbour's
                // image.  A sample more than half a pixel outside [-1, W] x [-1, H] is exactly zero with mask zero
                // (warp_sample), so it adds nothing to any sum and is not stored: decide that from cam.xy against
                // cam.z WITHOUT the four IEEE divisions of the exact coordinate (the products carry ~1e-6 relative
                // error against a margin of half a pixel; a NaN or a point at / behind the camera plane fails the
                // comparisons and takes the exact path).
                const float* P = ctx.P[j];
                const float c0 = vfd_dot4(P + 0, X[0], X[1], X[2], 1.0f), c1 = vfd_dot4(P + 4, X[0], X[1], X[2], 1.0f);
                const float den = VFD_ADD(vfd_dot4(P + 8, X[0], X[1], X[2], 1.0f), 1e-7f);
                if (den > 1e-6f && (c0 < -1.5f * den || c0 > ((float)W + 0.5f) * den || c1 < -1.5f * den ||
                                    c1 > ((float)H + 0.5f) * den))
                    continue;
            }
            const WarpSample s = warp_sample<false>(ctx.P[j], ctx.img[j], ctx.msk[j], X, W, H);
            if (s.v[0] != 0.f || s.v[1] != 0.f || s.v[2] != 0.f || s.m != 0.f) {
                raw[p] = make_float4(s.v[0], s.v[1], s.v[2], s.m);
                vb |= 1ull << (8 * k + j);
            }
            const bool m = (s_rmask[k][threadIdx.x] * s.m) != 0.f;   // (src_mask * warp_mask).bool()  (:46)
            const float w3 = s.v[0] + s.v[1] + s.v[2];
            sw1 += w3;
            sw2 += s.v[0] * s.v[0] + s.v[1] * s.v[1] + s.v[2] * s.v[2];
            if (m) {
                cnt += 1.f;
                swm += w3;
                srm += s_rsum[k][threadIdx.x];
            }
        }
        cnt = warp_sum(cnt); srm = warp_sum(srm); swm = warp_sum(swm); sw1 = warp_sum(sw1); sw2 = warp_sum(sw2);
        if (lane == 0) {
            atomicAdd(&sacc[j][0], (double)cnt);
            atomicAdd(&sacc[j][1], (double)srm);
            atomicAdd(&sacc[j][2], (double)swm);
            atomicAdd(&sacc[j][3], (double)sw1);
            atomicAdd(&sacc[j][4], (double)sw2);
        }
    }
    {
        uint8_t* vbits = a.vbits + (size_t)(b * d.n_call + ci) * HW;
#pragma unroll
        for (int k = 0; k < kStatPx; ++k) {
            const int v = v0 + k * kRowStep;
            if (u < W && v < H) vbits[v * W + u] = (uint8_t)(vb >> (8 * k));
        }
    }
    __syncthreads();
    double* out = a.stats + (size_t)(b * d.n_call + ci) * NW * 8;
    for (int i = threadIdx.x; i < NW * 8; i += kStatThreads) {
        const int j = i >> 3, q = i & 7;
        // the target-image sums were accumulated in row 0; every warp row gets a copy
        const double v = (q == 5 || q == 6) ? sacc[0][q] : sacc[j][q];
        if (v != 0.0) atomicAdd(out + i, v);
    }
}

// One thread per (target camera, warp): the batch-wide "any sample without overlap" test of :50-53, then the
// per-sample normalisation parameters.
__global__ void warp_stats_finalize_kernel(const StatsArgs a) {
    const vfd_warp_desc& d = a.d;
    const int NW = warp_count(d);
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= d.n_call * NW) return;
    const int ci = t / NW, j = t % NW;
    bool apply = d.intensity_align != 0;
    for (int b = 0; b < d.B && apply; ++b)
        if (a.stats[((size_t)(b * d.n_call + ci) * NW + j) * 8] == 0.0) apply = false;
    const double n = 3.0 * d.H * d.W;
    for (int b = 0; b < d.B; ++b) {
        const double* s = a.stats + ((size_t)(b * d.n_call + ci) * NW + j) * 8;
        float* p = a.params + ((size_t)(b * d.n_call + ci) * NW + j) * 8;
        const double msum = 3.0 * s[0] + 1e-8;           // mask repeated over 3 channels (:47-48)
        const double s_mean = (double)(float)(s[1] / msum), w_mean = (double)(float)(s[2] / msum);
        const double s_var = (s[6] - 2.0 * s_mean * s[5] + n * s_mean * s_mean) / n;
        const double w_var = (s[4] - 2.0 * w_mean * s[3] + n * w_mean * w_mean) / n;
        const float s_std = (float)sqrt(fmax(s_var, 0.0) + 1e-16), w_std = (float)sqrt(fmax(w_var, 0.0) + 1e-16);
        p[0] = (float)w_mean;
        p[1] = w_std + 1e-8f;
        p[2] = s_std;
        p[3] = (float)s_mean;
        p[4] = apply ? 1.f : 0.f;
        p[5] = 1.0f / p[1];
        // all parameters finite: an all-zero (skipped) raw sample normalises to exactly 0, as the reference's masked
        // product does; otherwise the loss kernels run the arithmetic so that NaN / Inf propagate as in the reference
        p[6] = (isfinite(p[0]) && isfinite(p[1]) && isfinite(p[2]) && isfinite(p[3]) && isfinite(p[5])) ? 1.f : 0.f;
        p[7] = 0.f;
    }
}

int check_warp_desc(const vfd_warp_desc* d) {
    if (!d) return VFD_ERR_NULL;
    if (d->B < 1 || d->num_cams < 1 || d->num_cams > VFD_MAX_CAMS || d->H < 2 || d->W < 2) return VFD_ERR_DESC;
    if (d->n_src < 1 || d->n_src > VFD_WARP_SRC) return VFD_ERR_UNSUPPORTED;
    if (d->cam0 < 0 || d->n_call < 1 || d->cam0 + d->n_call > d->num_cams) return VFD_ERR_DESC;
    if ((long)d->H * d->W >= (1L << 22)) return VFD_ERR_DESC;      // pixel indices and coordinates stay exact in fp32
    for (int f = 0; f <= d->n_src; ++f)
        if (!d->color[f]) return VFD_ERR_NULL;
    if (!d->mask || !d->K || !d->invK) return VFD_ERR_NULL;
    return VFD_OK;
}

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_warp_stats(const vfd_warp_desc* d, const float* depth, const float* T, double* stats,
                              float* params, float* raw, void* stream) {
    int rc = check_warp_desc(d);
    if (rc) return rc;
    if (!depth || !T || !stats || !params || !raw) return VFD_ERR_NULL;
    if (reinterpret_cast<uintptr_t>(raw) & 15) return VFD_ERR_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    const int NW = d->n_src + (1 + d->n_src) * VFD_WARP_NBR;
    if (cudaMemsetAsync(stats, 0, (size_t)d->B * d->n_call * NW * 8 * sizeof(double), st) != cudaSuccess)
        return VFD_ERR_CUDA;
    static_assert(kMaxW <= 8, "one validity byte per pixel");
    StatsArgs a = {*d, depth, T, stats, params, reinterpret_cast<float4*>(raw),
                   reinterpret_cast<uint8_t*>(reinterpret_cast<float4*>(raw) + warp_raw_f4(*d))};
    dim3 grid(((d->W + kStatTW - 1) / kStatTW) * ((d->H + kStatTH - 1) / kStatTH), d->n_call, d->B);
    warp_stats_kernel<<<grid, kStatThreads, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    const int nt = d->n_call * NW;
    warp_stats_finalize_kernel<<<(nt + 63) / 64, 64, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" size_t vfd_warp_raw_bytes(const vfd_warp_desc* d) {
    if (check_warp_desc(d)) return 0;
    return warp_raw_bytes(*d);
}
