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

__global__ void __launch_bounds__(kStatThreads) warp_stats_kernel(const StatsArgs a) {
    const vfd_warp_desc& d = a.d;
    __shared__ WarpCtx ctx;
    __shared__ double sacc[kMaxW][8];
    const int b = blockIdx.z, ci = blockIdx.y, cam = d.cam0 + ci;
    const int W = d.W, H = d.H, HW = H * W;
    warp_ctx_init(ctx, d, a.T, nullptr, b, ci);
    for (int i = threadIdx.x; i < kMaxW * 8; i += kStatThreads) (&sacc[0][0])[i] = 0.0;
    __syncthreads();
    const int NW = ctx.NW;
    const int lane = threadIdx.x & 31;
    const float inv_w = 1.0f / (float)W;
    const int p0 = blockIdx.x * kStatThreads * kStatPx + threadIdx.x;
    const float* depth = a.depth + (size_t)(b * d.n_call + ci) * HW;
    const float* ref = d.color[0] + (size_t)(b * d.num_cams + cam) * 3 * HW;
    const float* rmask = d.mask + (size_t)(b * d.num_cams + cam) * HW;

    // target image: unmasked sum and sum of squares (shared by all warps of this camera)
    {
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < kStatPx; ++k) {
            const int p = p0 + k * kStatThreads;
            if (p < HW) {
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float r = __ldg(ref + c * HW + p);
                    s1 += r;
                    s2 = fmaf(r, r, s2);
                }
            }
        }
        s1 = warp_sum(s1);
        s2 = warp_sum(s2);
        if (lane == 0) {
            atomicAdd(&sacc[0][5], (double)s1);
            atomicAdd(&sacc[0][6], (double)s2);
        }
    }
    unsigned long long vb = 0ull;      // validity bits of this thread's kStatPx pixels, one byte each
    for (int j = 0; j < NW; ++j) {
        if (!ctx.active[j]) continue;
        float4* raw = a.raw + ((size_t)(b * d.n_call + ci) * NW + j) * HW;
        float cnt = 0.f, srm = 0.f, swm = 0.f, sw1 = 0.f, sw2 = 0.f;
#pragma unroll 2
        for (int k = 0; k < kStatPx; ++k) {
            const int p = p0 + k * kStatThreads;
            if (p >= HW) break;
            // row / column without an integer division: (p + 0.5) / W is at least 0.5 / W away from an integer, far more
            // than the fp32 error of the product for H * W < 2^22 (check_warp_desc bounds it); one fix-up step for safety
            int v = __float2int_rz(((float)p + 0.5f) * inv_w);
            int u = p - v * W;
            if (u < 0) { u += W; --v; }
            if (u >= W) { u -= W; ++v; }
            float X[3];
            vfd_backproject(ctx.invK, (float)u, (float)v, __ldg(depth + p), X);
            const WarpSample s = warp_sample<false>(ctx.P[j], ctx.img[j], ctx.msk[j], X, W, H);
            if (s.v[0] != 0.f || s.v[1] != 0.f || s.v[2] != 0.f || s.m != 0.f) {
                raw[p] = make_float4(s.v[0], s.v[1], s.v[2], s.m);
                vb |= 1ull << (8 * k + j);
            }
            const bool m = (__ldg(rmask + p) * s.m) != 0.f;   // (src_mask * warp_mask).bool()  (:46)
            const float w3 = s.v[0] + s.v[1] + s.v[2];
            sw1 += w3;
            sw2 += s.v[0] * s.v[0] + s.v[1] * s.v[1] + s.v[2] * s.v[2];
            if (m) {
                cnt += 1.f;
                swm += w3;
                srm += __ldg(ref + p) + __ldg(ref + HW + p) + __ldg(ref + 2 * HW + p);
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
            const int p = p0 + k * kStatThreads;
            if (p < HW) vbits[p] = (uint8_t)(vb >> (8 * k));
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
    if ((long)d->H * d->W >= (1L << 22)) return VFD_ERR_DESC;      // pixel indices are exact in fp32, see warp_stats_kernel
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
    const int HW = d->H * d->W, per = kStatThreads * kStatPx;
    dim3 grid((HW + per - 1) / per, d->n_call, d->B);
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
