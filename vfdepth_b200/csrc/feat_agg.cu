// Multi-level feature aggregation in front of the fusion (network/fusion_depthnet.py:57-66, fusion_posenet.py:62-70):
//   feats_agg = LeakyReLU(conv1x1(cat[f_lev, up(f_lev+1), up(f_lev+2), ...]) + b),   up = bilinear, align_corners=True
// A 1x1 convolution and a bilinear resize are both linear and act on different axes, so they commute:
//   conv1x1(cat[...]) = W_0 f_lev + up(W_1 f_lev+1) + up(W_2 f_lev+2) + ...
// The host projects every level at its OWN resolution (plain GEMMs on channel-last rows: 128 + 256/4 + 512/16 = 224
// instead of 896 input channels per output pixel at the shipped sizes, and neither the resized maps nor their
// concatenation ever exist), and this kernel does the rest in one pass: resize the coarse projections with ATen's
// align_corners=True taps, add, bias, LeakyReLU, and write the channel-last [img, h, w, C] map the voxel gather reads.
// Backward: dZ = d_out * slope(out) is the gradient of the finest projection; the coarse ones gather theirs from dZ
// (every coarse pixel sums the fine pixels whose taps touch it: deterministic, no atomics).
#include "common.cuh"

namespace vfd {

constexpr int kAggThreads = 256;
constexpr int kAggMaxBlocks = 148 * 8;     // grid-stride kernels: at most 8 CTAs per SM of a B200

struct AggArgs {
    vfd_agg_desc d;
    const float* y0;                       // [img,h,w,C]
    const float* yc[VFD_AGG_LEVELS];       // [img,ch,cw,C]
    const float* bias;                     // [C]
    float* out;                            // [img,h,w,C]
    const float* d_out;
    float* d_y0;
    float* d_yc;                           // the level of this launch, [img,lch,lcw,C]
    int lch, lcw;
};

struct AggTap {
    int i0, i1;
    float l0, l1;
};

// ATen upsample_bilinear2d, align_corners=True: scale = (in - 1) / (out - 1), src = scale * dst
__device__ __forceinline__ float agg_scale(int in, int out) { return out > 1 ? (float)(in - 1) / (float)(out - 1) : 0.f; }

__device__ __forceinline__ AggTap agg_tap(int dst, int in, float scale) {
    AggTap t;
    const float src = scale * (float)dst;
    t.i0 = (int)src;
    t.i1 = t.i0 + ((t.i0 < in - 1) ? 1 : 0);
    t.l1 = src - (float)t.i0;
    t.l0 = 1.f - t.l1;
    return t;
}

__global__ void __launch_bounds__(kAggThreads) feat_agg_fwd_kernel(const AggArgs a) {
    const vfd_agg_desc& d = a.d;
    const int q = d.C / 4, per = kAggThreads / q;          // threads per pixel, pixels per CTA pass
    const int c4 = (threadIdx.x % q) * 4, sub = threadIdx.x / q;
    if (sub >= per) return;
    const long long npix = (long long)d.n_img * d.h * d.w;
    const float4 b = *reinterpret_cast<const float4*>(a.bias + c4);
    for (long long p = (long long)blockIdx.x * per + sub; p < npix; p += (long long)gridDim.x * per) {
        const int x = (int)(p % d.w), y = (int)((p / d.w) % d.h), img = (int)(p / ((long long)d.w * d.h));
        float4 z = ld_stream_f4(a.y0 + p * d.C + c4);
#pragma unroll
        for (int l = 0; l < VFD_AGG_LEVELS; ++l) {
            if (l >= d.n_coarse) break;
            const int ch = d.ch[l], cw = d.cw[l];
            const AggTap ty = agg_tap(y, ch, agg_scale(ch, d.h)), tx = agg_tap(x, cw, agg_scale(cw, d.w));
            const float* s = a.yc[l] + (size_t)img * ch * cw * d.C + c4;
            const float4 v00 = ldg_f4(s + ((size_t)ty.i0 * cw + tx.i0) * d.C), v01 = ldg_f4(s + ((size_t)ty.i0 * cw + tx.i1) * d.C);
            const float4 v10 = ldg_f4(s + ((size_t)ty.i1 * cw + tx.i0) * d.C), v11 = ldg_f4(s + ((size_t)ty.i1 * cw + tx.i1) * d.C);
            float4 r0 = f4_scale(tx.l0, v00), r1 = f4_scale(tx.l0, v10);
            f4_fma(r0, tx.l1, v01);
            f4_fma(r1, tx.l1, v11);
            f4_fma(z, ty.l0, r0);
            f4_fma(z, ty.l1, r1);
        }
        z.x += b.x; z.y += b.y; z.z += b.z; z.w += b.w;
        z.x = z.x > 0.f ? z.x : d.slope * z.x;
        z.y = z.y > 0.f ? z.y : d.slope * z.y;
        z.z = z.z > 0.f ? z.z : d.slope * z.z;
        z.w = z.w > 0.f ? z.w : d.slope * z.w;
        *reinterpret_cast<float4*>(a.out + p * d.C + c4) = z;
    }
}

// dZ = d_out * (out > 0 ? 1 : slope)  -> the gradient of the finest projection (and of the bias, summed by the caller's
// GEMM with a column of ones)
__global__ void __launch_bounds__(kAggThreads) feat_agg_bwd_fine_kernel(const AggArgs a) {
    const vfd_agg_desc& d = a.d;
    const long long n4 = (long long)d.n_img * d.h * d.w * d.C / 4;
    for (long long i = (long long)blockIdx.x * kAggThreads + threadIdx.x; i < n4; i += (long long)gridDim.x * kAggThreads) {
        const float4 o = ld_stream_f4(a.out + i * 4), g = ld_stream_f4(a.d_out + i * 4);
        float4 z;
        z.x = o.x > 0.f ? g.x : d.slope * g.x;
        z.y = o.y > 0.f ? g.y : d.slope * g.y;
        z.z = o.z > 0.f ? g.z : d.slope * g.z;
        z.w = o.w > 0.f ? g.w : d.slope * g.w;
        *reinterpret_cast<float4*>(a.d_y0 + i * 4) = z;
    }
}

// fine indices whose taps can touch coarse index qi: scale * dst in (qi - 1, qi + 1)
__device__ __forceinline__ void agg_range(int qi, float scale, int out, int& lo, int& hi) {
    if (scale <= 0.f) { lo = 0; hi = out - 1; return; }
    lo = (int)floorf((float)(qi - 1) / scale) - 1;      // one spare index either side: rounding of scale * dst
    hi = (int)ceilf((float)(qi + 1) / scale) + 1;
    lo = lo < 0 ? 0 : lo;
    hi = hi > out - 1 ? out - 1 : hi;
}

__device__ __forceinline__ float agg_weight(int dst, int in, float scale, int qi) {
    const AggTap t = agg_tap(dst, in, scale);
    return (t.i0 == qi ? t.l0 : 0.f) + (t.i1 == qi ? t.l1 : 0.f);
}

__global__ void __launch_bounds__(kAggThreads) feat_agg_bwd_coarse_kernel(const AggArgs a) {
    const vfd_agg_desc& d = a.d;
    const int ch = a.lch, cw = a.lcw;
    const int q = d.C / 4, per = kAggThreads / q;
    const int c4 = (threadIdx.x % q) * 4, sub = threadIdx.x / q;
    if (sub >= per) return;
    const float sy = agg_scale(ch, d.h), sx = agg_scale(cw, d.w);
    const long long npix = (long long)d.n_img * ch * cw;
    for (long long p = (long long)blockIdx.x * per + sub; p < npix; p += (long long)gridDim.x * per) {
        const int qx = (int)(p % cw), qy = (int)((p / cw) % ch), img = (int)(p / ((long long)cw * ch));
        int y0, y1, x0, x1;
        agg_range(qy, sy, d.h, y0, y1);
        agg_range(qx, sx, d.w, x0, x1);
        const float* g = a.d_y0 + (size_t)img * d.h * d.w * d.C + c4;
        float4 acc = f4_zero();
        for (int y = y0; y <= y1; ++y) {
            const float wy = agg_weight(y, ch, sy, qy);
            if (wy == 0.f) continue;
            float4 row = f4_zero();
            for (int x = x0; x <= x1; ++x) {
                const float wx = agg_weight(x, cw, sx, qx);
                if (wx != 0.f) f4_fma(row, wx, ldg_f4(g + ((size_t)y * d.w + x) * d.C));
            }
            f4_fma(acc, wy, row);
        }
        *reinterpret_cast<float4*>(a.d_yc + p * d.C + c4) = acc;
    }
}

static int check_agg(const vfd_agg_desc* d) {
    if (!d) return VFD_ERR_NULL;
    if (d->n_img < 1 || d->h < 1 || d->w < 1 || d->C < 4 || d->C % 4 || d->C > 4 * kAggThreads) return VFD_ERR_DESC;
    if (d->n_coarse < 0 || d->n_coarse > VFD_AGG_LEVELS) return VFD_ERR_DESC;
    for (int l = 0; l < d->n_coarse; ++l)
        if (d->ch[l] < 1 || d->cw[l] < 1 || d->ch[l] > d->h || d->cw[l] > d->w) return VFD_ERR_DESC;
    return VFD_OK;
}

static int agg_grid(long long pixels, int C) {
    const int per = kAggThreads / (C / 4);
    long long blocks = (pixels + per - 1) / per;
    if (blocks > kAggMaxBlocks) blocks = kAggMaxBlocks;
    return blocks < 1 ? 1 : (int)blocks;
}

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_feat_agg_fwd(const vfd_agg_desc* d, const float* y0, const float* const* yc, const float* bias,
                                float* out, void* stream) {
    int rc = check_agg(d);
    if (rc) return rc;
    if (!y0 || !bias || !out || (d->n_coarse > 0 && !yc)) return VFD_ERR_NULL;
    AggArgs a = {};
    a.d = *d; a.y0 = y0; a.bias = bias; a.out = out;
    for (int l = 0; l < d->n_coarse; ++l) {
        if (!yc[l]) return VFD_ERR_NULL;
        if (reinterpret_cast<uintptr_t>(yc[l]) & 15) return VFD_ERR_ALIGN;
        a.yc[l] = yc[l];
    }
    if ((reinterpret_cast<uintptr_t>(y0) & 15) || (reinterpret_cast<uintptr_t>(out) & 15) ||
        (reinterpret_cast<uintptr_t>(bias) & 15))
        return VFD_ERR_ALIGN;
    feat_agg_fwd_kernel<<<agg_grid((long long)d->n_img * d->h * d->w, d->C), kAggThreads, 0, (cudaStream_t)stream>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_feat_agg_bwd(const vfd_agg_desc* d, const float* out, const float* d_out, float* d_y0,
                                float* const* d_yc, void* stream) {
    int rc = check_agg(d);
    if (rc) return rc;
    if (!out || !d_out || !d_y0 || (d->n_coarse > 0 && !d_yc)) return VFD_ERR_NULL;
    for (int l = 0; l < d->n_coarse; ++l) {
        if (!d_yc[l]) return VFD_ERR_NULL;
        if (reinterpret_cast<uintptr_t>(d_yc[l]) & 15) return VFD_ERR_ALIGN;
    }
    if ((reinterpret_cast<uintptr_t>(out) & 15) || (reinterpret_cast<uintptr_t>(d_out) & 15) ||
        (reinterpret_cast<uintptr_t>(d_y0) & 15))
        return VFD_ERR_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    AggArgs a = {};
    a.d = *d; a.out = const_cast<float*>(out); a.d_out = d_out; a.d_y0 = d_y0;
    const long long n4 = (long long)d->n_img * d->h * d->w * d->C / 4;
    long long blocks = (n4 + kAggThreads * 4 - 1) / (kAggThreads * 4);
    const long long cap = kAggMaxBlocks;
    feat_agg_bwd_fine_kernel<<<(int)(blocks > cap ? cap : (blocks < 1 ? 1 : blocks)), kAggThreads, 0, st>>>(a);
    for (int l = 0; l < d->n_coarse; ++l) {
        a.lch = d->ch[l];
        a.lcw = d->cw[l];
        a.d_yc = d_yc[l];
        feat_agg_bwd_coarse_kernel<<<agg_grid((long long)d->n_img * d->ch[l] * d->cw[l], d->C), kAggThreads, 0, st>>>(a);
    }
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
