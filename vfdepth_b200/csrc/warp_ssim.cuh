// SSIM + L1 photometric loss at one pixel from shared-memory tiles (models/losses/loss_util.py:43-78) for all eight
// images of a camera at once, on Blackwell's packed fp32 pipe.
//
// The eight images are kept as four float2 pairs per channel:
//   pair 0 = (color_-1, color_+1)   temporal warps                      single_cam_loss.py:23-30
//   pair 1 = (overlap_-1, overlap_+1) spatio-temporal                   multi_cam_loss.py:38-56
//   pair 2 = (overlap_0, frame -1)  spatial warp | identity image       multi_cam_loss.py:21-26, single_cam_loss.py:33-43
//   pair 3 = (frame +1, target)     identity image | the target itself
// so one LDS.64 fetches a tap of two images and add/mul.f32x2 (sm_100 packed fp32: two IEEE round-to-nearest results
// per instruction, bit-identical to the scalar ops) update two window sums.
//
// Rounding.  The 3x3 window sums run tap by tap in row-major order -- the order ATen's avg_pool2d uses -- and are divided
// by 9 with x / 9 and x / 3 computed as q = x*r; q += fma(-9, q, x) * r, the correctly rounded quotient for every fp32 x
// (checked exhaustively over all 2^23 mantissas, tools/check_div_const.py).  The sums of x^2 and x*y accumulate with
// FUSED multiply-adds (one rounding per tap where the reference rounds the product tensor first): ptxas contracts a
// mul.rn.f32x2 feeding an add.rn.f32x2 into FFMA2 whatever the rounding modifiers say (found in the SASS in round 2), so
// the source now says fma -- every kernel that evaluates the SSIM then carries the same bits (warp_loss_fwd / _bwd /
// _sets are compared with each other bit for bit), at 3 instead of 5 packed instructions per tap and pair.  The
// difference to the reference's un-fused sequence is below the fp32 noise floor of the SSIM itself (E[x^2] - mu^2 is a
// difference of nearly equal numbers: 5e-5 of the loss between ANY two fp32 evaluations, the reference on CPU and on
// CUDA included) and is covered by the parity bars (tests/test_warp_gpu.py).  Only the final ssim = num / den uses a
// fast reciprocal: it is 1-2 ulp (1e-7) on a quantity with that noise floor.
#pragma once
#include "warp_common.cuh"

namespace vfd {

// ---- packed fp32 (PTX ISA 8.6+, sm_100+) ----
__device__ __forceinline__ unsigned long long f2_bits(float2 a) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a.x), "f"(a.y));
    return r;
}
__device__ __forceinline__ float2 bits_f2(unsigned long long r) {
    float2 a;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a.x), "=f"(a.y) : "l"(r));
    return a;
}
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
    unsigned long long r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(f2_bits(a)), "l"(f2_bits(b)));
    return bits_f2(r);
}
__device__ __forceinline__ float2 sub2(float2 a, float2 b) {
    unsigned long long r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(f2_bits(a)), "l"(f2_bits(b)));
    return bits_f2(r);
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
    unsigned long long r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(f2_bits(a)), "l"(f2_bits(b)));
    return bits_f2(r);
}
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(f2_bits(a)), "l"(f2_bits(b)), "l"(f2_bits(c)));
    return bits_f2(r);
}
__device__ __forceinline__ float2 splat2(float v) { return make_float2(v, v); }
__device__ __forceinline__ float fast_rcp(float x) {   // MUFU.RCP, ~1 ulp
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// correctly rounded x / K for K = 9 or 3 (Markstein: r = RN(1/K), one residual correction)
template <int K>
__device__ __forceinline__ float2 div_const2(float2 x) {
    const float r = 1.0f / (float)K;
    float2 q = mul2(x, splat2(r));
    const float2 rem = fma2(splat2(-(float)K), q, x);
    return fma2(rem, splat2(r), q);
}
template <int K>
__device__ __forceinline__ float div_const(float x) {
    const float r = 1.0f / (float)K;
    const float q = __fmul_rn(x, r);
    return __fmaf_rn(__fmaf_rn(-(float)K, q, x), r, q);
}

constexpr float kC1 = (float)(0.01 * 0.01);   // python: 0.01**2 in double, cast to fp32 when it meets the tensor
constexpr float kC2 = (float)(0.03 * 0.03);
constexpr int kPairs = 4;                      // float2 image pairs per channel (see the table above)

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

// derivative coefficients of one image pair: d L / d (window mean of x, of x^2, of x*y), per channel
struct SsimGrad2 {
    float2 a[3], b[3], c[3];
};

// Photometric loss 0.85 * mean_c clamp((1 - ssim_c) / 2, 0, 1) + 0.15 * mean_c |t_c - p_c| of the 7 images against
// the target (loss_util.py:70-78).  img: float2 planes [3 channels][kPairs][plane].  L[q] = losses of pair q
// (L[3].y, target against itself, is meaningless).  g (GRAD): coefficients for pairs 0..2.
// SP = false: the tile holds no spatial / spatio-temporal warp at all (their images are zero and their masks too), so
// pair 1 is not evaluated (L[1] = 0, g[1] untouched) -- its losses would be multiplied by a zero mask anyway.
template <bool GRAD, bool SP = true>
__device__ __forceinline__ void photometric8(const float2* __restrict__ img, int plane, const int* idx, float2* L,
                                             SsimGrad2* g) {
    float2 ssum[kPairs], lsum[kPairs];
#pragma unroll
    for (int q = 0; q < kPairs; ++q) ssum[q] = lsum[q] = splat2(0.f);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float2* pc = img + c * kPairs * plane;
        float2 sx[kPairs], sxx[kPairs], sxy[kPairs], xc[kPairs];
#pragma unroll
        for (int q = 0; q < kPairs; ++q) sx[q] = sxx[q] = sxy[q] = splat2(0.f);
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            float2 x[kPairs];
#pragma unroll
            for (int q = 0; q < kPairs; ++q) x[q] = (SP || q != 1) ? pc[q * plane + idx[k]] : splat2(0.f);
            const float2 yy = splat2(x[3].y);
#pragma unroll
            for (int q = 0; q < kPairs; ++q) {
                if (!SP && q == 1) continue;
                sx[q] = add2(sx[q], x[q]);
                sxx[q] = fma2(x[q], x[q], sxx[q]);
                sxy[q] = fma2(x[q], yy, sxy[q]);
                if (k == 4) xc[q] = x[q];
            }
        }
        const float mu_y = div_const<9>(sx[3].y), e_yy = div_const<9>(sxx[3].y);
        const float musq_y = __fmul_rn(mu_y, mu_y), sig_y = __fsub_rn(e_yy, musq_y);
        const float2 yc = splat2(xc[3].y);
#pragma unroll
        for (int q = 0; q < kPairs; ++q) {
            if (!SP && q == 1) continue;
            const float2 mu = div_const2<9>(sx[q]);
            const float2 musq = mul2(mu, mu), mu_xy = mul2(mu, splat2(mu_y));
            const float2 sig_x = sub2(div_const2<9>(sxx[q]), musq);
            const float2 sig_xy = sub2(div_const2<9>(sxy[q]), mu_xy);
            // 2*v + C: 2*v is exact, so the fused form rounds once exactly like mul-then-add
            const float2 n1 = fma2(splat2(2.f), mu_xy, splat2(kC1)), n2 = fma2(splat2(2.f), sig_xy, splat2(kC2));
            const float2 d1 = add2(add2(musq, splat2(musq_y)), splat2(kC1));
            const float2 d2 = add2(add2(sig_x, splat2(sig_y)), splat2(kC2));
            const float2 den = add2(mul2(d1, d2), splat2(1e-8f));
            const float2 num = mul2(n1, n2);
            const float2 inv = make_float2(fast_rcp(den.x), fast_rcp(den.y));
            const float2 ssim = mul2(num, inv);
            const float2 raw = mul2(sub2(splat2(1.f), ssim), splat2(0.5f));
            const float2 l = make_float2(fminf(fmaxf(raw.x, 0.f), 1.f), fminf(fmaxf(raw.y, 0.f), 1.f));
            // torch.clamp keeps a NaN (loss_util.py:66) where fminf / fmaxf drop it: ssim * 0 is +-0 for every finite ssim
            // (l + +-0 == l exactly, l >= +0) and NaN for a NaN ssim, at the price of one packed fma
            ssum[q] = add2(ssum[q], fma2(ssim, splat2(0.f), l));
            const float2 df = sub2(yc, xc[q]);
            lsum[q] = add2(lsum[q], make_float2(fabsf(df.x), fabsf(df.y)));
            if (GRAD && q < 3) {
                // d l / d ssim = -1/2 inside the clamp (torch: gradient passes where 0 <= raw <= 1), then * 0.85 / 3
                const float gk = -0.5f * 0.85f / 3.f;
                const float2 gs = make_float2((raw.x >= 0.f && raw.x <= 1.f) ? gk : 0.f,
                                              (raw.y >= 0.f && raw.y <= 1.f) ? gk : 0.f);
                const float2 gi = mul2(gs, inv);
                // n2 = 2 (sxy - mu_x mu_y) + C2 ; d2 = (sxx - mu_x^2) + sig_y + C2
                const float2 dn = mul2(splat2(2.f * mu_y), sub2(n2, n1));
                const float2 dd = mul2(mul2(splat2(2.f), mu), sub2(d2, d1));
                g[q].a[c] = mul2(gi, sub2(dn, mul2(ssim, dd)));
                g[q].b[c] = mul2(gi, mul2(ssim, mul2(splat2(-1.f), d1)));
                g[q].c[c] = mul2(gi, mul2(splat2(2.f), n1));
            }
        }
    }
#pragma unroll
    for (int q = 0; q < kPairs; ++q) {
        if (!SP && q == 1) { L[q] = splat2(0.f); continue; }
        const float2 s_mean = div_const2<3>(ssum[q]), l_mean = div_const2<3>(lsum[q]);
        L[q] = add2(mul2(splat2(0.85f), s_mean), mul2(splat2(0.15f), l_mean));
    }
}

}  // namespace vfd
