// Stand-alone SSIM / photometric loss of two images (models/losses/loss_util.py:43-78): the free functions
// compute_ssim_loss(pred, target) -> [B,3,H,W] and compute_photometric_loss(pred, target) -> [B,1,H,W] of the
// reference, for callers outside the fused rig kernels (which evaluate the same arithmetic for eight images at once,
// warp_ssim.cuh).  Same rounding: 3x3 reflect-padded window sums tap by tap in row-major order, un-fused, correctly
// rounded division by 9 and 3.  Backward w.r.t. pred: every pixel scatters its window's coefficients (atomics).
#include "warp_ssim.cuh"

namespace vfd {

constexpr int kPhThreads = 256, kPhBlocks = 148 * 2;

struct PhotoArgs {
    int B, H, W;
    const float* pred;     // [B,3,H,W]
    const float* target;   // [B,3,H,W]
    float* ssim;           // [B,3,H,W] or NULL   clamp((1 - ssim) / 2, 0, 1)
    float* photo;          // [B,1,H,W] or NULL   0.85 * mean_c ssim + 0.15 * mean_c |t - p|
    const float* d_ssim;   // backward upstreams (either may be NULL)
    const float* d_photo;
    float* d_pred;         // [B,3,H,W], zero-filled by the entry point
};

struct SsimPix {
    float l, raw, ssim, inv, n1, n2, d1, d2, mu_x, mu_y;
};

__device__ __forceinline__ SsimPix ssim_pixel(const float* x, const float* y) {
    float sx = 0.f, sy = 0.f, sxx = 0.f, syy = 0.f, sxy = 0.f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        sx = __fadd_rn(sx, x[k]);
        sy = __fadd_rn(sy, y[k]);
        sxx = __fadd_rn(sxx, __fmul_rn(x[k], x[k]));
        syy = __fadd_rn(syy, __fmul_rn(y[k], y[k]));
        sxy = __fadd_rn(sxy, __fmul_rn(x[k], y[k]));
    }
    SsimPix s;
    s.mu_x = div_const<9>(sx);
    s.mu_y = div_const<9>(sy);
    const float musq_x = __fmul_rn(s.mu_x, s.mu_x), musq_y = __fmul_rn(s.mu_y, s.mu_y), mu_xy = __fmul_rn(s.mu_x, s.mu_y);
    const float sig_x = __fsub_rn(div_const<9>(sxx), musq_x), sig_y = __fsub_rn(div_const<9>(syy), musq_y);
    const float sig_xy = __fsub_rn(div_const<9>(sxy), mu_xy);
    s.n1 = __fadd_rn(__fmul_rn(2.f, mu_xy), kC1);
    s.n2 = __fadd_rn(__fmul_rn(2.f, sig_xy), kC2);
    s.d1 = __fadd_rn(__fadd_rn(musq_x, musq_y), kC1);
    s.d2 = __fadd_rn(__fadd_rn(sig_x, sig_y), kC2);
    const float den = __fadd_rn(__fmul_rn(s.d1, s.d2), 1e-8f);
    s.inv = __frcp_rn(den);
    s.ssim = __fdiv_rn(__fmul_rn(s.n1, s.n2), den);
    s.raw = __fmul_rn(__fsub_rn(1.f, s.ssim), 0.5f);
    s.l = nan_clamp01(s.raw);      // torch.clamp keeps NaN (loss_util.py:66)
    return s;
}

__device__ __forceinline__ void load_window(const float* plane, int x, int y, int W, int H, float* v, int* idx) {
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
        for (int dx = -1; dx <= 1; ++dx) {
            const int k = (dy + 1) * 3 + dx + 1;
            const int i = reflect1(y + dy, H) * W + reflect1(x + dx, W);
            if (idx) idx[k] = i;
            v[k] = __ldg(plane + i);
        }
}

__global__ void __launch_bounds__(kPhThreads) photometric_fwd_kernel(const PhotoArgs a) {
    const int b = blockIdx.y, HW = a.H * a.W;
    for (int p = blockIdx.x * kPhThreads + threadIdx.x; p < HW; p += gridDim.x * kPhThreads) {
        const int x = p % a.W, y = p / a.W;
        float ssum = 0.f, lsum = 0.f;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const size_t o = ((size_t)b * 3 + c) * HW;
            float xv[9], yv[9];
            load_window(a.pred + o, x, y, a.W, a.H, xv, nullptr);
            load_window(a.target + o, x, y, a.W, a.H, yv, nullptr);
            const SsimPix s = ssim_pixel(xv, yv);
            if (a.ssim) a.ssim[o + p] = s.l;
            ssum = __fadd_rn(ssum, s.l);
            lsum = __fadd_rn(lsum, fabsf(__fsub_rn(yv[4], xv[4])));
        }
        if (a.photo)
            a.photo[(size_t)b * HW + p] = __fadd_rn(__fmul_rn(0.85f, div_const<3>(ssum)), __fmul_rn(0.15f, div_const<3>(lsum)));
    }
}

__global__ void __launch_bounds__(kPhThreads) photometric_bwd_kernel(const PhotoArgs a) {
    const int b = blockIdx.y, HW = a.H * a.W;
    for (int p = blockIdx.x * kPhThreads + threadIdx.x; p < HW; p += gridDim.x * kPhThreads) {
        const int x = p % a.W, y = p / a.W;
        const float gp = a.d_photo ? __ldg(a.d_photo + (size_t)b * HW + p) : 0.f;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const size_t o = ((size_t)b * 3 + c) * HW;
            const float gl = (a.d_ssim ? __ldg(a.d_ssim + o + p) : 0.f) + gp * (0.85f / 3.f);   // upstream of l_c
            float xv[9], yv[9];
            int idx[9];
            load_window(a.pred + o, x, y, a.W, a.H, xv, idx);
            load_window(a.target + o, x, y, a.W, a.H, yv, nullptr);
            float* g = a.d_pred + o;
            if (gl != 0.f) {
                const SsimPix s = ssim_pixel(xv, yv);
                if (s.raw >= 0.f && s.raw <= 1.f) {          // torch.clamp passes the gradient on the closed interval
                    const float gi = -0.5f * gl * s.inv;
                    // d ssim * den / d (mean x, mean x^2, mean xy); n2 = 2 (E[xy] - mu_x mu_y) + C2, d2 = E[x^2] - mu_x^2 + ...
                    const float dn = 2.f * s.mu_y * (s.n2 - s.n1);
                    const float dd = 2.f * s.mu_x * (s.d2 - s.d1);
                    const float ca = gi * (dn - s.ssim * dd) / 9.f;
                    const float cb = gi * (-s.ssim * s.d1) / 9.f;
                    const float cc = gi * (2.f * s.n1) / 9.f;
#pragma unroll
                    for (int k = 0; k < 9; ++k) atomicAdd(g + idx[k], ca + 2.f * cb * xv[k] + cc * yv[k]);
                }
            }
            if (gp != 0.f) {
                const float df = yv[4] - xv[4];
                const float sg = df > 0.f ? -1.f : (df < 0.f ? 1.f : 0.f);
                if (sg != 0.f) atomicAdd(g + p, gp * (0.15f / 3.f) * sg);
            }
        }
    }
}

static int check_ph(int B, int H, int W) {
    if (B < 1 || B > 65535 || H < 2 || W < 2 || (long)H * W >= (1L << 30)) return VFD_ERR_DESC;
    return VFD_OK;
}

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_photometric_loss_fwd(int B, int H, int W, const float* pred, const float* target, float* ssim,
                                        float* photo, void* stream) {
    int rc = check_ph(B, H, W);
    if (rc) return rc;
    if (!pred || !target || (!ssim && !photo)) return VFD_ERR_NULL;
    PhotoArgs a = {};
    a.B = B; a.H = H; a.W = W; a.pred = pred; a.target = target; a.ssim = ssim; a.photo = photo;
    photometric_fwd_kernel<<<dim3(kPhBlocks, B), kPhThreads, 0, (cudaStream_t)stream>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_photometric_loss_bwd(int B, int H, int W, const float* pred, const float* target, const float* d_ssim,
                                        const float* d_photo, float* d_pred, void* stream) {
    int rc = check_ph(B, H, W);
    if (rc) return rc;
    if (!pred || !target || !d_pred || (!d_ssim && !d_photo)) return VFD_ERR_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(d_pred, 0, sizeof(float) * 3 * (size_t)B * H * W, st) != cudaSuccess) return VFD_ERR_CUDA;
    PhotoArgs a = {};
    a.B = B; a.H = H; a.W = W; a.pred = pred; a.target = target; a.d_ssim = d_ssim; a.d_photo = d_photo; a.d_pred = d_pred;
    photometric_bwd_kernel<<<dim3(kPhBlocks, B), kPhThreads, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
