// Stand-alone view-synthesis warp: the sub-method surface of the reference that the fused loss kernels bypass
// (SURVEY.md 8b):  Projection.backproject / .reproject / .forward  (models/geometry/geometry_util.py:52-82) and
// ViewRendering.get_virtual_image (models/geometry/view_rendering.py:61-82) for ONE (source, target) pair per batch
// element, with the same bit-exact coordinate helpers (vfd_coords.h) and sampler (warp_common.cuh) the rig kernels use.
// Backward: gradients w.r.t. the target depth and the pre-multiplied projection P = (K_src @ T)[:3] (the host turns dP
// into d_T / d_K with two 4x4 products); per-CTA partial sums of dP, added by the host in CTA order (deterministic).
#include "warp_common.cuh"

namespace vfd {

constexpr int kWiThreads = 256;

struct WarpImageArgs {
    int B, H, W;
    const float* src_img;   // [B,3,H,W] or NULL (grid only)
    const float* src_mask;  // [B,1,H,W] or NULL
    const float* depth;     // [B,1,H,W]
    const float* invK;      // [B,4,4] target
    const float* K;         // [B,4,4] source
    const float* T;         // [B,4,4] target -> source
    float* img;             // [B,3,H,W]
    float* mask;            // [B,1,H,W]
    float* grid;            // [B,H,W,2] or NULL
    // backward
    const float* d_img;     // [B,3,H,W] or NULL
    const float* d_grid;    // [B,H,W,2] or NULL
    float* d_depth;         // [B,1,H,W]
    float* dP_part;         // [B,gridDim.x,12]
};

__device__ __forceinline__ void wi_load_mats(const WarpImageArgs& a, int b, float* s_P, float* s_invK) {
    if (threadIdx.x == 0) {
        float K[16], T[16], P[12];
#pragma unroll
        for (int i = 0; i < 16; ++i) { K[i] = __ldg(a.K + (size_t)b * 16 + i); T[i] = __ldg(a.T + (size_t)b * 16 + i); }
        vfd_premul_KT(K, T, P);
#pragma unroll
        for (int i = 0; i < 12; ++i) s_P[i] = P[i];
    }
    if (threadIdx.x >= 32 && threadIdx.x < 44) s_invK[threadIdx.x - 32] = __ldg(a.invK + (size_t)b * 16 + threadIdx.x - 32);
}

__global__ void __launch_bounds__(kWiThreads) warp_image_fwd_kernel(const WarpImageArgs a) {
    __shared__ float s_P[12], s_invK[12];
    const int b = blockIdx.y, HW = a.H * a.W;
    wi_load_mats(a, b, s_P, s_invK);
    __syncthreads();
    const float* dep = a.depth + (size_t)b * HW;
    for (int p = blockIdx.x * kWiThreads + threadIdx.x; p < HW; p += gridDim.x * kWiThreads) {
        const int x = p % a.W, y = p / a.W;
        float X[3];
        vfd_backproject(s_invK, (float)x, (float)y, __ldg(dep + p), X);
        if (a.src_img) {
            const WarpSample s = warp_sample<false>(s_P, a.src_img + (size_t)b * 3 * HW, a.src_mask + (size_t)b * HW, X,
                                                    a.W, a.H);
            float* o = a.img + (size_t)b * 3 * HW + p;
            o[0] = s.v[0]; o[HW] = s.v[1]; o[2 * (size_t)HW] = s.v[2];
            a.mask[(size_t)b * HW + p] = s.m;
            if (a.grid) reinterpret_cast<float2*>(a.grid)[(size_t)b * HW + p] = make_float2(s.gx, s.gy);
        } else {
            float cam[3], g[2];
            vfd_reproject(s_P, X, a.W, a.H, cam, g);
            reinterpret_cast<float2*>(a.grid)[(size_t)b * HW + p] = make_float2(g[0], g[1]);
        }
    }
}

// points3D [B,4,HW] = depth * (invK[:3,:3] @ (u,v,1)) with a row of ones  (geometry_util.py:52-60)
__global__ void __launch_bounds__(kWiThreads) backproject_kernel(const WarpImageArgs a, float* pts) {
    __shared__ float s_invK[12];
    const int b = blockIdx.y, HW = a.H * a.W;
    if (threadIdx.x < 12) s_invK[threadIdx.x] = __ldg(a.invK + (size_t)b * 16 + threadIdx.x);
    __syncthreads();
    for (int p = blockIdx.x * kWiThreads + threadIdx.x; p < HW; p += gridDim.x * kWiThreads) {
        float X[3];
        vfd_backproject(s_invK, (float)(p % a.W), (float)(p / a.W), __ldg(a.depth + (size_t)b * HW + p), X);
        float* o = pts + (size_t)b * 4 * HW + p;
        o[0] = X[0]; o[HW] = X[1]; o[2 * (size_t)HW] = X[2]; o[3 * (size_t)HW] = 1.f;
    }
}

// pix_coords [B,H,W,2] from points3D [B,4,HW] (4th row = 1, as backproject writes it)  (geometry_util.py:62-77)
__global__ void __launch_bounds__(kWiThreads) reproject_kernel(const WarpImageArgs a, const float* pts) {
    __shared__ float s_P[12], s_invK[12];
    const int b = blockIdx.y, HW = a.H * a.W;
    wi_load_mats(a, b, s_P, s_invK);
    __syncthreads();
    for (int p = blockIdx.x * kWiThreads + threadIdx.x; p < HW; p += gridDim.x * kWiThreads) {
        const float* q = pts + (size_t)b * 4 * HW + p;
        const float X[3] = {__ldg(q), __ldg(q + HW), __ldg(q + 2 * (size_t)HW)};
        float cam[3], g[2];
        vfd_reproject(s_P, X, a.W, a.H, cam, g);
        reinterpret_cast<float2*>(a.grid)[(size_t)b * HW + p] = make_float2(g[0], g[1]);
    }
}

__global__ void __launch_bounds__(kWiThreads) warp_image_bwd_kernel(const WarpImageArgs a) {
    __shared__ float s_P[12], s_invK[12];
    __shared__ float s_red[kWiThreads / 32][12];
    const int b = blockIdx.y, HW = a.H * a.W, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    wi_load_mats(a, b, s_P, s_invK);
    __syncthreads();
    const float* dep = a.depth + (size_t)b * HW;
    float dP[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) dP[i] = 0.f;
    for (int p = blockIdx.x * kWiThreads + threadIdx.x; p < HW; p += gridDim.x * kWiThreads) {
        const int x = p % a.W, y = p / a.W;
        float ray[3], X[3];
        vfd_pixel_ray(s_invK, (float)x, (float)y, ray);
        const float d = __ldg(dep + p);
        X[0] = VFD_MUL(d, ray[0]); X[1] = VFD_MUL(d, ray[1]); X[2] = VFD_MUL(d, ray[2]);
        float gix = 0.f, giy = 0.f, c0, c1, den;      // gradient w.r.t. cam.x / den and cam.y / den
        bool live = true;
        if (a.d_img) {
            const WarpSample s = warp_sample<true>(s_P, a.src_img + (size_t)b * 3 * HW, a.src_mask + (size_t)b * HW, X,
                                                   a.W, a.H);
            c0 = s.cam[0]; c1 = s.cam[1]; den = s.den;
            if (!s.nan) {
                const float* g = a.d_img + (size_t)b * 3 * HW + p;
                const float up[3] = {__ldg(g), __ldg(g + HW), __ldg(g + 2 * (size_t)HW)};
#pragma unroll
                for (int c = 0; c < 3; ++c) { gix = fmaf(up[c], s.dvx[c], gix); giy = fmaf(up[c], s.dvy[c], giy); }
            }
            live = !isnan(s.gx) && !isnan(s.gy);
        } else {
            float cam[3], g[2];
            vfd_reproject(s_P, X, a.W, a.H, cam, g);
            c0 = cam[0]; c1 = cam[1]; den = VFD_ADD(cam[2], 1e-7f);
        }
        if (a.d_grid) {
            const float2 gg = __ldg(reinterpret_cast<const float2*>(a.d_grid) + (size_t)b * HW + p);
            gix = fmaf(gg.x, 2.f / (float)(a.W - 1), gix);
            giy = fmaf(gg.y, 2.f / (float)(a.H - 1), giy);
        }
        float gd = 0.f;
        if (live && (gix != 0.f || giy != 0.f)) {
            const float inv = 1.f / den;
            const float dc[3] = {gix * inv, giy * inv, -(gix * c0 + giy * c1) * inv * inv};
            const float* P = s_P;
            // the depth-proportional parts cancel exactly in d(ix)/d(depth) (see warp_loss_bwd.cu)
            const float a0 = P[0] * ray[0] + P[1] * ray[1] + P[2] * ray[2];
            const float a1 = P[4] * ray[0] + P[5] * ray[1] + P[6] * ray[2];
            const float a2 = P[8] * ray[0] + P[9] * ray[1] + P[10] * ray[2];
            const float t2 = P[11] + 1e-7f;
            gd = (gix * (a0 * t2 - P[3] * a2) + giy * (a1 * t2 - P[7] * a2)) * inv * inv;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                dP[r * 4 + 0] = fmaf(dc[r], X[0], dP[r * 4 + 0]);
                dP[r * 4 + 1] = fmaf(dc[r], X[1], dP[r * 4 + 1]);
                dP[r * 4 + 2] = fmaf(dc[r], X[2], dP[r * 4 + 2]);
                dP[r * 4 + 3] += dc[r];
            }
        }
        a.d_depth[(size_t)b * HW + p] = gd;
    }
#pragma unroll
    for (int i = 0; i < 12; ++i) {
        float v = dP[i];
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_red[warp][i] = v;
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < kWiThreads / 32; ++w) v += s_red[w][threadIdx.x];
        a.dP_part[((size_t)b * gridDim.x + blockIdx.x) * 12 + threadIdx.x] = v;
    }
}

constexpr int kWiBlocks = 148;      // CTAs per batch element (fixed: the size of dP_part depends on it)

static int check_wi(int B, int H, int W) {
    if (B < 1 || B > 65535 || H < 2 || W < 2 || (long)H * W >= (1L << 30)) return VFD_ERR_DESC;
    return VFD_OK;
}

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_warp_image_blocks(void) { return kWiBlocks; }

extern "C" int vfd_warp_image_fwd(int B, int H, int W, const float* src_img, const float* src_mask, const float* depth,
                                  const float* invK, const float* K, const float* T, float* img, float* mask,
                                  float* grid, void* stream) {
    int rc = check_wi(B, H, W);
    if (rc) return rc;
    if (!depth || !invK || !K || !T) return VFD_ERR_NULL;
    if (src_img ? (!src_mask || !img || !mask) : !grid) return VFD_ERR_NULL;
    if (grid && (reinterpret_cast<uintptr_t>(grid) & 7)) return VFD_ERR_ALIGN;
    WarpImageArgs a = {};
    a.B = B; a.H = H; a.W = W; a.src_img = src_img; a.src_mask = src_mask; a.depth = depth; a.invK = invK; a.K = K;
    a.T = T; a.img = img; a.mask = mask; a.grid = grid;
    warp_image_fwd_kernel<<<dim3(kWiBlocks, B), kWiThreads, 0, (cudaStream_t)stream>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_warp_image_bwd(int B, int H, int W, const float* src_img, const float* src_mask, const float* depth,
                                  const float* invK, const float* K, const float* T, const float* d_img,
                                  const float* d_grid, float* d_depth, float* dP_part, void* stream) {
    int rc = check_wi(B, H, W);
    if (rc) return rc;
    if (!depth || !invK || !K || !T || !d_depth || !dP_part || (!d_img && !d_grid)) return VFD_ERR_NULL;
    if (d_img && (!src_img || !src_mask)) return VFD_ERR_NULL;
    if (d_grid && (reinterpret_cast<uintptr_t>(d_grid) & 7)) return VFD_ERR_ALIGN;
    WarpImageArgs a = {};
    a.B = B; a.H = H; a.W = W; a.src_img = src_img; a.src_mask = src_mask; a.depth = depth; a.invK = invK; a.K = K;
    a.T = T; a.d_img = d_img; a.d_grid = d_grid; a.d_depth = d_depth; a.dP_part = dP_part;
    warp_image_bwd_kernel<<<dim3(kWiBlocks, B), kWiThreads, 0, (cudaStream_t)stream>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_projection_backproject(int B, int H, int W, const float* depth, const float* invK, float* points, void* stream) {
    int rc = check_wi(B, H, W);
    if (rc) return rc;
    if (!depth || !invK || !points) return VFD_ERR_NULL;
    WarpImageArgs a = {};
    a.B = B; a.H = H; a.W = W; a.depth = depth; a.invK = invK;
    backproject_kernel<<<dim3(kWiBlocks, B), kWiThreads, 0, (cudaStream_t)stream>>>(a, points);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_projection_reproject(int B, int H, int W, const float* points, const float* K, const float* T, float* grid,
                             void* stream) {
    int rc = check_wi(B, H, W);
    if (rc) return rc;
    if (!points || !K || !T || !grid) return VFD_ERR_NULL;
    if (reinterpret_cast<uintptr_t>(grid) & 7) return VFD_ERR_ALIGN;
    WarpImageArgs a = {};
    a.B = B; a.H = H; a.W = W; a.K = K; a.T = T; a.invK = K; a.grid = grid;
    reproject_kernel<<<dim3(kWiBlocks, B), kWiThreads, 0, (cudaStream_t)stream>>>(a, points);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

// ---------------------------------------------------------------------------------------------------------
// ViewRendering.get_norm_image_single (view_rendering.py:39-59) for one warped image per batch element:
//   mask = (src_mask * warp_mask) != 0 on 3 channels; if ANY batch element has an empty mask the warp is returned as is;
//   mean = masked mean, var = sum over ALL pixels of (x - mean)^2 / (3 H W), std = sqrt(var + 1e-16)   (get_mean_std :30-37)
//   out = ((warp - w_mean) / (w_std + 1e-8) * s_std + s_mean) * warp_mask.
// The statistics are constants for autograd (torch.no_grad in the reference); sums in f64.
// ---------------------------------------------------------------------------------------------------------
namespace vfd {

struct NormArgs {
    int B, H, W;
    const float* ref;        // [B,3,H,W] target image (the reference's src_img argument)
    const float* ref_mask;   // [B,1,H,W]
    const float* warp;       // [B,3,H,W]
    const float* warp_mask;  // [B,1,H,W]
    double* stats;           // [B,8]: count, sum ref*m, sum warp*m, sum warp, sum warp^2, sum ref, sum ref^2, -
    float* out;              // [B,3,H,W]
    const float* d_out;
    float* d_warp;
};

__global__ void __launch_bounds__(kWiThreads) norm_stats_kernel(const NormArgs a) {
    __shared__ double s_red[kWiThreads / 32][7];
    const int b = blockIdx.y, HW = a.H * a.W, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double acc[7] = {0, 0, 0, 0, 0, 0, 0};
    for (int p = blockIdx.x * kWiThreads + threadIdx.x; p < HW; p += gridDim.x * kWiThreads) {
        const float m = (__ldg(a.ref_mask + (size_t)b * HW + p) * __ldg(a.warp_mask + (size_t)b * HW + p)) != 0.f ? 1.f : 0.f;
        float sr = 0.f, sw = 0.f, sr2 = 0.f, sw2 = 0.f;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float r = __ldg(a.ref + ((size_t)b * 3 + c) * HW + p), w = __ldg(a.warp + ((size_t)b * 3 + c) * HW + p);
            sr += r; sw += w; sr2 = fmaf(r, r, sr2); sw2 = fmaf(w, w, sw2);
        }
        acc[0] += 3.0 * m; acc[1] += (double)(sr * m); acc[2] += (double)(sw * m);
        acc[3] += sw; acc[4] += sw2; acc[5] += sr; acc[6] += sr2;
    }
#pragma unroll
    for (int i = 0; i < 7; ++i) {
        double v = acc[i];
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_red[warp][i] = v;
    }
    __syncthreads();
    if (threadIdx.x < 7) {
        double v = 0;
#pragma unroll
        for (int w = 0; w < kWiThreads / 32; ++w) v += s_red[w][threadIdx.x];
        atomicAdd(a.stats + (size_t)b * 8 + threadIdx.x, v);
    }
}

// {w_mean, 1 / (w_std + 1e-8), s_std, s_mean} of batch element b; apply = 0 when any element of the batch has no overlap
__device__ __forceinline__ bool norm_params(const NormArgs& a, int b, float* prm) {
    bool apply = true;
    for (int i = 0; i < a.B; ++i) apply = apply && a.stats[(size_t)i * 8] > 0.0;
    const double* s = a.stats + (size_t)b * 8;
    const double n = 3.0 * a.H * a.W, cnt = s[0] + 1e-8;
    const double wm = s[2] / cnt, sm = s[1] / cnt;
    const double wv = (s[4] - 2.0 * wm * s[3] + n * wm * wm) / n, sv = (s[6] - 2.0 * sm * s[5] + n * sm * sm) / n;
    prm[0] = (float)wm;
    prm[1] = (float)(1.0 / (sqrt((wv > 0 ? wv : 0) + 1e-16) + 1e-8));
    prm[2] = (float)sqrt((sv > 0 ? sv : 0) + 1e-16);
    prm[3] = (float)sm;
    return apply;
}

template <bool BWD>
__global__ void __launch_bounds__(kWiThreads) norm_apply_kernel(const NormArgs a) {
    __shared__ float s_prm[4];
    __shared__ int s_apply;
    const int b = blockIdx.y, HW = a.H * a.W;
    if (threadIdx.x == 0) s_apply = norm_params(a, b, s_prm) ? 1 : 0;
    __syncthreads();
    const bool apply = s_apply != 0;
    const float wm = s_prm[0], iw = s_prm[1], ss = s_prm[2], sm = s_prm[3];
    for (int p = blockIdx.x * kWiThreads + threadIdx.x; p < HW; p += gridDim.x * kWiThreads) {
        const float m = __ldg(a.warp_mask + (size_t)b * HW + p);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const size_t i = ((size_t)b * 3 + c) * HW + p;
            if (BWD) a.d_warp[i] = apply ? __ldg(a.d_out + i) * (iw * ss) * m : __ldg(a.d_out + i);
            else a.out[i] = apply ? ((__ldg(a.warp + i) - wm) * iw * ss + sm) * m : __ldg(a.warp + i);
        }
    }
}

}  // namespace vfd

extern "C" int vfd_norm_image_fwd(int B, int H, int W, const float* ref_img, const float* ref_mask, const float* warp_img,
                                  const float* warp_mask, double* stats, float* out, void* stream) {
    int rc = check_wi(B, H, W);
    if (rc) return rc;
    if (!ref_img || !ref_mask || !warp_img || !warp_mask || !stats || !out) return VFD_ERR_NULL;
    if (reinterpret_cast<uintptr_t>(stats) & 7) return VFD_ERR_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(stats, 0, sizeof(double) * 8 * B, st) != cudaSuccess) return VFD_ERR_CUDA;
    NormArgs a = {};
    a.B = B; a.H = H; a.W = W; a.ref = ref_img; a.ref_mask = ref_mask; a.warp = warp_img; a.warp_mask = warp_mask;
    a.stats = stats; a.out = out;
    norm_stats_kernel<<<dim3(kWiBlocks, B), kWiThreads, 0, st>>>(a);
    norm_apply_kernel<false><<<dim3(kWiBlocks, B), kWiThreads, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_norm_image_bwd(int B, int H, int W, const float* warp_mask, const double* stats, const float* d_out,
                                  float* d_warp, void* stream) {
    int rc = check_wi(B, H, W);
    if (rc) return rc;
    if (!warp_mask || !stats || !d_out || !d_warp) return VFD_ERR_NULL;
    NormArgs a = {};
    a.B = B; a.H = H; a.W = W; a.warp_mask = warp_mask; a.stats = const_cast<double*>(stats); a.d_out = d_out;
    a.d_warp = d_warp;
    norm_apply_kernel<true><<<dim3(kWiBlocks, B), kWiThreads, 0, (cudaStream_t)stream>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
