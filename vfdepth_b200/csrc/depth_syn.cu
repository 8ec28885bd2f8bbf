// Depth-synthesis consistency loss of the aug_depth configuration, fused:
//   ViewRendering.get_virtual_depth  models/geometry/view_rendering.py:84-116  (called for the two neighbours and the
//                                    camera itself, :201-241)
//   DepthSynLoss.compute_aug_losses  models/losses/depth_synthesis_loss.py:15-33 (consistency part)
// For every pixel of camera c's augmented view (depth `aug`) and every source s in rel_cam_list[c] + [c]:
//   pix   = project(aug, inverse(T), K_c^-1, K_s)                 T = inverse(ext_aug[c]) @ ext[s]  (source -> novel view)
//   z_s   = (T[:3,:] @ backproject(K_s^-1, depth_s))_z            source depth expressed in the novel view, per source pixel
//   dw    = bilinear(z_s, pix), m = nearest(mask_s, pix); NaN -> 2.0 / 0; range handling against [min_depth, max_depth]
//   loss += clamp(|aug - dw| / (aug + dw + 1e-8), 0, 1) * m,  den += m          (compute_masked_loss over the sources)
// The reference materialises 3 x (points, warped depth, grid, masks) per camera; here nothing but the two sums per
// camera leaves the kernel (optionally the warped depths / masks the logger reads).  The backward recomputes the
// forward and returns d/d aug (one write per pixel) and d/d depth_s (bilinear scatter, red.global.add.f32).
#include "common.cuh"

namespace vfd {

constexpr int kSynThreads = 256;
constexpr int kSynSrc = VFD_SYN_SRC;

struct SynArgs {
    vfd_syn_desc d;
    const float* depth;   // [B,cams,H,W]
    const float* aug;     // [B,n_call,H,W]
    const float* T;       // [B,n_call,3,4,4]
    const float* Tinv;    // [B,n_call,3,4,4]
    double* sums;         // [n_call][2]
    float* out_depth;     // [B,n_call,3,H,W] or NULL
    float* out_mask;      // [B,n_call,3,H,W] or NULL
    // backward
    const float* wcam;    // [n_call] upstream gradient of num (already divided by den + 1e-8)
    float* d_aug;         // [B,n_call,H,W]
    float* d_depth;       // [B,cams,H,W]
};

struct SynCtx {
    float P[kSynSrc][12];     // (K_src @ Tinv)[:3,:]
    float Tz[kSynSrc][4];     // row 2 of T
    float invKs[kSynSrc][12]; // K_src^-1 rows 0..2
    float invK[12];           // K_c^-1
    int src[kSynSrc];
};

__device__ __forceinline__ void syn_ctx_init(SynCtx& c, const SynArgs& a, int b, int ci) {
    const vfd_syn_desc& d = a.d;
    const int cam = d.cam0 + ci;
    if (threadIdx.x < 12) c.invK[threadIdx.x] = __ldg(d.invK + (size_t)(b * d.num_cams + cam) * 16 + threadIdx.x);
    if (threadIdx.x < kSynSrc) {
        const int i = threadIdx.x;
        int s = d.src[cam][i];
        if (s < 0 || s >= d.num_cams) s = -1;
        c.src[i] = s;
        if (s >= 0) {
            float K[16], Ti[16];
            const float* Kp = d.K + (size_t)(b * d.num_cams + s) * 16;
            const float* Tp = a.Tinv + ((size_t)(b * d.n_call + ci) * kSynSrc + i) * 16;
#pragma unroll
            for (int k = 0; k < 16; ++k) { K[k] = __ldg(Kp + k); Ti[k] = __ldg(Tp + k); }
            float P[12];
            vfd_premul_KT(K, Ti, P);
#pragma unroll
            for (int k = 0; k < 12; ++k) c.P[i][k] = P[k];
            const float* Tf = a.T + ((size_t)(b * d.n_call + ci) * kSynSrc + i) * 16;
#pragma unroll
            for (int k = 0; k < 4; ++k) c.Tz[i][k] = __ldg(Tf + 8 + k);
            const float* iK = d.invK + (size_t)(b * d.num_cams + s) * 16;
#pragma unroll
            for (int k = 0; k < 12; ++k) c.invKs[i][k] = __ldg(iK + k);
        }
    }
}

// z of T[:3,:] @ [d * K^-1 (u,v,1); 1] at source pixel (u, v)
__device__ __forceinline__ float syn_z(const float* Tz, const float* invKs, const float* dsrc, int W, int u, int v,
                                       float* slope) {
    float ray[3];
    vfd_pixel_ray(invKs, (float)u, (float)v, ray);
    const float dd = __ldg(dsrc + v * W + u);
    if (slope) *slope = Tz[0] * ray[0] + Tz[1] * ray[1] + Tz[2] * ray[2];
    return vfd_dot4(Tz, VFD_MUL(dd, ray[0]), VFD_MUL(dd, ray[1]), VFD_MUL(dd, ray[2]), 1.0f);
}

struct SynSample {
    float dw, m;          // warped depth after NaN / range handling, final mask
    bool diff;            // dw still depends on the inputs (not replaced)
    // bilinear bookkeeping for the backward
    int x0, y0;
    float w[4], z[4], slope[4];
    bool ok[4];
    float tx0, tx1, ty0, ty1;
    float cam[3], den;
};

template <bool GRAD>
__device__ __forceinline__ SynSample syn_sample(const SynCtx& c, int i, const float* X, const float* dsrc,
                                                const float* msrc, int W, int H, float min_depth, float max_depth) {
    SynSample s;
    float g[2];
    vfd_reproject(c.P[i], X, W, H, s.cam, g);
    s.den = VFD_ADD(s.cam[2], 1e-7f);
    const float ix = vfd_unnormalize(g[0], W), iy = vfd_unnormalize(g[1], H);
    const bool inb = vfd_in_bounds2(g[0], g[1]);
    s.diff = false;
    float dw, mval = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) { s.ok[k] = false; s.w[k] = 0.f; s.z[k] = 0.f; s.slope[k] = 0.f; }
    if (isnan(ix) || isnan(iy)) {
        dw = 2.0f;                 // :104-105 (NaN sample), mask :106-107
    } else {
        const float rx = rintf(ix), ry = rintf(iy);
        if (rx >= 0.f && rx <= (float)(W - 1) && ry >= 0.f && ry <= (float)(H - 1))
            mval = __ldg(msrc + (size_t)(int)ry * W + (int)rx);
        const float x0f = floorf(ix), y0f = floorf(iy);
        dw = 0.f;
        if (x0f >= -1.f && x0f <= (float)(W - 1) && y0f >= -1.f && y0f <= (float)(H - 1)) {
            const int x0 = (int)x0f, y0 = (int)y0f;
            s.x0 = x0; s.y0 = y0;
            s.tx1 = (x0f + 1.f) - ix; s.tx0 = ix - x0f; s.ty1 = (y0f + 1.f) - iy; s.ty0 = iy - y0f;
            const bool vx0 = x0 >= 0, vx1 = x0 + 1 <= W - 1, vy0 = y0 >= 0, vy1 = y0 + 1 <= H - 1;
            s.ok[0] = vx0 && vy0; s.ok[1] = vx1 && vy0; s.ok[2] = vx0 && vy1; s.ok[3] = vx1 && vy1;
            s.w[0] = s.tx1 * s.ty1; s.w[1] = s.tx0 * s.ty1; s.w[2] = s.tx1 * s.ty0; s.w[3] = s.tx0 * s.ty0;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (s.ok[k])
                    s.z[k] = syn_z(c.Tz[i], c.invKs[i], dsrc, W, x0 + (k & 1), y0 + (k >> 1), GRAD ? &s.slope[k] : nullptr);
            dw = s.z[0] * s.w[0];
            dw = fmaf(s.z[1], s.w[1], dw);
            dw = fmaf(s.z[2], s.w[2], dw);
            dw = fmaf(s.z[3], s.w[3], dw);
            s.diff = true;
        }
        if (isnan(dw)) { dw = 2.0f; s.diff = false; }
    }
    const bool vmin = dw > min_depth;
    if (!vmin) { dw = min_depth; s.diff = false; }
    const bool vmax = dw < max_depth;
    if (!vmax) { dw = max_depth; s.diff = false; }
    s.dw = dw;
    s.m = (inb ? 1.f : 0.f) * mval * (vmin ? 1.f : 0.f) * (vmax ? 1.f : 0.f);
    return s;
}

__device__ __forceinline__ double syn_block_sum(double v, double* s_red) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) s_red[wid] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0)
        for (int i = 0; i < kSynThreads / 32; ++i) t += s_red[i];
    return t;
}

__global__ void __launch_bounds__(kSynThreads) depth_syn_fwd_kernel(const SynArgs a) {
    __shared__ SynCtx ctx;
    __shared__ double s_red[kSynThreads / 32];
    const vfd_syn_desc& d = a.d;
    const int bc = blockIdx.y, b = bc / d.n_call, ci = bc % d.n_call;
    const int W = d.W, H = d.H, HW = H * W;
    syn_ctx_init(ctx, a, b, ci);
    __syncthreads();
    const float* aug = a.aug + (size_t)bc * HW;
    float num = 0.f, den = 0.f;
    for (int p = blockIdx.x * kSynThreads + threadIdx.x; p < HW; p += gridDim.x * kSynThreads) {
        const int x = p % W, y = p / W;
        const float av = __ldg(aug + p);
        float X[3];
        vfd_backproject(ctx.invK, (float)x, (float)y, av, X);
#pragma unroll
        for (int i = 0; i < kSynSrc; ++i) {
            const int s = ctx.src[i];
            if (s < 0) continue;
            const SynSample sm = syn_sample<false>(ctx, i, X, a.depth + (size_t)(b * d.num_cams + s) * HW,
                                                   d.mask + (size_t)(b * d.num_cams + s) * HW, W, H, d.min_depth, d.max_depth);
            float l = fabsf(av - sm.dw) / ((av + sm.dw) + 1e-8f);
            l = nan_clamp01(l);      // torch.clamp keeps NaN (depth_synthesis_loss.py:27)
            num += l * sm.m;
            den += sm.m;
            if (a.out_depth) a.out_depth[((size_t)bc * kSynSrc + i) * HW + p] = sm.dw;
            if (a.out_mask) a.out_mask[((size_t)bc * kSynSrc + i) * HW + p] = sm.m;
        }
    }
    const double tn = syn_block_sum((double)num, s_red);
    const double td = syn_block_sum((double)den, s_red);
    if (threadIdx.x == 0) {
        atomicAdd(a.sums + ci * 2 + 0, tn);
        atomicAdd(a.sums + ci * 2 + 1, td);
    }
}

__global__ void __launch_bounds__(kSynThreads) depth_syn_bwd_kernel(const SynArgs a) {
    __shared__ SynCtx ctx;
    const vfd_syn_desc& d = a.d;
    const int bc = blockIdx.y, b = bc / d.n_call, ci = bc % d.n_call;
    const int W = d.W, H = d.H, HW = H * W;
    syn_ctx_init(ctx, a, b, ci);
    __syncthreads();
    const float* aug = a.aug + (size_t)bc * HW;
    const float wc = __ldg(a.wcam + ci);
    for (int p = blockIdx.x * kSynThreads + threadIdx.x; p < HW; p += gridDim.x * kSynThreads) {
        const int x = p % W, y = p / W;
        const float av = __ldg(aug + p);
        float ray[3], X[3];
        vfd_pixel_ray(ctx.invK, (float)x, (float)y, ray);
        X[0] = VFD_MUL(av, ray[0]); X[1] = VFD_MUL(av, ray[1]); X[2] = VFD_MUL(av, ray[2]);
        float ga = 0.f;
#pragma unroll
        for (int i = 0; i < kSynSrc; ++i) {
            const int s = ctx.src[i];
            if (s < 0) continue;
            const SynSample sm = syn_sample<true>(ctx, i, X, a.depth + (size_t)(b * d.num_cams + s) * HW,
                                                  d.mask + (size_t)(b * d.num_cams + s) * HW, W, H, d.min_depth, d.max_depth);
            if (sm.m == 0.f) continue;
            const float r = av - sm.dw, q = (av + sm.dw) + 1e-8f;
            const float l = fabsf(r) / q;
            if (!(l >= 0.f && l <= 1.f)) continue;          // outside the clamp: no gradient
            const float sg = r > 0.f ? 1.f : (r < 0.f ? -1.f : 0.f);
            const float up = wc * sm.m;
            const float common = fabsf(r) / (q * q);
            ga += up * (sg / q - common);
            if (!sm.diff) continue;
            const float gdw = up * (-sg / q - common);
            // source depths
            float* gd = a.d_depth + (size_t)(b * d.num_cams + s) * HW;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (sm.ok[k]) atomicAdd(gd + (sm.y0 + (k >> 1)) * W + sm.x0 + (k & 1), gdw * sm.w[k] * sm.slope[k]);
            // sampling coordinate -> aug depth (same cancellation-free form as warp_loss_bwd)
            const float gix = gdw * ((sm.z[1] - sm.z[0]) * sm.ty1 + (sm.z[3] - sm.z[2]) * sm.ty0);
            const float giy = gdw * ((sm.z[2] - sm.z[0]) * sm.tx1 + (sm.z[3] - sm.z[1]) * sm.tx0);
            const float* P = ctx.P[i];
            const float inv = 1.f / sm.den;
            const float a0 = P[0] * ray[0] + P[1] * ray[1] + P[2] * ray[2];
            const float a1 = P[4] * ray[0] + P[5] * ray[1] + P[6] * ray[2];
            const float a2 = P[8] * ray[0] + P[9] * ray[1] + P[10] * ray[2];
            const float t2 = P[11] + 1e-7f;
            ga += (gix * (a0 * t2 - P[3] * a2) + giy * (a1 * t2 - P[7] * a2)) * inv * inv;
        }
        a.d_aug[(size_t)bc * HW + p] = ga;
    }
}

static int check_syn(const vfd_syn_desc* d) {
    if (!d) return VFD_ERR_NULL;
    if (d->B < 1 || d->num_cams < 1 || d->num_cams > VFD_MAX_CAMS || d->H < 2 || d->W < 2 || d->cam0 < 0 ||
        d->n_call < 1 || d->cam0 + d->n_call > d->num_cams)
        return VFD_ERR_DESC;
    if ((long)d->H * d->W >= (1L << 30)) return VFD_ERR_DESC;
    if (!(d->min_depth > 0.f) || !(d->max_depth > d->min_depth)) return VFD_ERR_DESC;
    if (!d->mask || !d->K || !d->invK) return VFD_ERR_NULL;
    return VFD_OK;
}

static dim3 syn_grid(const vfd_syn_desc* d) {
    int bx = (d->H * d->W + kSynThreads * 2 - 1) / (kSynThreads * 2);
    if (bx > 512) bx = 512;
    return dim3(bx, d->B * d->n_call);
}

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_depth_syn_fwd(const vfd_syn_desc* d, const float* depth, const float* aug, const float* T,
                                 const float* Tinv, double* sums, float* out_depth, float* out_mask, void* stream) {
    int rc = check_syn(d);
    if (rc) return rc;
    if (!depth || !aug || !T || !Tinv || !sums) return VFD_ERR_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(sums, 0, sizeof(double) * d->n_call * 2, st) != cudaSuccess) return VFD_ERR_CUDA;
    SynArgs a = {};
    a.d = *d; a.depth = depth; a.aug = aug; a.T = T; a.Tinv = Tinv; a.sums = sums; a.out_depth = out_depth; a.out_mask = out_mask;
    depth_syn_fwd_kernel<<<syn_grid(d), kSynThreads, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_depth_syn_bwd(const vfd_syn_desc* d, const float* depth, const float* aug, const float* T,
                                 const float* Tinv, const float* wcam, float* d_aug, float* d_depth, void* stream) {
    int rc = check_syn(d);
    if (rc) return rc;
    if (!depth || !aug || !T || !Tinv || !wcam || !d_aug || !d_depth) return VFD_ERR_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t bytes = sizeof(float) * (size_t)d->B * d->num_cams * d->H * d->W;
    if (cudaMemsetAsync(d_depth, 0, bytes, st) != cudaSuccess) return VFD_ERR_CUDA;
    SynArgs a = {};
    a.d = *d; a.depth = depth; a.aug = aug; a.T = T; a.Tinv = Tinv; a.wcam = wcam; a.d_aug = d_aug; a.d_depth = d_depth;
    depth_syn_bwd_kernel<<<syn_grid(d), kSynThreads, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
