// Pass 2 of the view-synthesis path, forward: intensity alignment of the raw warps pass 1 left + SSIM/L1 maps + min
// over frames + auto-mask + masked sums in ONE kernel -- no normalised image, overlap sum or loss map goes to HBM
// unless the caller asks for the materialised views (logger / parity tests).
//   models/geometry/view_rendering.py:118-198, models/losses/single_cam_loss.py:17-55, multi_cam_loss.py:16-59,
//   loss_util.py:6-25, 43-78.
// Grid: (tiles_x, tiles_y, B * n_call); CTA = 32 x 8 pixel tile with a 1-pixel halo (the 3x3 SSIM window).
#include "warp_tile.cuh"

namespace vfd {

struct LossFwdArgs {
    vfd_warp_desc d;
    const float* depth;
    const float* T;
    const float* params;
    const float4* raw;
    const uint8_t* vbits;
    const float* noise;
    double* sums;
    vfd_warp_views views;
    int has_views;
};

extern __shared__ __align__(16) float vfd_warp_smem[];

__global__ void __launch_bounds__(kTileThreads, 3) warp_loss_fwd_kernel(const LossFwdArgs a) {
    using Rg = Region<1>;
    const vfd_warp_desc& d = a.d;
    __shared__ WarpCtx ctx;
    __shared__ double sacc[8];
    __shared__ unsigned s_flags;
    float* sm = vfd_warp_smem;
    const int b = blockIdx.z / d.n_call, ci = blockIdx.z % d.n_call, cam = d.cam0 + ci;
    const int W = d.W, H = d.H, HW = H * W, n_src = d.n_src;
    const int tx0 = blockIdx.x * kTileW, ty0 = blockIdx.y * kTileH, ox = tx0 - 1, oy = ty0 - 1;
    warp_ctx_init(ctx, d, a.T, a.params, b, ci);
    if (threadIdx.x < 8) sacc[threadIdx.x] = 0.0;
    if (threadIdx.x == 0) s_flags = 0u;
    __syncthreads();
    const float* depth = a.depth + (size_t)(b * d.n_call + ci) * HW;
    const unsigned fl = fill_region<1>(ctx, d, b, ci, ox, oy, a.raw + (size_t)(b * d.n_call + ci) * ctx.NW * HW,
                                       a.vbits + (size_t)(b * d.n_call + ci) * HW, sm);
    const bool sp = (cta_or(fl, &s_flags) & 1u) != 0;     // any spatial / spatio-temporal warp in this tile's region?

    const int x = tx0 + (threadIdx.x % kTileW), y = ty0 + (threadIdx.x / kTileW);
    float acc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (threadIdx.x < kTilePixels && x < W && y < H) {
        const int p = y * W + x;
        const float rmask = __ldg(d.mask + (size_t)(b * d.num_cams + cam) * HW + p);
        float nz[VFD_WARP_SRC] = {0.f, 0.f};
        {   // static indices: a run-time index would put nz[] in local memory
            const float* np = a.noise + (size_t)(b * d.n_call + ci) * n_src * HW + p;
            nz[0] = __ldg(np);
            if (n_src > 1) nz[1] = __ldg(np + HW);
        }
        const PixelLoss o = sp ? pixel_loss<1, false, true>(sm, d, x, y, ox, oy, rmask, nz, nullptr)
                               : pixel_loss<1, false, false>(sm, d, x, y, ox, oy, rmask, nz, nullptr);
        acc[0] = VFD_MUL(o.reproj, o.automask); acc[1] = o.automask;
        acc[2] = VFD_MUL(o.spatial, o.sp_mask); acc[3] = o.sp_mask;
        acc[4] = VFD_MUL(o.st, o.st_mask);      acc[5] = o.st_mask;
        if (a.has_views) {
            const vfd_warp_views& v = a.views;
            const size_t bc = (size_t)(b * d.n_call + ci);
            const int ctr = (y - oy) * Rg::RW + (x - ox);
            const float2* img = reinterpret_cast<const float2*>(sm + Rg::kImg) + ctr;
            for (int c = 0; c < 3; ++c) {
                const float2 tp = img[(c * kPairs + 0) * Rg::R], st = img[(c * kPairs + 1) * Rg::R];
                const float2 sp = img[(c * kPairs + 2) * Rg::R];
                if (v.color) {
                    v.color[((bc * n_src + 0) * 3 + c) * HW + p] = tp.x;
                    if (n_src > 1) v.color[((bc * n_src + 1) * 3 + c) * HW + p] = tp.y;
                }
                if (v.overlap) {
                    v.overlap[((bc * (1 + n_src) + 0) * 3 + c) * HW + p] = sp.x;
                    v.overlap[((bc * (1 + n_src) + 1) * 3 + c) * HW + p] = st.x;
                    if (n_src > 1) v.overlap[((bc * (1 + n_src) + 2) * 3 + c) * HW + p] = st.y;
                }
            }
            for (int k = 0; k < n_src; ++k)
                if (v.color_mask) v.color_mask[(bc * n_src + k) * HW + p] = sm[Rg::kCmask + k * Rg::R + ctr];
            for (int f = 0; f <= n_src; ++f)
                if (v.overlap_mask) v.overlap_mask[(bc * (1 + n_src) + f) * HW + p] = sm[Rg::kOmask + f * Rg::R + ctr];
            if (v.reproj_loss) v.reproj_loss[bc * HW + p] = acc[0];
            if (v.reproj_mask) v.reproj_mask[bc * HW + p] = o.automask;
            if (v.selection) v.selection[bc * HW + p] = (float)(o.sel_reproj + 2 * o.sel_st);
            if (v.grid) {
                float X[3];
                vfd_backproject(ctx.invK, (float)x, (float)y, __ldg(depth + p), X);
                for (int j = 0; j < ctx.NW; ++j) {
                    float cm[3], g[2] = {0.f, 0.f};
                    if (ctx.active[j]) vfd_reproject(ctx.P[j], X, W, H, cm, g);
                    reinterpret_cast<float2*>(v.grid)[(bc * ctx.NW + j) * HW + p] = make_float2(g[0], g[1]);
                }
            }
        }
    }
    // CTA reduction: fp32 inside a warp (32 values), fp64 above
#pragma unroll
    for (int q = 0; q < 6; ++q) {
        float s = acc[q];
#pragma unroll
        for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if ((threadIdx.x & 31) == 0 && s != 0.f) atomicAdd(&sacc[q], (double)s);
    }
    __syncthreads();
    if (threadIdx.x < 6 && sacc[threadIdx.x] != 0.0) atomicAdd(a.sums + ci * 8 + threadIdx.x, sacc[threadIdx.x]);
}

int check_warp_desc(const vfd_warp_desc* d);   // warp_stats.cu

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_warp_loss_fwd(const vfd_warp_desc* d, const float* depth, const float* T, const float* params,
                                 const float* raw, const float* noise, double* sums, const vfd_warp_views* views, void* stream) {
    int rc = check_warp_desc(d);
    if (rc) return rc;
    if (!depth || !T || !params || !raw || !noise || !sums) return VFD_ERR_NULL;
    if (reinterpret_cast<uintptr_t>(raw) & 15) return VFD_ERR_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(sums, 0, (size_t)d->n_call * 8 * sizeof(double), st) != cudaSuccess) return VFD_ERR_CUDA;
    LossFwdArgs a = {};
    a.d = *d; a.depth = depth; a.T = T; a.params = params; a.raw = reinterpret_cast<const float4*>(raw); a.noise = noise; a.sums = sums;
    a.vbits = reinterpret_cast<const uint8_t*>(a.raw + warp_raw_f4(*d));
    if (views) { a.views = *views; a.has_views = 1; }
    const size_t smem = Region<1>::kFloats * sizeof(float);
    dim3 grid((d->W + kTileW - 1) / kTileW, (d->H + kTileH - 1) / kTileH, d->B * d->n_call);
    warp_loss_fwd_kernel<<<grid, kTileThreads, smem, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
