// Stand-alone forms of the remaining free functions of models/losses/loss_util.py, for callers outside the fused rig
// kernels (which evaluate the same arithmetic inside vfd_warp_loss_* / vfd_smooth_loss_*):
//   compute_auto_masks   (:6-18)   first arg-min over cat[reprojection, identity] == 0        -> float mask, no gradient
//   compute_masked_loss  (:21-25)  sum(loss * mask) / (sum(mask) + 1e-8)                      -> two f64 sums; d_loss
//   compute_edg_smooth_loss (:28-40) mean(|dx disp| exp(-mean_c |dx rgb|)) + the y term       -> two f64 sums; d_disp
// All three are HBM-bound single passes; sums are accumulated in f64 (one atomic per CTA), the products and the
// channel mean in un-fused fp32 in the reference's order.
#include "common.cuh"

namespace vfd {

constexpr int kLfThreads = 256, kLfBlocks = 148 * 4;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// two f64 partial sums of a CTA -> out[0], out[1]
__device__ __forceinline__ void block_add2(double a, double b, double* out) {
    __shared__ double sh[2][kLfThreads / 32];
    a = warp_sum(a);
    b = warp_sum(b);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) { sh[0][wid] = a; sh[1][wid] = b; }
    __syncthreads();
    if (wid == 0) {
        a = lane < kLfThreads / 32 ? sh[0][lane] : 0.0;
        b = lane < kLfThreads / 32 ? sh[1][lane] : 0.0;
        a = warp_sum(a);
        b = warp_sum(b);
        if (lane == 0) {
            if (a != 0.0) atomicAdd(out, a);
            if (b != 0.0) atomicAdd(out + 1, b);
        }
    }
}

// torch.argmin semantics: first minimum wins, a NaN is smaller than everything (first NaN wins)
__global__ void __launch_bounds__(kLfThreads) auto_masks_kernel(int F1, int F2, int HW, const float* reproj,
                                                                const float* identity, float* mask) {
    const int b = blockIdx.y;
    const float* rp = reproj + (size_t)b * F1 * HW;
    const float* id = identity + (size_t)b * F2 * HW;
    for (int p = blockIdx.x * kLfThreads + threadIdx.x; p < HW; p += gridDim.x * kLfThreads) {
        float best = __ldg(rp + p);
        int arg = 0;
        for (int f = 1; f < F1 + F2; ++f) {
            const float v = f < F1 ? __ldg(rp + (size_t)f * HW + p) : __ldg(id + (size_t)(f - F1) * HW + p);
            if (!(best != best) && (v < best || v != v)) { best = v; arg = f; }
        }
        mask[(size_t)b * HW + p] = arg == 0 ? 1.f : 0.f;
    }
}

__global__ void __launch_bounds__(kLfThreads) masked_loss_fwd_kernel(size_t n, const float* loss, const float* mask,
                                                                     double* sums) {
    double num = 0.0, den = 0.0;
    for (size_t i = (size_t)blockIdx.x * kLfThreads + threadIdx.x; i < n; i += (size_t)gridDim.x * kLfThreads) {
        const float m = __ldg(mask + i);
        num += (double)__fmul_rn(__ldg(loss + i), m);
        den += (double)m;
    }
    block_add2(num, den, sums);
}

// d_loss = d_out * mask / (sum(mask) + 1e-8)   (fp32 denominator as the reference forms it)
__global__ void __launch_bounds__(kLfThreads) masked_loss_bwd_kernel(size_t n, const float* mask, const double* sums,
                                                                     const float* d_out, float* d_loss) {
    const float den = __fadd_rn((float)sums[1], 1e-8f);
    const float g = __fdiv_rn(__ldg(d_out), den);
    for (size_t i = (size_t)blockIdx.x * kLfThreads + threadIdx.x; i < n; i += (size_t)gridDim.x * kLfThreads)
        d_loss[i] = __fmul_rn(g, __ldg(mask + i));
}

// exp(-mean_c |rgb[p] - rgb[q]|) ; channel sum in order c = 0..C-1, correctly rounded division by C
__device__ __forceinline__ float edge_weight(const float* rgb, int C, size_t HW, size_t p, size_t q) {
    float s = 0.f;
    for (int c = 0; c < C; ++c) s = __fadd_rn(s, fabsf(__fsub_rn(__ldg(rgb + c * HW + p), __ldg(rgb + c * HW + q))));
    return expf(-__fdiv_rn(s, (float)C));
}

__device__ __forceinline__ float sgn(float v) { return v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f); }

// forward (d_disp == NULL): sums[0] += sum_x |disp[x] - disp[x+1]| w_x, sums[1] += the y term.
// backward (d_disp != NULL): every pixel gathers its four edges, d_disp written once per pixel.
__global__ void __launch_bounds__(kLfThreads) edg_smooth_kernel(int B, int C, int H, int W, const float* rgb,
                                                                const float* disp, double* sums, const float* d_out,
                                                                float* d_disp) {
    const int HWi = H * W, b = blockIdx.y;
    const size_t HW = (size_t)HWi;
    float gx = 0.f, gy = 0.f;
    if (d_disp) {
        const float g = __ldg(d_out);
        gx = g / (float)((size_t)B * H * (W - 1));
        gy = g / (float)((size_t)B * (H - 1) * W);
    }
    const float* im = rgb + (size_t)b * C * HW;
    const float* dp = disp + (size_t)b * HW;
    double sx = 0.0, sy = 0.0;
    for (int p = blockIdx.x * kLfThreads + threadIdx.x; p < HWi; p += gridDim.x * kLfThreads) {
        const int y = p / W, x = p - y * W;
        const float d0 = __ldg(dp + p);
        float acc = 0.f;
        if (x + 1 < W) {
            const float df = __fsub_rn(d0, __ldg(dp + p + 1)), w = edge_weight(im, C, HW, p, p + 1);
            sx += (double)__fmul_rn(fabsf(df), w);
            acc += gx * sgn(df) * w;
        }
        if (y + 1 < H) {
            const float df = __fsub_rn(d0, __ldg(dp + p + W)), w = edge_weight(im, C, HW, p, p + W);
            sy += (double)__fmul_rn(fabsf(df), w);
            acc += gy * sgn(df) * w;
        }
        if (d_disp) {
            if (x > 0) acc -= gx * sgn(__fsub_rn(__ldg(dp + p - 1), d0)) * edge_weight(im, C, HW, p - 1, p);
            if (y > 0) acc -= gy * sgn(__fsub_rn(__ldg(dp + p - W), d0)) * edge_weight(im, C, HW, p - W, p);
            d_disp[(size_t)b * HW + p] = acc;
        }
    }
    if (!d_disp) block_add2(sx, sy, sums);
}

static int lf_blocks(size_t n) {
    const size_t need = (n + kLfThreads - 1) / kLfThreads;
    return (int)(need < (size_t)kLfBlocks ? (need ? need : 1) : kLfBlocks);
}

// one image per blockIdx.y; enough CTAs in x to fill the machine (4 per SM over the whole grid) without exceeding the image
static dim3 lf_grid(int B, int HW) {
    const int need = (HW + kLfThreads - 1) / kLfThreads;
    int per = (kLfBlocks + B - 1) / B;
    if (per > need) per = need;
    return dim3(per < 1 ? 1 : per, B);
}

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_auto_masks(int B, int F1, int F2, int HW, const float* reproj, const float* identity, float* mask,
                              void* stream) {
    if (B < 1 || B > 65535 || F1 < 1 || F2 < 0 || HW < 1 || HW >= (1 << 30)) return VFD_ERR_DESC;
    if (!reproj || !mask || (F2 > 0 && !identity)) return VFD_ERR_NULL;
    auto_masks_kernel<<<lf_grid(B, HW), kLfThreads, 0, (cudaStream_t)stream>>>(F1, F2, HW, reproj, identity, mask);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_masked_loss_fwd(size_t n, const float* loss, const float* mask, double* sums, void* stream) {
    if (n < 1) return VFD_ERR_DESC;
    if (!loss || !mask || !sums) return VFD_ERR_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(sums, 0, 2 * sizeof(double), st) != cudaSuccess) return VFD_ERR_CUDA;
    masked_loss_fwd_kernel<<<lf_blocks(n), kLfThreads, 0, st>>>(n, loss, mask, sums);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_masked_loss_bwd(size_t n, const float* mask, const double* sums, const float* d_out, float* d_loss,
                                   void* stream) {
    if (n < 1) return VFD_ERR_DESC;
    if (!mask || !sums || !d_out || !d_loss) return VFD_ERR_NULL;
    masked_loss_bwd_kernel<<<lf_blocks(n), kLfThreads, 0, (cudaStream_t)stream>>>(n, mask, sums, d_out, d_loss);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_edg_smooth_loss_fwd(int B, int C, int H, int W, const float* rgb, const float* disp, double* sums,
                                       void* stream) {
    if (B < 1 || B > 65535 || C < 1 || H < 2 || W < 2 || (long)H * W >= (1L << 30)) return VFD_ERR_DESC;
    if (!rgb || !disp || !sums) return VFD_ERR_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(sums, 0, 2 * sizeof(double), st) != cudaSuccess) return VFD_ERR_CUDA;
    edg_smooth_kernel<<<lf_grid(B, H * W), kLfThreads, 0, st>>>(B, C, H, W, rgb, disp, sums, nullptr, nullptr);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_edg_smooth_loss_bwd(int B, int C, int H, int W, const float* rgb, const float* disp,
                                       const float* d_out, float* d_disp, void* stream) {
    if (B < 1 || B > 65535 || C < 1 || H < 2 || W < 2 || (long)H * W >= (1L << 30)) return VFD_ERR_DESC;
    if (!rgb || !disp || !d_out || !d_disp) return VFD_ERR_NULL;
    edg_smooth_kernel<<<lf_grid(B, H * W), kLfThreads, 0, (cudaStream_t)stream>>>(B, C, H, W, rgb, disp, nullptr, d_out,
                                                                               d_disp);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
