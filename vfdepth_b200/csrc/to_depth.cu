// Disparity -> metric depth for all cameras of a rig batch (models/vfdepth.py:277-288 to_depth, called at :266-275):
//   d     = bilinear upsample of the decoder's sigmoid output to the training size (F.interpolate, align_corners=False;
//           the identity at scale 0, the only scale the shipped configs train)
//   disp  = 1/max_depth + (1/min_depth - 1/max_depth) * d
//   depth = (1 / disp) * K[0,0] / focal_length_scale
// The reference spends 5 elementwise launches per camera and scale on it; here one launch covers the rig batch, and the
// backward (depth -> decoder output) is one more.  Source index / weights exactly as ATen's upsample_bilinear2d forms
// them (area_pixel_compute_source_index: scale * (dst + 0.5) - 0.5 clamped at 0, scale = in / out).
#include "common.cuh"

namespace vfd {

constexpr int kDepthThreads = 256;

struct ToDepthArgs {
    int B, cams, hs, ws, H, W;
    const float* disp;     // [B,cams,hs,ws]
    const float* K;        // [B,cams,4,4]  intrinsics of scale 0
    float min_disp, range, focal;
    float* depth;          // [B,cams,H,W]
    const float* d_depth;  // [B,cams,H,W]
    float* d_disp;         // [B,cams,hs,ws]
};

struct Tap {
    int i0, i1;
    float l0, l1;
};

__device__ __forceinline__ Tap make_tap(int dst, int in, int out) {
    Tap t;
    if (in == out) {
        t.i0 = t.i1 = dst;
        t.l0 = 1.f;
        t.l1 = 0.f;
        return t;
    }
    const float scale = (float)in / (float)out;
    float src = scale * ((float)dst + 0.5f) - 0.5f;
    src = src < 0.f ? 0.f : src;
    t.i0 = (int)src;
    t.i1 = t.i0 + ((t.i0 < in - 1) ? 1 : 0);
    t.l1 = src - (float)t.i0;
    t.l0 = 1.f - t.l1;
    return t;
}

// Same-size case (scale 0, every shipped config): the bilinear taps are (1, 0) on the pixel itself, so the interpolation
// returns the input exactly for finite disparities (a sigmoid's output) and the pass is a float4 stream: 8 B per pixel
// forward, 12 B backward.  VEC = 4 needs H*W % 4 == 0 (checked by the launcher; 16-B alignment follows from the
// 256-B aligned tensors the host mirror passes -- the launcher checks the pointers too).
__device__ __forceinline__ float depth_of(float d, float k00, const ToDepthArgs& a) {
    const float disp = VFD_ADD(a.min_disp, VFD_MUL(a.range, d));
    const float depth = __fdiv_rn(1.f, disp);
    return __fdiv_rn(VFD_MUL(depth, k00), a.focal);
}
// the quotient is formed on the product -g * (k00 / focal) * range as in to_depth_bwd_kernel.  A zero numerator (pixels
// without a gradient) would send the whole warp down the slow path of the division; 0 / x is that signed zero itself,
// so those lanes divide a stand-in and keep the zero.
__device__ __forceinline__ float d_disp_of(float g, float d, float kf, const ToDepthArgs& a) {
    const float disp = a.min_disp + a.range * d;
    const float num = -g * kf * a.range;
    const bool zero = num == 0.f;
    const float q = (zero ? 1.f : num) / (disp * disp);
    return zero ? num : q;
}

__global__ void __launch_bounds__(kDepthThreads) to_depth_same_fwd_kernel(const ToDepthArgs a) {
    const int bc = blockIdx.y, n4 = (a.H * a.W) >> 2;
    const float4* s = reinterpret_cast<const float4*>(a.disp + (size_t)bc * a.H * a.W);
    float4* o = reinterpret_cast<float4*>(a.depth + (size_t)bc * a.H * a.W);
    const float k00 = __ldg(a.K + (size_t)bc * 16);
    for (int i = blockIdx.x * kDepthThreads + threadIdx.x; i < n4; i += gridDim.x * kDepthThreads) {
        const float4 v = __ldg(s + i);
        o[i] = make_float4(depth_of(v.x, k00, a), depth_of(v.y, k00, a), depth_of(v.z, k00, a), depth_of(v.w, k00, a));
    }
}

__global__ void __launch_bounds__(kDepthThreads) to_depth_same_bwd_kernel(const ToDepthArgs a) {
    const int bc = blockIdx.y, n4 = (a.H * a.W) >> 2;
    const float4* s = reinterpret_cast<const float4*>(a.disp + (size_t)bc * a.H * a.W);
    const float4* g = reinterpret_cast<const float4*>(a.d_depth + (size_t)bc * a.H * a.W);
    float4* gs = reinterpret_cast<float4*>(a.d_disp + (size_t)bc * a.H * a.W);
    const float kf = __ldg(a.K + (size_t)bc * 16) / a.focal;
    for (int i = blockIdx.x * kDepthThreads + threadIdx.x; i < n4; i += gridDim.x * kDepthThreads) {
        const float4 v = __ldg(s + i), u = __ldg(g + i);
        gs[i] = make_float4(d_disp_of(u.x, v.x, kf, a), d_disp_of(u.y, v.y, kf, a), d_disp_of(u.z, v.z, kf, a),
                            d_disp_of(u.w, v.w, kf, a));
    }
}

__global__ void __launch_bounds__(kDepthThreads) to_depth_fwd_kernel(const ToDepthArgs a) {
    const int bc = blockIdx.y, HW = a.H * a.W;
    const float* s = a.disp + (size_t)bc * a.hs * a.ws;
    float* o = a.depth + (size_t)bc * HW;
    const float k00 = __ldg(a.K + (size_t)bc * 16);
    for (int p = blockIdx.x * kDepthThreads + threadIdx.x; p < HW; p += gridDim.x * kDepthThreads) {
        const int x = p % a.W, y = p / a.W;
        const Tap tx = make_tap(x, a.ws, a.W), ty = make_tap(y, a.hs, a.H);
        const float v00 = __ldg(s + ty.i0 * a.ws + tx.i0), v01 = __ldg(s + ty.i0 * a.ws + tx.i1);
        const float v10 = __ldg(s + ty.i1 * a.ws + tx.i0), v11 = __ldg(s + ty.i1 * a.ws + tx.i1);
        const float d = VFD_ADD(VFD_MUL(ty.l0, VFD_ADD(VFD_MUL(tx.l0, v00), VFD_MUL(tx.l1, v01))),
                                VFD_MUL(ty.l1, VFD_ADD(VFD_MUL(tx.l0, v10), VFD_MUL(tx.l1, v11))));
        const float disp = VFD_ADD(a.min_disp, VFD_MUL(a.range, d));
        const float depth = __fdiv_rn(1.f, disp);
        o[p] = __fdiv_rn(VFD_MUL(depth, k00), a.focal);
    }
}

// d depth / d d = -(K00 / focal) * range / disp^2; scattered back through the bilinear weights
__global__ void __launch_bounds__(kDepthThreads) to_depth_bwd_kernel(const ToDepthArgs a) {
    const int bc = blockIdx.y, HW = a.H * a.W;
    const float* s = a.disp + (size_t)bc * a.hs * a.ws;
    const float* g = a.d_depth + (size_t)bc * HW;
    float* gs = a.d_disp + (size_t)bc * a.hs * a.ws;
    const float k00 = __ldg(a.K + (size_t)bc * 16);
    const bool same = a.hs == a.H && a.ws == a.W;
    for (int p = blockIdx.x * kDepthThreads + threadIdx.x; p < HW; p += gridDim.x * kDepthThreads) {
        const int x = p % a.W, y = p / a.W;
        const Tap tx = make_tap(x, a.ws, a.W), ty = make_tap(y, a.hs, a.H);
        const float v00 = __ldg(s + ty.i0 * a.ws + tx.i0), v01 = __ldg(s + ty.i0 * a.ws + tx.i1);
        const float v10 = __ldg(s + ty.i1 * a.ws + tx.i0), v11 = __ldg(s + ty.i1 * a.ws + tx.i1);
        const float d = ty.l0 * (tx.l0 * v00 + tx.l1 * v01) + ty.l1 * (tx.l0 * v10 + tx.l1 * v11);
        const float disp = a.min_disp + a.range * d;
        const float gd = -__ldg(g + p) * (k00 / a.focal) * a.range / (disp * disp);
        if (same) {
            gs[p] = gd;
        } else {
            atomicAdd(gs + ty.i0 * a.ws + tx.i0, gd * ty.l0 * tx.l0);
            atomicAdd(gs + ty.i0 * a.ws + tx.i1, gd * ty.l0 * tx.l1);
            atomicAdd(gs + ty.i1 * a.ws + tx.i0, gd * ty.l1 * tx.l0);
            atomicAdd(gs + ty.i1 * a.ws + tx.i1, gd * ty.l1 * tx.l1);
        }
    }
}

static int check_to_depth(int B, int cams, int hs, int ws, int H, int W, float min_depth, float max_depth, float focal) {
    if (B < 1 || cams < 1 || hs < 1 || ws < 1 || H < hs || W < ws) return VFD_ERR_DESC;
    if ((long)H * W >= (1L << 30)) return VFD_ERR_DESC;
    if (!(min_depth > 0.f) || !(max_depth > min_depth) || !(focal > 0.f)) return VFD_ERR_DESC;
    return VFD_OK;
}

static dim3 to_depth_grid(int B, int cams, int H, int W) {
    int bx = (H * W + kDepthThreads * 4 - 1) / (kDepthThreads * 4);
    if (bx > 256) bx = 256;
    return dim3(bx, B * cams);
}

static ToDepthArgs to_depth_args(int B, int cams, int hs, int ws, int H, int W, float min_depth, float max_depth,
                                 float focal) {
    ToDepthArgs a = {};
    a.B = B; a.cams = cams; a.hs = hs; a.ws = ws; a.H = H; a.W = W;
    // python: min_disp = 1/max_depth, max_disp = 1/min_depth, range = max_disp - min_disp in double, cast when they
    // meet the fp32 tensor
    const double min_disp = 1.0 / (double)max_depth, max_disp = 1.0 / (double)min_depth;
    a.min_disp = (float)min_disp;
    a.range = (float)(max_disp - min_disp);
    a.focal = focal;
    return a;
}

// the float4 same-size path: no resampling, whole float4s per image, 16-B aligned tensors
static bool to_depth_same(const ToDepthArgs& a, const void* p0, const void* p1, const void* p2) {
    const uintptr_t bits = (uintptr_t)p0 | (uintptr_t)p1 | (uintptr_t)p2;
    return a.hs == a.H && a.ws == a.W && ((a.H * a.W) & 3) == 0 && (bits & 15) == 0;
}

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_to_depth_fwd(int B, int cams, int hs, int ws, int H, int W, const float* disp, const float* K,
                                float min_depth, float max_depth, float focal_length_scale, float* depth, void* stream) {
    int rc = check_to_depth(B, cams, hs, ws, H, W, min_depth, max_depth, focal_length_scale);
    if (rc) return rc;
    if (!disp || !K || !depth) return VFD_ERR_NULL;
    ToDepthArgs a = to_depth_args(B, cams, hs, ws, H, W, min_depth, max_depth, focal_length_scale);
    a.disp = disp; a.K = K; a.depth = depth;
    if (to_depth_same(a, disp, depth, nullptr))
        to_depth_same_fwd_kernel<<<to_depth_grid(B, cams, H, W), kDepthThreads, 0, (cudaStream_t)stream>>>(a);
    else
        to_depth_fwd_kernel<<<to_depth_grid(B, cams, H, W), kDepthThreads, 0, (cudaStream_t)stream>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_to_depth_bwd(int B, int cams, int hs, int ws, int H, int W, const float* disp, const float* K,
                                float min_depth, float max_depth, float focal_length_scale, const float* d_depth,
                                float* d_disp, void* stream) {
    int rc = check_to_depth(B, cams, hs, ws, H, W, min_depth, max_depth, focal_length_scale);
    if (rc) return rc;
    if (!disp || !K || !d_depth || !d_disp) return VFD_ERR_NULL;
    ToDepthArgs a = to_depth_args(B, cams, hs, ws, H, W, min_depth, max_depth, focal_length_scale);
    a.disp = disp; a.K = K; a.d_depth = d_depth; a.d_disp = d_disp;
    cudaStream_t st = (cudaStream_t)stream;
    if (hs != H || ws != W) {
        if (cudaMemsetAsync(d_disp, 0, sizeof(float) * (size_t)B * cams * hs * ws, st) != cudaSuccess) return VFD_ERR_CUDA;
    }
    if (to_depth_same(a, disp, d_depth, d_disp))
        to_depth_same_bwd_kernel<<<to_depth_grid(B, cams, H, W), kDepthThreads, 0, st>>>(a);
    else
        to_depth_bwd_kernel<<<to_depth_grid(B, cams, H, W), kDepthThreads, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
