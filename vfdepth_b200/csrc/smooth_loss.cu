// Edge-aware disparity smoothness for all cameras of a rig batch in two launches per direction
// (models/losses/single_cam_loss.py:57-65 compute_smooth_loss + loss_util.py:28-40 compute_edg_smooth_loss, scale 0):
//   nd   = disp / (mean_HW(disp) + 1e-8)                                   per (rig, camera) image
//   gx   = |nd[x] - nd[x+1]| * exp(-mean_c |rgb[x] - rgb[x+1]|),  gy likewise along y
//   loss = mean(gx) + mean(gy)                                              means over the whole local batch of one camera
// The reference spends ~25 elementwise / reduction launches per camera on it; here the image means are one reduction
// launch and the loss one stencil launch (fp32 per thread, fp64 across threads); the backward to disp (the only
// differentiable input) is a stencil launch that also accumulates sum_i g_i * disp_i per image for the mean's chain
// rule, and a finishing launch.
#include <atomic>

#include "common.cuh"

namespace vfd {

constexpr int kSmoothThreads = 256;

struct SmoothArgs {
    int B, cams, H, W;
    const float* color;    // [B,cams,3,H,W]
    const float* disp;     // [B,cams,H,W]
    double* img_sum;       // [B*cams]   sum of disp per image
    double* sums;          // [cams][2]  sum gx, sum gy over the batch
    // backward
    const float* d_smooth; // [cams]
    double* gdot;          // [B*cams]   sum_i g_i * disp_i
    float* d_disp;         // [B,cams,H,W]
};

__device__ __forceinline__ double block_sum(double v, double* s_red) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) s_red[wid] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0)
        for (int i = 0; i < kSmoothThreads / 32; ++i) t += s_red[i];
    return t;   // valid in thread 0
}

__global__ void __launch_bounds__(kSmoothThreads) smooth_mean_kernel(const SmoothArgs a) {
    __shared__ double s_red[kSmoothThreads / 32];
    const int bc = blockIdx.y, HW = a.H * a.W;
    const float* d = a.disp + (size_t)bc * HW;
    double acc = 0.0;
    for (int p = blockIdx.x * kSmoothThreads + threadIdx.x; p < HW; p += gridDim.x * kSmoothThreads) acc += (double)__ldg(d + p);
    const double t = block_sum(acc, s_red);
    if (threadIdx.x == 0) atomicAdd(a.img_sum + bc, t);
}

// edge weight exp(-mean_c |rgb[p] - rgb[q]|); the channel mean as sum * (1/3), torch's CUDA mean (a division by the
// constant would send every warp holding one zero sum through the slow path of the division)
constexpr float kThird = 1.0f / 3.0f;
__device__ __forceinline__ float edge_weight(const float* col, int HW, int p, int q) {
    if (!col) return 1.f;      // plain smoothness (depth_synthesis_loss.py:35-41)
    const float d0 = fabsf(__ldg(col + p) - __ldg(col + q));
    const float d1 = fabsf(__ldg(col + HW + p) - __ldg(col + HW + q));
    const float d2 = fabsf(__ldg(col + 2 * HW + p) - __ldg(col + 2 * HW + q));
    return expf(-((d0 + d1 + d2) * kThird));
}

__device__ __forceinline__ float image_scale(const SmoothArgs& a, int bc) {
    const float mean = (float)(a.img_sum[bc] / (double)(a.H * a.W));
    return mean + 1e-8f;
}

__global__ void __launch_bounds__(kSmoothThreads) smooth_loss_kernel(const SmoothArgs a) {
    __shared__ double s_red[kSmoothThreads / 32];
    const int bc = blockIdx.y, cam = bc % a.cams, HW = a.H * a.W, W = a.W, H = a.H;
    const float* d = a.disp + (size_t)bc * HW;
    const float* col = a.color ? a.color + (size_t)bc * 3 * HW : nullptr;
    const float den = image_scale(a, bc);
    float sx = 0.f, sy = 0.f;
    for (int p = blockIdx.x * kSmoothThreads + threadIdx.x; p < HW; p += gridDim.x * kSmoothThreads) {
        const int x = p % W, y = p / W;
        const float n0 = __ldg(d + p) / den;
        if (x + 1 < W) sx += fabsf(n0 - __ldg(d + p + 1) / den) * edge_weight(col, HW, p, p + 1);
        if (y + 1 < H) sy += fabsf(n0 - __ldg(d + p + W) / den) * edge_weight(col, HW, p, p + W);
    }
    const double tx = block_sum((double)sx, s_red);
    const double ty = block_sum((double)sy, s_red);
    if (threadIdx.x == 0) {
        atomicAdd(a.sums + cam * 2 + 0, tx);
        atomicAdd(a.sums + cam * 2 + 1, ty);
    }
}

__device__ __forceinline__ float sgn(float v) { return v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f); }

// g_i = d loss / d nd_i (written to d_disp), and sum_i g_i * disp_i per image.
// TRAIN: unit upstream, and the loss sums of smooth_loss_kernel from the same taps -- the forward and the backward of a
// training step share one pass over the images; smooth_bwd_finish_kernel applies the upstream afterwards.
//
// Mapping.  A warp walks kSmoothRun rows downwards over 32 adjacent columns of which it OWNS the inner kSmoothOwn = 30
// (lanes 0 and 31 are halo lanes that only feed the shuffles); CTA = 8 such warps side by side.  Every edge of the
// stencil is then evaluated ONCE: the edge below a pixel (weight, normalised neighbour, neighbour colours) is carried
// in registers and becomes the edge above the next row's pixel, the edge to the right is handed to the lane on the right
// as its left edge by shuffle.  Per pixel that is one division and two expf instead of five and four, and 4 loads
// instead of 20, with no lane-divergent loads at the warp's edges -- and with the very same operands in the very same
// expressions, so d_disp is bit-identical to the pixel-per-thread form.  The row after next is loaded while the current
// one is evaluated (the walk is otherwise one dependent load chain per warp), and the sums leave per warp (no barrier).
constexpr int kSmoothRun = 16;
constexpr int kSmoothOwn = 30;

struct EdgeSide {      // what a pixel contributes to an edge: its normalised disparity and colours
    float n, c0, c1, c2;
};
struct RawSide {       // the same as loaded (disparity not yet normalised)
    float d, c0, c1, c2;
};
__device__ __forceinline__ RawSide load_raw(const float* d, const float* col, int HW, int p) {
    RawSide s;
    s.d = __ldg(d + p);
    s.c0 = s.c1 = s.c2 = 0.f;
    if (col) { s.c0 = __ldg(col + p); s.c1 = __ldg(col + HW + p); s.c2 = __ldg(col + 2 * HW + p); }
    return s;
}
__device__ __forceinline__ EdgeSide to_side(const RawSide& r, float den) {
    EdgeSide s;
    s.n = r.d / den;
    s.c0 = r.c0; s.c1 = r.c1; s.c2 = r.c2;
    return s;
}
// exp(-mean_c |rgb[p] - rgb[q]|) with p the left / upper pixel, as edge_weight() forms it
__device__ __forceinline__ float side_weight(bool has_color, const EdgeSide& p, const EdgeSide& q) {
    if (!has_color) return 1.f;
    const float d0 = fabsf(p.c0 - q.c0), d1 = fabsf(p.c1 - q.c1), d2 = fabsf(p.c2 - q.c2);
    return expf(-((d0 + d1 + d2) * kThird));
}
__device__ __forceinline__ double warp_sum(double v) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <bool TRAIN>
__global__ void __launch_bounds__(kSmoothThreads) smooth_bwd_kernel(const SmoothArgs a) {
    const int bc = blockIdx.z, cam = bc % a.cams, HW = a.H * a.W, W = a.W, H = a.H;
    const float* d = a.disp + (size_t)bc * HW;
    const float* col = a.color ? a.color + (size_t)bc * 3 * HW : nullptr;
    const bool hc = col != nullptr;
    float* g = a.d_disp + (size_t)bc * HW;
    const float den = image_scale(a, bc);
    const float up = TRAIN ? 1.f : __ldg(a.d_smooth + cam);
    const float ux = up / (float)((double)a.B * H * (W - 1)), uy = up / (float)((double)a.B * (H - 1) * W);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x = (blockIdx.x * (kSmoothThreads / 32) + warp) * kSmoothOwn + lane - 1;
    const int y0 = blockIdx.y * kSmoothRun, y1 = min(H, y0 + kSmoothRun);
    if (x - lane + 1 >= W) return;                 // the warp's first own column is outside (no CTA-wide barrier below)
    const bool in = x >= 0 && x < W;
    const bool own = in && lane >= 1 && lane <= kSmoothOwn;
    const bool has_r = own && x + 1 < W, has_l = own && x > 0;
    double acc = 0.0;
    float sx = 0.f, sy = 0.f;
    EdgeSide cur = {0.f, 0.f, 0.f, 0.f}, above = cur;
    RawSide next = {0.f, 0.f, 0.f, 0.f};           // row y + 1, in flight while row y is evaluated
    float w_top = 0.f, d_cur = 0.f;
    if (in) {
        const RawSide r0 = load_raw(d, col, HW, y0 * W + x);
        if (y0 + 1 < H) next = load_raw(d, col, HW, (y0 + 1) * W + x);
        d_cur = r0.d;
        cur = to_side(r0, den);
        if (y0 > 0) {
            above = to_side(load_raw(d, col, HW, (y0 - 1) * W + x), den);
            w_top = side_weight(hc, above, cur);
        }
    }
    for (int y = y0; y < y1; ++y) {
        const int p = y * W + x;
        const bool has_b = in && y + 1 < H;
        const RawSide below_raw = next;
        if (in && y + 2 < H && y + 1 < y1) next = load_raw(d, col, HW, p + 2 * W);
        // right neighbour: the lane on the right holds it; its edge weight is the left edge of that lane
        EdgeSide right;
        right.n = __shfl_down_sync(0xffffffffu, cur.n, 1);
        right.c0 = __shfl_down_sync(0xffffffffu, cur.c0, 1);
        right.c1 = __shfl_down_sync(0xffffffffu, cur.c1, 1);
        right.c2 = __shfl_down_sync(0xffffffffu, cur.c2, 1);
        const float w_r = side_weight(hc, cur, right);
        const float w_l = __shfl_up_sync(0xffffffffu, w_r, 1);
        const float n_l = __shfl_up_sync(0xffffffffu, cur.n, 1);
        EdgeSide below = cur;
        float w_b = 0.f;
        if (has_b) {
            below = to_side(below_raw, den);
            w_b = side_weight(hc, cur, below);
        }
        if (own) {
            float gi = 0.f;
            if (has_r) {
                const float df = cur.n - right.n;
                gi += ux * sgn(df) * w_r;
                if (TRAIN) sx += fabsf(df) * w_r;
            }
            if (has_l) gi -= ux * sgn(n_l - cur.n) * w_l;
            if (has_b) {
                const float df = cur.n - below.n;
                gi += uy * sgn(df) * w_b;
                if (TRAIN) sy += fabsf(df) * w_b;
            }
            if (y > 0) gi -= uy * sgn(above.n - cur.n) * w_top;
            g[p] = gi;
            acc += (double)gi * (double)d_cur;
        }
        above = cur;
        cur = below;
        d_cur = below_raw.d;
        w_top = w_b;
    }
    const double t = warp_sum(acc);
    if (lane == 0) atomicAdd(a.gdot + bc, t);
    if (TRAIN) {
        const double tx = warp_sum((double)sx), ty = warp_sum((double)sy);
        if (lane == 0) {
            atomicAdd(a.sums + cam * 2 + 0, tx);
            atomicAdd(a.sums + cam * 2 + 1, ty);
        }
    }
}

static dim3 smooth_bwd_grid(int B, int cams, int H, int W) {
    const int per_cta = (kSmoothThreads / 32) * kSmoothOwn;
    return dim3((W + per_cta - 1) / per_cta, (H + kSmoothRun - 1) / kSmoothRun, B * cams);
}

// d_disp_j = up * (g_j / den - (sum_i g_i disp_i) / (den^2 * HW))      (nd = disp / den, den = mean + 1e-8); up = 1
// unless the gradient map carries unit upstream (training mode: src = that map, d_smooth = the upstream)
__global__ void __launch_bounds__(kSmoothThreads) smooth_bwd_finish_kernel(const SmoothArgs a, const float* src) {
    const int bc = blockIdx.y, HW = a.H * a.W;
    float* g = a.d_disp + (size_t)bc * HW;
    const float* gs = src ? src + (size_t)bc * HW : g;
    const float den = image_scale(a, bc);
    const float up = src ? __ldg(a.d_smooth + bc % a.cams) : 1.f;
    const float shift = (float)(a.gdot[bc] / ((double)den * (double)den * (double)HW));
    for (int p = blockIdx.x * kSmoothThreads + threadIdx.x; p < HW; p += gridDim.x * kSmoothThreads)
        g[p] = up * (gs[p] / den - shift);
}

static int check_smooth(int B, int cams, int H, int W) {
    if (B < 1 || cams < 1 || H < 2 || W < 2) return VFD_ERR_DESC;
    if ((long)H * W >= (1L << 30)) return VFD_ERR_DESC;
    return VFD_OK;
}

static dim3 smooth_grid(int B, int cams, int H, int W) {
    int bx = (H * W + kSmoothThreads * 4 - 1) / (kSmoothThreads * 4);
    if (bx > 256) bx = 256;
    return dim3(bx, B * cams);
}

// ---- no-allocation variant of the forward: the per-image disparity sums live in library-owned device memory ----
// File-scope device array: every GPU that loads the library holds these VFD_SMOOTH_STATIC_MAX_IMAGES * 8 bytes (32 KB)
// until the process exits.  In HOST code the name below is only a stand-in for the symbol, not a device address; the
// address is looked up once per device with cudaGetSymbolAddress (static_scratch) and that pointer goes to the kernels.
static __device__ __align__(256) double vfd_smooth_static_img_sum[VFD_SMOOTH_STATIC_MAX_IMAGES];

constexpr int kMaxDevices = 64;
// device address of the array on the CURRENT device, nullptr on failure.  One look-up per device and process; the
// pointer is published with release / acquire so that concurrent host threads either see it or look it up again (the
// look-up is idempotent).
static double* static_scratch() {
    static std::atomic<double*> cache[kMaxDevices];
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
    double* p = cache[dev].load(std::memory_order_acquire);
    if (p) return p;
    void* addr = nullptr;
    if (cudaGetSymbolAddress(&addr, vfd_smooth_static_img_sum) != cudaSuccess || !addr) {
        (void)cudaGetLastError();      // do not leave the failure behind for the caller's next launch check
        return nullptr;
    }
    p = static_cast<double*>(addr);
    cache[dev].store(p, std::memory_order_release);
    return p;
}

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_smooth_loss_fwd(int B, int cams, int H, int W, const float* color, const float* disp,
                                   double* img_sum, double* sums, void* stream) {
    int rc = check_smooth(B, cams, H, W);
    if (rc) return rc;
    if (!disp || !img_sum || !sums) return VFD_ERR_NULL;      // color may be NULL
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(img_sum, 0, sizeof(double) * B * cams, st) != cudaSuccess) return VFD_ERR_CUDA;
    if (cudaMemsetAsync(sums, 0, sizeof(double) * cams * 2, st) != cudaSuccess) return VFD_ERR_CUDA;
    SmoothArgs a = {};
    a.B = B; a.cams = cams; a.H = H; a.W = W; a.color = color; a.disp = disp; a.img_sum = img_sum; a.sums = sums;
    const dim3 grid = smooth_grid(B, cams, H, W);
    smooth_mean_kernel<<<grid, kSmoothThreads, 0, st>>>(a);
    smooth_loss_kernel<<<grid, kSmoothThreads, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

// Same two launches as vfd_smooth_loss_fwd with img_sum in the library's static device array: no cudaMalloc / cudaFree /
// allocating routine on this path (cudaGetSymbolAddress on a device's first call, cudaMemsetAsync, two launches).  The
// array is shared by every call on the device and keeps the last call's contents, so its first B*cams entries are
// cleared here, on `stream`; two calls that overlap on the GPU (different streams) race on it.
extern "C" int vfd_smooth_loss_fwd_static(int B, int cams, int H, int W, const float* color, const float* disp,
                                          double* sums, void* stream) {
    int rc = check_smooth(B, cams, H, W);
    if (rc) return rc;
    if ((long)B * cams > VFD_SMOOTH_STATIC_MAX_IMAGES) return VFD_ERR_STATIC_LIMIT;
    if (!disp || !sums) return VFD_ERR_NULL;      // color may be NULL
    double* img_sum = static_scratch();
    if (!img_sum) return VFD_ERR_CUDA;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(img_sum, 0, sizeof(double) * B * cams, st) != cudaSuccess) return VFD_ERR_CUDA;
    if (cudaMemsetAsync(sums, 0, sizeof(double) * cams * 2, st) != cudaSuccess) return VFD_ERR_CUDA;
    SmoothArgs a = {};
    a.B = B; a.cams = cams; a.H = H; a.W = W; a.color = color; a.disp = disp; a.img_sum = img_sum; a.sums = sums;
    const dim3 grid = smooth_grid(B, cams, H, W);
    smooth_mean_kernel<<<grid, kSmoothThreads, 0, st>>>(a);
    smooth_loss_kernel<<<grid, kSmoothThreads, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_smooth_loss_bwd(int B, int cams, int H, int W, const float* color, const float* disp,
                                   const double* img_sum, const float* d_smooth, double* gdot, float* d_disp,
                                   void* stream) {
    int rc = check_smooth(B, cams, H, W);
    if (rc) return rc;
    if (!disp || !img_sum || !d_smooth || !gdot || !d_disp) return VFD_ERR_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(gdot, 0, sizeof(double) * B * cams, st) != cudaSuccess) return VFD_ERR_CUDA;
    SmoothArgs a = {};
    a.B = B; a.cams = cams; a.H = H; a.W = W; a.color = color; a.disp = disp;
    a.img_sum = const_cast<double*>(img_sum); a.d_smooth = d_smooth; a.gdot = gdot; a.d_disp = d_disp;
    const dim3 grid = smooth_grid(B, cams, H, W);
    smooth_bwd_kernel<false><<<smooth_bwd_grid(B, cams, H, W), kSmoothThreads, 0, st>>>(a);
    smooth_bwd_finish_kernel<<<grid, kSmoothThreads, 0, st>>>(a, nullptr);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

// Training mode: loss sums + unit-upstream gradient map g [B,cams,H,W] + gdot in one stencil pass (after the means);
// vfd_smooth_loss_train_bwd turns g into d_disp once the upstream d_smooth [cams] is known.
extern "C" int vfd_smooth_loss_train_fwd(int B, int cams, int H, int W, const float* color, const float* disp,
                                         double* img_sum, double* sums, double* gdot, float* g, void* stream) {
    int rc = check_smooth(B, cams, H, W);
    if (rc) return rc;
    if (!disp || !img_sum || !sums || !gdot || !g) return VFD_ERR_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(img_sum, 0, sizeof(double) * B * cams, st) != cudaSuccess) return VFD_ERR_CUDA;
    if (cudaMemsetAsync(sums, 0, sizeof(double) * cams * 2, st) != cudaSuccess) return VFD_ERR_CUDA;
    if (cudaMemsetAsync(gdot, 0, sizeof(double) * B * cams, st) != cudaSuccess) return VFD_ERR_CUDA;
    SmoothArgs a = {};
    a.B = B; a.cams = cams; a.H = H; a.W = W; a.color = color; a.disp = disp; a.img_sum = img_sum; a.sums = sums;
    a.gdot = gdot; a.d_disp = g;
    const dim3 grid = smooth_grid(B, cams, H, W);
    smooth_mean_kernel<<<grid, kSmoothThreads, 0, st>>>(a);
    smooth_bwd_kernel<true><<<smooth_bwd_grid(B, cams, H, W), kSmoothThreads, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_smooth_loss_train_bwd(int B, int cams, int H, int W, const double* img_sum, const double* gdot,
                                         const float* g, const float* d_smooth, float* d_disp, void* stream) {
    int rc = check_smooth(B, cams, H, W);
    if (rc) return rc;
    if (!img_sum || !gdot || !g || !d_smooth || !d_disp) return VFD_ERR_NULL;
    SmoothArgs a = {};
    a.B = B; a.cams = cams; a.H = H; a.W = W; a.img_sum = const_cast<double*>(img_sum); a.gdot = const_cast<double*>(gdot);
    a.d_smooth = d_smooth; a.d_disp = d_disp;
    smooth_bwd_finish_kernel<<<smooth_grid(B, cams, H, W), kSmoothThreads, 0, (cudaStream_t)stream>>>(a, g);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
