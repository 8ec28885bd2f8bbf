// Ray sampling: VFNet.project_voxel_into_image without reduce_dim (network/volumetric_fusionnet.py:232-262),
// forward and backward.
//
// Layout.  Samples leave the kernel as out[b, cam, v, u, d, c] (channel-last, c fastest): the 64 channels of one
// (pixel, depth bin) sample are one contiguous 256 B run, written by one half-warp with 128-bit stores and read
// back the same way by the backward.  The reference's channel order c*D + d (:262) is a fixed permutation of
// reduce_dim's input channels; the host layer applies it to reduce_dim.0.weight (29 MB, once per step) instead
// of transposing 295 MB of samples per rig.
//
// Mapping.  One warp owns one feature-map pixel of one camera (its whole ray of D depth bins).
//   phase 1  lane <-> depth bin: the bin's 3-D point in ego space with the reference's exact fp32 operation
//            order (vfd_coords.h), un-normalised as ATen does, floor + 3 fractions + an 8-bit corner mask,
//            kept in shared memory (20 B per bin).
//   phase 2  vectorised kernel (Cv % 4 == 0, Cv/4 divides 32; Cv = 64 in both reference configs): the Cv/4 lanes
//            of a sub-warp hold 4 channels each of ONE bin, so a warp interpolates 32/(Cv/4) bins per iteration
//            with LDG.128 (256 B contiguous per corner).  Generic kernel (any Cv): lane <-> channel.
// The volume of one rig (51.2 MB at DDAD size) is L2-resident; consecutive bins of a ray and neighbouring
// pixels share voxel cells, so most corner lines hit in L1.  Warps walk pixels with a grid stride.
#include <limits.h>

#include "common.cuh"

namespace vfd {

constexpr int kRayWarps = 8;
// persistent-style grid: 148 SMs x resident CTAs x a few waves of work items
constexpr long kRayGrid = 148 * 8 * 2;      // generic (non-vectorised) kernels
// Column kernels: ONE resident wave of persistent CTAs, i.e. CTAs per SM = what the registers allow (forward: 64
// registers -> 4, backward: 128 -> 2).  Measured at B = 8 -- forward 4 / 6 / 8 / 12 / 16 / 32 CTAs per SM: 0.545 / 0.668 /
// 0.581 / 0.588 / 0.603 / 0.601 ms (6 = one and a half waves); backward 2 / 8 / 32 / 64 / 128: 0.688 / 0.725 / 0.706 /
// 0.708 / 0.709 ms.
constexpr long kRayGridFwd = 148 * 4;
constexpr long kRayGridBwd = 148 * 2;

struct RayArgs {
    vfd_fusion_desc d;
    const float* vol;
    const float* invK;
    const float* E;
    float* out;
    float* grid;
    const float* d_out;
    float* d_vol;
};

struct RaySlot {
    int base;         // ((iz0*Y + iy0)*X + ix0), may address outside the volume: guarded by `corners`
    float fx, fy, fz; // fractions
    unsigned corners; // bit t set <=> corner t (x + 2y + 4z) is inside the volume
    int pk;           // (ix0+1) | (iy0+1) << 10 | (iz0+1) << 20: two cells are face neighbours <=> pk differs by +-1, +-2^10, +-2^20
};
constexpr int kPkY = 1 << 10, kPkZ = 1 << 20;

__device__ __forceinline__ void ray_phase1(const RayArgs& a, int b, int c, int p, int lane, RaySlot* slot) {
    const vfd_fusion_desc& d = a.d;
    const float* invK = a.invK + (size_t)(b * d.num_cams + c) * 16;
    const float* E = a.E + (size_t)(b * d.num_cams + c) * 16;
    float Em[12], iK[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) { Em[i] = __ldg(E + i); iK[i] = __ldg(invK + i); }
    const int u = p % d.w, v = p / d.w;
    float ray[3];
    vfd_pixel_ray(iK, (float)u, (float)v, ray);
    for (int dd = lane; dd < d.D; dd += VFD_WARP) {
        float g[3];
        vfd_ray_point_to_grid(Em, ray, __ldg(d.bins + dd), d.vox_start, d.vox_length, g);
        if (a.grid) {
            float* go = a.grid + ((((size_t)(b * d.num_cams + c) * d.D + dd) * d.h * d.w) + p) * 3;
            go[0] = g[0]; go[1] = g[1]; go[2] = g[2];
        }
        const float ix = vfd_unnormalize(g[0], d.X), iy = vfd_unnormalize(g[1], d.Y), iz = vfd_unnormalize(g[2], d.Z);
        const float x0 = floorf(ix), y0 = floorf(iy), z0 = floorf(iz);
        RaySlot s;
        s.corners = 0;
        s.base = 0;
        s.pk = 0;
        s.fx = ix - x0; s.fy = iy - y0; s.fz = iz - z0;
        // anything that touches the volume has floor in [-1, size-1]; NaN fails every comparison
        if (x0 >= -1.f && x0 <= (float)(d.X - 1) && y0 >= -1.f && y0 <= (float)(d.Y - 1) && z0 >= -1.f &&
            z0 <= (float)(d.Z - 1)) {
            const int xi = (int)x0, yi = (int)y0, zi = (int)z0;
            s.base = (zi * d.Y + yi) * d.X + xi;
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                const int xx = xi + (t & 1), yy = yi + ((t >> 1) & 1), zz = zi + (t >> 2);
                if (xx >= 0 && xx < d.X && yy >= 0 && yy < d.Y && zz >= 0 && zz < d.Z) s.corners |= 1u << t;
            }
        }
        slot[dd] = s;
    }
}

__device__ __forceinline__ void corner_weights(const RaySlot& s, float* w) {
    const float ex = 1.f - s.fx, ey = 1.f - s.fy, ez = 1.f - s.fz;
    const float a = ex * ey, b = s.fx * ey, c = ex * s.fy, d = s.fx * s.fy;
    w[0] = a * ez; w[1] = b * ez; w[2] = c * ez; w[3] = d * ez;
    w[4] = a * s.fz; w[5] = b * s.fz; w[6] = c * s.fz; w[7] = d * s.fz;
}

// the same eight weights, each splatted into both halves of a packed register (products formed exactly as above)
__device__ __forceinline__ void corner_weights2(const RaySlot& s, f32x2* w) {
    const float ex = 1.f - s.fx, ey = 1.f - s.fy, ez = 1.f - s.fz;
    const f32x2 ex2 = pack2(ex, ex), fx2 = pack2(s.fx, s.fx), ey2 = pack2(ey, ey), fy2 = pack2(s.fy, s.fy);
    const f32x2 ez2 = pack2(ez, ez), fz2 = pack2(s.fz, s.fz);
    const f32x2 a = mul2p(ex2, ey2), b = mul2p(fx2, ey2), c = mul2p(ex2, fy2), d = mul2p(fx2, fy2);
    w[0] = mul2p(a, ez2); w[1] = mul2p(b, ez2); w[2] = mul2p(c, ez2); w[3] = mul2p(d, ez2);
    w[4] = mul2p(a, fz2); w[5] = mul2p(b, fz2); w[6] = mul2p(c, fz2); w[7] = mul2p(d, fz2);
}

extern __shared__ __align__(16) unsigned char vfd_ray_smem[];

static size_t ray_smem_bytes(int D) { return (size_t)kRayWarps * D * sizeof(RaySlot); }

struct Task {
    int b, c, p;
};
__device__ __forceinline__ Task decode_task(long task, int P, int cams) {
    Task t;
    t.p = (int)(task % P);
    const int bc = (int)(task / P);
    t.c = bc % cams;
    t.b = bc / cams;
    return t;
}

// ------------------------------------------------------------------------------------------------------------
// generic kernels: lane <-> channel (any Cv)
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kRayWarps* VFD_WARP) ray_sample_fwd_kernel(const RayArgs a) {
    const vfd_fusion_desc& d = a.d;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int P = d.h * d.w, D = d.D, Cv = d.Cv;
    const long ntasks = (long)d.B * d.num_cams * P;
    RaySlot* slot = reinterpret_cast<RaySlot*>(vfd_ray_smem) + warp * D;
    const int XC = d.X * Cv, YXC = d.Y * XC;
    for (long task = (long)blockIdx.x * kRayWarps + warp; task < ntasks; task += (long)gridDim.x * kRayWarps) {
        const Task t = decode_task(task, P, d.num_cams);
        ray_phase1(a, t.b, t.c, t.p, lane, slot);
        __syncwarp();
        const float* vol = a.vol + (size_t)t.b * d.X * d.Y * d.Z * Cv;
        float* orow = a.out + (size_t)task * D * Cv;
        for (int dd = 0; dd < D; ++dd) {
            const RaySlot s = slot[dd];
            float w[8];
            corner_weights(s, w);
            for (int ch = lane; ch < Cv; ch += VFD_WARP) {
                float val = 0.f;
                const float* p0 = vol + (long)s.base * Cv + ch;
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (s.corners & (1u << k))
                        val = fmaf(w[k], __ldg(p0 + (k & 1) * Cv + ((k >> 1) & 1) * XC + (k >> 2) * YXC), val);
                orow[dd * Cv + ch] = val;
            }
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(kRayWarps* VFD_WARP) ray_sample_bwd_kernel(const RayArgs a) {
    const vfd_fusion_desc& d = a.d;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int P = d.h * d.w, D = d.D, Cv = d.Cv;
    const long ntasks = (long)d.B * d.num_cams * P;
    RaySlot* slot = reinterpret_cast<RaySlot*>(vfd_ray_smem) + warp * D;
    const int XC = d.X * Cv, YXC = d.Y * XC;
    for (long task = (long)blockIdx.x * kRayWarps + warp; task < ntasks; task += (long)gridDim.x * kRayWarps) {
        const Task t = decode_task(task, P, d.num_cams);
        ray_phase1(a, t.b, t.c, t.p, lane, slot);
        __syncwarp();
        float* dvol = a.d_vol + (size_t)t.b * d.X * d.Y * d.Z * Cv;
        const float* grow = a.d_out + (size_t)task * D * Cv;
        for (int dd = 0; dd < D; ++dd) {
            const RaySlot s = slot[dd];
            if (!s.corners) continue;
            float w[8];
            corner_weights(s, w);
            for (int ch = lane; ch < Cv; ch += VFD_WARP) {
                const float g = __ldg(grow + dd * Cv + ch);
                float* p0 = dvol + (long)s.base * Cv + ch;
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (s.corners & (1u << k))
                        red_add_f1(p0 + (k & 1) * Cv + ((k >> 1) & 1) * XC + (k >> 2) * YXC, w[k] * g);
            }
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------------------
// vectorised kernels (CV % 4 == 0, CV/4 divides 32): COLUMN WALK.
//   One warp owns one image column u of one camera and kBins = 32/(CV/4) consecutive depth bins; the CV/4 lanes of
//   a sub-warp hold 4 channels each and walk the h pixels of the column top to bottom at a fixed bin.
//   Consecutive pixels of a column at the same depth differ by depth/f metres in ego height (0.1 .. 1.1 m) against
//   1.5 m voxels, so runs of pixels share the same 2x2x2 voxel cell:
//     forward   the 8 corner rows (8 x LDG.128) stay in registers while the cell does not change;
//     backward  the 8 corner contributions are accumulated in registers and leave as 8 vector reductions
//               (red.global.add.v4.f32) only when the cell changes -- ncu showed the one-reduction-per-sample
//               version saturating the L2 atomic units (profiles/).
//   Slots (cell index, fractions, corner mask) of the kBins x h samples are computed lane-parallel first.
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ RaySlot make_slot(const vfd_fusion_desc& d, const float* Em, const float* iK, int u, int v,
                                             float dep, float* g /*3, out*/) {
    float ray[3];
    vfd_pixel_ray(iK, (float)u, (float)v, ray);
    vfd_ray_point_to_grid(Em, ray, dep, d.vox_start, d.vox_length, g);
    const float ix = vfd_unnormalize(g[0], d.X), iy = vfd_unnormalize(g[1], d.Y), iz = vfd_unnormalize(g[2], d.Z);
    const float x0 = floorf(ix), y0 = floorf(iy), z0 = floorf(iz);
    RaySlot s;
    s.corners = 0;
    s.base = 0;
    s.pk = 0;
    s.fx = ix - x0; s.fy = iy - y0; s.fz = iz - z0;
    if (x0 >= -1.f && x0 <= (float)(d.X - 1) && y0 >= -1.f && y0 <= (float)(d.Y - 1) && z0 >= -1.f &&
        z0 <= (float)(d.Z - 1)) {
        const int xi = (int)x0, yi = (int)y0, zi = (int)z0;
        s.base = (zi * d.Y + yi) * d.X + xi;
        s.pk = (xi + 1) | ((yi + 1) << 10) | ((zi + 1) << 20);
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            const int xx = xi + (t & 1), yy = yi + ((t >> 1) & 1), zz = zi + (t >> 2);
            if (xx >= 0 && xx < d.X && yy >= 0 && yy < d.Y && zz >= 0 && zz < d.Z) s.corners |= 1u << t;
        }
    }
    return s;
}

// slot of the column kernels: the eight trilinear weights are formed once by the lane that builds the slot instead
// of by every one of the 16 lanes that later consume it (48 bytes = three 128-bit shared loads)
struct __align__(16) RaySlotW {
    int base;
    unsigned corners;
    float w[8];
    int pk;        // packed cell coordinates (RaySlot::pk)
    float pad;
};

// axis bit (1 = x, 2 = y, 4 = z) and direction of the step from cell `from` to cell `to` when they share a face, else 0
__device__ __forceinline__ int face_step(int from, int to, bool* pos) {
    const int dlt = to - from;
    *pos = dlt > 0;
    const int a = dlt > 0 ? dlt : -dlt;
    return a == 1 ? 1 : (a == kPkY ? 2 : (a == kPkZ ? 4 : 0));
}

struct ColTask {
    int b, c, u, d0;
};
template <int KBINS>
__device__ __forceinline__ ColTask decode_col(long task, const vfd_fusion_desc& d) {
    const int nd = (d.D + KBINS - 1) / KBINS;
    ColTask t;
    t.d0 = (int)(task % nd) * KBINS;
    long r = task / nd;
    t.u = (int)(r % d.w);
    r /= d.w;
    t.c = (int)(r % d.num_cams);
    t.b = (int)(r / d.num_cams);
    return t;
}

template <int KBINS>
__device__ __forceinline__ void col_phase1(const RayArgs& a, const ColTask& t, int lane, RaySlotW* slot) {
    const vfd_fusion_desc& d = a.d;
    const float* invK = a.invK + (size_t)(t.b * d.num_cams + t.c) * 16;
    const float* E = a.E + (size_t)(t.b * d.num_cams + t.c) * 16;
    float Em[12], iK[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) { Em[i] = __ldg(E + i); iK[i] = __ldg(invK + i); }
    for (int idx = lane; idx < KBINS * d.h; idx += VFD_WARP) {
        const int q = idx / d.h, v = idx % d.h, dd = t.d0 + q;
        RaySlot s;
        s.corners = 0; s.base = 0; s.pk = 0; s.fx = s.fy = s.fz = 0.f;
        if (dd < d.D) {
            float g[3];
            s = make_slot(d, Em, iK, t.u, v, __ldg(d.bins + dd), g);
            if (a.grid) {
                float* go = a.grid + ((((size_t)(t.b * d.num_cams + t.c) * d.D + dd) * d.h + v) * d.w + t.u) * 3;
                go[0] = g[0]; go[1] = g[1]; go[2] = g[2];
            }
        }
        RaySlotW sw;
        sw.base = s.base;
        sw.corners = s.corners;
        corner_weights(s, sw.w);
        sw.pk = s.pk;
        sw.pad = 0.f;
        slot[idx] = sw;
    }
}

template <int CV>
__global__ void __launch_bounds__(kRayWarps* VFD_WARP) ray_sample_fwd_vec_kernel(const RayArgs a) {
    constexpr int kLanes = CV / 4, kBins = VFD_WARP / kLanes;
    const vfd_fusion_desc& d = a.d;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = lane / kLanes, l = lane % kLanes;
    const int H = d.h, W = d.w, D = d.D;
    const long ntasks = (long)d.B * d.num_cams * W * ((D + kBins - 1) / kBins);
    RaySlotW* slot = reinterpret_cast<RaySlotW*>(vfd_ray_smem) + warp * kBins * H;
    const int XC = d.X * CV, YXC = d.Y * XC;
    for (long task = (long)blockIdx.x * kRayWarps + warp; task < ntasks; task += (long)gridDim.x * kRayWarps) {
        const ColTask t = decode_col<kBins>(task, d);
        col_phase1<kBins>(a, t, lane, slot);
        __syncwarp();
        const int dd = t.d0 + q;
        if (dd < D) {
            const float* vol = a.vol + (size_t)t.b * d.X * d.Y * d.Z * CV + 4 * l;
            float* o = a.out + ((((size_t)(t.b * d.num_cams + t.c) * H) * W + t.u) * D + dd) * CV + 4 * l;
            const size_t ostride = (size_t)W * D * CV;
            int cur_base = INT_MIN;
            float4 vv[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) vv[k] = f4_zero();
            for (int v = 0; v < H; ++v, o += ostride) {      // (output row pointer advanced by addition, as in the backward)
                const RaySlotW s = slot[q * H + v];
                float4 r = f4_zero();
                if (s.corners) {
                    // (carrying the four corner rows a face step shares, as the backward does with its accumulators, was
                    // measured at 1.18 ms against 0.65 ms: 106 instead of 64 registers, and the reloads hit L1 anyway;
                    // requesting the next pixel's cell into L1 one row ahead with prefetch.global.L1 -- half of the stall
                    // samples are these eight loads -- 0.69 / 0.71 ms against 0.60 ms: the extra instructions cost more;
                    // carrying only the lower face on the one step direction that occurs (z - 1, see the backward) with
                    // four loads instead of eight -- 0.85 ms at 80 registers, 0.84 ms forced to 64)
                    if (s.base != cur_base) {
                        cur_base = s.base;
                        const float* p0 = vol + (long)s.base * CV;
                        if (s.corners == 0xffu) {      // interior cell (the usual case): eight unconditional loads
#pragma unroll
                            for (int k = 0; k < 8; ++k) vv[k] = ldg_f4(p0 + (k & 1) * CV + ((k >> 1) & 1) * XC + (k >> 2) * YXC);
                        } else {
#pragma unroll
                            for (int k = 0; k < 8; ++k)
                                vv[k] = (s.corners & (1u << k))
                                            ? ldg_f4(p0 + (k & 1) * CV + ((k >> 1) & 1) * XC + (k >> 2) * YXC)
                                            : f4_zero();
                        }
                    }
                    // packed fma.rn.f32x2: the same bits as four fmaf per corner at half the issue slots and half the
                    // fma-pipe cycles (a three-register FFMA and an FFMA2 both take the pipe for two cycles)
#pragma unroll
                    for (int k = 0; k < 8; ++k) f4_fma_w2(r, pack2(s.w[k], s.w[k]), vv[k]);
                }
                st_stream_f4(o, r);
            }
        }
        __syncwarp();
    }
}

template <int CV>
__global__ void __launch_bounds__(kRayWarps* VFD_WARP, 2) ray_sample_bwd_vec_kernel(const RayArgs a) {
    constexpr int kLanes = CV / 4, kBins = VFD_WARP / kLanes;
    const vfd_fusion_desc& d = a.d;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = lane / kLanes, l = lane % kLanes;
    const int H = d.h, W = d.w, D = d.D;
    const long ntasks = (long)d.B * d.num_cams * W * ((D + kBins - 1) / kBins);
    RaySlotW* slot = reinterpret_cast<RaySlotW*>(vfd_ray_smem) + warp * kBins * H;
    const int XC = d.X * CV, YXC = d.Y * XC;
    const bool carry = d.X < 1023 && d.Y < 1023 && d.Z < 1023;      // packed cell coordinates hold 10 bits per axis
    for (long task = (long)blockIdx.x * kRayWarps + warp; task < ntasks; task += (long)gridDim.x * kRayWarps) {
        const ColTask t = decode_col<kBins>(task, d);
        col_phase1<kBins>(a, t, lane, slot);
        __syncwarp();
        const int dd = t.d0 + q;
        if (dd < D) {
            float* dvol = a.d_vol + (size_t)t.b * d.X * d.Y * d.Z * CV + 4 * l;
            const float* gp = a.d_out + ((((size_t)(t.b * d.num_cams + t.c) * H) * W + t.u) * D + dd) * CV + 4 * l;
            const size_t gstride = (size_t)W * D * CV;
            int cur_base = INT_MIN, cur_pk = 0;
            unsigned cur_corners = 0;
            float4 acc[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] = f4_zero();
            // the gradient column is read once, straight from HBM: keep kPf rows per lane in flight (next chunk is
            // requested before the current one is consumed).
            // Measured alternatives, all slower than this form's 0.85 ms at B=8 (that kernel is in the history: commit 51c8f75, which ran the register form at bench size because of a shared-memory cap): the column
            // staged through a per-warp shared-memory ring by cp.async.bulk with mbarrier completion, 512-byte rows,
            // 16 / 12 / 8 rows in flight, issue cursor running ahead across tasks -- 1.31 / 1.33 / 1.24 ms at 2 and 3
            // CTAs per SM (one wait + warp sync + re-arm per 512-byte row costs more than the registers it frees);
            // kPf = 2 or 1 at 3 CTAs per SM -- 0.87 / 1.04 ms; a task order that spreads the resident warps over the
            // six cameras and distant columns -- 0.95 ms (the reductions want the L2 locality of neighbouring
            // columns, not more parallelism).
            constexpr int kPf = 6;      // rows in flight per lane; 3 / 4 / 6 / 8: 0.775 / 0.743 / 0.722 / 0.740 ms
            // Instruction count matters as much as the memory system here (ncu: 507 M warp instructions per launch, only
            // 106 M of them FFMA2 / FMUL2; issue 58 % at 16 warps/SM): advancing the row pointer by addition instead of a
            // 64-bit multiply per row took 0.843 -> 0.805 ms, carrying the face for the one step direction that occurs
            // (below) 0.806 -> 0.743 ms.  While the loop body still held three face-step variants it sat at the
            // instruction cache's limit: unrolling it 8 times to swap two register buffers without the copy below ran at
            // 1.14 ms, refilling g[i] in place right after row i is consumed 1.15 / 1.94 / 2.84 ms at kPf = 4 / 6 / 8.
            // Three CTAs per SM (80 registers): 0.77 / 0.98 ms at kPf = 2 / 4.
            float4 g[kPf], gn[kPf];
#pragma unroll
            for (int i = 0; i < kPf; ++i) g[i] = (i < H) ? ld_stream_f4(gp + i * gstride) : f4_zero();
            const float* gnext = gp + kPf * gstride;
            for (int v0 = 0; v0 < H; v0 += kPf) {
#pragma unroll
                for (int i = 0; i < kPf; ++i)
                    gn[i] = (v0 + kPf + i < H) ? ld_stream_f4(gnext + i * gstride) : f4_zero();
                gnext += kPf * gstride;
#pragma unroll
                for (int i = 0; i < kPf; ++i) {
                    const int v = v0 + i;
                    if (v >= H) break;
                    const RaySlotW s = slot[q * H + v];
                    if (s.corners) {
                        if (s.base != cur_base) {
                            float* p0 = dvol + (long)cur_base * CV;
                            // Walking down an image column the ray end drops: counted on the CPU at DDAD geometry, 58 % of the
                            // steps stay in the cell, 26.4 % enter the FACE neighbour below (z - 1) and every other kind of
                            // step together is 0.5 %.  That one face step keeps the four accumulators on the shared face
                            // (the old cell's lower corners are the new cell's upper ones): only the four corners that
                            // leave are reduced into global memory, which halves the L2 reduction traffic the kernel is
                            // bound by.  (A form that carried the face for all six directions was three times this code
                            // with a select per register: 0.806 ms.)
                            if (cur_corners && carry && s.pk == cur_pk - kPkZ) {
#pragma unroll
                                for (int lo = 0; lo < 4; ++lo) {
                                    const int hi = lo | 4;
                                    if (cur_corners & (1u << hi))
                                        red_add_f4(p0 + (hi & 1) * CV + ((hi >> 1) & 1) * XC + YXC, acc[hi]);
                                    acc[hi] = acc[lo];
                                    f4_fma_w2(acc[hi], pack2(s.w[hi], s.w[hi]), g[i]);
                                    f4_mul_w2(acc[lo], pack2(s.w[lo], s.w[lo]), g[i]);
                                }
                            } else {
                                if (cur_corners) {
#pragma unroll
                                    for (int k = 0; k < 8; ++k)
                                        if (cur_corners & (1u << k))
                                            red_add_f4(p0 + (k & 1) * CV + ((k >> 1) & 1) * XC + (k >> 2) * YXC, acc[k]);
                                }
                                // new cell: the accumulators start from the product (same value as 0 + w * g)
#pragma unroll
                                for (int k = 0; k < 8; ++k) f4_mul_w2(acc[k], pack2(s.w[k], s.w[k]), g[i]);
                            }
                            cur_base = s.base;
                            cur_pk = s.pk;
                            cur_corners = s.corners;
                        } else {
#pragma unroll
                            for (int k = 0; k < 8; ++k) f4_fma_w2(acc[k], pack2(s.w[k], s.w[k]), g[i]);
                        }
                    }
                }
#pragma unroll
                for (int i = 0; i < kPf; ++i) g[i] = gn[i];
            }
            if (cur_corners) {
                float* p0 = dvol + (long)cur_base * CV;
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (cur_corners & (1u << k))
                        red_add_f4(p0 + (k & 1) * CV + ((k >> 1) & 1) * XC + (k >> 2) * YXC, acc[k]);
            }
        }
        __syncwarp();
    }
}

static int check_ray_desc(const vfd_fusion_desc* d) {
    if (!d) return VFD_ERR_NULL;
    if (d->B < 1 || d->num_cams < 1 || d->num_cams > VFD_MAX_CAMS || d->h < 1 || d->w < 1 || d->X < 1 ||
        d->Y < 1 || d->Z < 1 || d->D < 1 || d->Cv < 1)
        return VFD_ERR_DESC;
    if (!d->bins) return VFD_ERR_NULL;
    if ((long)d->X * d->Y * d->Z * d->Cv >= (1L << 31)) return VFD_ERR_DESC;
    if (ray_smem_bytes(d->D) > 48 * 1024) return VFD_ERR_UNSUPPORTED;
    return VFD_OK;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

static int launch_ray(const RayArgs& a, bool bwd, cudaStream_t st) {
    const vfd_fusion_desc& d = a.d;
    const long tasks = (long)d.B * d.num_cams * d.h * d.w;
    const long blocks = (tasks + kRayWarps - 1) / kRayWarps;
    const unsigned grid = (unsigned)(blocks < kRayGrid ? blocks : kRayGrid);
    const size_t smem = ray_smem_bytes(d.D);
    const bool vec_ok = bwd ? (aligned16(a.d_out) && aligned16(a.d_vol)) : (aligned16(a.out) && aligned16(a.vol));
#define VFD_RAY_VEC(CVAL)                                                                                   \
    if (d.Cv == CVAL && vec_ok) {                                                                           \
        constexpr int kb = VFD_WARP / (CVAL / 4);                                                           \
        const size_t vsmem = (size_t)kRayWarps * kb * d.h * sizeof(RaySlotW);                               \
        const long vtasks = (long)d.B * d.num_cams * d.w * ((d.D + kb - 1) / kb);                           \
        const long vblocks = (vtasks + kRayWarps - 1) / kRayWarps;                                          \
        const long vcap = bwd ? kRayGridBwd : kRayGridFwd;                                                  \
        const unsigned vgrid = (unsigned)(vblocks < vcap ? vblocks : vcap);                                 \
        if (vsmem <= 48 * 1024) {                                                                           \
            if (bwd) ray_sample_bwd_vec_kernel<CVAL><<<vgrid, kRayWarps * VFD_WARP, vsmem, st>>>(a);        \
            else ray_sample_fwd_vec_kernel<CVAL><<<vgrid, kRayWarps * VFD_WARP, vsmem, st>>>(a);            \
            VFD_CHECK_LAUNCH();                                                                             \
            return VFD_OK;                                                                                  \
        }                                                                                                   \
    }
    VFD_RAY_VEC(64)
    VFD_RAY_VEC(32)
    VFD_RAY_VEC(128)
#undef VFD_RAY_VEC
    if (bwd) ray_sample_bwd_kernel<<<grid, kRayWarps * VFD_WARP, smem, st>>>(a);
    else ray_sample_fwd_kernel<<<grid, kRayWarps * VFD_WARP, smem, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_ray_sample_fwd(const vfd_fusion_desc* d, const float* vol, const float* invK, const float* E,
                                  float* out, float* grid, void* stream) {
    int rc = check_ray_desc(d);
    if (rc) return rc;
    if (!vol || !invK || !E || !out) return VFD_ERR_NULL;
    RayArgs a = {};
    a.d = *d; a.vol = vol; a.invK = invK; a.E = E; a.out = out; a.grid = grid;
    return launch_ray(a, false, (cudaStream_t)stream);
}

extern "C" int vfd_ray_sample_bwd(const vfd_fusion_desc* d, const float* invK, const float* E, const float* d_out,
                                  float* d_vol, void* stream) {
    int rc = check_ray_desc(d);
    if (rc) return rc;
    if (!invK || !E || !d_out || !d_vol) return VFD_ERR_NULL;
    RayArgs a = {};
    a.d = *d; a.invK = invK; a.E = E; a.d_out = d_out; a.d_vol = d_vol;
    const size_t bytes = (size_t)d->B * d->X * d->Y * d->Z * d->Cv * sizeof(float);
    if (cudaMemsetAsync(d_vol, 0, bytes, (cudaStream_t)stream) != cudaSuccess) return VFD_ERR_CUDA;
    return launch_ray(a, true, (cudaStream_t)stream);
}
