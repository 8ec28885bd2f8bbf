// Ray sampling: VFNet.project_voxel_into_image without reduce_dim (network/volumetric_fusionnet.py:232-262),
// forward and backward.
//
// Mapping.  One warp owns one feature-map pixel of one camera (its whole ray of D depth bins).
//   phase 1  lane <-> depth bin: the bin's 3-D point in ego space with the reference's exact fp32 operation
//            order (vfd_coords.h), un-normalised as ATen does, floor + 3 fractions + an 8-bit corner mask,
//            kept in shared memory.
//   phase 2  lane <-> channel (32 channels = one 128 B line of the channel-last volume per corner): walk the
//            D bins, trilinear-interpolate, park the result in a [channel][D+1] shared tile (conflict-free
//            both ways), then stream the tile out as the contiguous run out[pixel][c*D + d] with 128-bit
//            stores.  The reference layout channel = c*D + d (:262) is produced directly, channel-last.
// The volume of one rig (51.2 MB at DDAD size) is L2-resident; consecutive bins of a ray and neighbouring
// pixels share voxel cells, so most corner lines hit in L1.
#include "common.cuh"

namespace vfd {

constexpr int kRayWarps = 8;

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
};

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

__device__ __forceinline__ float corner_weight(const RaySlot& s, int t) {
    const float wx = (t & 1) ? s.fx : 1.f - s.fx;
    const float wy = (t & 2) ? s.fy : 1.f - s.fy;
    const float wz = (t & 4) ? s.fz : 1.f - s.fz;
    return wx * wy * wz;
}

extern __shared__ __align__(16) unsigned char vfd_ray_smem[];

__device__ __forceinline__ void ray_smem_carve(int D, int warp, RaySlot*& slot, float*& tile) {
    RaySlot* slots = reinterpret_cast<RaySlot*>(vfd_ray_smem);
    float* tiles = reinterpret_cast<float*>(slots + kRayWarps * D);
    slot = slots + warp * D;
    tile = tiles + (size_t)warp * VFD_WARP * (D + 1);
}

static size_t ray_smem_bytes(int D) {
    return (size_t)kRayWarps * D * sizeof(RaySlot) + (size_t)kRayWarps * VFD_WARP * (D + 1) * sizeof(float);
}

__global__ void __launch_bounds__(kRayWarps* VFD_WARP) ray_sample_fwd_kernel(const RayArgs a) {
    const vfd_fusion_desc& d = a.d;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int P = d.h * d.w, D = d.D, Cv = d.Cv;
    const long task = (long)blockIdx.x * kRayWarps + warp;  // (b*cams + c)*P + p
    if (task >= (long)d.B * d.num_cams * P) return;
    const int p = (int)(task % P), bc = (int)(task / P), c = bc % d.num_cams, b = bc / d.num_cams;
    RaySlot* slot;
    float* tile;
    ray_smem_carve(D, warp, slot, tile);
    ray_phase1(a, b, c, p, lane, slot);
    __syncwarp();
    const float* vol = a.vol + (size_t)b * d.X * d.Y * d.Z * Cv;
    const int XC = d.X * Cv, YXC = d.Y * XC;
    float* orow = a.out + (size_t)task * Cv * D;
    for (int cg = 0; cg < Cv; cg += VFD_WARP) {
        const int ch = cg + lane;
        const int nch = min(VFD_WARP, Cv - cg);
        for (int dd = 0; dd < D; ++dd) {
            const RaySlot s = slot[dd];
            float val = 0.f;
            if (s.corners && ch < Cv) {
                const float* p0 = vol + (long)s.base * Cv + ch;
#pragma unroll
                for (int t = 0; t < 8; ++t)
                    if (s.corners & (1u << t))
                        val = fmaf(corner_weight(s, t), __ldg(p0 + (t & 1) * Cv + ((t >> 1) & 1) * XC + (t >> 2) * YXC), val);
            }
            tile[lane * (D + 1) + dd] = val;
        }
        __syncwarp();
        // the group's outputs are the contiguous run orow[cg*D .. cg*D + nch*D)
        float* run = orow + (size_t)cg * D;
        const int len = nch * D;
        if (((reinterpret_cast<uintptr_t>(run) & 15) == 0) && (len % 4 == 0)) {
            for (int e = lane * 4; e < len; e += VFD_WARP * 4) {
                float4 v;
                v.x = tile[((e + 0) / D) * (D + 1) + (e + 0) % D];
                v.y = tile[((e + 1) / D) * (D + 1) + (e + 1) % D];
                v.z = tile[((e + 2) / D) * (D + 1) + (e + 2) % D];
                v.w = tile[((e + 3) / D) * (D + 1) + (e + 3) % D];
                st_stream_f4(run + e, v);
            }
        } else {
            for (int e = lane; e < len; e += VFD_WARP) run[e] = tile[(e / D) * (D + 1) + e % D];
        }
        __syncwarp();
    }
}

// Backward: d_vol[corner, ch] += w_corner * d_out[pixel][ch*D + d]; the run of one channel group is first
// brought into the shared tile with coalesced 128-bit loads, then scattered bin by bin (one 128 B reduction per
// corner per warp instruction).
__global__ void __launch_bounds__(kRayWarps* VFD_WARP) ray_sample_bwd_kernel(const RayArgs a) {
    const vfd_fusion_desc& d = a.d;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int P = d.h * d.w, D = d.D, Cv = d.Cv;
    const long task = (long)blockIdx.x * kRayWarps + warp;
    if (task >= (long)d.B * d.num_cams * P) return;
    const int p = (int)(task % P), bc = (int)(task / P), c = bc % d.num_cams, b = bc / d.num_cams;
    RaySlot* slot;
    float* tile;
    ray_smem_carve(D, warp, slot, tile);
    RayArgs a2 = a;
    a2.grid = nullptr;
    ray_phase1(a2, b, c, p, lane, slot);
    __syncwarp();
    float* dvol = a.d_vol + (size_t)b * d.X * d.Y * d.Z * Cv;
    const int XC = d.X * Cv, YXC = d.Y * XC;
    const float* grow = a.d_out + (size_t)task * Cv * D;
    for (int cg = 0; cg < Cv; cg += VFD_WARP) {
        const int ch = cg + lane;
        const int nch = min(VFD_WARP, Cv - cg);
        const float* run = grow + (size_t)cg * D;
        const int len = nch * D;
        if (((reinterpret_cast<uintptr_t>(run) & 15) == 0) && (len % 4 == 0)) {
            for (int e = lane * 4; e < len; e += VFD_WARP * 4) {
                const float4 v = ld_stream_f4(run + e);
                tile[((e + 0) / D) * (D + 1) + (e + 0) % D] = v.x;
                tile[((e + 1) / D) * (D + 1) + (e + 1) % D] = v.y;
                tile[((e + 2) / D) * (D + 1) + (e + 2) % D] = v.z;
                tile[((e + 3) / D) * (D + 1) + (e + 3) % D] = v.w;
            }
        } else {
            for (int e = lane; e < len; e += VFD_WARP) tile[(e / D) * (D + 1) + e % D] = __ldg(run + e);
        }
        __syncwarp();
        for (int dd = 0; dd < D; ++dd) {
            const RaySlot s = slot[dd];
            if (!s.corners || ch >= Cv) continue;
            const float g = tile[lane * (D + 1) + dd];
            if (g == 0.f) continue;
            float* p0 = dvol + (long)s.base * Cv + ch;
#pragma unroll
            for (int t = 0; t < 8; ++t)
                if (s.corners & (1u << t))
                    red_add_f1(p0 + (t & 1) * Cv + ((t >> 1) & 1) * XC + (t >> 2) * YXC, corner_weight(s, t) * g);
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------
// Vectorised variant for CV channels with CV % 4 == 0 and CV/4 a divisor of 32 (CV = 64 in both reference configs).
//   lane = (sub-warp q, l): the CV/4 lanes of a sub-warp hold 4 channels each of ONE depth bin, so a warp
//   interpolates 32/(CV/4) bins per iteration with LDG.128 (256 B contiguous per corner at CV=64) -- 4x fewer
//   issued instructions than the generic kernel, which ncu showed to be issue-bound (profiles/).
//   The [channel][D] shared tile is exactly the output run out[pixel][c*D+d], so it leaves the SM as ONE bulk
//   asynchronous copy (cp.async.bulk, the TMA engine) per pixel, overlapped with the next pixel's coordinates.
// Warps walk pixels with a grid stride so the tile, the slots and the mbarrier are set up once.
// ---------------------------------------------------------------------------------------------
template <int CV>
struct VecCfg {
    static constexpr int kLanes = CV / 4;            // lanes per depth bin
    static constexpr int kBins = VFD_WARP / kLanes;  // bins per warp iteration
};

template <int CV>
__device__ __forceinline__ void vec_smem_carve(int D, int warp, float*& tile, RaySlot*& slot, uint64_t*& bar) {
    float* tiles = reinterpret_cast<float*>(vfd_ray_smem);
    RaySlot* slots = reinterpret_cast<RaySlot*>(tiles + (size_t)kRayWarps * CV * D);
    uint64_t* bars = reinterpret_cast<uint64_t*>(slots + (size_t)kRayWarps * D + (((size_t)kRayWarps * D) & 1));
    tile = tiles + (size_t)warp * CV * D;
    slot = slots + (size_t)warp * D;
    bar = bars + warp;
}

template <int CV>
static size_t vec_smem_bytes(int D) {
    return (size_t)kRayWarps * CV * D * sizeof(float) + ((size_t)kRayWarps * D + 1) * sizeof(RaySlot) +
           kRayWarps * sizeof(uint64_t) + 16;
}

__device__ __forceinline__ void corner_weights(const RaySlot& s, float* w) {
    const float ex = 1.f - s.fx, ey = 1.f - s.fy, ez = 1.f - s.fz;
    const float a = ex * ey, b = s.fx * ey, c = ex * s.fy, d = s.fx * s.fy;
    w[0] = a * ez; w[1] = b * ez; w[2] = c * ez; w[3] = d * ez;
    w[4] = a * s.fz; w[5] = b * s.fz; w[6] = c * s.fz; w[7] = d * s.fz;
}

template <int CV>
__global__ void __launch_bounds__(kRayWarps* VFD_WARP) ray_sample_fwd_vec_kernel(const RayArgs a) {
    using Cfg = VecCfg<CV>;
    const vfd_fusion_desc& d = a.d;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = lane / Cfg::kLanes, l = lane % Cfg::kLanes;
    const int P = d.h * d.w, D = d.D;
    const long ntasks = (long)d.B * d.num_cams * P;
    float* tile;
    RaySlot* slot;
    uint64_t* bar;
    vec_smem_carve<CV>(D, warp, tile, slot, bar);
    const int XC = d.X * CV, YXC = d.Y * XC;
    for (long task = (long)blockIdx.x * kRayWarps + warp; task < ntasks; task += (long)gridDim.x * kRayWarps) {
        const int p = (int)(task % P), bc = (int)(task / P), c = bc % d.num_cams, b = bc / d.num_cams;
        ray_phase1(a, b, c, p, lane, slot);
        if (lane == 0) bulk_wait_read0();   // the previous pixel's tile has left shared memory
        __syncwarp();
        const float* vol = a.vol + (size_t)b * d.X * d.Y * d.Z * CV + 4 * l;
        for (int d0 = 0; d0 < D; d0 += Cfg::kBins) {
            const int dd = d0 + q;
            if (dd < D) {
                const RaySlot s = slot[dd];
                float4 r = f4_zero();
                if (s.corners) {
                    float w[8];
                    corner_weights(s, w);
                    const float* p0 = vol + (long)s.base * CV;
                    if (s.corners == 0xFFu) {
                        float4 v[8];
#pragma unroll
                        for (int t = 0; t < 8; ++t) v[t] = ldg_f4(p0 + (t & 1) * CV + ((t >> 1) & 1) * XC + (t >> 2) * YXC);
#pragma unroll
                        for (int t = 0; t < 8; ++t) f4_fma(r, w[t], v[t]);
                    } else {
#pragma unroll
                        for (int t = 0; t < 8; ++t)
                            if (s.corners & (1u << t))
                                f4_fma(r, w[t], ldg_f4(p0 + (t & 1) * CV + ((t >> 1) & 1) * XC + (t >> 2) * YXC));
                    }
                }
                float* t0 = tile + (4 * l) * D + dd;
                t0[0] = r.x; t0[D] = r.y; t0[2 * D] = r.z; t0[3 * D] = r.w;
            }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            bulk_store_s2g(a.out + (size_t)task * CV * D, tile, (uint32_t)(CV * D * sizeof(float)));
            bulk_commit();
        }
    }
    if (lane == 0) bulk_wait_read0();
}

template <int CV>
__global__ void __launch_bounds__(kRayWarps* VFD_WARP) ray_sample_bwd_vec_kernel(const RayArgs a) {
    using Cfg = VecCfg<CV>;
    const vfd_fusion_desc& d = a.d;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = lane / Cfg::kLanes, l = lane % Cfg::kLanes;
    const int P = d.h * d.w, D = d.D;
    const long ntasks = (long)d.B * d.num_cams * P;
    float* tile;
    RaySlot* slot;
    uint64_t* bar;
    vec_smem_carve<CV>(D, warp, tile, slot, bar);
    if (lane == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const int XC = d.X * CV, YXC = d.Y * XC;
    const uint32_t bytes = (uint32_t)(CV * D * sizeof(float));
    uint32_t phase = 0;
    for (long task = (long)blockIdx.x * kRayWarps + warp; task < ntasks; task += (long)gridDim.x * kRayWarps) {
        const int p = (int)(task % P), bc = (int)(task / P), c = bc % d.num_cams, b = bc / d.num_cams;
        if (lane == 0) {   // the pixel's upstream-gradient run arrives as one bulk copy while phase 1 runs
            mbar_expect_tx(bar, bytes);
            bulk_load_g2s(tile, a.d_out + (size_t)task * CV * D, bytes, bar);
        }
        ray_phase1(a, b, c, p, lane, slot);
        __syncwarp();
        mbar_wait(bar, phase);
        phase ^= 1;
        float* dvol = a.d_vol + (size_t)b * d.X * d.Y * d.Z * CV + 4 * l;
        for (int d0 = 0; d0 < D; d0 += Cfg::kBins) {
            const int dd = d0 + q;
            if (dd < D) {
                const RaySlot s = slot[dd];
                if (s.corners) {
                    const float* t0 = tile + (4 * l) * D + dd;
                    const float4 g = make_float4(t0[0], t0[D], t0[2 * D], t0[3 * D]);
                    float w[8];
                    corner_weights(s, w);
                    float* p0 = dvol + (long)s.base * CV;
#pragma unroll
                    for (int t = 0; t < 8; ++t)
                        if (s.corners & (1u << t))
                            red_add_f4(p0 + (t & 1) * CV + ((t >> 1) & 1) * XC + (t >> 2) * YXC, f4_scale(w[t], g));
                }
            }
        }
        __syncwarp();   // every lane is done with the tile before the next bulk load overwrites it
    }
}

static int check_ray_desc(const vfd_fusion_desc* d) {
    if (!d) return VFD_ERR_NULL;
    if (d->B < 1 || d->num_cams < 1 || d->num_cams > VFD_MAX_CAMS || d->h < 1 || d->w < 1 || d->X < 1 ||
        d->Y < 1 || d->Z < 1 || d->D < 1 || d->Cv < 1)
        return VFD_ERR_DESC;
    if (!d->bins) return VFD_ERR_NULL;
    if ((long)d->X * d->Y * d->Z * d->Cv >= (1L << 31)) return VFD_ERR_DESC;
    if (ray_smem_bytes(d->D) > 200 * 1024) return VFD_ERR_UNSUPPORTED;
    return VFD_OK;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
// persistent-style grid for the vectorised kernels: 148 SMs x 2 resident CTAs x 4 waves of work items
constexpr long kVecGrid = 148 * 2 * 4;

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_ray_sample_fwd(const vfd_fusion_desc* d, const float* vol, const float* invK, const float* E,
                                  float* out, float* grid, void* stream) {
    int rc = check_ray_desc(d);
    if (rc) return rc;
    if (!vol || !invK || !E || !out) return VFD_ERR_NULL;
    RayArgs a = {};
    a.d = *d; a.vol = vol; a.invK = invK; a.E = E; a.out = out; a.grid = grid;
    const long tasks = (long)d->B * d->num_cams * d->h * d->w;
    if (d->Cv == 64 && vec_smem_bytes<64>(d->D) <= 110 * 1024 && aligned16(out) && aligned16(vol)) {
        const size_t smem = vec_smem_bytes<64>(d->D);
        if (cudaFuncSetAttribute(ray_sample_fwd_vec_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem) != cudaSuccess)
            return VFD_ERR_CUDA;
        const long blocks = (tasks + kRayWarps - 1) / kRayWarps;
        ray_sample_fwd_vec_kernel<64><<<(unsigned)(blocks < kVecGrid ? blocks : kVecGrid), kRayWarps * VFD_WARP, smem,
                                        (cudaStream_t)stream>>>(a);
        VFD_CHECK_LAUNCH();
        return VFD_OK;
    }
    const size_t smem = ray_smem_bytes(d->D);
    if (cudaFuncSetAttribute(ray_sample_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
        cudaSuccess)
        return VFD_ERR_CUDA;
    ray_sample_fwd_kernel<<<(unsigned)((tasks + kRayWarps - 1) / kRayWarps), kRayWarps * VFD_WARP, smem,
                            (cudaStream_t)stream>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
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
    const long tasks = (long)d->B * d->num_cams * d->h * d->w;
    if (d->Cv == 64 && vec_smem_bytes<64>(d->D) <= 110 * 1024 && aligned16(d_out) && aligned16(d_vol)) {
        const size_t smem = vec_smem_bytes<64>(d->D);
        if (cudaFuncSetAttribute(ray_sample_bwd_vec_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem) != cudaSuccess)
            return VFD_ERR_CUDA;
        const long blocks = (tasks + kRayWarps - 1) / kRayWarps;
        ray_sample_bwd_vec_kernel<64><<<(unsigned)(blocks < kVecGrid ? blocks : kVecGrid), kRayWarps * VFD_WARP, smem,
                                        (cudaStream_t)stream>>>(a);
        VFD_CHECK_LAUNCH();
        return VFD_OK;
    }
    const size_t smem = ray_smem_bytes(d->D);
    if (cudaFuncSetAttribute(ray_sample_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
        cudaSuccess)
        return VFD_ERR_CUDA;
    ray_sample_bwd_kernel<<<(unsigned)((tasks + kRayWarps - 1) / kRayWarps), kRayWarps * VFD_WARP, smem,
                            (cudaStream_t)stream>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
