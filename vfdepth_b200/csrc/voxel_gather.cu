// Voxel gather: VFNet.backproject_into_voxel (network/volumetric_fusionnet.py:116-164) forward and backward.
//
// Mapping.  One warp owns 32 consecutive voxels (x fastest) of one rig.
//   phase 1  lane <-> voxel: project the voxel into every camera with the exact fp32 operation order of
//            the reference (vfd_coords.h), decide validity, keep (texel index, 2 weights, relative depth) per
//            camera in shared memory and the camera bit-mask in a register.
//   phase 2  lane <-> 4 channels: for each of the 32 voxels the warp reads the 4 bilinear corners of every
//            VALID camera as contiguous channel-last rows (C floats = 1 KB at C=256, 512 B per LDG.128 warp
//            instruction), accumulates the two overlap-group sums in registers and streams the finished rows
//            out with 128-bit stores.  The per-camera [B,257,N] tensors of the reference are never formed.
// The feature maps of one rig (23.6 MB at DDAD size) stay L2-resident while its 200 000 voxels are processed
// (blockIdx.y = rig), so HBM traffic is the 617 MB/rig (depth) or 206 MB/rig (pose) of output rows.
#include "gather_common.cuh"

namespace vfd {

constexpr int kGatherWarps = 8;

struct GatherArgs {
    vfd_fusion_desc d;
    const float* feats;
    const float* mask_small;
    const float* K;
    const float* Einv;
    float* x_no;      // depth: x_no, pose: avg
    float* x_ov;
    uint8_t* count;
    float* grid;
    uint8_t* valid;
    // backward
    const float* d_x_no;
    const float* d_x_ov;
    float* d_feats;
    // compact rows (only voxels with count 1 / 2 own a row); NULL = dense rows, one per voxel
    const int32_t* slot;
    // pose: 1 = rows ordered (y, x, z) -- [B,Y,X,Z*ld] IS the channels-last BEV map reduce_dim reads (:339-342)
    int bev_rows;
};

// dense row of voxel n = (z*Y + y)*X + x: the reference's flatten order, or z innermost for the BEV hand-off
__device__ __forceinline__ size_t dense_row(const GatherArgs& a, int b, long n) {
    const vfd_fusion_desc& d = a.d;
    const size_t N = (size_t)d.X * d.Y * d.Z;
    if (!a.bev_rows) return (size_t)b * N + n;
    const long xy = (long)d.X * d.Y;
    const long z = n / xy, r = n - z * xy;
    return (size_t)b * N + (size_t)r * d.Z + z;
}

// phase 1: returns the camera bit-mask of this lane's voxel and fills slot[c][lane]
__device__ __forceinline__ unsigned project_voxel(const GatherArgs& a, const float* s_mat, int b, long n, int lane,
                                                  bool in_range, CamSlot (*slot)[VFD_WARP]) {
    const vfd_fusion_desc& d = a.d;
    const int N = d.X * d.Y * d.Z;
    unsigned vmask = 0;
    float x = 0.f, y = 0.f, z = 0.f;
    if (in_range) {
        int xi = (int)(n % d.X), yi = (int)((n / d.X) % d.Y), zi = (int)(n / ((long)d.X * d.Y));
        x = __ldg(d.axis_x + xi);
        y = __ldg(d.axis_y + yi);
        z = __ldg(d.axis_z + zi);
    }
    for (int c = 0; c < d.num_cams; ++c) {
        const float* Einv = s_mat + c * 32;
        float g[2];
        CamSlot s;
        const bool ok = project_cam(d, a.mask_small + (size_t)(b * d.num_cams + c) * d.h * d.w, Einv, Einv + 16, x, y,
                                    z, s, g) && in_range;
        if (in_range) {
            size_t o = (size_t)(b * d.num_cams + c) * N + n;
            if (a.grid) reinterpret_cast<float2*>(a.grid)[o] = make_float2(g[0], g[1]);
            if (a.valid) a.valid[o] = ok ? 1 : 0;
        }
        slot[c][lane] = s;
        vmask |= (ok ? 1u : 0u) << c;
    }
    return vmask;
}

__device__ __forceinline__ void load_mats(const GatherArgs& a, int b, float* s_mat) {
    // per camera 32 floats: Einv 4x4 (16) then K 4x4 (16)
    const int n = a.d.num_cams * 32;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        int c = i >> 5, r = i & 31;
        const float* src = (r < 16 ? a.Einv : a.K) + (size_t)(b * a.d.num_cams + c) * 16 + (r & 15);
        s_mat[i] = __ldg(src);
    }
}

template <int MODE /*0 depth, 1 pose*/, int ITERS>
__global__ void __launch_bounds__(kGatherWarps* VFD_WARP) voxel_gather_fwd_kernel(const GatherArgs a) {
    const vfd_fusion_desc& d = a.d;
    __shared__ float s_mat[VFD_MAX_CAMS * 32];
    __shared__ CamSlot s_slot[kGatherWarps][VFD_MAX_CAMS][VFD_WARP];
    const int b = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int N = d.X * d.Y * d.Z;
    load_mats(a, b, s_mat);
    __syncthreads();
    const long n0 = ((long)blockIdx.x * kGatherWarps + warp) * VFD_WARP;
    if (n0 >= N) return;
    const long n = n0 + lane;
    unsigned vmask = project_voxel(a, s_mat, b, n, lane, n < N, s_slot[warp]);
    if (n < N && a.count) a.count[(size_t)b * N + n] = (uint8_t)__popc(vmask);
    int my_slot = -1;
    if (MODE == 0 && a.slot && n < N) my_slot = __ldg(a.slot + (size_t)b * N + n);
    __syncwarp();

    const int C = d.C, hw = d.h * d.w;
    const int jmax = min(VFD_WARP, (int)(N - n0));
    for (int j = 0; j < jmax; ++j) {
        unsigned m = __shfl_sync(0xffffffffu, vmask, j);
        const int cnt = __popc(m);
        const int sl = __shfl_sync(0xffffffffu, my_slot, j);
        if (MODE == 0 && a.slot && sl < 0) continue;   // compact rows: count 0 or >= 3 feeds neither MLP
        float4 acc0[ITERS], acc1[ITERS];
#pragma unroll
        for (int it = 0; it < ITERS; ++it) { acc0[it] = f4_zero(); acc1[it] = f4_zero(); }
        float z0 = 0.f, z1 = 0.f;
        while (m) {
            const int c = __ffs(m) - 1;
            m &= m - 1;
            const CamSlot s = s_slot[warp][c][j];
            const Corners q = make_corners(s, d.w, d.h, C);
            const float* base = a.feats + (size_t)(b * d.num_cams + c) * hw * C;
            const bool g1 = (MODE == 0) && d.group[c];
#pragma unroll
            for (int it = 0; it < ITERS; ++it) {
                const int k = it * 128 + lane * 4;
                if (k < C) {
                    float4 v0 = ldg_f4(base + q.off[0] + k), v1 = ldg_f4(base + q.off[1] + k);
                    float4 v2 = ldg_f4(base + q.off[2] + k), v3 = ldg_f4(base + q.off[3] + k);
                    float4 r = f4_scale(q.wgt[0], v0);
                    f4_fma(r, q.wgt[1], v1);
                    f4_fma(r, q.wgt[2], v2);
                    f4_fma(r, q.wgt[3], v3);
                    if (g1) f4_add(acc1[it], r); else f4_add(acc0[it], r);
                }
            }
            if (g1) z1 += s.zrel; else z0 += s.zrel;
        }
        const size_t row = a.slot ? (size_t)sl : dense_row(a, b, n0 + j);
        if (MODE == 0) {
            // dense: both rows of every voxel; compact: the x_no row of a count-1 voxel, the x_ov row of a count-2 one
            const bool wr_no = !a.slot || cnt == 1, wr_ov = !a.slot || cnt == 2;
            float* rno = a.x_no + row * d.ld_no;
            float* rov = a.x_ov + row * d.ld_ov;
            const float keep = (cnt == 1) ? 1.f : 0.f;
#pragma unroll
            for (int it = 0; it < ITERS; ++it) {
                const int k = it * 128 + lane * 4;
                if (k < C) {
                    float4 s = acc0[it];
                    f4_add(s, acc1[it]);
                    if (wr_no) st_stream_f4(rno + k, f4_scale(keep, s));
                    if (wr_ov) {
                        st_stream_f4(rov + k, acc0[it]);
                        st_stream_f4(rov + d.half_ov + k, acc1[it]);
                    }
                }
            }
            // tail columns: relative depth at column C, then zero padding
            if (wr_no)
                for (int k = C + lane * 4; k < d.ld_no; k += 128)
                    st_stream_f4(rno + k, make_float4(k == C ? keep * (z0 + z1) : 0.f, 0.f, 0.f, 0.f));
            if (wr_ov) {
                for (int k = C + lane * 4; k < d.half_ov; k += 128)
                    st_stream_f4(rov + k, make_float4(k == C ? z0 : 0.f, 0.f, 0.f, 0.f));
                for (int k = d.half_ov + C + lane * 4; k < d.ld_ov; k += 128)
                    st_stream_f4(rov + k, make_float4(k == d.half_ov + C ? z1 : 0.f, 0.f, 0.f, 0.f));
            }
        } else {
            float* ravg = a.x_no + row * d.ld_no;
            const float den = VFD_ADD((float)cnt, 1e-7f);  // count + 1e-7 (:162)
#pragma unroll
            for (int it = 0; it < ITERS; ++it) {
                const int k = it * 128 + lane * 4;
                if (k < C) {
                    float4 s = acc0[it];
                    st_stream_f4(ravg + k, make_float4(VFD_DIV(s.x, den), VFD_DIV(s.y, den), VFD_DIV(s.z, den),
                                                       VFD_DIV(s.w, den)));
                }
            }
            for (int k = C + lane * 4; k < d.ld_no; k += 128)
                st_stream_f4(ravg + k, make_float4(k == C ? VFD_DIV(z0, den) : 0.f, 0.f, 0.f, 0.f));
        }
    }
}

// Backward: d_feats[b,c,texel,:] += w_corner * g(voxel, c) for the 4 corners, with 128-bit vector reductions
// (red.global.add.v4.f32).  g = [count==1] * d_x_no + d_x_ov[group(c)]  (depth) or d_avg / (count+1e-7) (pose).
template <int MODE, int ITERS>
__global__ void __launch_bounds__(kGatherWarps* VFD_WARP) voxel_gather_bwd_kernel(const GatherArgs a) {
    const vfd_fusion_desc& d = a.d;
    __shared__ float s_mat[VFD_MAX_CAMS * 32];
    __shared__ CamSlot s_slot[kGatherWarps][VFD_MAX_CAMS][VFD_WARP];
    const int b = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int N = d.X * d.Y * d.Z;
    load_mats(a, b, s_mat);
    __syncthreads();
    const long n0 = ((long)blockIdx.x * kGatherWarps + warp) * VFD_WARP;
    if (n0 >= N) return;
    const long n = n0 + lane;
    unsigned vmask = project_voxel(a, s_mat, b, n, lane, n < N, s_slot[warp]);
    int my_slot = -1;
    if (MODE == 0 && a.slot && n < N) my_slot = __ldg(a.slot + (size_t)b * N + n);
    __syncwarp();

    const int C = d.C, hw = d.h * d.w;
    const int jmax = min(VFD_WARP, (int)(N - n0));
    for (int j = 0; j < jmax; ++j) {
        unsigned m = __shfl_sync(0xffffffffu, vmask, j);
        const int sl = __shfl_sync(0xffffffffu, my_slot, j);
        if (!m) continue;
        if (MODE == 0 && a.slot && sl < 0) continue;
        const int cnt = __popc(m);
        const size_t row = a.slot ? (size_t)sl : dense_row(a, b, n0 + j);
        float4 gno[ITERS];
        if (MODE == 0) {
#pragma unroll
            for (int it = 0; it < ITERS; ++it) {
                const int k = it * 128 + lane * 4;
                gno[it] = (cnt == 1 && k < C) ? ld_stream_f4(a.d_x_no + row * d.ld_no + k) : f4_zero();
            }
        } else {
            const float den = VFD_ADD((float)cnt, 1e-7f);
#pragma unroll
            for (int it = 0; it < ITERS; ++it) {
                const int k = it * 128 + lane * 4;
                float4 g = (k < C) ? ld_stream_f4(a.d_x_no + row * d.ld_no + k) : f4_zero();
                gno[it] = make_float4(g.x / den, g.y / den, g.z / den, g.w / den);
            }
        }
        while (m) {
            const int c = __ffs(m) - 1;
            m &= m - 1;
            const CamSlot s = s_slot[warp][c][j];
            const Corners q = make_corners(s, d.w, d.h, C);
            float* base = a.d_feats + (size_t)(b * d.num_cams + c) * hw * C;
#pragma unroll
            for (int it = 0; it < ITERS; ++it) {
                const int k = it * 128 + lane * 4;
                if (k < C) {
                    float4 g = gno[it];
                    if (MODE == 0 && (!a.slot || cnt == 2)) {
                        float4 o = ld_stream_f4(a.d_x_ov + row * d.ld_ov + (d.group[c] ? d.half_ov : 0) + k);
                        f4_add(g, o);
                    }
                    if (g.x == 0.f && g.y == 0.f && g.z == 0.f && g.w == 0.f) continue;
#pragma unroll
                    for (int t = 0; t < 4; ++t)
                        if (q.wgt[t] != 0.f) red_add_f4(base + q.off[t] + k, f4_scale(q.wgt[t], g));
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Classification: which voxels feed which MLP.  preprocess_non_overlap (:197-207) multiplies its input AND its
// output by [count==1], preprocess_overlap (:209-230) by [count==2]; a voxel seen by 0 or >=3 cameras feeds
// neither.  Three small launches give every count-1 / count-2 voxel a row ("slot") in a compact matrix, in voxel
// order (deterministic): count + per-block totals, an exclusive scan of the block totals, slots + index lists.
// ---------------------------------------------------------------------------------------------------------
struct ClassifyArgs {
    vfd_fusion_desc d;
    const float* mask_small;
    const float* K;
    const float* Einv;
    uint8_t* count;
    int32_t* slot;
    int64_t* idx_no;
    int64_t* idx_ov;
    int32_t* totals;
    int2* block_tot;   // [B * blocks_per_rig]
    void* table;       // optional corner table (gather_common.cuh), B*N rows
};

__global__ void __launch_bounds__(kGatherWarps* VFD_WARP) voxel_count_kernel(const ClassifyArgs c) {
    const vfd_fusion_desc& d = c.d;
    __shared__ float s_mat[VFD_MAX_CAMS * 32];
    __shared__ int s_n1[kGatherWarps], s_n2[kGatherWarps];
    const int b = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int N = d.X * d.Y * d.Z;
    GatherArgs a = {};
    a.d = d; a.mask_small = c.mask_small; a.K = c.K; a.Einv = c.Einv;
    load_mats(a, b, s_mat);
    __syncthreads();
    const long n = ((long)blockIdx.x * kGatherWarps + warp) * VFD_WARP + lane;
    int cnt = 0;
    if (n < N) {
        const int xi = (int)(n % d.X), yi = (int)((n / d.X) % d.Y), zi = (int)(n / ((long)d.X * d.Y));
        const float x = __ldg(d.axis_x + xi), y = __ldg(d.axis_y + yi), z = __ldg(d.axis_z + zi);
        CamSlot ent[2];
        int ecam[2] = {-1, -1};
        for (int cam = 0; cam < d.num_cams; ++cam) {
            const float* Einv = s_mat + cam * 32;
            float g[2];
            CamSlot cs;
            if (project_cam(d, c.mask_small + (size_t)(b * d.num_cams + cam) * d.h * d.w, Einv, Einv + 16, x, y, z, cs,
                            g)) {
                if (cnt == 0) { ent[0] = cs; ecam[0] = cam; }
                if (cnt == 1) { ent[1] = cs; ecam[1] = cam; }
                ++cnt;
            }
        }
        const size_t row = (size_t)b * N + n;
        c.count[row] = (uint8_t)cnt;
        if (c.table && (cnt == 1 || cnt == 2)) {
            const size_t rows = (size_t)d.B * N;
            const CornerTable t = carve_table(c.table, rows);
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                if (e < cnt) {
                    const Corners q = make_corners(ent[e], d.w, d.h, 1);     // pixel indices inside the camera's map
                    t.off[e * rows + row] = make_int4(q.off[0], q.off[1], q.off[2], q.off[3]);
                    t.wgt[e * rows + row] = make_float4(q.wgt[0], q.wgt[1], q.wgt[2], q.wgt[3]);
                    t.meta[e * rows + row] = make_int2(b * d.num_cams + ecam[e], __float_as_int(ent[e].zrel));
                } else {
                    t.meta[e * rows + row] = make_int2(-1, 0);
                }
            }
        }
    }
    const unsigned b1 = __ballot_sync(0xffffffffu, cnt == 1), b2 = __ballot_sync(0xffffffffu, cnt == 2);
    if (lane == 0) { s_n1[warp] = __popc(b1); s_n2[warp] = __popc(b2); }
    __syncthreads();
    if (threadIdx.x == 0) {
        int t1 = 0, t2 = 0;
        for (int i = 0; i < kGatherWarps; ++i) { t1 += s_n1[i]; t2 += s_n2[i]; }
        c.block_tot[(size_t)b * gridDim.x + blockIdx.x] = make_int2(t1, t2);
    }
}

// single CTA: exclusive scan of the per-block totals (in place), grand totals to totals[0..1]
__global__ void __launch_bounds__(1024) voxel_scan_kernel(int2* block_tot, int n, int32_t* totals) {
    __shared__ int2 s_warp[32];
    __shared__ int2 s_carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_carry = make_int2(0, 0);
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + threadIdx.x;
        const int2 v = i < n ? block_tot[i] : make_int2(0, 0);
        int2 inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int ux = __shfl_up_sync(0xffffffffu, inc.x, o), uy = __shfl_up_sync(0xffffffffu, inc.y, o);
            if (lane >= o) { inc.x += ux; inc.y += uy; }
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int2 w = s_warp[lane], winc = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int ux = __shfl_up_sync(0xffffffffu, winc.x, o), uy = __shfl_up_sync(0xffffffffu, winc.y, o);
                if (lane >= o) { winc.x += ux; winc.y += uy; }
            }
            s_warp[lane] = make_int2(winc.x - w.x, winc.y - w.y);   // exclusive warp offsets
        }
        __syncthreads();
        const int2 carry = s_carry, wo = s_warp[warp];
        if (i < n) block_tot[i] = make_int2(carry.x + wo.x + inc.x - v.x, carry.y + wo.y + inc.y - v.y);
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = make_int2(carry.x + wo.x + inc.x, carry.y + wo.y + inc.y);
        __syncthreads();
    }
    if (threadIdx.x == 0) { totals[0] = s_carry.x; totals[1] = s_carry.y; }
}

__global__ void __launch_bounds__(kGatherWarps* VFD_WARP) voxel_slot_kernel(const ClassifyArgs c) {
    const vfd_fusion_desc& d = c.d;
    __shared__ int s_n1[kGatherWarps], s_n2[kGatherWarps];
    const int b = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int N = d.X * d.Y * d.Z;
    const long n = ((long)blockIdx.x * kGatherWarps + warp) * VFD_WARP + lane;
    const int cnt = n < N ? c.count[(size_t)b * N + n] : 0;
    const unsigned b1 = __ballot_sync(0xffffffffu, cnt == 1), b2 = __ballot_sync(0xffffffffu, cnt == 2);
    if (lane == 0) { s_n1[warp] = __popc(b1); s_n2[warp] = __popc(b2); }
    __syncthreads();
    int2 off = c.block_tot[(size_t)b * gridDim.x + blockIdx.x];
    for (int i = 0; i < warp; ++i) { off.x += s_n1[i]; off.y += s_n2[i]; }
    const unsigned below = (1u << lane) - 1u;
    if (n < N) {
        const size_t row = (size_t)b * N + n;
        int sl = -1;
        if (cnt == 1) { sl = off.x + __popc(b1 & below); c.idx_no[sl] = (int64_t)row; }
        if (cnt == 2) { sl = off.y + __popc(b2 & below); c.idx_ov[sl] = (int64_t)row; }
        c.slot[row] = sl;
    }
}

static int check_desc(const vfd_fusion_desc* d) {
    if (!d) return VFD_ERR_NULL;
    if (d->B < 1 || d->num_cams < 1 || d->num_cams > VFD_MAX_CAMS || d->C < 4 || d->h < 2 || d->w < 2 ||
        d->X < 1 || d->Y < 1 || d->Z < 1)
        return VFD_ERR_DESC;
    if (d->C % 4 || d->C > 512) return VFD_ERR_UNSUPPORTED;
    if (!d->axis_x || !d->axis_y || !d->axis_z) return VFD_ERR_NULL;
    if ((long)d->h * d->w * d->C >= (1L << 31)) return VFD_ERR_DESC;
    return VFD_OK;
}

static int check_rows(const vfd_fusion_desc* d, bool depth) {
    if (d->ld_no < d->C + 1 || d->ld_no % 4) return VFD_ERR_ALIGN;
    if (depth && (d->half_ov < d->C + 1 || d->half_ov % 4 || d->ld_ov < d->half_ov + d->C + 1 || d->ld_ov % 4))
        return VFD_ERR_ALIGN;
    return VFD_OK;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <int MODE>
static int launch(const GatherArgs& a, bool bwd, cudaStream_t st) {
    const vfd_fusion_desc& d = a.d;
    const long N = (long)d.X * d.Y * d.Z;
    dim3 grid((unsigned)((N + kGatherWarps * VFD_WARP - 1) / (kGatherWarps * VFD_WARP)), d.B);
    dim3 block(kGatherWarps * VFD_WARP);
    const int iters = (d.C + 127) / 128;
#define VFD_GATHER_CASE(I)                                                            \
    case I:                                                                           \
        if (bwd) voxel_gather_bwd_kernel<MODE, I><<<grid, block, 0, st>>>(a);         \
        else voxel_gather_fwd_kernel<MODE, I><<<grid, block, 0, st>>>(a);             \
        break;
    switch (iters) {
        VFD_GATHER_CASE(1)
        VFD_GATHER_CASE(2)
        VFD_GATHER_CASE(3)
        VFD_GATHER_CASE(4)
        default: return VFD_ERR_UNSUPPORTED;
    }
#undef VFD_GATHER_CASE
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_voxel_gather_depth_fwd(const vfd_fusion_desc* d, const float* feats, const float* mask_small,
                                          const float* K, const float* Einv, float* x_no, float* x_ov,
                                          uint8_t* count, float* grid, uint8_t* valid, void* stream) {
    int rc = check_desc(d);
    if (rc) return rc;
    if ((rc = check_rows(d, true))) return rc;
    if (!feats || !mask_small || !K || !Einv || !x_no || !x_ov) return VFD_ERR_NULL;
    if (!aligned16(feats) || !aligned16(x_no) || !aligned16(x_ov) || (grid && !aligned16(grid))) return VFD_ERR_ALIGN;
    GatherArgs a = {};
    a.d = *d; a.feats = feats; a.mask_small = mask_small; a.K = K; a.Einv = Einv;
    a.x_no = x_no; a.x_ov = x_ov; a.count = count; a.grid = grid; a.valid = valid;
    return launch<0>(a, false, (cudaStream_t)stream);
}

extern "C" int vfd_voxel_gather_pose_fwd(const vfd_fusion_desc* d, const float* feats, const float* mask_small,
                                         const float* K, const float* Einv, float* avg, uint8_t* count,
                                         void* stream) {
    int rc = check_desc(d);
    if (rc) return rc;
    if ((rc = check_rows(d, false))) return rc;
    if (!feats || !mask_small || !K || !Einv || !avg) return VFD_ERR_NULL;
    if (!aligned16(feats) || !aligned16(avg)) return VFD_ERR_ALIGN;
    GatherArgs a = {};
    a.d = *d; a.feats = feats; a.mask_small = mask_small; a.K = K; a.Einv = Einv; a.x_no = avg; a.count = count;
    return launch<1>(a, false, (cudaStream_t)stream);
}

extern "C" int vfd_voxel_gather_pose_bev_fwd(const vfd_fusion_desc* d, const float* feats, const float* mask_small,
                                             const float* K, const float* Einv, float* bev, uint8_t* count,
                                             void* stream) {
    int rc = check_desc(d);
    if (rc) return rc;
    if ((rc = check_rows(d, false))) return rc;
    if (!feats || !mask_small || !K || !Einv || !bev) return VFD_ERR_NULL;
    if (!aligned16(feats) || !aligned16(bev)) return VFD_ERR_ALIGN;
    GatherArgs a = {};
    a.d = *d; a.feats = feats; a.mask_small = mask_small; a.K = K; a.Einv = Einv; a.x_no = bev; a.count = count;
    a.bev_rows = 1;
    return launch<1>(a, false, (cudaStream_t)stream);
}

static int bwd_common(const vfd_fusion_desc* d, GatherArgs& a, float* d_feats, cudaStream_t st) {
    const size_t bytes = (size_t)d->B * d->num_cams * d->h * d->w * d->C * sizeof(float);
    if (cudaMemsetAsync(d_feats, 0, bytes, st) != cudaSuccess) return VFD_ERR_CUDA;
    a.d_feats = d_feats;
    return VFD_OK;
}

extern "C" int vfd_voxel_gather_depth_bwd(const vfd_fusion_desc* d, const float* mask_small, const float* K,
                                          const float* Einv, const float* d_x_no, const float* d_x_ov,
                                          float* d_feats, void* stream) {
    int rc = check_desc(d);
    if (rc) return rc;
    if ((rc = check_rows(d, true))) return rc;
    if (!mask_small || !K || !Einv || !d_x_no || !d_x_ov || !d_feats) return VFD_ERR_NULL;
    if (!aligned16(d_feats) || !aligned16(d_x_no) || !aligned16(d_x_ov)) return VFD_ERR_ALIGN;
    GatherArgs a = {};
    a.d = *d; a.mask_small = mask_small; a.K = K; a.Einv = Einv; a.d_x_no = d_x_no; a.d_x_ov = d_x_ov;
    if ((rc = bwd_common(d, a, d_feats, (cudaStream_t)stream))) return rc;
    return launch<0>(a, true, (cudaStream_t)stream);
}

extern "C" int vfd_voxel_gather_pose_bwd(const vfd_fusion_desc* d, const float* mask_small, const float* K,
                                         const float* Einv, const float* d_avg, float* d_feats, void* stream) {
    int rc = check_desc(d);
    if (rc) return rc;
    if ((rc = check_rows(d, false))) return rc;
    if (!mask_small || !K || !Einv || !d_avg || !d_feats) return VFD_ERR_NULL;
    if (!aligned16(d_feats) || !aligned16(d_avg)) return VFD_ERR_ALIGN;
    GatherArgs a = {};
    a.d = *d; a.mask_small = mask_small; a.K = K; a.Einv = Einv; a.d_x_no = d_avg;
    if ((rc = bwd_common(d, a, d_feats, (cudaStream_t)stream))) return rc;
    return launch<1>(a, true, (cudaStream_t)stream);
}

extern "C" int vfd_voxel_gather_pose_bev_bwd(const vfd_fusion_desc* d, const float* mask_small, const float* K,
                                             const float* Einv, const float* d_bev, float* d_feats, void* stream) {
    int rc = check_desc(d);
    if (rc) return rc;
    if ((rc = check_rows(d, false))) return rc;
    if (!mask_small || !K || !Einv || !d_bev || !d_feats) return VFD_ERR_NULL;
    if (!aligned16(d_feats) || !aligned16(d_bev)) return VFD_ERR_ALIGN;
    GatherArgs a = {};
    a.d = *d; a.mask_small = mask_small; a.K = K; a.Einv = Einv; a.d_x_no = d_bev;
    a.bev_rows = 1;
    if ((rc = bwd_common(d, a, d_feats, (cudaStream_t)stream))) return rc;
    return launch<1>(a, true, (cudaStream_t)stream);
}

extern "C" size_t vfd_voxel_classify_workspace_bytes(const vfd_fusion_desc* d) {
    if (!d || d->B < 1) return 0;
    const long N = (long)d->X * d->Y * d->Z;
    const long blocks = (N + kGatherWarps * VFD_WARP - 1) / (kGatherWarps * VFD_WARP);
    return (size_t)d->B * blocks * sizeof(int2);
}

extern "C" size_t vfd_voxel_table_bytes(const vfd_fusion_desc* d) {
    if (!d || d->B < 1) return 0;
    return corner_table_bytes((size_t)d->B * d->X * d->Y * d->Z);
}

extern "C" int vfd_voxel_classify(const vfd_fusion_desc* d, const float* mask_small, const float* K,
                                  const float* Einv, uint8_t* count, int32_t* slot, int64_t* idx_no,
                                  int64_t* idx_ov, int32_t* totals, void* table, void* ws, size_t ws_bytes,
                                  void* stream) {
    int rc = check_desc(d);
    if (rc) return rc;
    if (!mask_small || !K || !Einv || !count || !ws) return VFD_ERR_NULL;
    // the compact lists are optional as a group: the fused path (vfd_voxel_fuse_*) needs only count + table
    const bool lists = slot || idx_no || idx_ov || totals;
    if (lists && (!slot || !idx_no || !idx_ov || !totals)) return VFD_ERR_NULL;
    if (ws_bytes < vfd_voxel_classify_workspace_bytes(d)) return VFD_ERR_WORKSPACE;
    if (reinterpret_cast<uintptr_t>(ws) & 7) return VFD_ERR_ALIGN;
    const long N = (long)d->X * d->Y * d->Z;
    if ((long)d->B * N >= (1L << 31)) return VFD_ERR_DESC;
    ClassifyArgs c = {};
    c.d = *d; c.mask_small = mask_small; c.K = K; c.Einv = Einv; c.count = count; c.slot = slot;
    c.idx_no = idx_no; c.idx_ov = idx_ov; c.totals = totals; c.block_tot = static_cast<int2*>(ws);
    if (table && (reinterpret_cast<uintptr_t>(table) & 15)) return VFD_ERR_ALIGN;
    c.table = table;
    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid((unsigned)((N + kGatherWarps * VFD_WARP - 1) / (kGatherWarps * VFD_WARP)), d->B);
    voxel_count_kernel<<<grid, kGatherWarps * VFD_WARP, 0, st>>>(c);
    if (lists) {
        voxel_scan_kernel<<<1, 1024, 0, st>>>(c.block_tot, (int)(grid.x * grid.y), totals);
        voxel_slot_kernel<<<grid, kGatherWarps * VFD_WARP, 0, st>>>(c);
    }
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" int vfd_voxel_gather_depth_compact_fwd(const vfd_fusion_desc* d, const float* feats,
                                                  const float* mask_small, const float* K, const float* Einv,
                                                  const int32_t* slot, float* x_no, float* x_ov, void* stream) {
    int rc = check_desc(d);
    if (rc) return rc;
    if ((rc = check_rows(d, true))) return rc;
    if (!feats || !mask_small || !K || !Einv || !slot || !x_no || !x_ov) return VFD_ERR_NULL;
    if (!aligned16(feats) || !aligned16(x_no) || !aligned16(x_ov)) return VFD_ERR_ALIGN;
    GatherArgs a = {};
    a.d = *d; a.feats = feats; a.mask_small = mask_small; a.K = K; a.Einv = Einv;
    a.x_no = x_no; a.x_ov = x_ov; a.slot = slot;
    return launch<0>(a, false, (cudaStream_t)stream);
}

extern "C" int vfd_voxel_gather_depth_compact_bwd(const vfd_fusion_desc* d, const float* mask_small,
                                                  const float* K, const float* Einv, const int32_t* slot,
                                                  const float* d_x_no, const float* d_x_ov, float* d_feats,
                                                  void* stream) {
    int rc = check_desc(d);
    if (rc) return rc;
    if ((rc = check_rows(d, true))) return rc;
    if (!mask_small || !K || !Einv || !slot || !d_x_no || !d_x_ov || !d_feats) return VFD_ERR_NULL;
    if (!aligned16(d_feats) || !aligned16(d_x_no) || !aligned16(d_x_ov)) return VFD_ERR_ALIGN;
    GatherArgs a = {};
    a.d = *d; a.mask_small = mask_small; a.K = K; a.Einv = Einv; a.d_x_no = d_x_no; a.d_x_ov = d_x_ov;
    a.slot = slot;
    if ((rc = bwd_common(d, a, d_feats, (cudaStream_t)stream))) return rc;
    return launch<0>(a, true, (cudaStream_t)stream);
}
