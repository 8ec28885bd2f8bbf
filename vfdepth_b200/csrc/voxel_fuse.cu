// Fused voxel gather + the two 1x1 MLPs of VFNet.backproject_into_voxel, model='depth'
// (network/volumetric_fusionnet.py:116-159 with preprocess_non_overlap :197-207 and preprocess_overlap :209-230),
// in "project, then gather" form.
//
// The reference materialises six [B,257,N] camera tensors and two masked GEMM inputs (617 MB/rig) and runs two dense
// 1x1 convolutions over ALL voxels, although only voxels seen by exactly one camera reach the first MLP and only voxels
// seen by exactly two reach the second.  Both MLPs are linear in the gathered features up to their LeakyReLU, and so
// is the bilinear gather, so the two commute (dense_proj.cu):
//   1. vfd_voxel_fuse_fwd multiplies every feature PIXEL once with the weight blocks its camera can need
//      (P [B*cams, h*w, 2 Cv]; tcgen05 + TMA GEMM, 3xTF32),
//   2. voxel_fuse_gather_kernel: Cv/4 lanes per voxel read the corner table of vfd_voxel_classify, gather the 4 bilinear
//      corners of the (one or two) cameras that see the voxel as Cv-float rows of P, add the relative-depth column and
//      the bias, apply LeakyReLU(0.1) and write the voxel row -- every row of the channel-last volume is written exactly
//      once, in order (zeros where count is 0 or >= 3), so there is no memset pass and no scattered store.
// L2 -> SM traffic per (voxel, camera): 4 x 256 B of P instead of 4 x 1 KB of feature rows; the GEMM has 184 320 rows
// (pixels, B=8) instead of 1 490 000 (voxel, camera) rows.  HBM: features in, volume out (P is L2-sized: 94 MB at B=8).
//
// Backward: voxel_fuse_scatter_kernel forms dZ = d_vol * LeakyReLU'(vol) per voxel, scatters w_k dZ into dP
// (red.global.add.v4.f32, 256 B rows) and accumulates the bias / relative-depth-column gradients; then
//   d_feats = dP . Wcat_g  (the same tcgen05 + TMA kernel, K = 2 Cv, N = C) and dWcat_g = dP^T . feats (dense_dw_tc_kernel:
// tcgen05 with both operands MN-major, TMA-staged)
// with a fixed-order reduction into d_w_no / d_w_ov.
#include "gather_common.cuh"

namespace vfd {

// dense_proj.cu
bool dense_proj_tc_ok(int K, int N);
int dense_proj(int n_img, int hw, int K, int N, int cams, const int* group, const float* X, float* Y, const float* Wfull,
               const float* Whi, const float* Wlo, cudaStream_t st);
int dense_dw_slabs(int n_img);
size_t dense_dw_part_floats(int n_img, int C, int N2);
int dense_dw(int n_img, int hw, int C, int N2, int cams, const int* group, const float* dP, const float* feats, float* part,
             int* n_parts, int* split, cudaStream_t st);
__global__ void wcat_prep_kernel(const float* __restrict__ w_no, const float* __restrict__ w_ov, int C, int Cv,
                                 int transpose, float* full, float* hi, float* lo);

constexpr int kFgWarps = 8, kFgThreads = kFgWarps * VFD_WARP;
constexpr int kFgMaxCv = 64;
constexpr int kFgCtasPerSm = 5;    // scatter kernel: one resident wave (48 registers -> 5 CTAs per SM); 4 / 5 / 8 / 10: 0.641 / 0.637 / 0.675 / 0.640 ms for the whole backward call
constexpr int kFgRun = 64;          // consecutive voxels one warp of the gather kernel walks

struct FuseArgs {
    vfd_fusion_desc d;
    const uint8_t* count;   // [B*N]
    const void* table;      // corner table of vfd_voxel_classify (gather_common.cuh): offsets are PIXEL indices
    const float* P;         // [B*cams][h*w][2 Cv]
    const float* w_no;      // [Cv][C+1]      conv_non_overlap.0.weight
    const float* b_no;      // [Cv]
    const float* w_ov;      // [Cv][2(C+1)]   conv_overlap.0.weight
    const float* b_ov;      // [Cv]
    float* vol;             // [B*N][Cv]
    // backward
    const float* d_vol;     // [B*N][Cv]
    float* dP;              // [B*cams][h*w][2 Cv], zero-filled by the caller
    float* partial;         // [gridDim.x][5][Cv]: db_no, dwz_no, db_ov, dwz_ov group 0, dwz_ov group 1
    int lpv;                // lanes per voxel: smallest power of two >= Cv / 4
};

__device__ __forceinline__ float leaky(float v) { return v > 0.f ? v : 0.1f * v; }

// one (voxel, camera) entry of the table
struct Entry {
    int4 pix;
    float4 wgt;
    int img;      // b * num_cams + c
    float zrel;
};
__device__ __forceinline__ Entry load_entry(const CornerTable& tb, size_t rows, int e, size_t v) {
    Entry en;
    en.pix = __ldg(tb.off + e * rows + v);
    en.wgt = __ldg(tb.wgt + e * rows + v);
    const int2 m = __ldg(tb.meta + e * rows + v);
    en.img = m.x;
    en.zrel = __int_as_float(m.y);
    return en;
}

// ---------------------------------------------------------------------------------------------------------
// forward: vol[v] = LeakyReLU(sum_e sum_k w_k P[img_e, pix_k, blk] + zrel_e * wz + bias)   blk = count - 1
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kFgThreads, 4) voxel_fuse_gather_kernel(const FuseArgs a) {
    const vfd_fusion_desc& d = a.d;
    __shared__ __align__(16) float s_c[5][kFgMaxCv];     // b_no, wz_no, b_ov, wz_ov[0], wz_ov[1]
    const int Cv = d.Cv, C1 = d.C + 1, hw = d.h * d.w, ldp = 2 * Cv;
    for (int i = threadIdx.x; i < 5 * kFgMaxCv; i += kFgThreads) {
        const int q = i / kFgMaxCv, n = i - q * kFgMaxCv;
        float v = 0.f;
        if (n < Cv) {
            if (q == 0) v = __ldg(a.b_no + n);
            else if (q == 1) v = __ldg(a.w_no + (size_t)n * C1 + d.C);
            else if (q == 2) v = __ldg(a.b_ov + n);
            else v = __ldg(a.w_ov + (size_t)n * 2 * C1 + (q - 3) * C1 + d.C);
        }
        s_c[q][n] = v;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int lpv = a.lpv, vpw = VFD_WARP / lpv, sub = lane / lpv, q4 = (lane % lpv) * 4;
    const bool chan = q4 < Cv;
    const size_t rows = (size_t)d.B * d.X * d.Y * d.Z;
    const CornerTable tb = carve_table(const_cast<void*>(a.table), rows);
    // neighbouring voxels (x fastest) project to neighbouring pixels and share bilinear corners: a warp walks a run of
    // kFgRun consecutive voxels, the CTA a block of kFgWarps runs, so the shared rows of P are served by L1
    const size_t v0 = ((size_t)blockIdx.x * kFgWarps + warp) * kFgRun;
    const size_t v1 = v0 + kFgRun < rows ? v0 + kFgRun : rows;
    for (size_t v = v0 + sub; v < v1; v += vpw) {
        const int cnt = __ldg(a.count + v);
        float4 acc = f4_zero();
        if ((cnt == 1 || cnt == 2) && chan) {
            const int blk = cnt - 1;
            Entry en[2];
            en[0] = load_entry(tb, rows, 0, v);
            if (cnt == 2) en[1] = load_entry(tb, rows, 1, v);
            float4 c[2][4];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                if (e < cnt) {
                    const float* base = a.P + (size_t)en[e].img * hw * ldp + blk * Cv + q4;
                    c[e][0] = ldg_f4(base + (size_t)en[e].pix.x * ldp);
                    c[e][1] = ldg_f4(base + (size_t)en[e].pix.y * ldp);
                    c[e][2] = ldg_f4(base + (size_t)en[e].pix.z * ldp);
                    c[e][3] = ldg_f4(base + (size_t)en[e].pix.w * ldp);
                }
            }
            acc = *reinterpret_cast<const float4*>(&s_c[2 * blk][q4]);                 // bias
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                if (e < cnt) {
                    float4 s = f4_scale(en[e].wgt.x, c[e][0]);
                    f4_fma(s, en[e].wgt.y, c[e][1]);
                    f4_fma(s, en[e].wgt.z, c[e][2]);
                    f4_fma(s, en[e].wgt.w, c[e][3]);
                    const int g = d.group[en[e].img % d.num_cams] ? 1 : 0;
                    const float4 wz = *reinterpret_cast<const float4*>(&s_c[blk ? 3 + g : 1][q4]);
                    f4_fma(s, en[e].zrel, wz);
                    f4_add(acc, s);
                }
            }
            acc = make_float4(leaky(acc.x), leaky(acc.y), leaky(acc.z), leaky(acc.w));
        }
        if (chan) st_stream_f4(a.vol + v * Cv + q4, acc);
    }
}

// ---------------------------------------------------------------------------------------------------------
// backward: dZ = d_vol * LeakyReLU'(vol);  dP[img_e, pix_k, blk] += w_k dZ;  bias / depth-column partial sums
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kFgThreads) voxel_fuse_scatter_kernel(const FuseArgs a) {
    const vfd_fusion_desc& d = a.d;
    __shared__ __align__(16) float s_part[kFgWarps][5][kFgMaxCv];
    const int Cv = d.Cv, hw = d.h * d.w, ldp = 2 * Cv;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int lpv = a.lpv, vpw = VFD_WARP / lpv, sub = lane / lpv, q4 = (lane % lpv) * 4;
    const bool chan = q4 < Cv;
    const size_t rows = (size_t)d.B * d.X * d.Y * d.Z;
    const CornerTable tb = carve_table(const_cast<void*>(a.table), rows);
    const size_t stride = (size_t)gridDim.x * kFgWarps * vpw;
    float4 s[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) s[i] = f4_zero();
    for (size_t v = ((size_t)blockIdx.x * kFgWarps + warp) * vpw + sub; v < rows; v += stride) {
        const int cnt = __ldg(a.count + v);
        if (!((cnt == 1 || cnt == 2) && chan)) continue;
        const int blk = cnt - 1;
        const float4 y = ld_stream_f4(a.vol + v * Cv + q4), g = ld_stream_f4(a.d_vol + v * Cv + q4);
        const float4 dz = make_float4(g.x * (y.x > 0.f ? 1.f : 0.1f), g.y * (y.y > 0.f ? 1.f : 0.1f),
                                      g.z * (y.z > 0.f ? 1.f : 0.1f), g.w * (y.w > 0.f ? 1.f : 0.1f));
        f4_add(s[2 * blk], dz);
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            if (e < cnt) {
                const Entry en = load_entry(tb, rows, e, v);
                float* base = a.dP + (size_t)en.img * hw * ldp + blk * Cv + q4;
                if (en.wgt.x != 0.f) red_add_f4(base + (size_t)en.pix.x * ldp, f4_scale(en.wgt.x, dz));
                if (en.wgt.y != 0.f) red_add_f4(base + (size_t)en.pix.y * ldp, f4_scale(en.wgt.y, dz));
                if (en.wgt.z != 0.f) red_add_f4(base + (size_t)en.pix.z * ldp, f4_scale(en.wgt.z, dz));
                if (en.wgt.w != 0.f) red_add_f4(base + (size_t)en.pix.w * ldp, f4_scale(en.wgt.w, dz));
                const int gq = d.group[en.img % d.num_cams] ? 1 : 0;
                f4_fma(s[blk ? 3 + gq : 1], en.zrel, dz);
            }
        }
    }
    // fixed-order reduction: over the warp's voxel slots (shuffles), over the CTA's warps (shared memory), over the CTAs
    // (voxel_fuse_reduce_kernel)
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        for (int o = lpv; o < VFD_WARP; o <<= 1) {
            s[i].x += __shfl_xor_sync(0xffffffffu, s[i].x, o);
            s[i].y += __shfl_xor_sync(0xffffffffu, s[i].y, o);
            s[i].z += __shfl_xor_sync(0xffffffffu, s[i].z, o);
            s[i].w += __shfl_xor_sync(0xffffffffu, s[i].w, o);
        }
        if (sub == 0 && chan) *reinterpret_cast<float4*>(&s_part[warp][i][q4]) = s[i];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 5 * Cv; i += kFgThreads) {
        const int q = i / Cv, n = i - q * Cv;
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kFgWarps; ++w) t += s_part[w][q][n];
        a.partial[((size_t)blockIdx.x * 5 + q) * Cv + n] = t;
    }
}

// d_w_no [Cv][C1], d_b_no [Cv], d_w_ov [Cv][2 C1], d_b_ov [Cv] from the dense partial products (dense_dw: n_parts blocks
// [2 Cv][C], each belonging to one overlap group) and the scatter kernel's per-CTA bias / depth-column sums, added in
// a fixed order.
struct ReduceArgs {
    const float* dw_part;   // [n_parts][2 Cv][C]
    const float* sc_part;   // [ctas][5][Cv]
    int n_parts, split, slabs, ctas, C, Cv, cams;
    int group[VFD_MAX_CAMS];
    float *d_w_no, *d_b_no, *d_w_ov, *d_b_ov;
};

__device__ __forceinline__ int part_group(const ReduceArgs& r, int i) {
    return r.split >= 0 ? (i >= r.split ? 1 : 0) : (r.group[(i / r.slabs) % r.cams] ? 1 : 0);
}

// grid.x covers the (n, c) elements 64 at a time, the block's 4 thread rows split the partial blocks; every thread sums
// its quarter in index order, the quarters are added in order: a fixed summation tree (deterministic given the partials)
constexpr int kRedCols = 64, kRedRows = 4;
constexpr int kBiasCols = 16, kBiasRows = kRedCols * kRedRows / kBiasCols;    // 16 x 16 <= the 2 x 4 x 64 floats of s_acc
__global__ void __launch_bounds__(kRedCols * kRedRows) voxel_fuse_reduce_kernel(const ReduceArgs r) {
    __shared__ float s_acc[kRedRows][2][kRedCols];
    const int C = r.C, Cv = r.Cv, C1 = C + 1, N2 = 2 * Cv;
    const int n_mat = N2 * C;            // (n, c)
    const int col = threadIdx.x % kRedCols, row = threadIdx.x / kRedCols;
    const int i = blockIdx.x * kRedCols + col;
    if (blockIdx.x * kRedCols < n_mat) {           // block-uniform: matrix part
        float acc[2] = {0.f, 0.f};
        if (i < n_mat) {
            const float* p = r.dw_part + i;
            const int per = (r.n_parts + kRedRows - 1) / kRedRows;
            const int k0 = row * per, k1 = min(r.n_parts, k0 + per);
            if (r.split >= 0) {          // two contiguous runs of blocks: independent loads, many in flight
                const int m = min(max(r.split, k0), k1);
#pragma unroll 8
                for (int k = k0; k < m; ++k) acc[0] += p[(size_t)k * n_mat];
#pragma unroll 8
                for (int k = m; k < k1; ++k) acc[1] += p[(size_t)k * n_mat];
            } else {
                for (int k = k0; k < k1; ++k) {
                    const float v = p[(size_t)k * n_mat];
                    if (part_group(r, k)) acc[1] += v; else acc[0] += v;
                }
            }
        }
        s_acc[row][0][col] = acc[0];
        s_acc[row][1][col] = acc[1];
        __syncthreads();
        if (row == 0 && i < n_mat) {
            float a0 = 0.f, a1 = 0.f;
#pragma unroll
            for (int q = 0; q < kRedRows; ++q) { a0 += s_acc[q][0][col]; a1 += s_acc[q][1][col]; }
            const int n = i / C, c = i - n * C;
            if (n >= Cv) {
                r.d_w_ov[(size_t)(n - Cv) * 2 * C1 + c] = a0;
                r.d_w_ov[(size_t)(n - Cv) * 2 * C1 + C1 + c] = a1;
            } else {
                r.d_w_no[(size_t)n * C1 + c] = a0 + a1;     // W_no serves both camera groups
            }
        }
        return;
    }
    // bias / relative-depth-column part: kBiasCols elements per block, the block's kBiasRows thread rows each add every
    // kBiasRows-th CTA partial in index order, the rows are then added in order (fixed tree; one thread per element
    // walking all ~1200 partials was most of this kernel's time)
    const int bcol = threadIdx.x % kBiasCols, brow = threadIdx.x / kBiasCols;
    const int j = (blockIdx.x - (n_mat + kRedCols - 1) / kRedCols) * kBiasCols + bcol;
    float* s_b = &s_acc[0][0][0];
    float acc = 0.f;
    if (j < 5 * Cv) {
#pragma unroll 8
        for (int c = brow; c < r.ctas; c += kBiasRows) acc += r.sc_part[(size_t)c * 5 * Cv + j];
    }
    s_b[brow * kBiasCols + bcol] = acc;
    __syncthreads();
    if (brow != 0 || j >= 5 * Cv) return;
    acc = 0.f;
#pragma unroll
    for (int q = 0; q < kBiasRows; ++q) acc += s_b[q * kBiasCols + bcol];
    const int q = j / Cv, n = j - q * Cv;
    if (q == 0) r.d_b_no[n] = acc;
    else if (q == 1) r.d_w_no[(size_t)n * C1 + C] = acc;
    else if (q == 2) r.d_b_ov[n] = acc;
    else r.d_w_ov[(size_t)n * 2 * C1 + (q - 3) * C1 + C] = acc;
}

static int check_fuse(const vfd_fusion_desc* d) {
    if (!d) return VFD_ERR_NULL;
    if (d->B < 1 || d->num_cams < 1 || d->num_cams > VFD_MAX_CAMS || d->C < 4 || d->h < 2 || d->w < 2 || d->X < 1 ||
        d->Y < 1 || d->Z < 1 || d->Cv < 4)
        return VFD_ERR_DESC;
    if (d->C % 4 || d->C > 256 || d->Cv > kFgMaxCv || d->Cv % 4) return VFD_ERR_UNSUPPORTED;
    if ((long)d->h * d->w * d->C >= (1L << 31)) return VFD_ERR_DESC;
    if ((long)d->B * d->X * d->Y * d->Z >= (1L << 31)) return VFD_ERR_DESC;
    return VFD_OK;
}

static int sm_count() {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    return n;
}

static int lanes_per_voxel(int Cv) {
    int l = 1;
    while (l * 4 < Cv) l <<= 1;
    return l;
}

static size_t align256(size_t b) { return (b + 255) / 256 * 256; }

// workspace carving (floats, every block 256-byte aligned)
struct FuseWs {
    size_t p, w_full, w_hi, w_lo, dw_part, sc_part, total;
};
static FuseWs carve_ws(const vfd_fusion_desc* d, bool bwd, int ctas) {
    FuseWs w = {};
    const size_t n_img = (size_t)d->B * d->num_cams, hw = (size_t)d->h * d->w, N2 = 2 * (size_t)d->Cv;
    size_t o = 0;
    w.p = o;        o += align256(n_img * hw * N2 * sizeof(float));
    w.w_full = o;   o += align256(2 * N2 * d->C * sizeof(float));
    w.w_hi = o;     o += align256(2 * N2 * d->C * sizeof(float));
    w.w_lo = o;     o += align256(2 * N2 * d->C * sizeof(float));
    if (bwd) {
        w.dw_part = o;  o += align256(dense_dw_part_floats((int)n_img, d->C, (int)N2) * sizeof(float));
        w.sc_part = o;  o += align256((size_t)ctas * 5 * d->Cv * sizeof(float));
    }
    w.total = o;
    return w;
}

}  // namespace vfd

using namespace vfd;

extern "C" size_t vfd_voxel_fuse_fwd_workspace_bytes(const vfd_fusion_desc* d) {
    if (check_fuse(d)) return 0;
    return carve_ws(d, false, 0).total;
}

extern "C" int vfd_voxel_fuse_fwd(const vfd_fusion_desc* d, const float* feats, const uint8_t* count, const void* table,
                                  const float* w_no, const float* b_no, const float* w_ov, const float* b_ov, float* vol,
                                  void* ws, size_t ws_bytes, void* stream) {
    int rc = check_fuse(d);
    if (rc) return rc;
    if (!feats || !count || !table || !w_no || !b_no || !w_ov || !b_ov || !vol || !ws) return VFD_ERR_NULL;
    if ((reinterpret_cast<uintptr_t>(feats) & 15) || (reinterpret_cast<uintptr_t>(vol) & 15) ||
        (reinterpret_cast<uintptr_t>(table) & 15) || (reinterpret_cast<uintptr_t>(ws) & 255))
        return VFD_ERR_ALIGN;
    const FuseWs w = carve_ws(d, false, 0);
    if (ws_bytes < w.total) return VFD_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const int sms = sm_count();
    if (sms <= 0) return VFD_ERR_CUDA;
    char* wsb = static_cast<char*>(ws);
    float* P = reinterpret_cast<float*>(wsb + w.p);
    float* wf = reinterpret_cast<float*>(wsb + w.w_full);
    float* wh = reinterpret_cast<float*>(wsb + w.w_hi);
    float* wl = reinterpret_cast<float*>(wsb + w.w_lo);
    const int N2 = 2 * d->Cv, n_img = d->B * d->num_cams, hw = d->h * d->w;
    wcat_prep_kernel<<<(2 * N2 * d->C + 255) / 256, 256, 0, st>>>(w_no, w_ov, d->C, d->Cv, 0, wf, wh, wl);
    VFD_CHECK_LAUNCH();
    if ((rc = dense_proj(n_img, hw, d->C, N2, d->num_cams, d->group, feats, P, wf, wh, wl, st))) return rc;
    FuseArgs a = {};
    a.d = *d; a.count = count; a.table = table; a.P = P;
    a.w_no = w_no; a.b_no = b_no; a.w_ov = w_ov; a.b_ov = b_ov; a.vol = vol;
    a.lpv = lanes_per_voxel(d->Cv);
    const size_t n_vox = (size_t)d->B * d->X * d->Y * d->Z;
    voxel_fuse_gather_kernel<<<(unsigned)((n_vox + kFgWarps * kFgRun - 1) / (kFgWarps * kFgRun)), kFgThreads, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

static int bwd_ctas(int sms) { return sms * kFgCtasPerSm; }

extern "C" size_t vfd_voxel_fuse_bwd_workspace_bytes(const vfd_fusion_desc* d) {
    if (check_fuse(d)) return 0;
    const int sms = sm_count();
    return carve_ws(d, true, bwd_ctas(sms > 0 ? sms : 256)).total;
}

extern "C" int vfd_voxel_fuse_bwd(const vfd_fusion_desc* d, const float* feats, const uint8_t* count, const void* table,
                                  const float* w_no, const float* w_ov, const float* vol, const float* d_vol,
                                  float* d_feats, float* d_w_no, float* d_b_no, float* d_w_ov, float* d_b_ov, void* ws,
                                  size_t ws_bytes, void* stream) {
    int rc = check_fuse(d);
    if (rc) return rc;
    if (!feats || !count || !table || !w_no || !w_ov || !vol || !d_vol || !d_feats || !d_w_no || !d_b_no || !d_w_ov ||
        !d_b_ov || !ws)
        return VFD_ERR_NULL;
    if ((reinterpret_cast<uintptr_t>(feats) & 15) || (reinterpret_cast<uintptr_t>(d_feats) & 15) ||
        (reinterpret_cast<uintptr_t>(vol) & 15) || (reinterpret_cast<uintptr_t>(d_vol) & 15) ||
        (reinterpret_cast<uintptr_t>(table) & 15) || (reinterpret_cast<uintptr_t>(ws) & 255))
        return VFD_ERR_ALIGN;
    const int sms = sm_count();
    if (sms <= 0) return VFD_ERR_CUDA;
    const int ctas = bwd_ctas(sms);
    const FuseWs w = carve_ws(d, true, ctas);
    if (ws_bytes < w.total) return VFD_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    char* wsb = static_cast<char*>(ws);
    float* dP = reinterpret_cast<float*>(wsb + w.p);
    float* wf = reinterpret_cast<float*>(wsb + w.w_full);
    float* wh = reinterpret_cast<float*>(wsb + w.w_hi);
    float* wl = reinterpret_cast<float*>(wsb + w.w_lo);
    float* dwp = reinterpret_cast<float*>(wsb + w.dw_part);
    float* scp = reinterpret_cast<float*>(wsb + w.sc_part);
    const int N2 = 2 * d->Cv, n_img = d->B * d->num_cams, hw = d->h * d->w;
    if (cudaMemsetAsync(dP, 0, (size_t)n_img * hw * N2 * sizeof(float), st) != cudaSuccess) return VFD_ERR_CUDA;
    wcat_prep_kernel<<<(2 * N2 * d->C + 255) / 256, 256, 0, st>>>(w_no, w_ov, d->C, d->Cv, 1, wf, wh, wl);
    VFD_CHECK_LAUNCH();
    FuseArgs a = {};
    a.d = *d; a.count = count; a.table = table; a.vol = const_cast<float*>(vol); a.d_vol = d_vol; a.dP = dP;
    a.partial = scp; a.lpv = lanes_per_voxel(d->Cv);
    voxel_fuse_scatter_kernel<<<ctas, kFgThreads, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    // d_feats[img, p, :] = dP[img, p, :] . Wcat_g   (K = 2 Cv, N = C; weight planes transposed: [2][C][2 Cv])
    if ((rc = dense_proj(n_img, hw, N2, d->C, d->num_cams, d->group, dP, d_feats, wf, wh, wl, st))) return rc;
    int n_parts = 0, split = -1;
    if ((rc = dense_dw(n_img, hw, d->C, N2, d->num_cams, d->group, dP, feats, dwp, &n_parts, &split, st))) return rc;
    ReduceArgs r = {};
    r.dw_part = dwp; r.sc_part = scp; r.n_parts = n_parts; r.split = split; r.slabs = dense_dw_slabs(n_img); r.ctas = ctas;
    r.C = d->C; r.Cv = d->Cv; r.cams = d->num_cams;
    for (int c = 0; c < VFD_MAX_CAMS; ++c) r.group[c] = c < d->num_cams ? d->group[c] : 0;
    r.d_w_no = d_w_no; r.d_b_no = d_b_no; r.d_w_ov = d_w_ov; r.d_b_ov = d_b_ov;
    const int blocks = (N2 * d->C + kRedCols - 1) / kRedCols + (5 * d->Cv + kBiasCols - 1) / kBiasCols;
    voxel_fuse_reduce_kernel<<<blocks, kRedCols * kRedRows, 0, st>>>(r);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
