// Fused voxel gather + the two 1x1 MLPs of VFNet.backproject_into_voxel, model='depth'
// (network/volumetric_fusionnet.py:116-159 with preprocess_non_overlap :197-207 and preprocess_overlap :209-230).
//
// the reference materialises six [B,257,N] camera tensors, two masked GEMM inputs (617 MB/rig) and
// runs two dense 1x1 convolutions over ALL voxels although only voxels seen by exactly one camera reach the first MLP
// and only voxels seen by exactly two reach the second.  Here a CTA takes a tile of 64 such voxels from the compact
// lists of vfd_voxel_classify, gathers their bilinear feature rows straight into shared memory (the feature maps are
// L2-resident), multiplies them with the MLP weight that also sits in shared memory and writes LeakyReLU(x W^T + b)
// into the channel-last voxel volume -- the GEMM rows never exist in HBM.  Algorithmic HBM bytes: 23.6 MB of feature
// maps + 51.2 MB of volume per rig instead of 640 MB.
//
// Both directions run as three passes over the lists, one per (C+1) x Cv weight block:
//   pass 0: count-1 voxels x W_no;  pass 1: count-2 voxels, camera group 0 x W_ov[:, :C+1];  pass 2: group 1 x W_ov[:, C+1:]
// (forward pass 1 parks its raw accumulators in the volume, pass 2 adds its own, the bias and the LeakyReLU).
//
// The products run on the tensor cores through mma.sync.m16n8k8 TF32 with the error-compensated 3xTF32 split
// (x = hi + lo, x*w ~ hi*hi + hi*lo + lo*hi, fp32 accumulate), which keeps the result within fp32 round-off of the
// reference's fp32 GEMM (tests: 1e-4 of the tensor's scale; measured ~3e-6).  The weight block is split once per pass
// into TF32 hi/lo planes in shared memory.
//
// Persistent grid (one CTA per SM, 16 warps); the list lengths are read from device memory (vfd_voxel_classify's
// totals), so nothing here needs a host round trip.
#include <stdlib.h>

#include "gather_common.cuh"

namespace vfd {

constexpr int kFuseRows = 64;       // voxels per tile
constexpr int kFuseWarps = 16;
constexpr int kFuseThreads = kFuseWarps * VFD_WARP;
constexpr int kRowsPerWarp = kFuseRows / kFuseWarps;
constexpr int kWS = 72;             // row stride of [k][n] planes (8k+n banks: conflict-free B fragments)
constexpr int kMaxEnt = 2;          // cameras of one group that can see a listed voxel (count <= 2)

struct FuseArgs {
    vfd_fusion_desc d;
    const float* feats;
    const float* mask_small;
    const float* K;
    const float* Einv;
    const int64_t* idx_no;
    const int64_t* idx_ov;
    const int32_t* totals;
    const float* w_no;   // [Cv][C+1]      conv_non_overlap.0.weight
    const float* b_no;   // [Cv]
    const float* w_ov;   // [Cv][2(C+1)]   conv_overlap.0.weight
    const float* b_ov;   // [Cv]
    float* vol;          // [B*N][Cv]
    // backward
    const float* d_vol;  // [B*N][Cv]
    float* d_feats;      // [B,cams,h,w,C]
    float* partial;      // [gridDim.x][3][Cv*(C+2)] per-CTA weight-gradient partial sums
    const void* table;   // corner table of vfd_voxel_classify (gather_common.cuh), B*N rows
    int KP;              // round_up(C+1, 8)
    int NT;              // ceil(Cv / 8)
};

// a group of warps that works on one tile: the whole CTA (barrier 0) or half of it (named barriers 1, 2)
struct Group {
    int tid, size, bar;
};
__device__ __forceinline__ void group_sync(const Group& G) {
    asm volatile("bar.sync %0, %1;" ::"r"(G.bar), "r"(G.size) : "memory");
}

// per-tile bookkeeping shared by forward and backward: for every row of the tile the (at most two) cameras of the
// current pass that see the voxel, with their bilinear corners precomputed.  ROWS rows per tile.
struct TileMeta {
    long long* row;   // [ROWS]     b*N+n, -1 past the end of the list
    int* cnt;         // [ROWS]     entries of this row in this pass
    int4* off;        // [2][ROWS]  corner offsets (floats) inside the camera's feature map
    float4* wgt;      // [2][ROWS]  corner weights
    int2* meta;       // [2][ROWS]  {b*num_cams + c, bits of zrel}
    unsigned char* ok;  // [8][ROWS]  validity of (camera, row) while the entries are being numbered
};
__host__ __device__ constexpr size_t meta_bytes(int rows) {
    return (size_t)rows * (8 + 4 + kMaxEnt * (16 + 16 + 8) + VFD_MAX_CAMS);
}

template <int ROWS>
__device__ __forceinline__ TileMeta carve_meta(unsigned char* base) {
    TileMeta m;
    m.row = reinterpret_cast<long long*>(base);
    base += ROWS * 8;
    m.off = reinterpret_cast<int4*>(base);
    base += kMaxEnt * ROWS * 16;
    m.wgt = reinterpret_cast<float4*>(base);
    base += kMaxEnt * ROWS * 16;
    m.meta = reinterpret_cast<int2*>(base);
    base += kMaxEnt * ROWS * 8;
    m.cnt = reinterpret_cast<int*>(base);
    base += ROWS * 4;
    m.ok = base;
    return m;
}

// One thread's share of a tile's metadata, loaded from the corner table of vfd_voxel_classify one tile ahead of its
// use (the index -> table chain of two dependent global loads is hidden behind the previous tile's work).
// Thread (part = tid / ROWS, row r = tid % ROWS): parts 0..3 carry the corner offsets / weights of entry 0 / 1, part 4
// the two meta words, the row id and the entry count; parts 5..7 idle.
struct TilePf {
    long long row;
    int2 m0, m1;
    uint4 piece;
};

__device__ __forceinline__ bool entry_wanted(int2 m, unsigned want, int num_cams) {
    return m.x >= 0 && ((want >> (m.x % num_cams)) & 1u);
}

template <int ROWS>
__device__ __forceinline__ TilePf prefetch_tile(const FuseArgs& a, const CornerTable& tb, size_t rows_total,
                                                const Group& G, const int64_t* idx, int base, int M, unsigned want,
                                                bool tile_ok) {
    TilePf p;
    p.row = -1;
    p.m0 = p.m1 = make_int2(-1, 0);
    p.piece = make_uint4(0u, 0u, 0u, 0u);
    const int r = G.tid & (ROWS - 1), part = G.tid / ROWS;
    if (tile_ok && part < 5 && base + r < M) {
        p.row = idx[base + r];
        p.m0 = __ldg(tb.meta + p.row);
        p.m1 = __ldg(tb.meta + rows_total + p.row);
        if (part < 4) {
            const int e = part >> 1;
            if (entry_wanted(e ? p.m1 : p.m0, want, a.d.num_cams)) {
                const uint4* src = (part & 1) ? reinterpret_cast<const uint4*>(tb.wgt) : reinterpret_cast<const uint4*>(tb.off);
                p.piece = __ldg(src + e * rows_total + p.row);
            }
        }
    }
    return p;
}

// write the prefetched share into the tile metadata: the wanted entries of a row, in camera order.  The caller
// synchronises the group before the metadata is read.
template <int ROWS>
__device__ __forceinline__ void commit_tile(const TileMeta& tm, const Group& G, const TilePf& p, unsigned want,
                                            int num_cams) {
    const int r = G.tid & (ROWS - 1), part = G.tid / ROWS;
    const bool w0 = entry_wanted(p.m0, want, num_cams), w1 = entry_wanted(p.m1, want, num_cams);
    if (part < 4) {
        const int e = part >> 1;
        const int dest = e ? (w0 ? 1 : 0) : 0;
        if (e ? w1 : w0) {
            uint4* dst = (part & 1) ? reinterpret_cast<uint4*>(tm.wgt) : reinterpret_cast<uint4*>(tm.off);
            dst[dest * ROWS + r] = p.piece;
        }
    } else if (part == 4) {
        if (w0) tm.meta[r] = p.m0;
        if (w1) tm.meta[(w0 ? 1 : 0) * ROWS + r] = p.m1;
        tm.row[r] = p.row;
        tm.cnt[r] = (w0 ? 1 : 0) + (w1 ? 1 : 0);
    }
}

// warp-cooperative gather of NR consecutive rows into shared memory: arow[0..C) = sum over the row's entries of the
// bilinear sample, arow[C] = sum of the relative depths, arow[C+1..KP) = 0.  NCH = ceil(C / 128) chunks of 4 floats per
// lane.  The corner rows of all NR voxels (first entry each) are requested before any is consumed: NR * NCH * 4 128-bit
// loads in flight per lane; a second entry (two cameras of one group, rare) is added afterwards.
template <int NCH, int ROWS, int NR>
__device__ __forceinline__ void gather_rows(const FuseArgs& a, const TileMeta& tm, int r0, int lane, float* arow0,
                                            int stride) {
    const int C = a.d.C;
    const size_t map = (size_t)a.d.h * a.d.w * C;
    int cnt[NR];
    float4 wg[NR], v[NR][NCH][4];
    float zsum[NR];
#pragma unroll
    for (int n = 0; n < NR; ++n) {
        const int r = r0 + n;
        cnt[n] = tm.cnt[r];
        const int4 off = tm.off[r];
        wg[n] = tm.wgt[r];
        const int2 me = tm.meta[r];
        const float* base = a.feats + (size_t)me.x * map + lane * 4;
        zsum[n] = cnt[n] > 0 ? __int_as_float(me.y) : 0.f;
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
            const bool on = cnt[n] > 0 && i * 128 + lane * 4 < C;
            v[n][i][0] = on ? ldg_f4(base + off.x + i * 128) : f4_zero();
            v[n][i][1] = on ? ldg_f4(base + off.y + i * 128) : f4_zero();
            v[n][i][2] = on ? ldg_f4(base + off.z + i * 128) : f4_zero();
            v[n][i][3] = on ? ldg_f4(base + off.w + i * 128) : f4_zero();
        }
    }
#pragma unroll
    for (int n = 0; n < NR; ++n) {
        const int r = r0 + n;
        float* arow = arow0 + (size_t)n * stride;
        float4 acc[NCH];
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
            acc[i] = f4_scale(wg[n].x, v[n][i][0]);
            f4_fma(acc[i], wg[n].y, v[n][i][1]);
            f4_fma(acc[i], wg[n].z, v[n][i][2]);
            f4_fma(acc[i], wg[n].w, v[n][i][3]);
        }
        for (int e = 1; e < cnt[n]; ++e) {
            const int4 off = tm.off[e * ROWS + r];
            const float4 w2 = tm.wgt[e * ROWS + r];
            const int2 me = tm.meta[e * ROWS + r];
            const float* base = a.feats + (size_t)me.x * map + lane * 4;
            zsum[n] += __int_as_float(me.y);
#pragma unroll
            for (int i = 0; i < NCH; ++i) {
                if (i * 128 + lane * 4 < C) {
                    float4 t = f4_scale(w2.x, ldg_f4(base + off.x + i * 128));
                    f4_fma(t, w2.y, ldg_f4(base + off.y + i * 128));
                    f4_fma(t, w2.z, ldg_f4(base + off.z + i * 128));
                    f4_fma(t, w2.w, ldg_f4(base + off.w + i * 128));
                    f4_add(acc[i], t);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < NCH; ++i)
            if (i * 128 + lane * 4 < C) *reinterpret_cast<float4*>(arow + i * 128 + lane * 4) = acc[i];
        for (int k = C + lane; k < a.KP; k += VFD_WARP) arow[k] = (k == C) ? zsum[n] : 0.f;
    }
}

// ---------------------------------------------------------------------------------------------------------
// Forward.  The CTA's 16 warps form two groups of 8 that share the weight planes but otherwise run independently
// (named barriers), each on its own 32-voxel tiles: while one group waits for its L2 gathers the other keeps the tensor
// pipe busy.  Inside a group: 2 row blocks of 16 x 4 quarters of the K range, partial sums reduced through shared memory.
// ---------------------------------------------------------------------------------------------------------
constexpr int kFwdRows = 32;
constexpr int kFwdGroups = 2;
constexpr int kFwdGroupThreads = kFuseThreads / kFwdGroups;

__host__ __device__ inline size_t fuse_fwd_group_bytes(int KP) {
    const size_t a_tile = (size_t)kFwdRows * (KP + 4) * 4, partials = (size_t)4 * kFwdRows * kWS * 4;
    return meta_bytes(kFwdRows) + (a_tile > partials ? a_tile : partials);
}
__host__ __device__ inline size_t fuse_fwd_smem_bytes(int KP) {
    return (size_t)2 * KP * kWS * 4 + kFwdGroups * fuse_fwd_group_bytes(KP);
}

template <int NCH>
__global__ void __launch_bounds__(kFuseThreads, 1) voxel_fuse_fwd_kernel(const FuseArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const vfd_fusion_desc& d = a.d;
    const int KP = a.KP, AS = KP + 4, NT = a.NT, C1 = d.C + 1, Cv = d.Cv;
    uint32_t* Whi = reinterpret_cast<uint32_t*>(smem_raw);                // [KP][kWS] tf32 hi plane, [k][n]
    uint32_t* Wlo = Whi + (size_t)KP * kWS;
    const int grp = threadIdx.x / kFwdGroupThreads;
    const Group G = {(int)threadIdx.x - grp * kFwdGroupThreads, kFwdGroupThreads, 1 + grp};
    unsigned char* gbase = smem_raw + (size_t)2 * KP * kWS * 4 + (size_t)grp * fuse_fwd_group_bytes(KP);
    const TileMeta tm = carve_meta<kFwdRows>(gbase);
    float* sA = reinterpret_cast<float*>(gbase + meta_bytes(kFwdRows));   // [32][AS]; later the K-split partials
    const int warp = G.tid >> 5, lane = G.tid & 31, g = lane >> 2, t = lane & 3;
    const int wm = warp & 1, wk = warp >> 1;      // 16-row block, quarter of the K range
    const int M1 = a.totals[0], M2 = a.totals[1];
    const int T1 = (M1 + kFwdRows - 1) / kFwdRows, T2 = (M2 + kFwdRows - 1) / kFwdRows;
    unsigned gmask[2] = {0u, 0u};
    for (int c = 0; c < d.num_cams; ++c) gmask[d.group[c] ? 1 : 0] |= 1u << c;
    const int nk = KP / 8, kq = (nk + 3) / 4;
    const int ks0 = wk * kq, ks1 = min(nk, ks0 + kq);
    const size_t rows_total = (size_t)d.B * d.X * d.Y * d.Z;
    const CornerTable tb = carve_table(const_cast<void*>(a.table), rows_total);

    for (int pass = 0; pass < 3; ++pass) {
        const int M = pass ? M2 : M1, T = pass ? T2 : T1;
        const int64_t* idx = pass ? a.idx_ov : a.idx_no;
        const unsigned want = pass ? gmask[pass - 1] : 0xffffffffu;
        const float* bias = pass ? a.b_ov : a.b_no;
        __syncthreads();
        {   // weight block [k][n] = w[n][koff + k], split into TF32 hi / lo planes, zero padded
            const float* w = pass ? a.w_ov : a.w_no;
            const int ld = pass ? 2 * C1 : C1, koff = pass == 2 ? C1 : 0;
            for (int i = threadIdx.x; i < KP * kWS; i += blockDim.x) { Whi[i] = 0u; Wlo[i] = 0u; }
            __syncthreads();
            for (int i = threadIdx.x; i < Cv * C1; i += blockDim.x) {
                const int n = i / C1, k = i - n * C1;
                uint32_t hi, lo;
                split_tf32(__ldg(w + (size_t)n * ld + koff + k), hi, lo);
                Whi[k * kWS + n] = hi;
                Wlo[k * kWS + n] = lo;
            }
        }
        __syncthreads();
        const int tstep = gridDim.x * kFwdGroups;
        int tile = blockIdx.x * kFwdGroups + grp;
        TilePf pf = prefetch_tile<kFwdRows>(a, tb, rows_total, G, idx, tile * kFwdRows, M, want, tile < T);
        for (; tile < T; tile += tstep) {
            commit_tile<kFwdRows>(tm, G, pf, want, d.num_cams);
            pf = prefetch_tile<kFwdRows>(a, tb, rows_total, G, idx, (tile + tstep) * kFwdRows, M, want, tile + tstep < T);
            group_sync(G);
#pragma unroll
            for (int rr = 0; rr < 4; rr += 2)
                gather_rows<NCH, kFwdRows, 2>(a, tm, warp * 4 + rr, lane, sA + (size_t)(warp * 4 + rr) * AS, AS);
            group_sync(G);

            // [16 rows x (this warp's K quarter)] x [K x Cv]: 3xTF32 on the tensor cores
            float acc[8][4];
#pragma unroll
            for (int j = 0; j < 8; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[j][i] = 0.f;
            {
                const float* A0 = sA + (size_t)(wm * 16 + g) * AS + t;
                const float* A1 = A0 + 8 * AS;
                const uint32_t* Bh = Whi + t * kWS + g;
                const uint32_t* Bl = Wlo + t * kWS + g;
                for (int ks = ks0; ks < ks1; ++ks) {
                    uint32_t ah[4], al[4];
                    split_tf32(A0[ks * 8], ah[0], al[0]);
                    split_tf32(A1[ks * 8], ah[1], al[1]);
                    split_tf32(A0[ks * 8 + 4], ah[2], al[2]);
                    split_tf32(A1[ks * 8 + 4], ah[3], al[3]);
                    const uint32_t* bh = Bh + ks * 8 * kWS;
                    const uint32_t* bl = Bl + ks * 8 * kWS;
#pragma unroll
                    for (int jg = 0; jg < 8; jg += 4) {
                        if (jg < NT) {      // four independent accumulators per round: no back-to-back dependent MMAs
                            uint32_t bh0[4], bh1[4], bl0[4], bl1[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                bh0[j] = bh[(jg + j) * 8];
                                bh1[j] = bh[4 * kWS + (jg + j) * 8];
                                bl0[j] = bl[(jg + j) * 8];
                                bl1[j] = bl[4 * kWS + (jg + j) * 8];
                            }
#pragma unroll
                            for (int j = 0; j < 4; ++j) mma_tf32(acc[jg + j], al, bh0[j], bh1[j]);
#pragma unroll
                            for (int j = 0; j < 4; ++j) mma_tf32(acc[jg + j], ah, bl0[j], bl1[j]);
#pragma unroll
                            for (int j = 0; j < 4; ++j) mma_tf32(acc[jg + j], ah, bh0[j], bh1[j]);
                        }
                    }
                }
            }
            group_sync(G);       // the group is done with the A tile: reuse it for the K-split partial sums
            float* P = sA;       // [4][32][kWS]
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (j < NT) {
                    float* p0 = P + ((size_t)wk * kFwdRows + wm * 16 + g) * kWS + j * 8 + 2 * t;
                    *reinterpret_cast<float2*>(p0) = make_float2(acc[j][0], acc[j][1]);
                    *reinterpret_cast<float2*>(p0 + 8 * kWS) = make_float2(acc[j][2], acc[j][3]);
                }
            }
            group_sync(G);
            {   // sum the four K quarters, finish the row and write it (one 16-byte store per quad)
                const int nq = Cv / 4;
                for (int i = G.tid; i < kFwdRows * nq; i += G.size) {
                    const int r = i / nq, q = (i - r * nq) * 4;
                    const long long row = tm.row[r];
                    if (row < 0) continue;
                    float4 y = *reinterpret_cast<const float4*>(P + (size_t)r * kWS + q);
#pragma unroll
                    for (int u = 1; u < 4; ++u)
                        f4_add(y, *reinterpret_cast<const float4*>(P + ((size_t)u * kFwdRows + r) * kWS + q));
                    float* o = a.vol + (size_t)row * Cv + q;
                    if (pass == 2) f4_add(y, *reinterpret_cast<const float4*>(o));   // the group-0 half parked by pass 1
                    if (pass != 1) {
                        f4_add(y, ldg_f4(bias + q));
                        y.x = y.x > 0.f ? y.x : 0.1f * y.x;
                        y.y = y.y > 0.f ? y.y : 0.1f * y.y;
                        y.z = y.z > 0.f ? y.z : 0.1f * y.z;
                        y.w = y.w > 0.f ? y.w : 0.1f * y.w;
                    }
                    *reinterpret_cast<float4*>(o) = y;
                }
            }
            group_sync(G);     // the metadata and the A tile are rewritten by the next tile
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Forward on the 5th-generation tensor cores (tcgen05.mma kind::tf32, accumulators in TMEM).  Used when C % 32 == 0 and
// Cv % 16 == 0 (the DDAD / nuScenes configurations); the mma.sync kernel above covers the rest.
//
// Tile = 128 (voxel, camera-group) rows of one pass x Cv outputs; K = C runs in chunks of 32 channels (one 128-byte
// swizzle atom row per voxel).  Roles, one CTA per SM, 13 warps:
//   warps 0-3   epilogue: tcgen05.ld of the finished accumulator (TMEM lane = row), + relative-depth column, bias,
//               LeakyReLU, 16-byte stores of the 256-byte voxel row
//   warp  4     one elected lane issues the MMAs: per chunk 4 K-steps x {A_lo W_hi, A_hi W_lo, A_hi W_hi} (3xTF32);
//               tcgen05.commit releases the A stage / publishes the accumulator through mbarriers
//   warps 5-12  producers: bilinear gather of the chunk's 4 corner segments (128 B each) per voxel straight from the
//               L2-resident feature maps, split into TF32 hi / lo, written in the K-major SWIZZLE_128B layout the
//               MMA descriptors expect (2 stages of 32 KB)
// The weight block of the pass sits in shared memory as hi / lo planes in the same layout (128 KB at C=256, Cv=64);
// two accumulators (2 x Cv TMEM columns) let the epilogue of tile i overlap the MMAs of tile i+1.  Descriptor encoding
// validated by tools/umma_probe.cu.
// ---------------------------------------------------------------------------------------------------------
constexpr int kTcRows = 128;
constexpr int kTcEpiWarps = 4, kTcProdWarps = 8;
constexpr int kTcProdThreads = kTcProdWarps * VFD_WARP;
constexpr int kTcThreads = (kTcEpiWarps + 1 + kTcProdWarps) * VFD_WARP;
constexpr int kTcStages = 2;
constexpr int kTcStageBytes = 2 * kTcRows * 128;       // hi + lo plane of one 32-channel chunk

struct TcMeta {          // per tile (double buffered)
    long long row[kTcRows];
    int4 off[kMaxEnt][kTcRows];
    float4 wgt[kMaxEnt][kTcRows];
    int cam[kMaxEnt][kTcRows];      // b*num_cams + c
    int cnt[kTcRows];
    float zrel[kTcRows];            // sum of the wanted entries' relative depths
};

struct TcBars {
    uint64_t full[kTcStages], empty[kTcStages], acc_full[2], acc_empty[2], meta_free[2];
    uint32_t tmem_base;
};

__host__ __device__ inline size_t fuse_tc_smem_bytes(int C, int Cv) {
    return 1024 /*alignment slack*/ + (size_t)2 * (C / 32) * Cv * 128 + (size_t)kTcStages * kTcStageBytes +
           2 * sizeof(TcMeta) + 3 * 64 * sizeof(float) + sizeof(TcBars);
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t* bar) {   // arrives on `bar` when all MMAs issued so far have finished
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// shared-memory matrix descriptor: K-major, SWIZZLE_128B, 8-row atoms of 1024 bytes (tools/umma_probe.cu)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
           ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
        "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
        : "memory");
}
// byte offset of (row r, 16-byte chunk c16) inside a [rows][128 B] SWIZZLE_128B tile
__device__ __forceinline__ uint32_t sw128_off(int r, int c16) {
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c16 ^ (r & 7)) << 4));
}

__global__ void __launch_bounds__(kTcThreads, 1) voxel_fuse_fwd_tc_kernel(const FuseArgs a) {
    extern __shared__ unsigned char smem_raw[];
    const vfd_fusion_desc& d = a.d;
    const int C = d.C, C1 = C + 1, Cv = d.Cv, NCH = C / 32;
    unsigned char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // swizzle atoms: 1024-byte aligned
    const uint32_t wplane = (uint32_t)NCH * Cv * 128;      // bytes of one weight plane: [NCH][Cv rows][128 B]
    unsigned char* sW = base;
    unsigned char* sA = sW + 2 * wplane;                   // [stage][hi, lo][128 rows][128 B]
    TcMeta* meta = reinterpret_cast<TcMeta*>(sA + kTcStages * kTcStageBytes);
    float* sWz = reinterpret_cast<float*>(meta + 2);       // [64] relative-depth column of the pass' weight block
    float* sBias = sWz + 64;                               // [64]
    TcBars* bars = reinterpret_cast<TcBars*>(sBias + 128);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int M1 = a.totals[0], M2 = a.totals[1];
    const int T1 = (M1 + kTcRows - 1) / kTcRows, T2 = (M2 + kTcRows - 1) / kTcRows;
    unsigned gmask[2] = {0u, 0u};
    for (int c = 0; c < d.num_cams; ++c) gmask[d.group[c] ? 1 : 0] |= 1u << c;
    const size_t rows_total = (size_t)d.B * d.X * d.Y * d.Z;
    const CornerTable tb = carve_table(const_cast<void*>(a.table), rows_total);
    const size_t map = (size_t)d.h * d.w * C;

    if (tid == 0) {
        for (int s = 0; s < kTcStages; ++s) { mbar_init(&bars->full[s], kTcProdThreads); mbar_init(&bars->empty[s], 1); }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&bars->acc_full[i], 1);
            mbar_init(&bars->acc_empty[i], kTcEpiWarps * VFD_WARP);
            mbar_init(&bars->meta_free[i], kTcEpiWarps * VFD_WARP);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kTcEpiWarps) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(smem_u32(&bars->tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = bars->tmem_base;
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(Cv >> 3) << 17) | ((uint32_t)(kTcRows >> 4) << 24);

    uint32_t stage = 0, phase = 0;      // A-stage ring position of this role (producers and the MMA warp)
    uint32_t ntile = 0;                 // tiles this role has started (same sequence in every role)

    for (int pass = 0; pass < 3; ++pass) {
        const int M = pass ? M2 : M1, T = pass ? T2 : T1;
        const int64_t* idx = pass ? a.idx_ov : a.idx_no;
        const unsigned want = pass ? gmask[pass - 1] : 0xffffffffu;
        __syncthreads();                // every role has drained the previous pass: the weight planes are free
        {
            const float* w = pass ? a.w_ov : a.w_no;
            const int ld = pass ? 2 * C1 : C1, koff = pass == 2 ? C1 : 0;
            for (int i = tid; i < Cv * C; i += kTcThreads) {
                const int n = i / C, k = i - n * C;
                uint32_t hi, lo;
                split_tf32(__ldg(w + (size_t)n * ld + koff + k), hi, lo);
                const uint32_t o = (uint32_t)(k >> 5) * Cv * 128 + sw128_off(n, (k & 31) >> 2) + (k & 3) * 4;
                *reinterpret_cast<uint32_t*>(sW + o) = hi;
                *reinterpret_cast<uint32_t*>(sW + wplane + o) = lo;
            }
            if (tid < 64) {
                sWz[tid] = tid < Cv ? __ldg(w + (size_t)tid * ld + koff + C) : 0.f;
                sBias[tid] = tid < Cv ? __ldg((pass ? a.b_ov : a.b_no) + tid) : 0.f;
            }
        }
        fence_proxy_async_smem();
        __syncthreads();

        if (warp > kTcEpiWarps) {
            // ===================== producers =====================
            const int ptid = tid - (kTcEpiWarps + 1) * VFD_WARP;
            const int c16 = ptid & 7, r0 = ptid >> 3;        // item it: row r0 + 32 it, 16-byte chunk c16 of the 128-byte row
            for (int tile = blockIdx.x; tile < T; tile += gridDim.x, ++ntile) {
                TcMeta& tm = meta[ntile & 1];
                mbar_wait(&bars->meta_free[ntile & 1], ((ntile >> 1) & 1) ^ 1);
                {   // metadata of the tile from the corner table: thread (half, r)
                    const int r = ptid & (kTcRows - 1), half = ptid >> 7;
                    const int li = tile * kTcRows + r;
                    long long row = -1;
                    int2 m0 = make_int2(-1, 0), m1 = make_int2(-1, 0);
                    if (li < M) {
                        row = idx[li];
                        m0 = __ldg(tb.meta + row);
                        m1 = __ldg(tb.meta + rows_total + row);
                    }
                    const bool w0 = entry_wanted(m0, want, d.num_cams), w1 = entry_wanted(m1, want, d.num_cams);
                    if (half == 0) {
                        if (w0) {
                            tm.off[0][r] = __ldg(tb.off + row);
                            tm.wgt[0][r] = __ldg(tb.wgt + row);
                            tm.cam[0][r] = m0.x;
                        }
                        tm.row[r] = row;
                        tm.cnt[r] = (w0 ? 1 : 0) + (w1 ? 1 : 0);
                        tm.zrel[r] = (w0 ? __int_as_float(m0.y) : 0.f) + (w1 ? __int_as_float(m1.y) : 0.f);
                    } else if (w1) {
                        const int dst = w0 ? 1 : 0;
                        tm.off[dst][r] = __ldg(tb.off + rows_total + row);
                        tm.wgt[dst][r] = __ldg(tb.wgt + rows_total + row);
                        tm.cam[dst][r] = m1.x;
                    }
                }
                asm volatile("bar.sync 1, %0;" ::"n"(kTcProdThreads) : "memory");
                int4 off[4];
                float4 wg[4];
                const float* src[4];
                int cnt[4];
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int r = r0 + 32 * it;
                    cnt[it] = tm.cnt[r];
                    off[it] = tm.off[0][r];
                    wg[it] = tm.wgt[0][r];
                    src[it] = a.feats + (size_t)(cnt[it] > 0 ? tm.cam[0][r] : 0) * map + c16 * 4;
                }
                for (int kc = 0; kc < NCH; ++kc) {
                    mbar_wait(&bars->empty[stage], phase ^ 1);
                    float4 v[4][4];
#pragma unroll
                    for (int it = 0; it < 4; ++it) {
                        const float* p = src[it] + kc * 32;
                        const bool on = cnt[it] > 0;
                        v[it][0] = on ? ldg_f4(p + off[it].x) : f4_zero();
                        v[it][1] = on ? ldg_f4(p + off[it].y) : f4_zero();
                        v[it][2] = on ? ldg_f4(p + off[it].z) : f4_zero();
                        v[it][3] = on ? ldg_f4(p + off[it].w) : f4_zero();
                    }
                    unsigned char* st_hi = sA + stage * kTcStageBytes;
#pragma unroll
                    for (int it = 0; it < 4; ++it) {
                        const int r = r0 + 32 * it;
                        float4 x = f4_scale(wg[it].x, v[it][0]);
                        f4_fma(x, wg[it].y, v[it][1]);
                        f4_fma(x, wg[it].z, v[it][2]);
                        f4_fma(x, wg[it].w, v[it][3]);
                        if (cnt[it] > 1) {       // a second camera of the same group sees the voxel (rare)
                            const int4 o2 = tm.off[1][r];
                            const float4 w2 = tm.wgt[1][r];
                            const float* p = a.feats + (size_t)tm.cam[1][r] * map + c16 * 4 + kc * 32;
                            float4 y = f4_scale(w2.x, ldg_f4(p + o2.x));
                            f4_fma(y, w2.y, ldg_f4(p + o2.y));
                            f4_fma(y, w2.z, ldg_f4(p + o2.z));
                            f4_fma(y, w2.w, ldg_f4(p + o2.w));
                            f4_add(x, y);
                        }
                        uint4 hi, lo;
                        split_tf32(x.x, hi.x, lo.x);
                        split_tf32(x.y, hi.y, lo.y);
                        split_tf32(x.z, hi.z, lo.z);
                        split_tf32(x.w, hi.w, lo.w);
                        const uint32_t o = sw128_off(r, c16);
                        *reinterpret_cast<uint4*>(st_hi + o) = hi;
                        *reinterpret_cast<uint4*>(st_hi + kTcRows * 128 + o) = lo;
                    }
                    fence_proxy_async_smem();
                    mbar_arrive(&bars->full[stage]);
                    if (++stage == kTcStages) { stage = 0; phase ^= 1; }
                }
            }
        } else if (warp == kTcEpiWarps) {
            // ===================== MMA issue (lane 0) =====================
            for (int tile = blockIdx.x; tile < T; tile += gridDim.x, ++ntile) {
                const uint32_t acc = ntile & 1;
                mbar_wait(&bars->acc_empty[acc], ((ntile >> 1) & 1) ^ 1);
                tc_fence_after();
                for (int kc = 0; kc < NCH; ++kc) {
                    mbar_wait(&bars->full[stage], phase);
                    tc_fence_after();
                    if (lane == 0) {
                        const uint64_t a_hi = umma_desc(smem_u32(sA + stage * kTcStageBytes));
                        const uint64_t a_lo = umma_desc(smem_u32(sA + stage * kTcStageBytes + kTcRows * 128));
                        const uint64_t b_hi = umma_desc(smem_u32(sW + (uint32_t)kc * Cv * 128));
                        const uint64_t b_lo = umma_desc(smem_u32(sW + wplane + (uint32_t)kc * Cv * 128));
                        const uint32_t dcol = tmem + acc * 64;
#pragma unroll
                        for (int k = 0; k < 4; ++k) {     // 8 TF32 = 32 bytes per K-step: +2 in the 16-byte address field
                            umma_tf32(dcol, a_lo + 2 * k, b_hi + 2 * k, idesc, (kc | k) ? 1u : 0u);
                            umma_tf32(dcol, a_hi + 2 * k, b_lo + 2 * k, idesc, 1u);
                            umma_tf32(dcol, a_hi + 2 * k, b_hi + 2 * k, idesc, 1u);
                        }
                        tc_commit(&bars->empty[stage]);                    // frees the A stage when these MMAs are done
                        if (kc == NCH - 1) tc_commit(&bars->acc_full[acc]);   // ... and publishes the accumulator
                    }
                    __syncwarp();
                    if (++stage == kTcStages) { stage = 0; phase ^= 1; }
                }
            }
        } else {
            // ===================== epilogue (TMEM lane = row of the tile) =====================
            for (int tile = blockIdx.x; tile < T; tile += gridDim.x, ++ntile) {
                const uint32_t acc = ntile & 1;
                const TcMeta& tm = meta[ntile & 1];
                mbar_wait(&bars->acc_full[acc], (ntile >> 1) & 1);
                tc_fence_after();
                const long long row = tm.row[tid];
                const float z = tm.zrel[tid];
                mbar_arrive(&bars->meta_free[ntile & 1]);
                float* o = a.vol + (size_t)(row >= 0 ? row : 0) * Cv;
                for (int c0 = 0; c0 < Cv; c0 += 16) {
                    uint32_t v[16];
                    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + acc * 64 + (uint32_t)c0;
                    asm volatile(
                        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                        : "r"(taddr));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (row >= 0) {
#pragma unroll
                        for (int q = 0; q < 16; q += 4) {
                            float4 y = make_float4(__uint_as_float(v[q]), __uint_as_float(v[q + 1]),
                                                   __uint_as_float(v[q + 2]), __uint_as_float(v[q + 3]));
                            const float4 wz = *reinterpret_cast<const float4*>(sWz + c0 + q);
                            f4_fma(y, z, wz);
                            if (pass == 2) f4_add(y, *reinterpret_cast<const float4*>(o + c0 + q));   // parked by pass 1
                            if (pass != 1) {
                                f4_add(y, *reinterpret_cast<const float4*>(sBias + c0 + q));
                                y.x = y.x > 0.f ? y.x : 0.1f * y.x;
                                y.y = y.y > 0.f ? y.y : 0.1f * y.y;
                                y.z = y.z > 0.f ? y.z : 0.1f * y.z;
                                y.w = y.w > 0.f ? y.w : 0.1f * y.w;
                            }
                            *reinterpret_cast<float4*>(o + c0 + q) = y;
                        }
                    }
                }
                tc_fence_before();
                mbar_arrive(&bars->acc_empty[acc]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kTcEpiWarps) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tmem));
}

// ---------------------------------------------------------------------------------------------------------
// Backward.  Per pass (weight block) and per tile of 64 voxels:
//   dZ  = d_vol[rows] * LeakyReLU'(vol[rows])       (y > 0 <=> pre-activation > 0), split into TF32 hi/lo planes
//   A   = the gathered feature rows, recomputed (L2-resident feature maps)  [64 x (C+1)]
//   dW += A^T dZ    tensor cores, accumulators stay in registers for the whole pass (one 16-row block per warp);
//                   the relative-depth column and db are two SIMT dot products
//   dA  = dZ W      tensor cores -> shared memory -> bilinear scatter into d_feats with red.global.add.v4.f32
// Per-CTA weight-gradient partials are written to the workspace and summed in CTA order by a second kernel, so dW / db
// are deterministic; d_feats is accumulated with L2 atomics like vfd_voxel_gather_depth_bwd.
// ---------------------------------------------------------------------------------------------------------
constexpr int kZS = 72;   // row stride of the dZ planes

__host__ __device__ inline int fuse_sn(int KP) { return (KP - 8 + 31) / 32 * 32 + 8; }   // stride = 8 (mod 32)

__host__ __device__ inline size_t fuse_bwd_smem_bytes(int KP, int NT) {
    const int SN = fuse_sn(KP);
    return meta_bytes(kFuseRows) + (size_t)NT * 8 * SN * 4 + (size_t)kFuseRows * SN * 4 + (size_t)2 * kFuseRows * kZS * 4;
}

template <int NCH>
__global__ void __launch_bounds__(kFuseThreads, 1) voxel_fuse_bwd_kernel(const FuseArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const vfd_fusion_desc& d = a.d;
    const int KP = a.KP, SN = fuse_sn(KP), NT = a.NT, C = d.C, C1 = C + 1, Cv = d.Cv;
    const TileMeta tm = carve_meta<kFuseRows>(smem_raw);
    const Group G = {(int)threadIdx.x, kFuseThreads, 0};
    float* sW = reinterpret_cast<float*>(smem_raw + meta_bytes(kFuseRows));          // [NT*8][SN] natural layout [out][in]
    float* sA = sW + (size_t)NT * 8 * SN;                                 // [64][SN]: A, then dA
    uint32_t* Zhi = reinterpret_cast<uint32_t*>(sA + (size_t)kFuseRows * SN);   // [64][kZS]
    uint32_t* Zlo = Zhi + kFuseRows * kZS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int wm = warp & 3, wn = warp >> 2;
    const int M1 = a.totals[0], M2 = a.totals[1];
    const int T1 = (M1 + kFuseRows - 1) / kFuseRows, T2 = (M2 + kFuseRows - 1) / kFuseRows;
    unsigned gmask[2] = {0u, 0u};
    for (int c = 0; c < d.num_cams; ++c) gmask[d.group[c] ? 1 : 0] |= 1u << c;
    const int MT = (C + 15) / 16;          // 16-row blocks of dW^T handled by the tensor cores (<= 16: one per warp)
    const int NA = (C + 7) / 8;            // 8-column blocks of dA
    const int NAW = (NA + 3) / 4;          // ... per warp column group (<= 8)
    const size_t pstride = (size_t)Cv * (C1 + 1);     // one (CTA, pass) partial: [Cv][C1] weights then [Cv] bias
    const size_t map = (size_t)d.h * d.w * C;
    const size_t rows_total = (size_t)d.B * d.X * d.Y * d.Z;
    const CornerTable tb = carve_table(const_cast<void*>(a.table), rows_total);

    for (int pass = 0; pass < 3; ++pass) {
        const int M = pass ? M2 : M1, T = pass ? T2 : T1;
        const int64_t* idx = pass ? a.idx_ov : a.idx_no;
        const unsigned want = pass ? gmask[pass - 1] : 0xffffffffu;
        __syncthreads();
        {   // natural-layout weight block [out][in], zero padded
            const float* w = pass ? a.w_ov : a.w_no;
            const int ld = pass ? 2 * C1 : C1, koff = pass == 2 ? C1 : 0;
            for (int i = threadIdx.x; i < NT * 8 * SN; i += blockDim.x) {
                const int o = i / SN, k = i - o * SN;
                sW[i] = (o < Cv && k < C1) ? __ldg(w + (size_t)o * ld + koff + k) : 0.f;
            }
        }
        float accw[8][4];
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) accw[j][i] = 0.f;
        float db = 0.f, dwz = 0.f;           // threads o < Cv: bias gradient and the relative-depth column of dW

        int tile = blockIdx.x;
        TilePf pf = prefetch_tile<kFuseRows>(a, tb, rows_total, G, idx, tile * kFuseRows, M, want, tile < T);
        for (; tile < T; tile += gridDim.x) {
            commit_tile<kFuseRows>(tm, G, pf, want, d.num_cams);
            pf = prefetch_tile<kFuseRows>(a, tb, rows_total, G, idx, (tile + (int)gridDim.x) * kFuseRows, M, want,
                                          tile + (int)gridDim.x < T);
            __syncthreads();
            // dZ tile, split once
            for (int i = threadIdx.x; i < kFuseRows * (kZS / 4); i += blockDim.x) {
                const int r = i / (kZS / 4), q = (i - r * (kZS / 4)) * 4;
                const long long row = tm.row[r];
                float4 z = f4_zero();
                if (row >= 0 && q < Cv) {
                    const float4 gy = ld_stream_f4(a.d_vol + (size_t)row * Cv + q);
                    const float4 y = ld_stream_f4(a.vol + (size_t)row * Cv + q);
                    z.x = y.x > 0.f ? gy.x : 0.1f * gy.x;
                    z.y = y.y > 0.f ? gy.y : 0.1f * gy.y;
                    z.z = y.z > 0.f ? gy.z : 0.1f * gy.z;
                    z.w = y.w > 0.f ? gy.w : 0.1f * gy.w;
                }
                uint4 hi, lo;
                split_tf32(z.x, hi.x, lo.x);
                split_tf32(z.y, hi.y, lo.y);
                split_tf32(z.z, hi.z, lo.z);
                split_tf32(z.w, hi.w, lo.w);
                *reinterpret_cast<uint4*>(Zhi + r * kZS + q) = hi;
                *reinterpret_cast<uint4*>(Zlo + r * kZS + q) = lo;
            }
            // recompute the gathered rows
#pragma unroll
            for (int rr = 0; rr < kRowsPerWarp; rr += 2)
                gather_rows<NCH, kFuseRows, 2>(a, tm, warp * kRowsPerWarp + rr, lane,
                                               sA + (size_t)(warp * kRowsPerWarp + rr) * SN, SN);
            __syncthreads();

            // dW^T block [16 in x Cv out] += A^T [16 x 64] dZ [64 x Cv]
            if (warp < MT) {
                const float* At = sA + t * SN + warp * 16 + g;
                const uint32_t* Bh = Zhi + t * kZS + g;
                const uint32_t* Bl = Zlo + t * kZS + g;
#pragma unroll 2
                for (int ks = 0; ks < kFuseRows; ks += 8) {
                    uint32_t ah[4], al[4];
                    split_tf32(At[ks * SN], ah[0], al[0]);
                    split_tf32(At[ks * SN + 8], ah[1], al[1]);
                    split_tf32(At[(ks + 4) * SN], ah[2], al[2]);
                    split_tf32(At[(ks + 4) * SN + 8], ah[3], al[3]);
#pragma unroll
                    for (int jg = 0; jg < 8; jg += 4) {
                        if (jg < NT) {
                            uint32_t bh0[4], bh1[4], bl0[4], bl1[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                bh0[j] = Bh[ks * kZS + (jg + j) * 8];
                                bh1[j] = Bh[(ks + 4) * kZS + (jg + j) * 8];
                                bl0[j] = Bl[ks * kZS + (jg + j) * 8];
                                bl1[j] = Bl[(ks + 4) * kZS + (jg + j) * 8];
                            }
#pragma unroll
                            for (int j = 0; j < 4; ++j) mma_tf32(accw[jg + j], al, bh0[j], bh1[j]);
#pragma unroll
                            for (int j = 0; j < 4; ++j) mma_tf32(accw[jg + j], ah, bl0[j], bl1[j]);
#pragma unroll
                            for (int j = 0; j < 4; ++j) mma_tf32(accw[jg + j], ah, bh0[j], bh1[j]);
                        }
                    }
                }
            }
            if (threadIdx.x < Cv) {
                const int o = threadIdx.x;
                float sb = 0.f, sz = 0.f;
                for (int r = 0; r < kFuseRows; ++r) {
                    const float z = __uint_as_float(Zhi[r * kZS + o]) + __uint_as_float(Zlo[r * kZS + o]);
                    sb += z;
                    sz = fmaf(z, sA[r * SN + C], sz);
                }
                db += sb;
                dwz += sz;
            }

            // dA [64 x C] = dZ [64 x Cv] W [Cv x C]
            float acca[8][4];
#pragma unroll
            for (int j = 0; j < 8; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) acca[j][i] = 0.f;
            {
                const uint32_t* Z0h = Zhi + (wm * 16 + g) * kZS + t;
                const uint32_t* Z0l = Zlo + (wm * 16 + g) * kZS + t;
                const float* Wb = sW + t * SN + wn * NAW * 8 + g;
                for (int ks = 0; ks < NT * 8; ks += 8) {
                    uint32_t ah[4], al[4];
                    ah[0] = Z0h[ks];            al[0] = Z0l[ks];
                    ah[1] = Z0h[ks + 8 * kZS];  al[1] = Z0l[ks + 8 * kZS];
                    ah[2] = Z0h[ks + 4];        al[2] = Z0l[ks + 4];
                    ah[3] = Z0h[ks + 4 + 8 * kZS];  al[3] = Z0l[ks + 4 + 8 * kZS];
#pragma unroll
                    for (int jg = 0; jg < 8; jg += 4) {
                        if (jg < NAW) {      // columns past NA read zero-padded / unused weight columns: harmless
                            uint32_t bh0[4], bh1[4], bl0[4], bl1[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                split_tf32(Wb[ks * SN + (jg + j) * 8], bh0[j], bl0[j]);
                                split_tf32(Wb[(ks + 4) * SN + (jg + j) * 8], bh1[j], bl1[j]);
                            }
#pragma unroll
                            for (int j = 0; j < 4; ++j) mma_tf32(acca[jg + j], al, bh0[j], bh1[j]);
#pragma unroll
                            for (int j = 0; j < 4; ++j) mma_tf32(acca[jg + j], ah, bl0[j], bl1[j]);
#pragma unroll
                            for (int j = 0; j < 4; ++j) mma_tf32(acca[jg + j], ah, bh0[j], bh1[j]);
                        }
                    }
                }
            }
            __syncthreads();          // every warp is done reading A: overwrite it with dA
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (j < NAW && wn * NAW + j < NA) {
                    const int col = (wn * NAW + j) * 8 + 2 * t;
                    float* r0 = sA + (size_t)(wm * 16 + g) * SN + col;
                    *reinterpret_cast<float2*>(r0) = make_float2(acca[j][0], acca[j][1]);
                    *reinterpret_cast<float2*>(r0 + 8 * SN) = make_float2(acca[j][2], acca[j][3]);
                }
            }
            __syncthreads();
            // bilinear scatter of the dA rows
#pragma unroll
            for (int rr = 0; rr < kRowsPerWarp; ++rr) {
                const int r = warp * kRowsPerWarp + rr;
                const int cnt = tm.cnt[r];
                for (int e = 0; e < cnt; ++e) {
                    const int4 off = tm.off[e * kFuseRows + r];
                    const float4 wg = tm.wgt[e * kFuseRows + r];
                    float* fb = a.d_feats + (size_t)tm.meta[e * kFuseRows + r].x * map + lane * 4;
#pragma unroll
                    for (int i = 0; i < NCH; ++i) {
                        if (i * 128 + lane * 4 < C) {
                            const float4 gv = *reinterpret_cast<const float4*>(sA + (size_t)r * SN + i * 128 + lane * 4);
                            if (wg.x != 0.f) red_add_f4(fb + off.x + i * 128, f4_scale(wg.x, gv));
                            if (wg.y != 0.f) red_add_f4(fb + off.y + i * 128, f4_scale(wg.y, gv));
                            if (wg.z != 0.f) red_add_f4(fb + off.z + i * 128, f4_scale(wg.z, gv));
                            if (wg.w != 0.f) red_add_f4(fb + off.w + i * 128, f4_scale(wg.w, gv));
                        }
                    }
                }
            }
            __syncthreads();
        }

        // flush this pass' weight-gradient partials: partial[cta][pass][out][in], then [out] bias
        float* P = a.partial + ((size_t)blockIdx.x * 3 + pass) * pstride;
        if (warp < MT) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (j < NT) {
                    const int in0 = warp * 16 + g, o0 = j * 8 + 2 * t;
                    if (o0 < Cv) {
                        if (in0 < C) { P[(size_t)o0 * C1 + in0] = accw[j][0]; P[(size_t)(o0 + 1) * C1 + in0] = accw[j][1]; }
                        if (in0 + 8 < C) { P[(size_t)o0 * C1 + in0 + 8] = accw[j][2]; P[(size_t)(o0 + 1) * C1 + in0 + 8] = accw[j][3]; }
                    }
                }
            }
        }
        if (threadIdx.x < Cv) {
            P[(size_t)threadIdx.x * C1 + C] = dwz;
            P[(size_t)Cv * C1 + threadIdx.x] = db;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Backward, data half, on tcgen05: dA [128 rows x C] = dZ [128 x Cv] W [Cv x C] per tile (3xTF32, one N = C instruction
// per K-step, accumulator = C TMEM columns, double buffered), then the bilinear scatter of the dA rows into d_feats.
//   warps 0-3   drain: tcgen05.ld 32 columns at a time -> staging tile in shared memory -> 8 lanes per row push
//               128-byte contiguous red.global.add.v4.f32 to the row's 4 corners (x entries)
//   warp  4     MMA issue
//   warps 5-12  producers: dZ = d_vol * LeakyReLU'(vol) of the tile's rows, split hi / lo, K-major SWIZZLE_128B;
//               tile metadata (corner table) prefetched in registers and published once the previous tile is drained
// W^T of the pass ([in][out], hi / lo planes, 128 KB at C=256, Cv=64) stays in shared memory.
// ---------------------------------------------------------------------------------------------------------
struct TcBwdBars {
    uint64_t full, empty, acc_full[2], acc_empty[2], meta_free;
    uint32_t tmem_base;
};
constexpr int kStageCols = 32;     // dA columns drained per round

__host__ __device__ inline size_t fuse_tc_bwd_data_smem_bytes(int C, int Cv) {
    const int nkc = (Cv + 31) / 32;
    return 1024 + (size_t)2 * nkc * C * 128 + (size_t)2 * nkc * kTcRows * 128 + (size_t)kTcRows * kStageCols * 4 +
           sizeof(TcMeta) + sizeof(TcBwdBars);
}

__global__ void __launch_bounds__(kTcThreads, 1) voxel_fuse_bwd_data_tc_kernel(const FuseArgs a) {
    extern __shared__ unsigned char smem_raw[];
    const vfd_fusion_desc& d = a.d;
    const int C = d.C, C1 = C + 1, Cv = d.Cv, NKC = (Cv + 31) / 32, KS = Cv / 8;
    unsigned char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t wplane = (uint32_t)NKC * C * 128;           // W^T plane: [NKC][C rows (in)][128 B (32 out)]
    const uint32_t zplane = (uint32_t)NKC * kTcRows * 128;     // dZ plane:  [NKC][128 rows][128 B]
    unsigned char* sW = base;
    unsigned char* sZ = sW + 2 * wplane;
    float* sStage = reinterpret_cast<float*>(sZ + 2 * zplane);                 // [128][32]
    TcMeta* tm = reinterpret_cast<TcMeta*>(sStage + kTcRows * kStageCols);
    TcBwdBars* bars = reinterpret_cast<TcBwdBars*>(tm + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int M1 = a.totals[0], M2 = a.totals[1];
    const int T1 = (M1 + kTcRows - 1) / kTcRows, T2 = (M2 + kTcRows - 1) / kTcRows;
    unsigned gmask[2] = {0u, 0u};
    for (int c = 0; c < d.num_cams; ++c) gmask[d.group[c] ? 1 : 0] |= 1u << c;
    const size_t rows_total = (size_t)d.B * d.X * d.Y * d.Z;
    const CornerTable tb = carve_table(const_cast<void*>(a.table), rows_total);
    const size_t map = (size_t)d.h * d.w * C;

    if (tid == 0) {
        mbar_init(&bars->full, kTcProdThreads);
        mbar_init(&bars->empty, 1);
        mbar_init(&bars->meta_free, kTcEpiWarps * VFD_WARP);
        for (int i = 0; i < 2; ++i) { mbar_init(&bars->acc_full[i], 1); mbar_init(&bars->acc_empty[i], kTcEpiWarps * VFD_WARP); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kTcEpiWarps) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&bars->tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = bars->tmem_base;
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(C >> 3) << 17) | ((uint32_t)(kTcRows >> 4) << 24);
    uint32_t ntile = 0;

    for (int pass = 0; pass < 3; ++pass) {
        const int M = pass ? M2 : M1, T = pass ? T2 : T1;
        const int64_t* idx = pass ? a.idx_ov : a.idx_no;
        const unsigned want = pass ? gmask[pass - 1] : 0xffffffffu;
        __syncthreads();
        {   // W^T planes: element (in, out) = w[out][koff + in]
            const float* w = pass ? a.w_ov : a.w_no;
            const int ld = pass ? 2 * C1 : C1, koff = pass == 2 ? C1 : 0;
            for (int i = tid; i < Cv * C; i += kTcThreads) {
                const int o = i / C, k = i - o * C;          // coalesced read along in
                uint32_t hi, lo;
                split_tf32(__ldg(w + (size_t)o * ld + koff + k), hi, lo);
                const uint32_t off = (uint32_t)(o >> 5) * C * 128 + sw128_off(k, (o & 31) >> 2) + (o & 3) * 4;
                *reinterpret_cast<uint32_t*>(sW + off) = hi;
                *reinterpret_cast<uint32_t*>(sW + wplane + off) = lo;
            }
        }
        fence_proxy_async_smem();
        __syncthreads();

        if (warp > kTcEpiWarps) {
            // ===================== producers =====================
            const int ptid = tid - (kTcEpiWarps + 1) * VFD_WARP;
            const int nq = Cv / 4;                                   // float4 per row
            for (int tile = blockIdx.x; tile < T; tile += gridDim.x, ++ntile) {
                // this thread's share of the metadata, from the corner table (published below)
                const int r = ptid & (kTcRows - 1), half = ptid >> 7;
                const int li = tile * kTcRows + r;
                long long row = -1;
                int2 m0 = make_int2(-1, 0), m1 = make_int2(-1, 0);
                if (li < M) {
                    row = idx[li];
                    m0 = __ldg(tb.meta + row);
                    m1 = __ldg(tb.meta + rows_total + row);
                }
                const bool w0 = entry_wanted(m0, want, d.num_cams), w1 = entry_wanted(m1, want, d.num_cams);
                int4 e_off = make_int4(0, 0, 0, 0);
                float4 e_wgt = f4_zero();
                if (half == 0 ? w0 : w1) {
                    e_off = __ldg(tb.off + (half ? rows_total : 0) + row);
                    e_wgt = __ldg(tb.wgt + (half ? rows_total : 0) + row);
                }
                // dZ of the tile into the (single) operand stage
                mbar_wait(&bars->empty, (ntile & 1) ^ 1);
                for (int item = ptid; item < kTcRows * nq; item += kTcProdThreads) {
                    const int rr = item / nq, q4 = item - rr * nq;
                    const int l2 = tile * kTcRows + rr;
                    float4 z = f4_zero();
                    if (l2 < M) {
                        const long long rw = idx[l2];
                        const float4 gy = ld_stream_f4(a.d_vol + (size_t)rw * Cv + q4 * 4);
                        const float4 y = ld_stream_f4(a.vol + (size_t)rw * Cv + q4 * 4);
                        z.x = y.x > 0.f ? gy.x : 0.1f * gy.x;
                        z.y = y.y > 0.f ? gy.y : 0.1f * gy.y;
                        z.z = y.z > 0.f ? gy.z : 0.1f * gy.z;
                        z.w = y.w > 0.f ? gy.w : 0.1f * gy.w;
                    }
                    uint4 hi, lo;
                    split_tf32(z.x, hi.x, lo.x);
                    split_tf32(z.y, hi.y, lo.y);
                    split_tf32(z.z, hi.z, lo.z);
                    split_tf32(z.w, hi.w, lo.w);
                    const uint32_t o = (uint32_t)(q4 >> 3) * kTcRows * 128 + sw128_off(rr, q4 & 7);
                    *reinterpret_cast<uint4*>(sZ + o) = hi;
                    *reinterpret_cast<uint4*>(sZ + zplane + o) = lo;
                }
                // publish the metadata once the previous tile has been drained
                mbar_wait(&bars->meta_free, (ntile & 1) ^ 1);
                if (half == 0) {
                    if (w0) { tm->off[0][r] = e_off; tm->wgt[0][r] = e_wgt; tm->cam[0][r] = m0.x; }
                    tm->row[r] = row;
                    tm->cnt[r] = (w0 ? 1 : 0) + (w1 ? 1 : 0);
                } else if (w1) {
                    const int dst = w0 ? 1 : 0;
                    tm->off[dst][r] = e_off; tm->wgt[dst][r] = e_wgt; tm->cam[dst][r] = m1.x;
                }
                fence_proxy_async_smem();
                mbar_arrive(&bars->full);
            }
        } else if (warp == kTcEpiWarps) {
            // ===================== MMA issue =====================
            for (int tile = blockIdx.x; tile < T; tile += gridDim.x, ++ntile) {
                const uint32_t acc = ntile & 1;
                mbar_wait(&bars->acc_empty[acc], ((ntile >> 1) & 1) ^ 1);
                mbar_wait(&bars->full, ntile & 1);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t dcol = tmem + acc * 256;
                    for (int ks = 0; ks < KS; ++ks) {
                        const uint32_t kc = ks >> 2, kk = 2 * (ks & 3);
                        const uint64_t a_hi = umma_desc(smem_u32(sZ + kc * kTcRows * 128)) + kk;
                        const uint64_t a_lo = umma_desc(smem_u32(sZ + zplane + kc * kTcRows * 128)) + kk;
                        const uint64_t b_hi = umma_desc(smem_u32(sW + kc * C * 128)) + kk;
                        const uint64_t b_lo = umma_desc(smem_u32(sW + wplane + kc * C * 128)) + kk;
                        umma_tf32(dcol, a_lo, b_hi, idesc, ks ? 1u : 0u);
                        umma_tf32(dcol, a_hi, b_lo, idesc, 1u);
                        umma_tf32(dcol, a_hi, b_hi, idesc, 1u);
                    }
                    tc_commit(&bars->empty);
                    tc_commit(&bars->acc_full[acc]);
                }
                __syncwarp();
            }
        } else {
            // ===================== drain + scatter =====================
            for (int tile = blockIdx.x; tile < T; tile += gridDim.x, ++ntile) {
                const uint32_t acc = ntile & 1;
                mbar_wait(&bars->acc_full[acc], (ntile >> 1) & 1);
                tc_fence_after();
                for (int c0 = 0; c0 < C; c0 += kStageCols) {
                    uint32_t v[32];
                    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + acc * 256 + (uint32_t)c0;
                    asm volatile(
                        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
                        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                        : "r"(taddr));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    // staging: row tid, 32 columns; 16-byte chunks rotated by the row so that the column reads below
                    // and these row writes are both conflict-free
                    float* srow = sStage + tid * kStageCols;
#pragma unroll
                    for (int q = 0; q < 8; ++q)
                        *reinterpret_cast<uint4*>(srow + (((q + tid) & 7) << 2)) = make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                    __syncwarp();      // each warp drains the 32 rows it staged itself
                    const int sub = lane >> 3, l8 = lane & 7;         // 4 rows per instruction, 8 lanes x 16 B per row
                    for (int rb = 0; rb < 32; rb += 4) {
                        const int r = warp * 32 + rb + sub;
                        const int cnt = tm->cnt[r];
                        const float4 gv = *reinterpret_cast<const float4*>(sStage + r * kStageCols + (((l8 + r) & 7) << 2));
                        for (int e = 0; e < cnt; ++e) {
                            const int4 off = tm->off[e][r];
                            const float4 wg = tm->wgt[e][r];
                            float* fb = a.d_feats + (size_t)tm->cam[e][r] * map + c0 + l8 * 4;
                            if (wg.x != 0.f) red_add_f4(fb + off.x, f4_scale(wg.x, gv));
                            if (wg.y != 0.f) red_add_f4(fb + off.y, f4_scale(wg.y, gv));
                            if (wg.z != 0.f) red_add_f4(fb + off.z, f4_scale(wg.z, gv));
                            if (wg.w != 0.f) red_add_f4(fb + off.w, f4_scale(wg.w, gv));
                        }
                    }
                    __syncwarp();
                }
                tc_fence_before();
                mbar_arrive(&bars->acc_empty[acc]);
                mbar_arrive(&bars->meta_free);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kTcEpiWarps) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

// ---------------------------------------------------------------------------------------------------------
// Backward, weight half (companion of voxel_fuse_bwd_data_tc_kernel):  dW^T += A^T dZ, db += sum dZ over the rows of
// every pass.  Warp specialised so that the L2 gather and the tensor pipe overlap instead of alternating:
//   warps 0-7   producers: tile metadata from the corner table, dZ tile split into TF32 hi / lo planes, bilinear
//               gather of the tile's 32 feature rows -> one of kWtStages shared-memory stages
//   warps 8-15  consumers: each owns two 16-channel blocks of dW^T (accumulators in registers for the whole pass),
//               3xTF32 mma.sync over the stage; the relative-depth column and db are SIMT partial sums
// full / empty mbarriers per stage; the producers run ahead across pass boundaries.  Partials per CTA as in the
// single-kernel backward (summed in CTA order by voxel_fuse_reduce_kernel).
// ---------------------------------------------------------------------------------------------------------
constexpr int kWtRows = 32, kWtStages = 3, kWtHalf = kFuseThreads / 2;

__host__ __device__ inline size_t fuse_wt_stage_bytes(int KP) {
    return meta_bytes(kWtRows) + (size_t)kWtRows * fuse_sn(KP) * 4 + (size_t)2 * kWtRows * kZS * 4;
}
__host__ __device__ inline size_t fuse_wt_smem_bytes(int KP) { return kWtStages * fuse_wt_stage_bytes(KP); }

template <int NCH>
__global__ void __launch_bounds__(kFuseThreads, 1) voxel_fuse_bwd_weight_kernel(const FuseArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint64_t s_full[kWtStages], s_empty[kWtStages];
    __shared__ float s_part[2][4][64];
    const vfd_fusion_desc& d = a.d;
    const int KP = a.KP, SN = fuse_sn(KP), NT = a.NT, C = d.C, C1 = C + 1, Cv = d.Cv;
    const size_t stage_bytes = fuse_wt_stage_bytes(KP);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int M1 = a.totals[0], M2 = a.totals[1];
    const int T1 = (M1 + kWtRows - 1) / kWtRows, T2 = (M2 + kWtRows - 1) / kWtRows;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kWtStages; ++s) { mbar_init(&s_full[s], kWtHalf); mbar_init(&s_empty[s], kWtHalf); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t ntile = 0;

    if (warp < kFuseWarps / 2) {
        // ===================== producers =====================
        const Group G = {(int)threadIdx.x, kWtHalf, 1};
        unsigned gmask[2] = {0u, 0u};
        for (int c = 0; c < d.num_cams; ++c) gmask[d.group[c] ? 1 : 0] |= 1u << c;
        const size_t rows_total = (size_t)d.B * d.X * d.Y * d.Z;
        const CornerTable tb = carve_table(const_cast<void*>(a.table), rows_total);
        for (int pass = 0; pass < 3; ++pass) {
            const int M = pass ? M2 : M1, T = pass ? T2 : T1;
            const int64_t* idx = pass ? a.idx_ov : a.idx_no;
            const unsigned want = pass ? gmask[pass - 1] : 0xffffffffu;
            int tile = blockIdx.x;
            TilePf pf = prefetch_tile<kWtRows>(a, tb, rows_total, G, idx, tile * kWtRows, M, want, tile < T);
            for (; tile < T; tile += gridDim.x, ++ntile) {
                const uint32_t s = ntile % kWtStages, ph = (ntile / kWtStages) & 1;
                unsigned char* st = smem_raw + s * stage_bytes;
                const TileMeta tm = carve_meta<kWtRows>(st);
                float* sA = reinterpret_cast<float*>(st + meta_bytes(kWtRows));
                uint32_t* Zhi = reinterpret_cast<uint32_t*>(sA + (size_t)kWtRows * SN);
                uint32_t* Zlo = Zhi + kWtRows * kZS;
                mbar_wait(&s_empty[s], ph ^ 1);
                commit_tile<kWtRows>(tm, G, pf, want, d.num_cams);
                pf = prefetch_tile<kWtRows>(a, tb, rows_total, G, idx, (tile + (int)gridDim.x) * kWtRows, M, want,
                                            tile + (int)gridDim.x < T);
                group_sync(G);
                for (int i = G.tid; i < kWtRows * (kZS / 4); i += kWtHalf) {
                    const int r = i / (kZS / 4), q = (i - r * (kZS / 4)) * 4;
                    const long long row = tm.row[r];
                    float4 z = f4_zero();
                    if (row >= 0 && q < Cv) {
                        const float4 gy = ld_stream_f4(a.d_vol + (size_t)row * Cv + q);
                        const float4 y = ld_stream_f4(a.vol + (size_t)row * Cv + q);
                        z.x = y.x > 0.f ? gy.x : 0.1f * gy.x;
                        z.y = y.y > 0.f ? gy.y : 0.1f * gy.y;
                        z.z = y.z > 0.f ? gy.z : 0.1f * gy.z;
                        z.w = y.w > 0.f ? gy.w : 0.1f * gy.w;
                    }
                    uint4 hi, lo;
                    split_tf32(z.x, hi.x, lo.x);
                    split_tf32(z.y, hi.y, lo.y);
                    split_tf32(z.z, hi.z, lo.z);
                    split_tf32(z.w, hi.w, lo.w);
                    *reinterpret_cast<uint4*>(Zhi + r * kZS + q) = hi;
                    *reinterpret_cast<uint4*>(Zlo + r * kZS + q) = lo;
                }
                constexpr int kPer = kWtRows / (kFuseWarps / 2);     // rows per producer warp
#pragma unroll
                for (int rr = 0; rr < kPer; rr += 2)
                    gather_rows<NCH, kWtRows, 2>(a, tm, warp * kPer + rr, lane, sA + (size_t)(warp * kPer + rr) * SN, SN);
                mbar_arrive(&s_full[s]);
            }
        }
    } else {
        // ===================== consumers =====================
        const int cw = warp - kFuseWarps / 2, ctid = threadIdx.x - kWtHalf, g = lane >> 2, t = lane & 3;
        const int MT = (C + 15) / 16;
        const int so = ctid & 63, sp = ctid >> 6;                  // SIMT sums: output so, rows sp*8 .. sp*8+7 of the tile
        const size_t pstride = (size_t)Cv * (C1 + 1);
        for (int pass = 0; pass < 3; ++pass) {
            const int T = pass ? T2 : T1;
            float accw[2][8][4];
#pragma unroll
            for (int b = 0; b < 2; ++b)
#pragma unroll
                for (int j = 0; j < 8; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) accw[b][j][i] = 0.f;
            float db = 0.f, dwz = 0.f;
            for (int tile = blockIdx.x; tile < T; tile += gridDim.x, ++ntile) {
                const uint32_t s = ntile % kWtStages, ph = (ntile / kWtStages) & 1;
                unsigned char* st = smem_raw + s * stage_bytes;
                const float* sA = reinterpret_cast<const float*>(st + meta_bytes(kWtRows));
                const uint32_t* Zhi = reinterpret_cast<const uint32_t*>(sA + (size_t)kWtRows * SN);
                const uint32_t* Zlo = Zhi + kWtRows * kZS;
                mbar_wait(&s_full[s], ph);
                if (cw * 2 < MT) {
                    const bool two = cw * 2 + 1 < MT;
                    const float* At = sA + t * SN + cw * 32 + g;
                    const uint32_t* Bh = Zhi + t * kZS + g;
                    const uint32_t* Bl = Zlo + t * kZS + g;
#pragma unroll 1
                    for (int ks = 0; ks < kWtRows; ks += 8) {
                        uint32_t ah[2][4], al[2][4];
#pragma unroll
                        for (int b = 0; b < 2; ++b) {
                            if (b == 0 || two) {
                                split_tf32(At[ks * SN + b * 16], ah[b][0], al[b][0]);
                                split_tf32(At[ks * SN + b * 16 + 8], ah[b][1], al[b][1]);
                                split_tf32(At[(ks + 4) * SN + b * 16], ah[b][2], al[b][2]);
                                split_tf32(At[(ks + 4) * SN + b * 16 + 8], ah[b][3], al[b][3]);
                            }
                        }
#pragma unroll
                        for (int jg = 0; jg < 8; jg += 4) {
                            if (jg < NT) {
                                uint32_t bh0[4], bh1[4], bl0[4], bl1[4];
#pragma unroll
                                for (int j = 0; j < 4; ++j) {
                                    bh0[j] = Bh[ks * kZS + (jg + j) * 8];
                                    bh1[j] = Bh[(ks + 4) * kZS + (jg + j) * 8];
                                    bl0[j] = Bl[ks * kZS + (jg + j) * 8];
                                    bl1[j] = Bl[(ks + 4) * kZS + (jg + j) * 8];
                                }
#pragma unroll
                                for (int b = 0; b < 2; ++b) {
                                    if (b == 0 || two) {
#pragma unroll
                                        for (int j = 0; j < 4; ++j) mma_tf32(accw[b][jg + j], al[b], bh0[j], bh1[j]);
#pragma unroll
                                        for (int j = 0; j < 4; ++j) mma_tf32(accw[b][jg + j], ah[b], bl0[j], bl1[j]);
#pragma unroll
                                        for (int j = 0; j < 4; ++j) mma_tf32(accw[b][jg + j], ah[b], bh0[j], bh1[j]);
                                    }
                                }
                            }
                        }
                    }
                }
                if (so < Cv) {
#pragma unroll
                    for (int r = sp * 8; r < sp * 8 + 8; ++r) {
                        const float z = __uint_as_float(Zhi[r * kZS + so]) + __uint_as_float(Zlo[r * kZS + so]);
                        db += z;
                        dwz = fmaf(z, sA[r * SN + C], dwz);
                    }
                }
                mbar_arrive(&s_empty[s]);
            }
            // flush this pass' partials: partial[cta][pass][out][in], then [out] bias
            float* P = a.partial + ((size_t)blockIdx.x * 3 + pass) * pstride;
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                if (cw * 2 + b < MT) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (j < NT) {
                            const int in0 = (cw * 2 + b) * 16 + g, o0 = j * 8 + 2 * t;
                            if (o0 < Cv) {
                                if (in0 < C) { P[(size_t)o0 * C1 + in0] = accw[b][j][0]; P[(size_t)(o0 + 1) * C1 + in0] = accw[b][j][1]; }
                                if (in0 + 8 < C) { P[(size_t)o0 * C1 + in0 + 8] = accw[b][j][2]; P[(size_t)(o0 + 1) * C1 + in0 + 8] = accw[b][j][3]; }
                            }
                        }
                    }
                }
            }
            s_part[0][sp][so] = db;
            s_part[1][sp][so] = dwz;
            asm volatile("bar.sync 2, %0;" ::"n"(kWtHalf) : "memory");
            if (sp == 0 && so < Cv) {
                P[(size_t)so * C1 + C] = (s_part[1][0][so] + s_part[1][1][so]) + (s_part[1][2][so] + s_part[1][3][so]);
                P[(size_t)Cv * C1 + so] = (s_part[0][0][so] + s_part[0][1][so]) + (s_part[0][2][so] + s_part[0][3][so]);
            }
            asm volatile("bar.sync 2, %0;" ::"n"(kWtHalf) : "memory");
        }
    }
}

// sum the per-CTA partials in CTA order -> d_w_no [Cv][C1], d_b_no, d_w_ov [Cv][2 C1], d_b_ov
__global__ void voxel_fuse_reduce_kernel(const float* partial, int ctas, int Cv, int C1, float* d_w_no, float* d_b_no,
                                         float* d_w_ov, float* d_b_ov) {
    const size_t pstride = (size_t)Cv * (C1 + 1);
    const int per = Cv * (C1 + 1);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 3 * per) return;
    const int pass = i / per, e = i - pass * per;
    float acc = 0.f;
    for (int c = 0; c < ctas; ++c) acc += partial[((size_t)c * 3 + pass) * pstride + e];
    if (e < Cv * C1) {
        const int o = e / C1, k = e - o * C1;
        if (pass == 0) d_w_no[(size_t)o * C1 + k] = acc;
        else d_w_ov[(size_t)o * 2 * C1 + (pass - 1) * C1 + k] = acc;
    } else {
        const int o = e - Cv * C1;
        if (pass == 0) d_b_no[o] = acc;
        if (pass == 1) d_b_ov[o] = acc;
    }
}

static int check_fuse(const vfd_fusion_desc* d) {
    if (!d) return VFD_ERR_NULL;
    if (d->B < 1 || d->num_cams < 1 || d->num_cams > VFD_MAX_CAMS || d->C < 4 || d->h < 2 || d->w < 2 || d->X < 1 ||
        d->Y < 1 || d->Z < 1 || d->Cv < 2)
        return VFD_ERR_DESC;
    if (d->C % 4 || d->C > 256 || d->Cv > 64 || d->Cv % 4) return VFD_ERR_UNSUPPORTED;
    if (!d->axis_x || !d->axis_y || !d->axis_z) return VFD_ERR_NULL;
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

template <typename Kern>
static int launch_fuse(Kern kern, const FuseArgs& a, size_t smem, int sms, cudaStream_t st) {
    if (smem > 227 * 1024) return VFD_ERR_UNSUPPORTED;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return VFD_ERR_CUDA;
    kern<<<sms, kFuseThreads, smem, st>>>(a);
    return VFD_OK;
}

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_voxel_fuse_fwd(const vfd_fusion_desc* d, const float* feats, const float* mask_small,
                                  const float* K, const float* Einv, const int64_t* idx_no, const int64_t* idx_ov,
                                  const int32_t* totals, const void* table, const float* w_no, const float* b_no,
                                  const float* w_ov, const float* b_ov, float* vol, void* stream) {
    int rc = check_fuse(d);
    if (rc) return rc;
    if (!feats || !mask_small || !K || !Einv || !idx_no || !idx_ov || !totals || !table || !w_no || !b_no || !w_ov ||
        !b_ov || !vol)
        return VFD_ERR_NULL;
    if ((reinterpret_cast<uintptr_t>(feats) & 15) || (reinterpret_cast<uintptr_t>(vol) & 15)) return VFD_ERR_ALIGN;
    FuseArgs a = {};
    a.d = *d; a.feats = feats; a.mask_small = mask_small; a.K = K; a.Einv = Einv;
    a.idx_no = idx_no; a.idx_ov = idx_ov; a.totals = totals;
    a.w_no = w_no; a.b_no = b_no; a.w_ov = w_ov; a.b_ov = b_ov; a.vol = vol; a.table = table;
    a.KP = (d->C + 1 + 7) / 8 * 8;
    a.NT = (d->Cv + 7) / 8;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t vol_bytes = (size_t)d->B * d->X * d->Y * d->Z * d->Cv * sizeof(float);
    if (cudaMemsetAsync(vol, 0, vol_bytes, st) != cudaSuccess) return VFD_ERR_CUDA;
    const int sms = sm_count();
    if (sms <= 0) return VFD_ERR_CUDA;
    // tcgen05 kernel for the production shapes; VFD_FUSE_LEGACY=1 forces the mma.sync kernel (tests compare the two)
    const char* legacy = getenv("VFD_FUSE_LEGACY");
    if (d->C % 32 == 0 && d->Cv % 16 == 0 && !(legacy && legacy[0] == '1')) {
        const size_t smem = fuse_tc_smem_bytes(d->C, d->Cv);
        if (smem > 227 * 1024) return VFD_ERR_UNSUPPORTED;
        if (cudaFuncSetAttribute(voxel_fuse_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
            cudaSuccess)
            return VFD_ERR_CUDA;
        voxel_fuse_fwd_tc_kernel<<<sms, kTcThreads, smem, st>>>(a);
        VFD_CHECK_LAUNCH();
        return VFD_OK;
    }
    const size_t smem = fuse_fwd_smem_bytes(a.KP);
    rc = d->C > 128 ? launch_fuse(voxel_fuse_fwd_kernel<2>, a, smem, sms, st)
                    : launch_fuse(voxel_fuse_fwd_kernel<1>, a, smem, sms, st);
    if (rc) return rc;
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

extern "C" size_t vfd_voxel_fuse_bwd_workspace_bytes(const vfd_fusion_desc* d) {
    if (!d || d->Cv < 1 || d->C < 1) return 0;
    const int sms = sm_count();
    return (size_t)(sms > 0 ? sms : 256) * 3 * d->Cv * (d->C + 2) * sizeof(float);
}

extern "C" int vfd_voxel_fuse_bwd(const vfd_fusion_desc* d, const float* feats, const float* mask_small,
                                  const float* K, const float* Einv, const int64_t* idx_no, const int64_t* idx_ov,
                                  const int32_t* totals, const void* table, const float* w_no, const float* w_ov,
                                  const float* vol, const float* d_vol, float* d_feats, float* d_w_no, float* d_b_no, float* d_w_ov,
                                  float* d_b_ov, void* ws, size_t ws_bytes, void* stream) {
    int rc = check_fuse(d);
    if (rc) return rc;
    if (!feats || !mask_small || !K || !Einv || !idx_no || !idx_ov || !totals || !table || !w_no || !w_ov || !vol || !d_vol ||
        !d_feats || !d_w_no || !d_b_no || !d_w_ov || !d_b_ov || !ws)
        return VFD_ERR_NULL;
    if ((reinterpret_cast<uintptr_t>(feats) & 15) || (reinterpret_cast<uintptr_t>(d_feats) & 15) ||
        (reinterpret_cast<uintptr_t>(vol) & 15) || (reinterpret_cast<uintptr_t>(d_vol) & 15) ||
        (reinterpret_cast<uintptr_t>(ws) & 3))
        return VFD_ERR_ALIGN;
    const int sms = sm_count();
    if (sms <= 0) return VFD_ERR_CUDA;
    if (ws_bytes < (size_t)sms * 3 * d->Cv * (d->C + 2) * sizeof(float)) return VFD_ERR_WORKSPACE;
    FuseArgs a = {};
    a.d = *d; a.feats = feats; a.mask_small = mask_small; a.K = K; a.Einv = Einv;
    a.idx_no = idx_no; a.idx_ov = idx_ov; a.totals = totals; a.w_no = w_no; a.w_ov = w_ov; a.table = table;
    a.vol = const_cast<float*>(vol); a.d_vol = d_vol; a.d_feats = d_feats; a.partial = static_cast<float*>(ws);
    a.KP = (d->C + 1 + 7) / 8 * 8;
    a.NT = (d->Cv + 7) / 8;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t fbytes = (size_t)d->B * d->num_cams * d->h * d->w * d->C * sizeof(float);
    if (cudaMemsetAsync(d_feats, 0, fbytes, st) != cudaSuccess) return VFD_ERR_CUDA;
    const size_t smem = fuse_bwd_smem_bytes(a.KP, a.NT);
    const char* legacy = getenv("VFD_FUSE_LEGACY");
    const size_t smem_tc = fuse_tc_bwd_data_smem_bytes(d->C, d->Cv);
    if (d->C % 32 == 0 && d->Cv % 16 == 0 && smem_tc <= 227 * 1024 && !(legacy && legacy[0] == '1')) {
        // production shapes: dA + scatter on tcgen05, weight gradients on the mma.sync kernel
        if (cudaFuncSetAttribute(voxel_fuse_bwd_data_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem_tc) != cudaSuccess)
            return VFD_ERR_CUDA;
        voxel_fuse_bwd_data_tc_kernel<<<sms, kTcThreads, smem_tc, st>>>(a);
        const size_t smem_wt = fuse_wt_smem_bytes(a.KP);
        rc = d->C > 128 ? launch_fuse(voxel_fuse_bwd_weight_kernel<2>, a, smem_wt, sms, st)
                        : launch_fuse(voxel_fuse_bwd_weight_kernel<1>, a, smem_wt, sms, st);
    } else {
        rc = d->C > 128 ? launch_fuse(voxel_fuse_bwd_kernel<2>, a, smem, sms, st)
                        : launch_fuse(voxel_fuse_bwd_kernel<1>, a, smem, sms, st);
    }
    if (rc) return rc;
    const int total = 3 * d->Cv * (d->C + 2);
    voxel_fuse_reduce_kernel<<<(total + 255) / 256, 256, 0, st>>>(a.partial, sms, d->Cv, d->C + 1, d_w_no, d_b_no,
                                                                  d_w_ov, d_b_ov);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
