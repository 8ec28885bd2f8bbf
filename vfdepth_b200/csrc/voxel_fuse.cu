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
// The product runs on the tensor cores through mma.sync.m16n8k8 TF32 with the error-compensated 3xTF32 split
// (x = hi + lo, x*w ~ hi*hi + hi*lo + lo*hi, fp32 accumulate), which keeps the result within fp32 round-off of the
// reference's fp32 GEMM (tests: 1e-4 of the tensor's scale; measured ~1e-6).
//
// Persistent grid (one CTA per SM, 16 warps); the list lengths are read from device memory (vfd_voxel_classify's
// totals), so nothing here needs a host round trip.
#include "gather_common.cuh"

namespace vfd {

constexpr int kFuseRows = 64;       // voxels per tile
constexpr int kFuseWarps = 16;
constexpr int kFuseThreads = kFuseWarps * VFD_WARP;
constexpr int kWS = 72;             // shared-memory row stride of the [k][n] weight (8k+n banks: conflict-free B fragments)

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
    float* partial;      // [gridDim.x][3][Cv][KP+... ] per-CTA weight-gradient partial sums
    int KP;              // round_up(C+1, 8)
    int NT;              // ceil(Cv / 8)
};

struct FuseSmem {
    float* W[2];     // [KP][kWS] each
    float* A;        // [kFuseRows][AS]
    CamSlot* slot;   // [VFD_MAX_CAMS][kFuseRows]
    unsigned* vmask; // [kFuseRows]
    long long* row;  // [kFuseRows]  b*N+n or -1
};

__host__ __device__ inline size_t fuse_fwd_smem_bytes(int KP) {
    return (size_t)2 * KP * kWS * 4 + (size_t)kFuseRows * (KP + 4) * 4 + sizeof(CamSlot) * VFD_MAX_CAMS * kFuseRows +
           kFuseRows * 4 + kFuseRows * 8;
}

__device__ __forceinline__ FuseSmem carve_fwd(unsigned char* base, int KP) {
    FuseSmem s;
    s.row = reinterpret_cast<long long*>(base);
    base += kFuseRows * 8;
    s.W[0] = reinterpret_cast<float*>(base);
    base += (size_t)KP * kWS * 4;
    s.W[1] = reinterpret_cast<float*>(base);
    base += (size_t)KP * kWS * 4;
    s.A = reinterpret_cast<float*>(base);
    base += (size_t)kFuseRows * (KP + 4) * 4;
    s.slot = reinterpret_cast<CamSlot*>(base);
    base += sizeof(CamSlot) * VFD_MAX_CAMS * kFuseRows;
    s.vmask = reinterpret_cast<unsigned*>(base);
    return s;
}

// sW[k][n] = w[n][koff + k] for k < C1, n < Cv; zero elsewhere (k < KP, n < kWS)
__device__ __forceinline__ void load_weight_kn(float* sW, const float* w, int ld, int koff, int C1, int Cv, int KP) {
    for (int i = threadIdx.x; i < KP * kWS; i += blockDim.x) sW[i] = 0.f;
    __syncthreads();
    for (int i = threadIdx.x; i < Cv * C1; i += blockDim.x) {
        const int n = i / C1, k = i - n * C1;
        sW[k * kWS + n] = __ldg(w + (size_t)n * ld + koff + k);
    }
}

// threads r < kFuseRows: project list entry base+r into every camera
__device__ __forceinline__ void project_tile(const FuseArgs& a, const FuseSmem& s, const int64_t* idx, int base,
                                             int M) {
    const vfd_fusion_desc& d = a.d;
    const int r = threadIdx.x;
    if (r >= kFuseRows) return;
    const long N = (long)d.X * d.Y * d.Z;
    long long row = -1;
    unsigned vm = 0;
    if (base + r < M) {
        row = idx[base + r];
        const int b = (int)(row / N);
        const long n = row - (long)b * N;
        const int xi = (int)(n % d.X), yi = (int)((n / d.X) % d.Y), zi = (int)(n / ((long)d.X * d.Y));
        const float x = __ldg(d.axis_x + xi), y = __ldg(d.axis_y + yi), z = __ldg(d.axis_z + zi);
        for (int c = 0; c < d.num_cams; ++c) {
            const size_t bc = (size_t)b * d.num_cams + c;
            float g[2];
            CamSlot cs;
            const bool ok = project_cam(d, a.mask_small + bc * d.h * d.w, a.Einv + bc * 16, a.K + bc * 16, x, y, z, cs, g);
            s.slot[c * kFuseRows + r] = cs;
            vm |= (ok ? 1u : 0u) << c;
        }
    }
    s.row[r] = row;
    s.vmask[r] = vm;
}

// warp-cooperative gather of one row: A[r][0..C) = sum over cameras in `m` of the bilinear sample, A[r][C] = sum zrel,
// A[r][C+1..KP) = 0
__device__ __forceinline__ void gather_row(const FuseArgs& a, const FuseSmem& s, int r, unsigned m, int lane, int AS) {
    const vfd_fusion_desc& d = a.d;
    const int C = d.C, hw = d.h * d.w;
    float* arow = s.A + (size_t)r * AS;
    const long long row = s.row[r];
    const long N = (long)d.X * d.Y * d.Z;
    const int b = row >= 0 ? (int)(row / N) : 0;
    float zsum = 0.f;
    for (int k0 = 0; k0 < C; k0 += 128) {
        const int k = k0 + lane * 4;
        float4 acc = f4_zero();
        unsigned mm = m;
        while (mm) {
            const int c = __ffs(mm) - 1;
            mm &= mm - 1;
            const CamSlot cs = s.slot[c * kFuseRows + r];
            if (k0 == 0) zsum += cs.zrel;
            if (k < C) {
                const Corners q = make_corners(cs, d.w, d.h, C);
                const float* base = a.feats + ((size_t)b * d.num_cams + c) * hw * C;
                const float4 v0 = ldg_f4(base + q.off[0] + k), v1 = ldg_f4(base + q.off[1] + k);
                const float4 v2 = ldg_f4(base + q.off[2] + k), v3 = ldg_f4(base + q.off[3] + k);
                float4 t = f4_scale(q.wgt[0], v0);
                f4_fma(t, q.wgt[1], v1);
                f4_fma(t, q.wgt[2], v2);
                f4_fma(t, q.wgt[3], v3);
                f4_add(acc, t);
            }
        }
        if (k < C) *reinterpret_cast<float4*>(arow + k) = acc;
    }
    for (int k = C + lane; k < a.KP; k += VFD_WARP) arow[k] = (k == C) ? zsum : 0.f;
}

__global__ void __launch_bounds__(kFuseThreads, 1) voxel_fuse_fwd_kernel(const FuseArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const vfd_fusion_desc& d = a.d;
    const int KP = a.KP, AS = KP + 4, NT = a.NT, C1 = d.C + 1, Cv = d.Cv;
    const FuseSmem s = carve_fwd(smem_raw, KP);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int wm = warp & 3, wn = warp >> 2;
    const int M1 = a.totals[0], M2 = a.totals[1];
    const int T1 = (M1 + kFuseRows - 1) / kFuseRows, T2 = (M2 + kFuseRows - 1) / kFuseRows;
    unsigned gmask[2] = {0u, 0u};
    for (int c = 0; c < d.num_cams; ++c) gmask[d.group[c] ? 1 : 0] |= 1u << c;

    int cur = -1;
    for (int tile = blockIdx.x; tile < T1 + T2; tile += gridDim.x) {
        const int type = tile < T1 ? 0 : 1;
        const int base = (type ? tile - T1 : tile) * kFuseRows;
        const int M = type ? M2 : M1;
        const int64_t* idx = type ? a.idx_ov : a.idx_no;
        if (type != cur) {
            __syncthreads();
            if (type == 0) {
                load_weight_kn(s.W[0], a.w_no, C1, 0, C1, Cv, KP);
            } else {
                load_weight_kn(s.W[0], a.w_ov, 2 * C1, 0, C1, Cv, KP);
                load_weight_kn(s.W[1], a.w_ov, 2 * C1, C1, C1, Cv, KP);
            }
            cur = type;
        }
        project_tile(a, s, idx, base, M);
        __syncthreads();

        float acc[2][4];
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[j][i] = 0.f;

        for (int half = 0; half <= type; ++half) {
            // gather: 4 rows per warp
#pragma unroll 2
            for (int rr = 0; rr < kFuseRows / kFuseWarps; ++rr) {
                const int r = warp * (kFuseRows / kFuseWarps) + rr;
                const unsigned m = s.vmask[r] & (type ? gmask[half] : 0xffffffffu);
                gather_row(a, s, r, m, lane, AS);
            }
            __syncthreads();
            // [64 x KP] x [KP x Cv] on the tensor cores, 3xTF32
            const float* A0 = s.A + (size_t)(wm * 16 + g) * AS + t;
            const float* A1 = A0 + 8 * AS;
            const float* Wb = s.W[half] + t * kWS + wn * 16 + g;
            if (wn * 2 < NT) {
                for (int ks = 0; ks < KP; ks += 8) {
                    uint32_t ah[4], al[4];
                    split_tf32(A0[ks], ah[0], al[0]);
                    split_tf32(A1[ks], ah[1], al[1]);
                    split_tf32(A0[ks + 4], ah[2], al[2]);
                    split_tf32(A1[ks + 4], ah[3], al[3]);
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        if (wn * 2 + j < NT) {
                            uint32_t bh0, bl0, bh1, bl1;
                            split_tf32(Wb[ks * kWS + j * 8], bh0, bl0);
                            split_tf32(Wb[(ks + 4) * kWS + j * 8], bh1, bl1);
                            mma_tf32(acc[j], al, bh0, bh1);
                            mma_tf32(acc[j], ah, bl0, bl1);
                            mma_tf32(acc[j], ah, bh0, bh1);
                        }
                    }
                }
            }
            __syncthreads();
        }
        // epilogue: + bias, LeakyReLU(0.1), scatter the rows into the volume
        const float* bias = type ? a.b_ov : a.b_no;
        const long long row0 = s.row[wm * 16 + g], row1 = s.row[wm * 16 + g + 8];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int col = (wn * 2 + j) * 8 + 2 * t;
            if (wn * 2 + j < NT && col < Cv) {
                const float b0 = __ldg(bias + col), b1 = __ldg(bias + col + 1);
                float y0 = acc[j][0] + b0, y1 = acc[j][1] + b1, y2 = acc[j][2] + b0, y3 = acc[j][3] + b1;
                y0 = y0 > 0.f ? y0 : 0.1f * y0;
                y1 = y1 > 0.f ? y1 : 0.1f * y1;
                y2 = y2 > 0.f ? y2 : 0.1f * y2;
                y3 = y3 > 0.f ? y3 : 0.1f * y3;
                if (row0 >= 0) *reinterpret_cast<float2*>(a.vol + (size_t)row0 * Cv + col) = make_float2(y0, y1);
                if (row1 >= 0) *reinterpret_cast<float2*>(a.vol + (size_t)row1 * Cv + col) = make_float2(y2, y3);
            }
        }
        __syncthreads();   // s.row / s.vmask are rewritten by the next tile's projection
    }
}

// ---------------------------------------------------------------------------------------------------------
// Backward.  Three passes over the compact lists, one per weight block (W_no, W_ov[:, :C+1], W_ov[:, C+1:]); a tile of
// 64 voxels per step:
//   dZ  = d_vol[rows] * LeakyReLU'(vol[rows])                         (y > 0 <=> pre-activation > 0)
//   A   = the gathered feature rows, recomputed (L2-resident feature maps)  [64 x (C+1)]
//   dW += A^T dZ    tensor cores, accumulators stay in registers for the whole pass (one 16-row block per warp);
//                   the relative-depth column and db are two SIMT dot products
//   dA  = dZ W      tensor cores -> shared memory -> bilinear scatter into d_feats with red.global.add.v4.f32
// Per-CTA weight-gradient partials are written to the workspace and summed in CTA order by a second kernel, so dW / db
// are deterministic; d_feats is accumulated with L2 atomics like vfd_voxel_gather_depth_bwd.
// ---------------------------------------------------------------------------------------------------------
constexpr int kZS = 72;   // row stride of the dZ tile

__host__ __device__ inline int fuse_sn(int KP) { return (KP - 8 + 31) / 32 * 32 + 8; }   // stride = 8 (mod 32)

__host__ __device__ inline size_t fuse_bwd_smem_bytes(int KP, int NT) {
    const int SN = fuse_sn(KP);
    return (size_t)kFuseRows * 8 + (size_t)NT * 8 * SN * 4 + (size_t)kFuseRows * SN * 4 + (size_t)kFuseRows * kZS * 4 +
           sizeof(CamSlot) * VFD_MAX_CAMS * kFuseRows + kFuseRows * 4;
}

struct FuseBwdSmem {
    FuseSmem g;      // A (stride SN), slot, vmask, row;  g.W[0] = natural-layout weight [NT*8][SN]
    float* Z;        // [kFuseRows][kZS]
};

__device__ __forceinline__ FuseBwdSmem carve_bwd(unsigned char* base, int KP, int NT) {
    const int SN = fuse_sn(KP);
    FuseBwdSmem s;
    s.g.row = reinterpret_cast<long long*>(base);
    base += kFuseRows * 8;
    s.g.W[0] = reinterpret_cast<float*>(base);
    s.g.W[1] = nullptr;
    base += (size_t)NT * 8 * SN * 4;
    s.g.A = reinterpret_cast<float*>(base);
    base += (size_t)kFuseRows * SN * 4;
    s.Z = reinterpret_cast<float*>(base);
    base += (size_t)kFuseRows * kZS * 4;
    s.g.slot = reinterpret_cast<CamSlot*>(base);
    base += sizeof(CamSlot) * VFD_MAX_CAMS * kFuseRows;
    s.g.vmask = reinterpret_cast<unsigned*>(base);
    return s;
}

__global__ void __launch_bounds__(kFuseThreads, 1) voxel_fuse_bwd_kernel(const FuseArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const vfd_fusion_desc& d = a.d;
    const int KP = a.KP, SN = fuse_sn(KP), NT = a.NT, C = d.C, C1 = C + 1, Cv = d.Cv;
    const FuseBwdSmem s = carve_bwd(smem_raw, KP, NT);
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

    for (int pass = 0; pass < 3; ++pass) {
        const int M = pass ? M2 : M1, T = pass ? T2 : T1;
        const int64_t* idx = pass ? a.idx_ov : a.idx_no;
        const unsigned want = pass ? gmask[pass - 1] : 0xffffffffu;
        // natural-layout weight block [out][in], zero padded
        __syncthreads();
        {
            const float* w = pass ? a.w_ov : a.w_no;
            const int ld = pass ? 2 * C1 : C1, koff = pass == 2 ? C1 : 0;
            for (int i = threadIdx.x; i < NT * 8 * SN; i += blockDim.x) {
                const int o = i / SN, k = i - o * SN;
                s.g.W[0][i] = (o < Cv && k < C1) ? __ldg(w + (size_t)o * ld + koff + k) : 0.f;
            }
        }
        float accw[8][4];
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) accw[j][i] = 0.f;
        float db = 0.f, dwz = 0.f;           // threads o < Cv: bias gradient and the relative-depth column of dW

        for (int tile = blockIdx.x; tile < T; tile += gridDim.x) {
            const int base = tile * kFuseRows;
            project_tile(a, s.g, idx, base, M);
            __syncthreads();
            // dZ tile
            for (int i = threadIdx.x; i < kFuseRows * (kZS / 4); i += blockDim.x) {
                const int r = i / (kZS / 4), q = (i - r * (kZS / 4)) * 4;
                const long long row = s.g.row[r];
                float4 z = f4_zero();
                if (row >= 0 && q < Cv) {
                    const float4 gy = ld_stream_f4(a.d_vol + (size_t)row * Cv + q);
                    const float4 y = ld_stream_f4(a.vol + (size_t)row * Cv + q);
                    z.x = y.x > 0.f ? gy.x : 0.1f * gy.x;
                    z.y = y.y > 0.f ? gy.y : 0.1f * gy.y;
                    z.z = y.z > 0.f ? gy.z : 0.1f * gy.z;
                    z.w = y.w > 0.f ? gy.w : 0.1f * gy.w;
                }
                *reinterpret_cast<float4*>(s.Z + r * kZS + q) = z;
            }
            // recompute the gathered rows
#pragma unroll 2
            for (int rr = 0; rr < kFuseRows / kFuseWarps; ++rr) {
                const int r = warp * (kFuseRows / kFuseWarps) + rr;
                gather_row(a, s.g, r, s.g.vmask[r] & want, lane, SN);
            }
            __syncthreads();

            // dW^T block [16 in x Cv out] += A^T [16 x 64] dZ [64 x Cv]
            if (warp < MT) {
                const float* At = s.g.A + t * SN + warp * 16 + g;
                const float* Zb = s.Z + t * kZS + g;
#pragma unroll 2
                for (int ks = 0; ks < kFuseRows; ks += 8) {
                    uint32_t ah[4], al[4];
                    split_tf32(At[ks * SN], ah[0], al[0]);
                    split_tf32(At[ks * SN + 8], ah[1], al[1]);
                    split_tf32(At[(ks + 4) * SN], ah[2], al[2]);
                    split_tf32(At[(ks + 4) * SN + 8], ah[3], al[3]);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (j < NT) {
                            uint32_t bh0, bl0, bh1, bl1;
                            split_tf32(Zb[ks * kZS + j * 8], bh0, bl0);
                            split_tf32(Zb[(ks + 4) * kZS + j * 8], bh1, bl1);
                            mma_tf32(accw[j], al, bh0, bh1);
                            mma_tf32(accw[j], ah, bl0, bl1);
                            mma_tf32(accw[j], ah, bh0, bh1);
                        }
                    }
                }
            }
            if (threadIdx.x < Cv) {
                const int o = threadIdx.x;
                float sb = 0.f, sz = 0.f;
                for (int r = 0; r < kFuseRows; ++r) {
                    const float z = s.Z[r * kZS + o];
                    sb += z;
                    sz = fmaf(z, s.g.A[r * SN + C], sz);
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
                const float* Z0 = s.Z + (wm * 16 + g) * kZS + t;
                const float* Z1 = Z0 + 8 * kZS;
                const float* Wb = s.g.W[0] + t * SN + wn * NAW * 8 + g;
                for (int ks = 0; ks < NT * 8; ks += 8) {
                    uint32_t ah[4], al[4];
                    split_tf32(Z0[ks], ah[0], al[0]);
                    split_tf32(Z1[ks], ah[1], al[1]);
                    split_tf32(Z0[ks + 4], ah[2], al[2]);
                    split_tf32(Z1[ks + 4], ah[3], al[3]);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (j < NAW && wn * NAW + j < NA) {
                            uint32_t bh0, bl0, bh1, bl1;
                            split_tf32(Wb[ks * SN + j * 8], bh0, bl0);
                            split_tf32(Wb[(ks + 4) * SN + j * 8], bh1, bl1);
                            mma_tf32(acca[j], al, bh0, bh1);
                            mma_tf32(acca[j], ah, bl0, bl1);
                            mma_tf32(acca[j], ah, bh0, bh1);
                        }
                    }
                }
            }
            __syncthreads();          // every warp is done reading A: overwrite it with dA
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (j < NAW && wn * NAW + j < NA) {
                    const int col = (wn * NAW + j) * 8 + 2 * t;
                    float* r0 = s.g.A + (size_t)(wm * 16 + g) * SN + col;
                    *reinterpret_cast<float2*>(r0) = make_float2(acca[j][0], acca[j][1]);
                    *reinterpret_cast<float2*>(r0 + 8 * SN) = make_float2(acca[j][2], acca[j][3]);
                }
            }
            __syncthreads();
            // bilinear scatter of the dA rows
            {
                const long N = (long)d.X * d.Y * d.Z;
                const int hw = d.h * d.w;
                for (int rr = 0; rr < kFuseRows / kFuseWarps; ++rr) {
                    const int r = warp * (kFuseRows / kFuseWarps) + rr;
                    const long long row = s.g.row[r];
                    if (row < 0) continue;
                    const int b = (int)(row / N);
                    unsigned mm = s.g.vmask[r] & want;
                    while (mm) {
                        const int c = __ffs(mm) - 1;
                        mm &= mm - 1;
                        const Corners q = make_corners(s.g.slot[c * kFuseRows + r], d.w, d.h, C);
                        float* fb = a.d_feats + ((size_t)b * d.num_cams + c) * hw * C;
                        for (int k = lane * 4; k < C; k += 128) {
                            const float4 gv = *reinterpret_cast<const float4*>(s.g.A + (size_t)r * SN + k);
#pragma unroll
                            for (int u = 0; u < 4; ++u)
                                if (q.wgt[u] != 0.f) red_add_f4(fb + q.off[u] + k, f4_scale(q.wgt[u], gv));
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

}  // namespace vfd

using namespace vfd;

extern "C" int vfd_voxel_fuse_fwd(const vfd_fusion_desc* d, const float* feats, const float* mask_small,
                                  const float* K, const float* Einv, const int64_t* idx_no, const int64_t* idx_ov,
                                  const int32_t* totals, const float* w_no, const float* b_no, const float* w_ov,
                                  const float* b_ov, float* vol, void* stream) {
    int rc = check_fuse(d);
    if (rc) return rc;
    if (!feats || !mask_small || !K || !Einv || !idx_no || !idx_ov || !totals || !w_no || !b_no || !w_ov || !b_ov ||
        !vol)
        return VFD_ERR_NULL;
    if ((reinterpret_cast<uintptr_t>(feats) & 15) || (reinterpret_cast<uintptr_t>(vol) & 7)) return VFD_ERR_ALIGN;
    FuseArgs a = {};
    a.d = *d; a.feats = feats; a.mask_small = mask_small; a.K = K; a.Einv = Einv;
    a.idx_no = idx_no; a.idx_ov = idx_ov; a.totals = totals;
    a.w_no = w_no; a.b_no = b_no; a.w_ov = w_ov; a.b_ov = b_ov; a.vol = vol;
    a.KP = (d->C + 1 + 7) / 8 * 8;
    a.NT = (d->Cv + 7) / 8;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t vol_bytes = (size_t)d->B * d->X * d->Y * d->Z * d->Cv * sizeof(float);
    if (cudaMemsetAsync(vol, 0, vol_bytes, st) != cudaSuccess) return VFD_ERR_CUDA;
    const size_t smem = fuse_fwd_smem_bytes(a.KP);
    if (smem > 227 * 1024) return VFD_ERR_UNSUPPORTED;
    if (cudaFuncSetAttribute(voxel_fuse_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
        cudaSuccess)
        return VFD_ERR_CUDA;
    const int sms = sm_count();
    if (sms <= 0) return VFD_ERR_CUDA;
    voxel_fuse_fwd_kernel<<<sms, kFuseThreads, smem, st>>>(a);
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
                                  const int32_t* totals, const float* w_no, const float* w_ov, const float* vol,
                                  const float* d_vol, float* d_feats, float* d_w_no, float* d_b_no, float* d_w_ov,
                                  float* d_b_ov, void* ws, size_t ws_bytes, void* stream) {
    int rc = check_fuse(d);
    if (rc) return rc;
    if (!feats || !mask_small || !K || !Einv || !idx_no || !idx_ov || !totals || !w_no || !w_ov || !vol || !d_vol ||
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
    a.idx_no = idx_no; a.idx_ov = idx_ov; a.totals = totals; a.w_no = w_no; a.w_ov = w_ov;
    a.vol = const_cast<float*>(vol); a.d_vol = d_vol; a.d_feats = d_feats; a.partial = static_cast<float*>(ws);
    a.KP = (d->C + 1 + 7) / 8 * 8;
    a.NT = (d->Cv + 7) / 8;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t fbytes = (size_t)d->B * d->num_cams * d->h * d->w * d->C * sizeof(float);
    if (cudaMemsetAsync(d_feats, 0, fbytes, st) != cudaSuccess) return VFD_ERR_CUDA;
    const size_t smem = fuse_bwd_smem_bytes(a.KP, a.NT);
    if (smem > 227 * 1024) return VFD_ERR_UNSUPPORTED;
    if (cudaFuncSetAttribute(voxel_fuse_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
        cudaSuccess)
        return VFD_ERR_CUDA;
    voxel_fuse_bwd_kernel<<<sms, kFuseThreads, smem, st>>>(a);
    const int total = 3 * d->Cv * (d->C + 2);
    voxel_fuse_reduce_kernel<<<(total + 255) / 256, 256, 0, st>>>(a.partial, sms, d->Cv, d->C + 1, d_w_no, d_b_no,
                                                                  d_w_ov, d_b_ov);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}
