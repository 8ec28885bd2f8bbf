// Dense products on either side of the fused voxel gather (voxel_fuse.cu), "project, then gather".
//
// The two 1x1 MLPs of VFNet.backproject_into_voxel (network/volumetric_fusionnet.py:197-230) are linear up to their
// LeakyReLU, and so is the bilinear gather in front of them (:143-146):
//     W . (sum_k w_k feat[corner_k])  ==  sum_k w_k (W . feat[corner_k]).
// Instead of gathering a 257-channel row per (voxel, camera) and multiplying 1.5 M such rows (B=8) with the MLP
// weights, every FEATURE PIXEL (184 320 at B=8) is multiplied once with both weight blocks its camera can need --
//     P[img, pixel, 0:Cv]    = W_no[:, :C]                  . feat[img, pixel, :]     (voxels seen by one camera)
//     P[img, pixel, Cv:2Cv]  = W_ov[:, g*(C+1) : g*(C+1)+C] . feat[img, pixel, :]     g = overlap group of the camera
// -- a dense, TMA-loadable GEMM with 8x fewer rows, and the voxel kernels gather 256-byte rows of P instead of 1 KB rows of
// the feature maps (4x less L2 traffic, no tensor work in the gather).  The backward mirrors it: the voxel kernel scatters
// d(pre-activation) into dP, then   d_feats = dP . Wcat_g   and   dWcat_g = dP^T . feats   are dense products again.
//
//   dense_proj_tc_kernel   Y[img, r, 0:N] = X[img, r, 0:K] . W_g[0:N, 0:K]^T on tcgen05 (kind::tf32, 3xTF32 split, fp32
//                          accuracy), operands staged by TMA (cp.async.bulk.tensor, SWIZZLE_128B), accumulators in TMEM.
//                          Forward: X = feats, K = C, N = 2 Cv.  Backward: X = dP, K = 2 Cv, N = C, W = Wcat^T.
//   dense_proj_simt_kernel the same product for shapes the tensor path does not take (C % 16, Cv % 8).
//   dense_dw_kernel        dWcat partial sums over slabs of pixels (fp32 FMA, packed f32x2), reduced in a fixed order.
#include "gather_common.cuh"
#include "umma.cuh"

namespace vfd {

struct ProjArgs {
    int n_img, hw, K, N, cams;
    int group[VFD_MAX_CAMS];
    const float* X;      // [n_img][hw][K]
    float* Y;            // [n_img][hw][N]
    const float* Wfull;  // [2][N][K] fp32 (SIMT path)
};

// ---------------------------------------------------------------------------------------------------------
// weight planes: Wcat_g [2 Cv][C] = [W_no[:, :C] ; W_ov[:, g (C+1) : g (C+1) + C]]  (g = 0, 1), as fp32 and as TF32
// hi / lo planes; transposed ([C][2 Cv]) for the backward.
// ---------------------------------------------------------------------------------------------------------
__global__ void wcat_prep_kernel(const float* __restrict__ w_no, const float* __restrict__ w_ov, int C, int Cv,
                                 int transpose, float* full, float* hi, float* lo) {
    const int N2 = 2 * Cv, C1 = C + 1;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 2 * N2 * C) return;
    const int g = idx / (N2 * C), r = idx - g * N2 * C;
    int n, k;
    if (transpose) { k = r / N2; n = r - k * N2; } else { n = r / C; k = r - n * C; }
    const float v = n < Cv ? __ldg(w_no + (size_t)n * C1 + k) : __ldg(w_ov + (size_t)(n - Cv) * 2 * C1 + g * C1 + k);
    uint32_t h, l;
    split_tf32(v, h, l);
    full[idx] = v;
    hi[idx] = __uint_as_float(h);
    lo[idx] = __uint_as_float(l);
}

// ---------------------------------------------------------------------------------------------------------
// tcgen05 + TMA product.  One CTA per SM, 10 warps:
//   warps 0-3  epilogue: tcgen05.ld of the finished accumulator (TMEM lane = row of the tile), 64-byte runs to Y
//   warp  4    TMA producer (one lane): per 32-channel K chunk the X tile [128 rows][128 B] and the chunk of both
//              weight planes [N rows][128 B] land in a stage of the ring, SWIZZLE_128B, completing on tma_full[stage]
//   warp  5    MMA issue (one lane): per chunk up to 4 K-steps x {X_lo W_hi, X_hi W_lo, X_hi W_hi} (3xTF32);
//              tcgen05.commit frees the stage / publishes the accumulator (2 accumulators of N TMEM columns)
//   warps 6-9  split: X arrives as fp32; the tile is rewritten in place as its TF32 hi part and its lo part goes to the
//              stage's second A buffer (same byte offsets, so the swizzle is untouched), then fence.proxy.async
// ---------------------------------------------------------------------------------------------------------
constexpr int kPjRows = 128;
constexpr int kPjThreads = 10 * VFD_WARP;
constexpr int kPjMaxStages = 4;

struct PjBars {
    uint64_t tma_full[kPjMaxStages], conv_full[kPjMaxStages], empty[kPjMaxStages], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

__host__ __device__ inline uint32_t pj_stage_bytes(int N) { return 2u * kPjRows * 128u + 2u * (uint32_t)N * 128u; }
static inline int pj_stages(int N) {
    const int s = (int)((200u * 1024u) / pj_stage_bytes(N));
    return s > kPjMaxStages ? kPjMaxStages : s;
}
static inline size_t pj_smem_bytes(int N) { return 1024 + (size_t)pj_stages(N) * pj_stage_bytes(N) + sizeof(PjBars); }

__global__ void __launch_bounds__(kPjThreads, 1)
    dense_proj_tc_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapWhi,
                         const __grid_constant__ CUtensorMap mapWlo, const ProjArgs a, int stages, int tmem_cols) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // swizzle atoms: 1024-byte aligned
    const int N = a.N, K = a.K, NCH = (K + 31) / 32;
    const uint32_t sbytes = pj_stage_bytes(N), a_bytes = kPjRows * 128u, b_bytes = (uint32_t)N * 128u;
    PjBars* bars = reinterpret_cast<PjBars*>(base + (size_t)stages * sbytes);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int tiles_per_img = (a.hw + kPjRows - 1) / kPjRows, n_tiles = a.n_img * tiles_per_img;

    if (tid == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&bars->tma_full[s], 1);
            mbar_init(&bars->conv_full[s], 4 * VFD_WARP);
            mbar_init(&bars->empty[s], 1);
        }
        for (int i = 0; i < 2; ++i) { mbar_init(&bars->acc_full[i], 1); mbar_init(&bars->acc_empty[i], 4 * VFD_WARP); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        if (lane == 0) { tma_prefetch_desc(&mapX); tma_prefetch_desc(&mapWhi); tma_prefetch_desc(&mapWlo); }
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars->tmem_base)),
                     "r"(tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = bars->tmem_base;

    uint32_t stage = 0, phase = 0, ntile = 0;
    if (warp == 4) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                const int img = tile / tiles_per_img, r0 = (tile - img * tiles_per_img) * kPjRows;
                const int g = a.group[img % a.cams] ? 1 : 0;
                for (int kc = 0; kc < NCH; ++kc) {
                    mbar_wait(&bars->empty[stage], phase ^ 1);
                    unsigned char* st = base + (size_t)stage * sbytes;
                    mbar_expect_tx(&bars->tma_full[stage], a_bytes + 2 * b_bytes);
                    tma_load_3d(st, &mapX, kc * 32, r0, img, &bars->tma_full[stage]);
                    tma_load_3d(st + 2 * a_bytes, &mapWhi, kc * 32, 0, g, &bars->tma_full[stage]);
                    tma_load_3d(st + 2 * a_bytes + b_bytes, &mapWlo, kc * 32, 0, g, &bars->tma_full[stage]);
                    if (++stage == (uint32_t)stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 5) {
        // ===================== MMA issue =====================
        const uint32_t idesc = umma_idesc_tf32(kPjRows, N);
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++ntile) {
            const uint32_t acc = ntile & 1;
            mbar_wait(&bars->acc_empty[acc], ((ntile >> 1) & 1) ^ 1);
            tc_fence_after();
            for (int kc = 0; kc < NCH; ++kc) {
                mbar_wait(&bars->tma_full[stage], phase);      // the weight chunk (written by the TMA engine)
                mbar_wait(&bars->conv_full[stage], phase);     // the split X chunk (written by the split warps)
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t s0 = smem_u32(base + (size_t)stage * sbytes);
                    const uint64_t a_hi = umma_desc(s0), a_lo = umma_desc(s0 + a_bytes);
                    const uint64_t b_hi = umma_desc(s0 + 2 * a_bytes), b_lo = umma_desc(s0 + 2 * a_bytes + b_bytes);
                    const uint32_t dcol = tmem + acc * (uint32_t)N;
                    const int ksteps = min(4, (K - kc * 32 + 7) / 8);
                    for (int k = 0; k < ksteps; ++k) {     // 8 TF32 = 32 bytes per K-step: +2 in the 16-byte address field
                        umma_tf32(dcol, a_lo + 2 * k, b_hi + 2 * k, idesc, (kc | k) ? 1u : 0u);
                        umma_tf32(dcol, a_hi + 2 * k, b_lo + 2 * k, idesc, 1u);
                        umma_tf32(dcol, a_hi + 2 * k, b_hi + 2 * k, idesc, 1u);
                    }
                    tc_commit(&bars->empty[stage]);
                    if (kc == NCH - 1) tc_commit(&bars->acc_full[acc]);
                }
                __syncwarp();
                if (++stage == (uint32_t)stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp >= 6) {
        // ===================== split X into TF32 hi (in place) + lo =====================
        const int ctid = tid - 6 * VFD_WARP;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            for (int kc = 0; kc < NCH; ++kc) {
                mbar_wait(&bars->tma_full[stage], phase);
                uint4* pa = reinterpret_cast<uint4*>(base + (size_t)stage * sbytes);
                uint4* pl = reinterpret_cast<uint4*>(base + (size_t)stage * sbytes + a_bytes);
#pragma unroll
                for (int i = 0; i < (int)(kPjRows * 128 / 16) / (4 * VFD_WARP); ++i) {
                    const int o = ctid + i * 4 * VFD_WARP;
                    const uint4 v = pa[o];
                    uint4 h, l;
                    split_tf32(__uint_as_float(v.x), h.x, l.x);
                    split_tf32(__uint_as_float(v.y), h.y, l.y);
                    split_tf32(__uint_as_float(v.z), h.z, l.z);
                    split_tf32(__uint_as_float(v.w), h.w, l.w);
                    pa[o] = h;
                    pl[o] = l;
                }
                fence_proxy_async_smem();
                mbar_arrive(&bars->conv_full[stage]);
                if (++stage == (uint32_t)stages) { stage = 0; phase ^= 1; }
            }
        }
    } else {
        // ===================== epilogue (TMEM lane = row of the tile) =====================
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++ntile) {
            const uint32_t acc = ntile & 1;
            const int img = tile / tiles_per_img, r = (tile - img * tiles_per_img) * kPjRows + tid;
            mbar_wait(&bars->acc_full[acc], (ntile >> 1) & 1);
            tc_fence_after();
            float* o = a.Y + ((size_t)img * a.hw + (r < a.hw ? r : 0)) * N;
            for (int c0 = 0; c0 < N; c0 += 16) {
                uint32_t v[16];
                tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + acc * (uint32_t)N + (uint32_t)c0, v);
                if (r < a.hw) {
#pragma unroll
                    for (int q = 0; q < 16; q += 4)
                        *reinterpret_cast<float4*>(o + c0 + q) =
                            make_float4(__uint_as_float(v[q]), __uint_as_float(v[q + 1]), __uint_as_float(v[q + 2]),
                                        __uint_as_float(v[q + 3]));
                }
            }
            tc_fence_before();
            mbar_arrive(&bars->acc_empty[acc]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(tmem_cols));
}

// fp32 FMA form of the same product for the shapes outside the tensor path: one thread per output element
__global__ void __launch_bounds__(256) dense_proj_simt_kernel(const ProjArgs a) {
    const size_t total = (size_t)a.n_img * a.hw * a.N;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
        const int n = (int)(i % a.N);
        const size_t row = i / a.N;
        const int img = (int)(row / a.hw);
        const int g = a.group[img % a.cams] ? 1 : 0;
        const float* x = a.X + row * a.K;
        const float* w = a.Wfull + ((size_t)g * a.N + n) * a.K;
        float acc = 0.f;
        for (int k = 0; k < a.K; k += 4) {
            const float4 xv = ldg_f4(x + k), wv = ldg_f4(w + k);
            acc = fmaf(xv.x, wv.x, acc);
            acc = fmaf(xv.y, wv.y, acc);
            acc = fmaf(xv.z, wv.z, acc);
            acc = fmaf(xv.w, wv.w, acc);
        }
        a.Y[i] = acc;
    }
}

static int sm_count_pj() {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    return n;
}

bool dense_proj_tc_ok(int K, int N) {
    return K % 16 == 0 && N % 16 == 0 && N >= 16 && N <= 256 && K >= 16 && K <= 256;
}

// Y[n_img][hw][N] = X[n_img][hw][K] . W_g^T with W planes [2][N][K] (full / hi / lo)
int dense_proj(int n_img, int hw, int K, int N, int cams, const int* group, const float* X, float* Y, const float* Wfull,
               const float* Whi, const float* Wlo, cudaStream_t st) {
    ProjArgs a = {};
    a.n_img = n_img; a.hw = hw; a.K = K; a.N = N; a.cams = cams;
    for (int c = 0; c < VFD_MAX_CAMS; ++c) a.group[c] = c < cams ? group[c] : 0;
    a.X = X; a.Y = Y; a.Wfull = Wfull;
    if (dense_proj_tc_ok(K, N)) {
        CUtensorMap mx, mh, ml;
        if (!make_map_3d(&mx, X, (uint64_t)K, (uint64_t)hw, (uint64_t)n_img, kPjRows) ||
            !make_map_3d(&mh, Whi, (uint64_t)K, (uint64_t)N, 2, (uint32_t)N) ||
            !make_map_3d(&ml, Wlo, (uint64_t)K, (uint64_t)N, 2, (uint32_t)N))
            return VFD_ERR_CUDA;
        const int sms = sm_count_pj();
        if (sms <= 0) return VFD_ERR_CUDA;
        const size_t smem = pj_smem_bytes(N);
        if (cudaFuncSetAttribute(dense_proj_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return VFD_ERR_CUDA;
        int cols = 32;
        while (cols < 2 * N) cols <<= 1;
        const int tiles = n_img * ((hw + kPjRows - 1) / kPjRows);
        dense_proj_tc_kernel<<<tiles < sms ? tiles : sms, kPjThreads, smem, st>>>(mx, mh, ml, a, pj_stages(N), cols);
    } else {
        const size_t total = (size_t)n_img * hw * N;
        size_t blocks = (total + 255) / 256;
        if (blocks > 148 * 16) blocks = 148 * 16;
        dense_proj_simt_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
    }
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

// ---------------------------------------------------------------------------------------------------------
// dWcat partial sums:  part[img, slab][n][c] = sum over the slab's pixels of dP[img, p, n] * feats[img, p, c].
// CTA = one slab of one image, 256 threads; thread (tn, tc) owns the 8 x 16 block n = 8 tn .. +7, c = 4 tc + 64 j .. +3
// (j = 0..3) of the (<= 128) x (<= 256) product in registers (packed fp32 FMAs), pixels stream through shared memory
// 16 at a time with the next chunk's global loads in flight during the FMAs.
// ---------------------------------------------------------------------------------------------------------
constexpr int kDwKT = 16, kDwN = 128, kDwC = 256, kDwThreads = 256;

struct DwArgs {
    int hw, C, N2, slabs, slab_len;
    const float* dP;     // [n_img][hw][N2]
    const float* feats;  // [n_img][hw][C]
    float* part;         // [n_img][slabs][N2][C]
};

__global__ void __launch_bounds__(kDwThreads, 1) dense_dw_kernel(const DwArgs a) {
    __shared__ __align__(16) float sP[kDwKT][kDwN];
    __shared__ __align__(16) float sF[kDwKT][kDwC];
    const int tid = threadIdx.x, tn = tid >> 4, tc = tid & 15;
    const int img = blockIdx.y, slab = blockIdx.x;
    const int p_begin = slab * a.slab_len, p_end = min(a.hw, p_begin + a.slab_len);
    const int N2 = a.N2, C = a.C;
    const int qP = N2 / 4, qF = C / 4;                       // float4 per row
    const int nP = kDwKT * qP, nF = kDwKT * qF;              // float4 per chunk
    const float* gP = a.dP + (size_t)img * a.hw * N2;
    const float* gF = a.feats + (size_t)img * a.hw * C;
    for (int i = tid; i < kDwKT * kDwN; i += kDwThreads) (&sP[0][0])[i] = 0.f;
    for (int i = tid; i < kDwKT * kDwC; i += kDwThreads) (&sF[0][0])[i] = 0.f;
    f32x2 acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = pack2(0.f, 0.f);
    float4 rP[2], rF[4];
    auto fetch = [&](int p0) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int e = tid + i * kDwThreads;
            rP[i] = f4_zero();
            if (e < nP) {
                const int k = e / qP, q = e - k * qP;
                if (p0 + k < p_end) rP[i] = ldg_f4(gP + (size_t)(p0 + k) * N2 + q * 4);
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int e = tid + i * kDwThreads;
            rF[i] = f4_zero();
            if (e < nF) {
                const int k = e / qF, q = e - k * qF;
                if (p0 + k < p_end) rF[i] = ldg_f4(gF + (size_t)(p0 + k) * C + q * 4);
            }
        }
    };
    auto stash = [&]() {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int e = tid + i * kDwThreads;
            if (e < nP) { const int k = e / qP, q = e - k * qP; *reinterpret_cast<float4*>(&sP[k][q * 4]) = rP[i]; }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int e = tid + i * kDwThreads;
            if (e < nF) { const int k = e / qF, q = e - k * qF; *reinterpret_cast<float4*>(&sF[k][q * 4]) = rF[i]; }
        }
    };
    fetch(p_begin);
    __syncthreads();          // zero fill done
    for (int p0 = p_begin; p0 < p_end; p0 += kDwKT) {
        stash();
        __syncthreads();
        if (p0 + kDwKT < p_end) fetch(p0 + kDwKT);
#pragma unroll 4
        for (int k = 0; k < kDwKT; ++k) {
            const float4 a0 = *reinterpret_cast<const float4*>(&sP[k][tn * 8]);
            const float4 a1 = *reinterpret_cast<const float4*>(&sP[k][tn * 8 + 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            f32x2 f[8];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 v = *reinterpret_cast<const float4*>(&sF[k][tc * 4 + j * 64]);
                f[2 * j] = pack2(v.x, v.y);
                f[2 * j + 1] = pack2(v.z, v.w);
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const f32x2 s = pack2(av[i], av[i]);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fma2p(s, f[j], acc[i][j]);
            }
        }
        __syncthreads();
    }
    float* out = a.part + ((size_t)img * a.slabs + slab) * N2 * C;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int n = tn * 8 + i;
        if (n >= N2) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = tc * 4 + j * 64;
            if (c >= C) continue;
            float4 v;
            unpack2(acc[i][2 * j], v.x, v.y);
            unpack2(acc[i][2 * j + 1], v.z, v.w);
            *reinterpret_cast<float4*>(out + (size_t)n * C + c) = v;
        }
    }
}

int dense_dw_slabs(int n_img) {
    int s = (2 * 148 + n_img - 1) / n_img;
    return s < 1 ? 1 : s;
}

// ---------------------------------------------------------------------------------------------------------
// dWcat on tcgen05 (N2 == 128, C % 16 == 0):  D[n][c] += sum over pixels dP[p][n] * feats[p][c].  The contraction runs
// over the ROWS of both tensors, so both operands are MN-major: TMA boxes {32 channels, 32 pixels} land as [pixel][128 B]
// blocks with SWIZZLE_128B_ATOM_32B, which is the one MN-major layout kind::tf32 takes (umma_desc_mn).  One CTA per SM
// accumulates the [128 x C] product of its share of one camera group's pixels in TMEM (C columns) and writes it once.
//   warps 0-3  epilogue (after the last chunk): tcgen05.ld -> partial [cta][128][C]
//   warp  4    TMA producer: per 32-pixel chunk N2/32 + C/32 boxes into a stage (2 stages x 96 KB at C = 256)
//   warp  5    MMA issue: 4 K-steps of 8 pixels x {A_lo B_hi, A_hi B_lo, A_hi B_hi}
//   warps 6-13 split both operand tiles into TF32 hi (in place) + lo (second half of the stage)
// ---------------------------------------------------------------------------------------------------------
constexpr int kDtKP = 32;                        // pixels per chunk
constexpr int kDtThreads = 14 * VFD_WARP, kDtConvThreads = 8 * VFD_WARP;
constexpr int kDtBlock = kDtKP * 128;            // bytes of one {32 channels, 32 pixels} box

struct DwTcArgs {
    int hw, C, N2, cams, chunks_per_img;
    int n_group_img[2];               // images per overlap group
    int ncam[2];                      // cameras per group
    int camlist[2][VFD_MAX_CAMS];     // their indices
    int cta0[3];                      // CTAs [cta0[g], cta0[g+1]) work on group g
    float* part;                      // [ctas][N2][C]
};

struct DtBars {
    uint64_t tma_full[2], conv_full[2], empty[2], acc_full;
    uint32_t tmem_base;
};

static inline uint32_t dt_stage_bytes(int C, int N2) { return 2u * (uint32_t)(N2 / 32 + C / 32) * kDtBlock; }
static inline size_t dt_smem_bytes(int C, int N2) { return 1024 + 2 * (size_t)dt_stage_bytes(C, N2) + sizeof(DtBars); }

__global__ void __launch_bounds__(kDtThreads, 1)
    dense_dw_tc_kernel(const __grid_constant__ CUtensorMap mapP, const __grid_constant__ CUtensorMap mapF, const DwTcArgs a,
                       int tmem_cols) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int C = a.C, N2 = a.N2, nbA = N2 / 32, nbB = C / 32;
    const uint32_t half = (uint32_t)(nbA + nbB) * kDtBlock, sbytes = 2 * half;   // [A hi | B hi | A lo | B lo]
    DtBars* bars = reinterpret_cast<DtBars*>(base + 2 * (size_t)sbytes);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = (int)blockIdx.x >= a.cta0[1] ? 1 : 0;
    const int nc = a.cta0[g + 1] - a.cta0[g], ci = (int)blockIdx.x - a.cta0[g];
    const long items = (long)a.n_group_img[g] * a.chunks_per_img;
    const long it0 = items * ci / nc, it1 = items * (ci + 1) / nc;

    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            mbar_init(&bars->tma_full[s], 1);
            mbar_init(&bars->conv_full[s], kDtConvThreads);
            mbar_init(&bars->empty[s], 1);
        }
        mbar_init(&bars->acc_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        if (lane == 0) { tma_prefetch_desc(&mapP); tma_prefetch_desc(&mapF); }
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars->tmem_base)),
                     "r"(tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = bars->tmem_base;

    uint32_t stage = 0, phase = 0;
    if (warp == 4) {
        if (lane == 0) {
            for (long it = it0; it < it1; ++it) {
                const int ki = (int)(it / a.chunks_per_img), p0 = (int)(it - (long)ki * a.chunks_per_img) * kDtKP;
                const int img = (ki / a.ncam[g]) * a.cams + a.camlist[g][ki % a.ncam[g]];
                mbar_wait(&bars->empty[stage], phase ^ 1);
                unsigned char* st = base + (size_t)stage * sbytes;
                mbar_expect_tx(&bars->tma_full[stage], half);
                for (int k = 0; k < nbA; ++k) tma_load_3d(st + k * kDtBlock, &mapP, k * 32, p0, img, &bars->tma_full[stage]);
                for (int k = 0; k < nbB; ++k)
                    tma_load_3d(st + (nbA + k) * kDtBlock, &mapF, k * 32, p0, img, &bars->tma_full[stage]);
                if (++stage == 2) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 5) {
        const uint32_t idesc = umma_idesc_tf32_mn(N2, C);
        for (long it = it0; it < it1; ++it) {
            mbar_wait(&bars->conv_full[stage], phase);
            tc_fence_after();
            if (lane == 0) {
                const uint32_t s0 = smem_u32(base + (size_t)stage * sbytes);
#pragma unroll
                for (int ks = 0; ks < kDtKP / 8; ++ks) {     // 8 pixels = 8 rows of 128 bytes per K-step
                    const uint32_t o = (uint32_t)ks * 1024u;
                    const uint64_t a_hi = umma_desc_mn(s0 + o, kDtBlock), a_lo = umma_desc_mn(s0 + half + o, kDtBlock);
                    const uint64_t b_hi = umma_desc_mn(s0 + nbA * kDtBlock + o, kDtBlock);
                    const uint64_t b_lo = umma_desc_mn(s0 + half + nbA * kDtBlock + o, kDtBlock);
                    umma_tf32(tmem, a_lo, b_hi, idesc, (it > it0 || ks) ? 1u : 0u);
                    umma_tf32(tmem, a_hi, b_lo, idesc, 1u);
                    umma_tf32(tmem, a_hi, b_hi, idesc, 1u);
                }
                tc_commit(&bars->empty[stage]);
                if (it == it1 - 1) tc_commit(&bars->acc_full);
            }
            __syncwarp();
            if (++stage == 2) { stage = 0; phase ^= 1; }
        }
    } else if (warp >= 6) {
        const int ctid = tid - 6 * VFD_WARP;
        const int n16 = (int)(half / 16);
        for (long it = it0; it < it1; ++it) {
            mbar_wait(&bars->tma_full[stage], phase);
            uint4* ph = reinterpret_cast<uint4*>(base + (size_t)stage * sbytes);
            uint4* pl = reinterpret_cast<uint4*>(base + (size_t)stage * sbytes + half);
            for (int o = ctid; o < n16; o += kDtConvThreads) {
                const uint4 v = ph[o];
                uint4 h, l;
                split_tf32(__uint_as_float(v.x), h.x, l.x);
                split_tf32(__uint_as_float(v.y), h.y, l.y);
                split_tf32(__uint_as_float(v.z), h.z, l.z);
                split_tf32(__uint_as_float(v.w), h.w, l.w);
                ph[o] = h;
                pl[o] = l;
            }
            fence_proxy_async_smem();
            mbar_arrive(&bars->conv_full[stage]);
            if (++stage == 2) { stage = 0; phase ^= 1; }
        }
    } else {
        // epilogue: TMEM lane = n
        float* o = a.part + ((size_t)blockIdx.x * N2 + tid) * C;
        if (it1 > it0) {
            mbar_wait(&bars->acc_full, 0);
            tc_fence_after();
            for (int c0 = 0; c0 < C; c0 += 16) {
                uint32_t v[16];
                tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
                for (int q = 0; q < 16; q += 4)
                    *reinterpret_cast<float4*>(o + c0 + q) = make_float4(__uint_as_float(v[q]), __uint_as_float(v[q + 1]),
                                                                       __uint_as_float(v[q + 2]), __uint_as_float(v[q + 3]));
            }
        } else {
            for (int c0 = 0; c0 < C; c0 += 4) *reinterpret_cast<float4*>(o + c0) = f4_zero();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(tmem_cols));
}

bool dense_dw_tc_ok(int C, int N2) { return N2 == 128 && C % 32 == 0 && C >= 32 && C <= 256; }

// number of partial [N2][C] blocks dense_dw writes, and which overlap group each belongs to
int dense_dw_parts(int n_img, int C, int N2) {
    if (dense_dw_tc_ok(C, N2)) {
        const int sms = sm_count_pj();
        return sms > 0 ? sms : 148;
    }
    return n_img * dense_dw_slabs(n_img);
}

size_t dense_dw_part_floats(int n_img, int C, int N2) {
    int p = dense_dw_parts(n_img, C, N2);
    if (p < 256) p = 256;          // sized before the SM count is known to the caller's allocator: never too small
    return (size_t)p * N2 * C;
}

// -> *n_parts partial blocks [N2][C] in `part`.  *split >= 0 (tensor path): blocks [0, split) belong to overlap group 0,
// the rest to group 1.  *split < 0 (fp32 FMA path): block i belongs to the group of image i / dense_dw_slabs(n_img).
int dense_dw(int n_img, int hw, int C, int N2, int cams, const int* group, const float* dP, const float* feats, float* part,
             int* n_parts, int* split, cudaStream_t st) {
    if (N2 > kDwN || C > kDwC || N2 % 4 || C % 4) return VFD_ERR_UNSUPPORTED;
    if (dense_dw_tc_ok(C, N2)) {
        DwTcArgs a = {};
        a.hw = hw; a.C = C; a.N2 = N2; a.cams = cams; a.chunks_per_img = (hw + kDtKP - 1) / kDtKP; a.part = part;
        for (int c = 0; c < cams; ++c) {
            const int g = group[c] ? 1 : 0;
            a.camlist[g][a.ncam[g]++] = c;
        }
        const int B = n_img / cams;
        a.n_group_img[0] = B * a.ncam[0];
        a.n_group_img[1] = B * a.ncam[1];
        const int ctas = dense_dw_parts(n_img, C, N2);
        int c0 = (int)((long)ctas * a.n_group_img[0] / n_img);
        if (a.n_group_img[0] > 0 && c0 < 1) c0 = 1;
        if (a.n_group_img[1] > 0 && c0 > ctas - 1) c0 = ctas - 1;
        if (a.n_group_img[1] == 0) c0 = ctas;
        a.cta0[0] = 0; a.cta0[1] = c0; a.cta0[2] = ctas;
        *n_parts = ctas;
        *split = c0;
        CUtensorMap mp, mf;
        if (!make_map_3d(&mp, dP, (uint64_t)N2, (uint64_t)hw, (uint64_t)n_img, kDtKP, true) ||
            !make_map_3d(&mf, feats, (uint64_t)C, (uint64_t)hw, (uint64_t)n_img, kDtKP, true))
            return VFD_ERR_CUDA;
        const size_t smem = dt_smem_bytes(C, N2);
        if (smem > 227 * 1024) return VFD_ERR_UNSUPPORTED;
        if (cudaFuncSetAttribute(dense_dw_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return VFD_ERR_CUDA;
        int cols = 32;
        while (cols < C) cols <<= 1;
        dense_dw_tc_kernel<<<ctas, kDtThreads, smem, st>>>(mp, mf, a, cols);
        VFD_CHECK_LAUNCH();
        return VFD_OK;
    }
    DwArgs a = {};
    a.hw = hw; a.C = C; a.N2 = N2; a.slabs = dense_dw_slabs(n_img);
    a.slab_len = ((hw + a.slabs - 1) / a.slabs + kDwKT - 1) / kDwKT * kDwKT;
    a.dP = dP; a.feats = feats; a.part = part;
    *n_parts = n_img * a.slabs;
    *split = -1;
    dense_dw_kernel<<<dim3(a.slabs, n_img), kDwThreads, 0, st>>>(a);
    VFD_CHECK_LAUNCH();
    return VFD_OK;
}

}  // namespace vfd
