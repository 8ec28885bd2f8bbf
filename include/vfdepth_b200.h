/* vfdepth_b200 -- C ABI of the B200 (sm_100a) implementation of VFDepth's volumetric-fusion and
 * view-synthesis hot path.
 *
 * The reference (tronglh241/VFDepth) is pure PyTorch and has no FFI of its own; the boundary it offers is
 * the Python method surface of VFNet / Projection / ViewRendering / the loss modules.  Each entry point
 * below replaces the geometry core of one of those methods (file:line cited per function); the Python
 * host layer in vfdepth_b200/ mirrors the reference signatures on top of this ABI (INTEGRATION.md shows the
 * ctypes binding a maintainer adds to the reference).
 *
 * Conventions
 *  - plain pointers + sizes, no torch types; all pointers are DEVICE pointers unless stated otherwise;
 *  - every function returns 0 on success or a negative vfd_status; vfd_last_error(code) names it;
 *  - nothing allocates, frees or retains device memory; scratch comes from the caller;
 *  - everything is launched on the cudaStream_t passed in (as void*), nothing synchronises the device,
 *    no global mutable state: entry points are re-entrant (autograd threads, DDP streams);
 *  - fp32 everywhere (the reference's dtype); layouts are channel-last and stated per function.
 */
#ifndef VFDEPTH_B200_H_
#define VFDEPTH_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VFD_VERSION 200 /* major*100 + minor */
#define VFD_MAX_CAMS 8

typedef enum vfd_status {
    VFD_OK = 0,
    VFD_ERR_NULL = -1,      /* a required pointer is NULL */
    VFD_ERR_DESC = -2,      /* descriptor field out of range */
    VFD_ERR_ALIGN = -3,     /* a pointer / leading dimension violates the stated alignment */
    VFD_ERR_WORKSPACE = -4, /* workspace too small */
    VFD_ERR_CUDA = -5,      /* cudaGetLastError() after launch was not cudaSuccess */
    VFD_ERR_UNSUPPORTED = -6,
    VFD_ERR_STATIC_LIMIT = -7 /* size exceeds the fixed capacity of a no-allocation (static scratch) entry point */
} vfd_status;

int vfd_version(void);
const char* vfd_last_error(int code);

/* --------------------------------------------------------------------------------------------------
 * Fusion geometry (network/volumetric_fusionnet.py).
 * Voxel n = (z*Y + y)*X + x, x fastest (create_voxel_grid :67-80, flattened :24-29).
 * ------------------------------------------------------------------------------------------------ */
typedef struct vfd_fusion_desc {
    int32_t B;        /* rigs in the batch */
    int32_t num_cams; /* 3 or 6 (preprocess_overlap :218-225); <= VFD_MAX_CAMS */
    int32_t C;        /* feature channels (fusion_feat_in_dim); multiple of 4 */
    int32_t h, w;     /* feature-map size = image size / 2^(fusion_level+1) (:32-34) */
    int32_t X, Y, Z;  /* voxel_size */
    int32_t D;        /* proj_d_bins */
    int32_t Cv;       /* channels of the voxel volume the rays sample (voxel_pre_dim[-1]) */
    int32_t ld_no;    /* row stride (floats) of x_no / pose rows, >= C+1, multiple of 4 */
    int32_t ld_ov;    /* row stride (floats) of x_ov rows, >= 2*half_ov, multiple of 4 */
    int32_t half_ov;  /* offset (floats) of the second camera group inside an x_ov row, >= C+1, multiple of 4 */
    int32_t group[VFD_MAX_CAMS]; /* overlap group (0/1) of every camera: {0,1,1,0,0,1} for 6 cams (:222-223) */
    float depth_div;      /* voxel_size[0]: the relative-depth channel is v.z / depth_div (:145) */
    float vox_start[3];   /* voxel_str_p as fp32 (:256-258) */
    float vox_length[3];  /* voxel_end_p - voxel_str_p, computed in double then cast (:257) */
    const float* axis_x;  /* device, [X] = torch.linspace(str,end,X)  (:72) */
    const float* axis_y;  /* device, [Y] */
    const float* axis_z;  /* device, [Z] */
    const float* bins;    /* device, [D] = torch.linspace(proj_d_str, proj_d_end, D)  (:39) */
} vfd_fusion_desc;

/* VFNet.backproject_into_voxel, model='depth' (:116-159) up to the inputs of the two 1x1 MLPs (:197-230):
 *   feats      [B,cams,h,w,C]      channel-last feature maps (feats_agg of :124)
 *   mask_small [B,cams,h,w]        self-occlusion mask already resized to h x w (:129)
 *   K          [B,cams,4,4]        intrinsics at the feature scale (inputs['K', fusion_level+1])
 *   Einv       [B,cams,4,4]        ego->camera (inputs['extrinsics_inv'])
 *   x_no       [B,N,ld_no]         (sum_c feat_c) * [count==1]; cols C+1.. are written as 0
 *   x_ov       [B,N,ld_ov]         [sum_{group 0} feat_c | pad | sum_{group 1} feat_c | pad]
 *   count      [B,N]  uint8        number of cameras that see the voxel (:152)
 *   grid  (optional) [B,cams,N,2]  normalised sampling coordinates (:166-181), bit-exact
 *   valid (optional) [B,cams,N] u8 validity mask (:183-195), bit-exact
 * feat_c = [bilinear(feats_c) ; v.z/depth_div] * valid_c  (:143-146). */
int vfd_voxel_gather_depth_fwd(const vfd_fusion_desc* d, const float* feats, const float* mask_small,
                               const float* K, const float* Einv, float* x_no, float* x_ov, uint8_t* count,
                               float* grid, uint8_t* valid, void* stream);

/* backward of the above w.r.t. feats: d_feats [B,cams,h,w,C] is zero-filled here, then accumulated. */
int vfd_voxel_gather_depth_bwd(const vfd_fusion_desc* d, const float* mask_small, const float* K,
                               const float* Einv, const float* d_x_no, const float* d_x_ov, float* d_feats,
                               void* stream);

/* Which voxels feed which MLP (preprocess_non_overlap :197-207 keeps count==1, preprocess_overlap :209-230 keeps
 * count==2; the rest of the volume is zero).  Gives every count-1 voxel a row in a compact x_no matrix and every
 * count-2 voxel a row in a compact x_ov matrix, in voxel order (b major, then n) -- deterministic.
 *   count   [B,N] u8      cameras that see the voxel (:152), bit-exact
 *   slot    [B,N] i32     row of the voxel in ITS compact matrix, -1 when count is 0 or >= 3
 *   idx_no  [>= B*N] i64  out: idx_no[r] = b*N+n of compact x_no row r  (first totals[0] entries are written)
 *   idx_ov  [>= B*N] i64  same for x_ov (first totals[1] entries)
 *   totals  [2] i32       device: number of count-1 and count-2 voxels; the caller reads it back to size the MLP GEMMs
 *   table   (optional)    vfd_voxel_table_bytes(d) bytes, 16-byte aligned: for every listed voxel the bilinear corner
 *                         pixel indices / weights, camera and relative depth of the one or two cameras that see it, so that
 *                         vfd_voxel_fuse_fwd / _bwd do not project again; NULL = not written
 *   ws                    scratch of vfd_voxel_classify_workspace_bytes(d) bytes, 8-byte aligned
 * slot / idx_no / idx_ov / totals may be NULL together: only count (+ table) is produced, in one launch. */
size_t vfd_voxel_classify_workspace_bytes(const vfd_fusion_desc* d);
size_t vfd_voxel_table_bytes(const vfd_fusion_desc* d);
int vfd_voxel_classify(const vfd_fusion_desc* d, const float* mask_small, const float* K, const float* Einv,
                       uint8_t* count, int32_t* slot, int64_t* idx_no, int64_t* idx_ov, int32_t* totals, void* table,
                       void* ws, size_t ws_bytes, void* stream);

/* vfd_voxel_gather_depth_fwd / _bwd on compact rows: x_no [totals[0], ld_no] and x_ov [totals[1], ld_ov] hold one
 * row per voxel that reaches the corresponding MLP (row = slot[b,n]); voxels with slot < 0 are skipped entirely. */
int vfd_voxel_gather_depth_compact_fwd(const vfd_fusion_desc* d, const float* feats, const float* mask_small,
                                       const float* K, const float* Einv, const int32_t* slot, float* x_no,
                                       float* x_ov, void* stream);
int vfd_voxel_gather_depth_compact_bwd(const vfd_fusion_desc* d, const float* mask_small, const float* K,
                                       const float* Einv, const int32_t* slot, const float* d_x_no,
                                       const float* d_x_ov, float* d_feats, void* stream);

/* Fused gather + both 1x1 MLPs: VFNet.backproject_into_voxel, model='depth', :116-159 INCLUDING
 * preprocess_non_overlap (:197-207) and preprocess_overlap (:209-230), in "project, then gather" form: the MLPs are
 * linear up to their LeakyReLU and so is the bilinear gather, so every feature PIXEL is multiplied once with the weight
 * blocks its camera can need (dense tcgen05 + TMA product, 3xTF32 = fp32-level accuracy; B*cams*h*w rows instead of one
 * row per (voxel, camera)) and the voxel kernel gathers Cv-float rows of that product, adds the relative-depth column
 * and the bias and applies LeakyReLU(0.1).  Neither the GEMM rows nor the gathered rows reach HBM.
 *   count, table                    from vfd_voxel_classify (table is required)
 *   w_no [Cv][C+1], b_no [Cv]       conv_non_overlap.0.{weight,bias}   (Conv1d weight with the k=1 axis squeezed)
 *   w_ov [Cv][2(C+1)], b_ov [Cv]    conv_overlap.0.{weight,bias}
 *   vol  [B,N,Cv]                   channel-last voxel volume (voxel_feat of :158 transposed); every row is written:
 *                                   LeakyReLU_0.1(x W^T + b) for count-1 / count-2 voxels, zeros elsewhere
 *   ws                              scratch of vfd_voxel_fuse_fwd_workspace_bytes(d) bytes, 256-byte aligned
 * C multiple of 4, <= 256; Cv multiple of 4, <= 64.  Deterministic. */
size_t vfd_voxel_fuse_fwd_workspace_bytes(const vfd_fusion_desc* d);
int vfd_voxel_fuse_fwd(const vfd_fusion_desc* d, const float* feats, const uint8_t* count, const void* table,
                       const float* w_no, const float* b_no, const float* w_ov, const float* b_ov, float* vol, void* ws,
                       size_t ws_bytes, void* stream);

/* Backward of vfd_voxel_fuse_fwd w.r.t. the feature maps and the four MLP parameters.
 *   vol    [B,N,Cv]  the forward output (the sign of the pre-activation is read from it)
 *   d_vol  [B,N,Cv]  upstream gradient
 *   d_feats [B,cams,h,w,C]  written (every element once, by the dense product dP . Wcat)
 *   d_w_no [Cv][C+1], d_b_no [Cv], d_w_ov [Cv][2(C+1)], d_b_ov [Cv]  written
 *   ws     scratch of vfd_voxel_fuse_bwd_workspace_bytes(d) bytes, 256-byte aligned
 * The per-voxel gradients are scattered into the pixel-domain buffer dP with L2 atomics (summation order varies
 * between runs at the 1e-7 level); everything after that is reduced in a fixed order. */
size_t vfd_voxel_fuse_bwd_workspace_bytes(const vfd_fusion_desc* d);
int vfd_voxel_fuse_bwd(const vfd_fusion_desc* d, const float* feats, const uint8_t* count, const void* table,
                       const float* w_no, const float* w_ov, const float* vol, const float* d_vol, float* d_feats,
                       float* d_w_no, float* d_b_no, float* d_w_ov, float* d_b_ov, void* ws, size_t ws_bytes,
                       void* stream);

/* VFNet.backproject_into_voxel, model='pose' (:160-162):  avg [B,N,ld_no] = sum_c feat_c / (count + 1e-7) */
int vfd_voxel_gather_pose_fwd(const vfd_fusion_desc* d, const float* feats, const float* mask_small,
                              const float* K, const float* Einv, float* avg, uint8_t* count, void* stream);
int vfd_voxel_gather_pose_bwd(const vfd_fusion_desc* d, const float* mask_small, const float* K,
                              const float* Einv, const float* d_avg, float* d_feats, void* stream);

/* The pose gather with rows ordered (y, x, z) instead of the reference's (z, y, x): bev [B,Y,X,Z,ld_no] read as
 * [B,Y,X,Z*ld_no] is the channels-last form of the BEV map VFNet.forward hands to reduce_dim (:339-342; channel
 * z*ld_no + c there is channel c*Z + z of the reference, pad columns are zero), so the [B,257*Z,Y,X] reshape copy
 * (411 MB of traffic per rig and call, again in backward) disappears.  count stays in voxel order. */
int vfd_voxel_gather_pose_bev_fwd(const vfd_fusion_desc* d, const float* feats, const float* mask_small,
                                  const float* K, const float* Einv, float* bev, uint8_t* count, void* stream);
int vfd_voxel_gather_pose_bev_bwd(const vfd_fusion_desc* d, const float* mask_small, const float* K,
                                  const float* Einv, const float* d_bev, float* d_feats, void* stream);

/* VFNet.project_voxel_into_image without reduce_dim (:245-262):
 *   vol   [B,Z,Y,X,Cv]       channel-last voxel volume (voxel_feat of :240)
 *   invK  [B,cams,4,4]       inputs['inv_K', fusion_level+1];   E [B,cams,4,4] camera->ego
 *   out   [B,cams,h,w,D,Cv]  channel-last samples, c fastest.  The reference's channel order c*D + d (the .view
 *                            of :262) is a fixed permutation of reduce_dim's INPUT channels; the host layer applies
 *                            it to reduce_dim.0.weight instead of transposing the samples
 *   grid (optional) [B,cams,D,h,w,3]  normalised sampling coordinates (:251-260), bit-exact */
int vfd_ray_sample_fwd(const vfd_fusion_desc* d, const float* vol, const float* invK, const float* E,
                       float* out, float* grid, void* stream);
/* d_vol [B,Z,Y,X,Cv] is zero-filled here, then accumulated. */
int vfd_ray_sample_bwd(const vfd_fusion_desc* d, const float* invK, const float* E, const float* d_out,
                       float* d_vol, void* stream);

/* --------------------------------------------------------------------------------------------------
 * View synthesis + photometric loss (models/geometry/geometry_util.py:33-82, view_rendering.py:30-82,118-198,
 * models/losses/loss_util.py:6-25,43-78, single_cam_loss.py:17-55, multi_cam_loss.py:16-59,94-138), scale 0.
 *
 * For a target camera `cam` there are NW = n_src + (1+n_src)*VFD_WARP_NBR warps of a source image into its view:
 *   j <  n_src                       temporal: frame 1+j of the SAME camera          (view_rendering.py:133-157)
 *   j =  n_src + fi*VFD_WARP_NBR+k   frame fi (0 = current) of NEIGHBOUR nbr[cam][k]  (:160-198), summed over k
 * "frame fi" indexes desc.color[]: fi = 0 is frame 0, fi >= 1 is frame_ids[fi].
 * ------------------------------------------------------------------------------------------------ */
#define VFD_WARP_SRC 2 /* max source frames besides frame 0 */
#define VFD_WARP_NBR 2 /* neighbours per camera (utils/misc.py:10) */
#define VFD_WARP_MAXW (VFD_WARP_SRC + (1 + VFD_WARP_SRC) * VFD_WARP_NBR)

typedef struct vfd_warp_desc {
    int32_t B, num_cams, H, W;
    int32_t n_src;           /* len(frame_ids) - 1: 1 or 2 */
    int32_t cam0, n_call;    /* this call processes target cameras [cam0, cam0 + n_call) */
    int32_t intensity_align; /* training.intensity_align (view_rendering.py:148) */
    int32_t nbr[VFD_MAX_CAMS][VFD_WARP_NBR]; /* rel_cam_list: neighbour camera or -1 (skipped, :167-168) */
    const float* color[1 + VFD_WARP_SRC];    /* device, per frame [B,cams,3,H,W]: inputs['color', f, 0] */
    const float* mask;                       /* device [B,cams,H,W]: inputs['mask'] */
    const float* K;                          /* device [B,cams,4,4]: inputs['K', 0] */
    const float* invK;                       /* device [B,cams,4,4]: inputs['inv_K', 0] */
} vfd_warp_desc;

/* optional materialised views of one call (any pointer may be NULL); [B,n_call,...] */
typedef struct vfd_warp_views {
    float* color;        /* [B,n_call,n_src,3,H,W]    ('color', f, 0)       view_rendering.py:156 */
    float* color_mask;   /* [B,n_call,n_src,H,W]      ('color_mask', f, 0)  :157 */
    float* overlap;      /* [B,n_call,1+n_src,3,H,W]  ('overlap', f, 0)     :197 */
    float* overlap_mask; /* [B,n_call,1+n_src,H,W]    ('overlap_mask', f, 0) :198 (before multi_cam_loss.py:28) */
    float* reproj_loss;  /* [B,n_call,H,W]            ('reproj_loss', 0)    single_cam_loss.py:53 */
    float* reproj_mask;  /* [B,n_call,H,W]            ('reproj_mask', 0)    :54 */
    float* grid;         /* [B,n_call,NW,H,W,2]  normalised sampling coordinates (geometry_util.py:62-77), bit-exact */
    float* selection;    /* [B,n_call,H,W]  which frame won the two torch.min over frames: (reprojection index)
                            + 2 * (spatio-temporal index)  (single_cam_loss.py:30, multi_cam_loss.py:56) */
} vfd_warp_views;

/* Pass 1 -- intensity-alignment statistics (view_rendering.py:30-59) of every warp.
 *   depth  [B,n_call,H,W]        target depth (('depth', 0) of each camera)
 *   T      [B,n_call,NW,4,4]     target->source transforms (cam_T_cam / rel_pose_dict, pose.py:66-96)
 *   stats  [B,n_call,NW,8] f64   scratch, zero-filled here: {count, sum ref*m, sum warp*m, sum warp, sum warp^2,
 *                                sum ref, sum ref^2, -}
 *   params [B,n_call,NW,8] f32   out: {w_mean, w_std+1e-8, s_std, s_mean, apply, 1/(w_std+1e-8), finite, -}; apply = 0
 *                                when intensity_align is off or the batch-wide early exit of :50-53 fires for warp j;
 *                                finite = 1 when every parameter is finite
 *   raw    vfd_warp_raw_bytes(d) bytes, 16-byte aligned, out: [B,n_call,NW,H,W,4] floats = the un-normalised warp
 *                                (r, g, b) and warp mask of every source image (get_virtual_image, :61-82), followed by
 *                                [B,n_call,H,W] bytes = which of a pixel's NW warps are not all-zero.  Only those
 *                                entries are written, and only those are read back by the loss passes (a spatial
 *                                warp leaves the neighbour camera's image for most pixels of the target). */
size_t vfd_warp_raw_bytes(const vfd_warp_desc* d);
int vfd_warp_stats(const vfd_warp_desc* d, const float* depth, const float* T, double* stats, float* params,
                   float* raw, void* stream);

/* Pass 2 -- normalise the raw warps, SSIM+L1 maps, min over frames, auto-mask, masked sums.
 *   noise  [B,n_call,n_src,H,W]  the torch.randn of single_cam_loss.py:45-46 (drawn by the caller)
 *   sums   [n_call,8] f64        zero-filled here: {sum reproj*auto, sum auto, sum spatial*m, sum m,
 *                                sum spatio-temporal*m, sum m, -, -} over the whole local batch (loss_util.py:21-25)
 *   views  optional materialised images (NULL for the fused path) */
int vfd_warp_loss_fwd(const vfd_warp_desc* d, const float* depth, const float* T, const float* params,
                      const float* raw, const float* noise, double* sums, const vfd_warp_views* views, void* stream);

size_t vfd_warp_bwd_workspace_bytes(const vfd_warp_desc* d);

/* Backward of pass 2 w.r.t. depth and T (statistics are constants, view_rendering.py:45).
 *   d_num   [n_call,3] f32   dL/d(sum reproj*auto), dL/d(sum spatial*m), dL/d(sum spatio-temporal*m)
 *   d_depth [B,n_call,H,W]   written (every pixel exactly once, no atomics)
 *   d_T     [B,n_call,NW,4,4] written; per-block partial sums go through `ws` and are reduced in a fixed
 *                            order (deterministic) */
int vfd_warp_loss_bwd(const vfd_warp_desc* d, const float* depth, const float* T, const float* params,
                      const float* raw, const float* noise, const float* d_num, float* d_depth, float* d_T, void* ws,
                      size_t ws_bytes, void* stream);

/* Training mode of pass 2: ONE call evaluates the loss sums of vfd_warp_loss_fwd and the gradients of
 * vfd_warp_loss_bwd with unit upstream, per loss set (the backward recomputes the forward anyway).  Two launches: the
 * temporal set (reprojection + auto-mask) on every tile, then the spatial + spatio-temporal sets on the tiles a
 * neighbour camera's warp reaches (csrc/warp_loss_sets.cu).
 *   sums [n_call][8] f64 as vfd_warp_loss_fwd;  g3 [B,n_call,3,H,W] = d loss_set / d depth;  ws
 *   (vfd_warp_bwd_workspace_bytes, 16-byte aligned) = the d_T partial sums -- to be kept until
 *   vfd_warp_loss_train_bwd -- followed by what the first launch hands to the second (a flag byte per tile, the
 *   auto-mask decision byte per pixel).  vfd_warp_loss_train_bwd applies
 *   d_num [n_call,3] = d total / d num_set:  d_depth = sum_set d_num * g3,  d_T likewise (fixed-order fp64 reduction). */
int vfd_warp_loss_train_fwd(const vfd_warp_desc* d, const float* depth, const float* T, const float* params,
                            const float* raw, const float* noise, double* sums, float* g3, void* ws, size_t ws_bytes,
                            void* stream);
int vfd_warp_loss_train_bwd(const vfd_warp_desc* d, const float* d_num, const float* g3, const void* ws, size_t ws_bytes,
                            float* d_depth, float* d_T, void* stream);

/* --------------------------------------------------------------------------------------------------
 * Edge-aware disparity smoothness (models/losses/single_cam_loss.py:57-65, loss_util.py:28-40), scale 0, all cameras
 * of the batch at once.
 *   color   [B,cams,3,H,W]  inputs['color', 0, 0]; NULL = no edge weighting (the depth_sm_loss of
 *                           depth_synthesis_loss.py:35-41 on ('disp', 0, 'aug'))
 *   disp    [B,cams,H,W]    ('disp', 0) of every camera
 *   img_sum [B*cams] f64    out: sum of disp per image (the mean of single_cam_loss.py:63), kept for the backward
 *   sums    [cams][2] f64   out: sum over the batch of |d nd / dx| * exp(-|d rgb / dx|) and of the y term; the loss of
 *                           camera c is sums[c][0] / (B*H*(W-1)) + sums[c][1] / (B*(H-1)*W)   (the two .mean() of :40)
 * Both buffers are zero-filled here. */
int vfd_smooth_loss_fwd(int B, int cams, int H, int W, const float* color, const float* disp, double* img_sum,
                        double* sums, void* stream);
/* No-allocation variant of vfd_smooth_loss_fwd.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  The per-image disparity sums -- the `img_sum` scratch the entry point above takes from the caller -- live in a
 * file-scope `static __device__` array inside the library (csrc/smooth_loss.cu), VFD_SMOOTH_STATIC_MAX_IMAGES doubles
 * = 32 KB that every GPU which loads the library holds until the process exits.  Same arguments and results otherwise:
 * color [B,cams,3,H,W] (may be NULL), disp [B,cams,H,W], sums [cams][2] f64 out (zero-filled here).
 *   - B * cams > VFD_SMOOTH_STATIC_MAX_IMAGES is rejected with VFD_ERR_STATIC_LIMIT before anything is launched.
 *   - Calls no cudaMalloc / cudaFree / allocating library routine: the array's device address is looked up with
 *     cudaGetSymbolAddress once per device (the device current at the call), then memset + two launches on `stream`.
 *   - Every call on a device shares the array, which keeps the previous call's contents: the entries a call uses are
 *     cleared on `stream` first.  Calls that overlap on the GPU (different streams, no ordering) race on it. */
#define VFD_SMOOTH_STATIC_MAX_IMAGES 4096
int vfd_smooth_loss_fwd_static(int B, int cams, int H, int W, const float* color, const float* disp, double* sums,
                               void* stream);
/* backward w.r.t. disp:  d_smooth [cams] upstream gradient of the per-camera loss, gdot [B*cams] f64 scratch
 * (zero-filled here), d_disp [B,cams,H,W] written. */
int vfd_smooth_loss_bwd(int B, int cams, int H, int W, const float* color, const float* disp, const double* img_sum,
                        const float* d_smooth, double* gdot, float* d_disp, void* stream);

/* Training mode of the smoothness: the means, then ONE stencil pass that yields the loss sums, the gradient map
 * g [B,cams,H,W] with unit upstream and gdot [B*cams] f64; vfd_smooth_loss_train_bwd applies d_smooth [cams]:
 * d_disp = d_smooth[c] * (g / den - gdot / (den^2 HW)).  color may be NULL as above. */
int vfd_smooth_loss_train_fwd(int B, int cams, int H, int W, const float* color, const float* disp, double* img_sum,
                              double* sums, double* gdot, float* g, void* stream);
int vfd_smooth_loss_train_bwd(int B, int cams, int H, int W, const double* img_sum, const double* gdot, const float* g,
                              const float* d_smooth, float* d_disp, void* stream);

/* --------------------------------------------------------------------------------------------------
 * Disparity -> metric depth (models/vfdepth.py:277-288 to_depth, called per camera and scale at :266-275), all cameras
 * of the batch at once:
 *   d = bilinear upsample (align_corners=False) of disp [B,cams,hs,ws] to [H,W]  (identity when hs==H, ws==W)
 *   depth [B,cams,H,W] = (1 / (1/max_depth + (1/min_depth - 1/max_depth) * d)) * K[b,c,0,0] / focal_length_scale
 *   K [B,cams,4,4] = inputs['K', 0]. */
int vfd_to_depth_fwd(int B, int cams, int hs, int ws, int H, int W, const float* disp, const float* K, float min_depth,
                     float max_depth, float focal_length_scale, float* depth, void* stream);
/* backward w.r.t. disp: d_depth [B,cams,H,W] -> d_disp [B,cams,hs,ws] (zero-filled here when it is accumulated). */
int vfd_to_depth_bwd(int B, int cams, int hs, int ws, int H, int W, const float* disp, const float* K, float min_depth,
                     float max_depth, float focal_length_scale, const float* d_depth, float* d_disp, void* stream);

/* --------------------------------------------------------------------------------------------------
 * Depth-synthesis consistency loss of the aug_depth configuration:
 * ViewRendering.get_virtual_depth (models/geometry/view_rendering.py:84-116) for the sources
 * rel_cam_list[cam] + [cam] (:201-241) fused with the consistency part of DepthSynLoss.compute_aug_losses
 * (models/losses/depth_synthesis_loss.py:15-33).
 *   depth  [B,cams,H,W]       ('depth', 0) of every camera (sources)
 *   aug    [B,n_call,H,W]     ('depth', 0, 'aug') of the target cameras [cam0, cam0 + n_call)
 *   T      [B,n_call,3,4,4]   inverse(extrinsics_aug[cam]) @ extrinsics[src]  (source camera -> novel view), slot i = src[cam][i]
 *   Tinv   [B,n_call,3,4,4]   inverse(T)  (torch.inverse, view_rendering.py:96)
 *   sums   [n_call][2] f64    out (zero-filled here): sum loss*mask, sum mask over the sources and the batch;
 *                             depth_con_loss = sums[c][0] / (sums[c][1] + 1e-8)  (loss_util.py:21-25)
 *   out_depth / out_mask [B,n_call,3,H,W] or NULL: ('tform_depth', 0) / ('tform_depth_mask', 0) for the logger */
#define VFD_SYN_SRC 3
typedef struct vfd_syn_desc {
    int32_t B, num_cams, H, W;
    int32_t cam0, n_call;
    int32_t src[VFD_MAX_CAMS][VFD_SYN_SRC]; /* source camera of slot i, -1 = skipped (:210-212) */
    float min_depth, max_depth;             /* training.min_depth / max_depth (:109-113) */
    const float* mask;                      /* device [B,cams,H,W]  inputs['mask'] */
    const float* K;                         /* device [B,cams,4,4]  inputs['K', 0] */
    const float* invK;                      /* device [B,cams,4,4]  inputs['inv_K', 0] */
} vfd_syn_desc;
int vfd_depth_syn_fwd(const vfd_syn_desc* d, const float* depth, const float* aug, const float* T, const float* Tinv,
                      double* sums, float* out_depth, float* out_mask, void* stream);
/* backward: wcam [n_call] = upstream gradient of sums[c][0]; d_aug [B,n_call,H,W] written, d_depth [B,cams,H,W]
 * zero-filled here and accumulated (bilinear scatter). */
int vfd_depth_syn_bwd(const vfd_syn_desc* d, const float* depth, const float* aug, const float* T, const float* Tinv,
                      const float* wcam, float* d_aug, float* d_depth, void* stream);

/* --------------------------------------------------------------------------------------------------
 * Stand-alone warp of ONE (source, target) pair per batch element: the sub-method surface of the reference that the
 * fused rig kernels bypass.  Same bit-exact coordinates and sampler as vfd_warp_loss_*.
 *   Projection.forward (models/geometry/geometry_util.py:79-82)      vfd_warp_image_fwd with src_img = NULL: grid only
 *   ViewRendering.get_virtual_image (view_rendering.py:61-82)        vfd_warp_image_fwd: img (NaN -> 2.0), valid * mask
 *   Projection.backproject / .reproject (:52-60 / :62-77)            vfd_projection_backproject / _reproject
 *   src_img [B,3,H,W], src_mask [B,1,H,W], depth [B,1,H,W] (target), invK [B,4,4] (target), K [B,4,4] (source),
 *   T [B,4,4] (target -> source); img [B,3,H,W], mask [B,1,H,W], grid [B,H,W,2] (optional with an image).
 * Backward (d_img and / or d_grid given): d_depth [B,1,H,W] written; dP_part [B, vfd_warp_image_blocks(), 12] per-CTA
 * partial sums of the gradient w.r.t. P = (K @ T)[:3,:] -- the caller adds them in order (deterministic) and forms
 * d_T = K[:3,:]^T dP, d_K[:3,:] = dP T^T.  No gradient to src_img (an input of the training step). */
int vfd_warp_image_blocks(void);
int vfd_warp_image_fwd(int B, int H, int W, const float* src_img, const float* src_mask, const float* depth,
                       const float* invK, const float* K, const float* T, float* img, float* mask, float* grid,
                       void* stream);
int vfd_warp_image_bwd(int B, int H, int W, const float* src_img, const float* src_mask, const float* depth,
                       const float* invK, const float* K, const float* T, const float* d_img, const float* d_grid,
                       float* d_depth, float* dP_part, void* stream);
/* points [B,4,H*W] = depth * (invK[:3,:3] @ (u,v,1)), row 3 = 1;  grid [B,H,W,2] from such points (row 3 taken as 1) */
int vfd_projection_backproject(int B, int H, int W, const float* depth, const float* invK, float* points, void* stream);
int vfd_projection_reproject(int B, int H, int W, const float* points, const float* K, const float* T, float* grid,
                             void* stream);

/* ViewRendering.get_norm_image_single (view_rendering.py:39-59) for one warped image per batch element: intensity
 * alignment of warp_img to the target image over the overlap (ref_mask * warp_mask), batch-wide early exit when any
 * element has no overlap.  stats [B,8] f64 scratch (zero-filled here, kept for the backward); out [B,3,H,W].
 * Backward w.r.t. warp_img only (the statistics are constants, as under torch.no_grad in the reference). */
int vfd_norm_image_fwd(int B, int H, int W, const float* ref_img, const float* ref_mask, const float* warp_img,
                       const float* warp_mask, double* stats, float* out, void* stream);
int vfd_norm_image_bwd(int B, int H, int W, const float* warp_mask, const double* stats, const float* d_out,
                       float* d_warp, void* stream);

/* loss_util.compute_ssim_loss / compute_photometric_loss (models/losses/loss_util.py:43-78) of two images:
 *   ssim  [B,3,H,W] (optional) = clamp((1 - SSIM_3x3,reflect) / 2, 0, 1) per channel
 *   photo [B,1,H,W] (optional) = 0.85 * mean_c ssim + 0.15 * mean_c |target - pred|
 * backward w.r.t. pred for upstreams d_ssim and / or d_photo; d_pred [B,3,H,W] zero-filled here, accumulated. */
int vfd_photometric_loss_fwd(int B, int H, int W, const float* pred, const float* target, float* ssim, float* photo,
                             void* stream);
int vfd_photometric_loss_bwd(int B, int H, int W, const float* pred, const float* target, const float* d_ssim,
                             const float* d_photo, float* d_pred, void* stream);

/* loss_util.compute_auto_masks (models/losses/loss_util.py:6-18): mask [B,1,HW] = 1 where the first arg-min over
 * cat[reproj [B,F1,HW], identity [B,F2,HW]] along the channel is channel 0 (torch.argmin order: first minimum, NaN
 * smallest), else 0.  F2 = 0 / identity = NULL compares the reprojection channels only.  No gradient. */
int vfd_auto_masks(int B, int F1, int F2, int HW, const float* reproj, const float* identity, float* mask, void* stream);

/* loss_util.compute_masked_loss (:21-25) over n elements: sums [2] f64 (zero-filled here) = sum(loss * mask), sum(mask);
 * the value is sums[0] / (sums[1] + 1e-8).  backward: d_loss [n] = d_out[0] * mask / (sums[1] + 1e-8), d_out a device
 * scalar. */
int vfd_masked_loss_fwd(size_t n, const float* loss, const float* mask, double* sums, void* stream);
int vfd_masked_loss_bwd(size_t n, const float* mask, const double* sums, const float* d_out, float* d_loss, void* stream);

/* loss_util.compute_edg_smooth_loss (:28-40) WITHOUT the mean normalisation of single_cam_loss.py:57-65 (which
 * vfd_smooth_loss_* fuses): rgb [B,C,H,W], disp [B,1,H,W]; sums [2] f64 (zero-filled here) = sum of
 * |disp[x] - disp[x+1]| exp(-mean_c |rgb[x] - rgb[x+1]|) and of the y term; the value is
 * sums[0] / (B*H*(W-1)) + sums[1] / (B*(H-1)*W).  backward w.r.t. disp (d_out a device scalar), one write per pixel. */
int vfd_edg_smooth_loss_fwd(int B, int C, int H, int W, const float* rgb, const float* disp, double* sums, void* stream);
int vfd_edg_smooth_loss_bwd(int B, int C, int H, int W, const float* rgb, const float* disp, const float* d_out,
                            float* d_disp, void* stream);

/* --------------------------------------------------------------------------------------------------
 * Multi-level feature aggregation in front of the fusion (network/fusion_depthnet.py:57-66,
 * network/fusion_posenet.py:62-70):  LeakyReLU(conv1x1(cat[f_lev, up(f_lev+1), ...]) + b), up = F.interpolate(bilinear,
 * align_corners=True).  conv1x1 and the resize commute, so the caller projects each level at its own resolution
 * (channel-last GEMMs with the column blocks of the conv1x1 weight) and this entry point resizes, adds, applies bias and
 * LeakyReLU and writes the channel-last map the voxel kernels read:
 *   y0  [n_img,h,w,C]            projection of the finest level (no bias)
 *   yc  host array of n_coarse device pointers, yc[l] = [n_img,ch[l],cw[l],C]
 *   out [n_img,h,w,C]            feats_agg, channel-last; C % 4 == 0, C <= 1024, 16-byte aligned pointers. */
#define VFD_AGG_LEVELS 3
typedef struct vfd_agg_desc {
    int32_t n_img, h, w, C;
    int32_t n_coarse;
    int32_t ch[VFD_AGG_LEVELS], cw[VFD_AGG_LEVELS];
    float slope;                            /* LeakyReLU negative slope (0.1, network/blocks.py:47) */
} vfd_agg_desc;
int vfd_feat_agg_fwd(const vfd_agg_desc* d, const float* y0, const float* const* yc, const float* bias, float* out,
                     void* stream);
/* backward: d_y0 [n_img,h,w,C] = d_out * (out > 0 ? 1 : slope) (its column sums are the bias gradient), d_yc[l] written
 * (gathered from d_y0 through the transposed resize; deterministic). */
int vfd_feat_agg_bwd(const vfd_agg_desc* d, const float* out, const float* d_out, float* d_y0, float* const* d_yc,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VFDEPTH_B200_H_ */
