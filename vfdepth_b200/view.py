"""View synthesis + photometric loss on the CUDA kernels: the host-side mirror of
/root/reference/models/geometry/view_rendering.py (ViewRendering), models/geometry/pose.py:66-96 (relative camera
poses) and models/losses/{single_cam,multi_cam}_loss.py + loss_util.py (MultiCamLoss), at scale 0.

Three kernels replace the ~1 000 ATen launches per rig of the reference (csrc/warp_stats.cu, warp_loss_fwd.cu,
warp_loss_bwd.cu): pass 1 warps every source image and reduces the intensity-alignment statistics, pass 2 warps
again and goes straight to the masked loss sums, the backward re-warps and writes d_depth / d_T.  No warped image
reaches HBM unless ``materialize`` asks for the ``('color', f, 0)`` / ``('overlap', f, 0)`` views (logger, tests).

There is no CPU path: tensors must be CUDA fp32, the library must load, non-zero status codes raise.
"""
import ctypes

import torch

from . import _lib
from . import config as vcfg
from .ops import _chk, _stream

NBR = _lib.VFD_WARP_NBR


def num_warps(n_src):
    return n_src + (1 + n_src) * NBR


class WarpInputs:
    """The rig-level tensors every warp call reads, validated once: colour frames, mask, K, inv_K (scale 0)."""

    def __init__(self, cfg, inputs):
        t, d = cfg['training'], cfg['data']
        self.frame_ids = list(t['frame_ids'])
        self.n_src = len(self.frame_ids) - 1
        if self.frame_ids[0] != 0 or not 1 <= self.n_src <= _lib.VFD_WARP_SRC:
            raise NotImplementedError('frame_ids must be [0, f1] or [0, f1, f2]')
        self.num_cams = d['num_cams']
        self.H, self.W = t['height'], t['width']
        self.intensity_align = bool(t.get('intensity_align', True))
        self.rel = {c: [n for n in d['rel_cam_list'][c]][:NBR] for c in range(self.num_cams)}
        color0 = inputs[('color', 0, 0)]
        self.B = color0.shape[0]
        shp = (self.B, self.num_cams, 3, self.H, self.W)
        self.color = [_chk(inputs[('color', f, 0)], "inputs['color', %d, 0]" % f, shp) for f in self.frame_ids]
        mask = inputs['mask']
        if mask.dtype != torch.float32:
            mask = mask.float()
        self.mask = _chk(mask.reshape(self.B, self.num_cams, self.H, self.W), "inputs['mask']")
        self.K = _chk(inputs[('K', 0)].float().contiguous(), "inputs['K', 0]", (self.B, self.num_cams, 4, 4))
        self.invK = _chk(inputs[('inv_K', 0)].float().contiguous(), "inputs['inv_K', 0]", (self.B, self.num_cams, 4, 4))
        self.device = color0.device

    def desc(self, cam0, n_call):
        d = _lib.WarpDesc()
        d.B, d.num_cams, d.H, d.W = self.B, self.num_cams, self.H, self.W
        d.n_src, d.cam0, d.n_call = self.n_src, cam0, n_call
        d.intensity_align = 1 if self.intensity_align else 0
        for c in range(_lib.VFD_MAX_CAMS):
            for k in range(NBR):
                nb = self.rel.get(c, [])
                d.nbr[c][k] = nb[k] if k < len(nb) and nb[k] < self.num_cams else -1
        for i, t in enumerate(self.color):
            d.color[i] = t.data_ptr()
        d.mask, d.K, d.invK = self.mask.data_ptr(), self.K.data_ptr(), self.invK.data_ptr()
        return d


def warp_stats(wi, depth, T, cam0=0):
    """pass 1 -> (params [B,n_call,NW,8], raw): the intensity-alignment constants (autograd constants:
    view_rendering.py:45 runs under no_grad) and the scratch the loss passes read back -- the raw warps (r, g, b, mask)
    [B,n_call,NW,H,W,4] followed by one validity byte per pixel [B,n_call,H,W] (only warps that are not all-zero at a
    pixel are stored and re-read; vfd_warp_raw_bytes)."""
    n_call = depth.shape[1]
    NW = num_warps(wi.n_src)
    _chk(depth, 'depth', (wi.B, n_call, wi.H, wi.W))
    _chk(T, 'T', (wi.B, n_call, NW, 4, 4))
    stats = torch.empty((wi.B, n_call, NW, 8), device=wi.device, dtype=torch.float64)
    params = torch.empty((wi.B, n_call, NW, 8), device=wi.device, dtype=torch.float32)
    d = wi.desc(cam0, n_call)
    raw = torch.empty((_lib.load().vfd_warp_raw_bytes(ctypes.byref(d)),), device=wi.device, dtype=torch.uint8)
    _lib.call('vfd_warp_stats', ctypes.byref(d), _lib.ptr(depth), _lib.ptr(T), _lib.ptr(stats), _lib.ptr(params),
              _lib.ptr(raw), _stream())
    return params, raw


VIEW_KEYS = ('color', 'color_mask', 'overlap', 'overlap_mask', 'reproj_loss', 'reproj_mask', 'grid', 'selection')


def _alloc_views(wi, n_call, want):
    n, H, W, B = wi.n_src, wi.H, wi.W, wi.B
    shapes = {'color': (B, n_call, n, 3, H, W), 'color_mask': (B, n_call, n, H, W),
              'overlap': (B, n_call, 1 + n, 3, H, W), 'overlap_mask': (B, n_call, 1 + n, H, W),
              'reproj_loss': (B, n_call, H, W), 'reproj_mask': (B, n_call, H, W),
              'grid': (B, n_call, num_warps(n), H, W, 2), 'selection': (B, n_call, H, W)}
    out = {k: torch.empty(shapes[k], device=wi.device, dtype=torch.float32) for k in want}
    v = _lib.WarpViews()
    for k, t in out.items():
        setattr(v, k, t.data_ptr())
    return v, out


class _WarpLoss(torch.autograd.Function):
    """(depth, T) -> masked-loss numerators [n_call,3] (differentiable) and denominators [n_call,3]."""

    fuse_training = True    # False: separate vfd_warp_loss_fwd / vfd_warp_loss_bwd launches (tests compare the two)

    @staticmethod
    def forward(ctx, depth, T, wi, params, raw, noise, cam0, views_struct):
        n_call = depth.shape[1]
        _chk(depth, 'depth', (wi.B, n_call, wi.H, wi.W))
        _chk(T, 'T', (wi.B, n_call, num_warps(wi.n_src), 4, 4))
        _chk(noise, 'noise', (wi.B, n_call, wi.n_src, wi.H, wi.W))
        sums = torch.empty((n_call, 8), device=wi.device, dtype=torch.float64)
        d = wi.desc(cam0, n_call)
        ctx.wi, ctx.cam0 = wi, cam0
        # training: one launch computes the sums and the unit-upstream gradients (vfd_warp_loss_train_fwd)
        ctx.fused = views_struct is None and (ctx.needs_input_grad[0] or ctx.needs_input_grad[1]) and _WarpLoss.fuse_training
        if ctx.fused:
            nbytes = _lib.load().vfd_warp_bwd_workspace_bytes(ctypes.byref(d))
            ws = torch.empty((max(nbytes, 4) // 4,), device=wi.device, dtype=torch.float32)
            g3 = torch.empty((wi.B, n_call, 3, wi.H, wi.W), device=wi.device)
            _lib.call('vfd_warp_loss_train_fwd', ctypes.byref(d), _lib.ptr(depth), _lib.ptr(T), _lib.ptr(params),
                      _lib.ptr(raw), _lib.ptr(noise), _lib.ptr(sums), _lib.ptr(g3), _lib.ptr(ws), ctypes.c_size_t(nbytes),
                      _stream())
            ctx.nbytes, ctx.shapes = nbytes, (depth.shape, T.shape)
            ctx.save_for_backward(g3, ws)
        else:
            vptr = ctypes.byref(views_struct) if views_struct is not None else None
            _lib.call('vfd_warp_loss_fwd', ctypes.byref(d), _lib.ptr(depth), _lib.ptr(T), _lib.ptr(params),
                      _lib.ptr(raw), _lib.ptr(noise), _lib.ptr(sums), vptr, _stream())
            ctx.save_for_backward(depth, T, params, raw, noise)
        s = sums.float()
        num, den = s[:, 0:6:2].contiguous(), s[:, 1:6:2].contiguous()
        ctx.mark_non_differentiable(den)
        return num, den

    @staticmethod
    def backward(ctx, d_num, _):
        wi = ctx.wi
        d_num = d_num.float().contiguous()
        if ctx.fused:
            g3, ws = ctx.saved_tensors
            n_call = g3.shape[1]
            d = wi.desc(ctx.cam0, n_call)
            d_depth = torch.empty(ctx.shapes[0], device=wi.device)
            d_T = torch.empty(ctx.shapes[1], device=wi.device)
            _lib.call('vfd_warp_loss_train_bwd', ctypes.byref(d), _lib.ptr(d_num), _lib.ptr(g3), _lib.ptr(ws),
                      ctypes.c_size_t(ctx.nbytes), _lib.ptr(d_depth), _lib.ptr(d_T), _stream())
            return d_depth, d_T, None, None, None, None, None, None
        depth, T, params, raw, noise = ctx.saved_tensors
        n_call = depth.shape[1]
        d = wi.desc(ctx.cam0, n_call)
        d_depth = torch.empty_like(depth)
        d_T = torch.empty_like(T)
        nbytes = _lib.load().vfd_warp_bwd_workspace_bytes(ctypes.byref(d))
        ws = torch.empty((max(nbytes, 4) // 4,), device=wi.device, dtype=torch.float32)
        _lib.call('vfd_warp_loss_bwd', ctypes.byref(d), _lib.ptr(depth), _lib.ptr(T), _lib.ptr(params),
                  _lib.ptr(raw), _lib.ptr(noise), _lib.ptr(d_num), _lib.ptr(d_depth), _lib.ptr(d_T), _lib.ptr(ws),
                  ctypes.c_size_t(nbytes), _stream())
        return d_depth, d_T, None, None, None, None, None, None


def warp_loss_sums(wi, depth, T, noise, cam0=0, stats=None, want_views=()):
    """-> (num [n_call,3], den [n_call,3], views dict).  Columns: reprojection, spatial, spatio-temporal
    (single_cam_loss.py:55, multi_cam_loss.py:29, :59 are num / (den + 1e-8))."""
    depth = depth.contiguous()
    T = T.contiguous()
    params, raw = stats if stats is not None else warp_stats(wi, depth.detach(), T.detach(), cam0)
    vstruct, views = (None, {})
    if want_views:
        vstruct, views = _alloc_views(wi, depth.shape[1], tuple(want_views))
    num, den = _WarpLoss.apply(depth, T, wi, params, raw, noise, cam0, vstruct)
    return num, den, views


# --------------------------------------------------------------------------------------------------
# pose plumbing (models/geometry/pose.py:66-96): 4x4 products on [B,4,4] -- stays torch, differentiable
# --------------------------------------------------------------------------------------------------
_INDEX_CACHE = {}


def _device_index(idx, device):
    """a Python index list as a cached device tensor (indexing with a list costs a pageable host->device copy per
    call, which also cannot be captured in a CUDA graph)"""
    key = (tuple(idx), str(device))
    if key not in _INDEX_CACHE:
        _INDEX_CACHE[key] = torch.tensor(list(idx), dtype=torch.long, device=device)
    return _INDEX_CACHE[key]


def rig_transforms(wi, inputs, cam_T_cam):
    """cam_T_cam {f: [B,cams,4,4]} -> T [B,cams,NW,4,4] in the kernel's warp order (include/vfdepth_b200.h):
    temporal warps first, then for every frame the two neighbours' relative poses ext_inv[nbr] @ ext[cam] (@ T_f)."""
    E, Einv = inputs['extrinsics'].float(), inputs['extrinsics_inv'].float()
    B, ncam = wi.B, wi.num_cams
    eye = torch.eye(4, device=E.device, dtype=E.dtype).expand(B, ncam, 4, 4)
    cols = [cam_T_cam[f] for f in wi.frame_ids[1:]]
    rel0 = []
    for k in range(NBR):
        idx = [wi.rel[c][k] if k < len(wi.rel[c]) and wi.rel[c][k] < ncam else c for c in range(ncam)]
        rel0.append(torch.matmul(Einv[:, _device_index(idx, E.device)], E))                      # pose.py:81
    for fi, f in enumerate(wi.frame_ids):
        for k in range(NBR):
            cols.append(rel0[k] if fi == 0 else torch.matmul(rel0[k], cam_T_cam[f]))   # pose.py:95
    del eye
    return torch.stack(cols, 2)


def _expected_rel_keys(wi, cam, spatio=True, spatio_temporal=True):
    """the keys Pose.compute_relative_cam_poses (pose.py:66-96) writes for camera ``cam``"""
    nbrs = [n for n in wi.rel[cam] if n < wi.num_cams]
    keys = [(0, n) for n in nbrs] if spatio else []
    if spatio_temporal:
        keys += [(f, n) for f in wi.frame_ids[1:] for n in nbrs]
    return keys


def camera_transforms(wi, target_view, rel_pose_dict, cam):
    """one camera: T [B,1,NW,4,4] from ('cam_T_cam', 0, f) and the reference's rel_pose_dict.  A neighbour pose the
    reference would read (view_rendering.py:176, :186) but the dict lacks raises KeyError, as it does there; slots of
    neighbours that do not exist (partial rigs, pose.py:78-79) hold the identity and are inactive in the kernels."""
    cols = [target_view[('cam_T_cam', 0, f)] for f in wi.frame_ids[1:]]
    eye = torch.eye(4, device=wi.device).expand(wi.B, 4, 4)
    for f in wi.frame_ids:
        for k in range(NBR):
            nb = wi.rel[cam][k] if k < len(wi.rel[cam]) else -1
            if 0 <= nb < wi.num_cams:
                if (f, nb) not in rel_pose_dict:
                    raise KeyError((f, nb))
                cols.append(rel_pose_dict[(f, nb)])
            else:
                cols.append(eye)
    return torch.stack(cols, 1).unsqueeze(1).float()


class _SmoothLoss(torch.autograd.Function):
    """vfd_smooth_loss_fwd / _bwd: edge-aware smoothness of every camera of the batch, gradient to disp."""

    fuse_training = True    # False: separate forward / backward stencil passes (tests compare the two)

    @staticmethod
    def forward(ctx, color, disp):
        B, n, H, W = disp.shape
        if color is not None:
            _chk(color, 'color', (B, n, 3, H, W))
        _chk(disp, 'disp', (B, n, H, W))
        dev = disp.device
        img_sum = torch.empty((B * n,), device=dev, dtype=torch.float64)
        sums = torch.empty((n, 2), device=dev, dtype=torch.float64)
        ctx.color = color
        ctx.train = bool(ctx.needs_input_grad[1]) and _SmoothLoss.fuse_training
        if ctx.train:      # loss sums and the unit-upstream gradient map from one stencil pass
            gdot = torch.empty((B * n,), device=dev, dtype=torch.float64)
            g = torch.empty_like(disp)
            _lib.call('vfd_smooth_loss_train_fwd', B, n, H, W, _lib.ptr(color), _lib.ptr(disp), _lib.ptr(img_sum),
                      _lib.ptr(sums), _lib.ptr(gdot), _lib.ptr(g), _stream())
            ctx.save_for_backward(img_sum, gdot, g)
        else:
            _lib.call('vfd_smooth_loss_fwd', B, n, H, W, _lib.ptr(color), _lib.ptr(disp), _lib.ptr(img_sum),
                      _lib.ptr(sums), _stream())
            ctx.save_for_backward(disp, img_sum)
        out = sums[:, 0] / float(B * H * (W - 1)) + sums[:, 1] / float(B * (H - 1) * W)
        return out.float()

    @staticmethod
    def backward(ctx, d_smooth):
        d_smooth = d_smooth.float().contiguous()
        if ctx.train:
            img_sum, gdot, g = ctx.saved_tensors
            B, n, H, W = g.shape
            d_disp = torch.empty_like(g)
            _lib.call('vfd_smooth_loss_train_bwd', B, n, H, W, _lib.ptr(img_sum), _lib.ptr(gdot), _lib.ptr(g),
                      _lib.ptr(d_smooth), _lib.ptr(d_disp), _stream())
            return None, d_disp
        disp, img_sum = ctx.saved_tensors
        color = ctx.color
        B, n, H, W = disp.shape
        gdot = torch.empty((B * n,), device=disp.device, dtype=torch.float64)
        d_disp = torch.empty_like(disp)
        _lib.call('vfd_smooth_loss_bwd', B, n, H, W, _lib.ptr(color), _lib.ptr(disp), _lib.ptr(img_sum),
                  _lib.ptr(d_smooth), _lib.ptr(gdot), _lib.ptr(d_disp), _stream())
        return None, d_disp


def smooth_loss(color, disp):
    """single_cam_loss.py:57-65 + loss_util.py:28-40 for all cameras at once.
    color [B,cams,3,H,W] (None: no edge weighting, the depth_sm_loss of depth_synthesis_loss.py:35-41),
    disp [B,cams,H,W] -> per-camera loss [cams]."""
    return _SmoothLoss.apply(None if color is None else color.contiguous(), disp.contiguous())


class _DepthSyn(torch.autograd.Function):
    """vfd_depth_syn_fwd / _bwd: masked consistency sums of the warped source depths against the novel-view depth."""

    @staticmethod
    def forward(ctx, depth, aug, T, Tinv, desc, keep, views):
        d = desc
        sums = torch.empty((d.n_call, 2), device=depth.device, dtype=torch.float64)
        out_d, out_m = views if views is not None else (None, None)
        _lib.call('vfd_depth_syn_fwd', ctypes.byref(d), _lib.ptr(depth), _lib.ptr(aug), _lib.ptr(T), _lib.ptr(Tinv),
                  _lib.ptr(sums), _lib.ptr(out_d), _lib.ptr(out_m), _stream())
        ctx.desc, ctx.keep = d, keep
        ctx.save_for_backward(depth, aug, T, Tinv)
        num, den = sums[:, 0].float(), sums[:, 1].float()
        ctx.mark_non_differentiable(den)
        return num, den

    @staticmethod
    def backward(ctx, d_num, _d_den):
        depth, aug, T, Tinv = ctx.saved_tensors
        d_aug, d_depth = torch.empty_like(aug), torch.empty_like(depth)
        _lib.call('vfd_depth_syn_bwd', ctypes.byref(ctx.desc), _lib.ptr(depth), _lib.ptr(aug), _lib.ptr(T),
                  _lib.ptr(Tinv), _lib.ptr(d_num.float().contiguous()), _lib.ptr(d_aug), _lib.ptr(d_depth), _stream())
        return d_depth, d_aug, None, None, None, None, None


def depth_syn_loss(cfg, inputs, depth, aug_depth, disp_aug, cam0=0, want_views=False):
    """The aug_depth losses for target cameras [cam0, cam0 + n_call) of the whole rig batch:
    view_rendering.py:201-241 (get_virtual_depth :84-116 for rel_cam_list[cam] + [cam]) fused with
    depth_synthesis_loss.py:15-43.
    depth [B,cams,H,W] = ('depth', 0) of every camera; aug_depth / disp_aug [B,n_call,H,W] = ('depth', 0, 'aug') /
    ('disp', 0, 'aug') of the targets; inputs needs 'mask', 'extrinsics', 'extrinsics_aug', ('K', 0), ('inv_K', 0).
    -> (depth_con_loss [n_call], depth_sm_loss [n_call], views {'tform_depth', 'tform_depth_mask'} [B,n_call,3,H,W])."""
    t = cfg['training']
    n = cfg['data']['num_cams']
    B, cams, H, W = depth.shape
    n_call = aug_depth.shape[1]
    _chk(depth, 'depth', (B, n, H, W))
    _chk(aug_depth, 'aug_depth', (B, n_call, H, W))
    mask = inputs['mask'].reshape(B, n, H, W).contiguous()
    K, invK = inputs[('K', 0)].contiguous(), inputs[('inv_K', 0)].contiguous()
    d = _lib.SynDesc()
    d.B, d.num_cams, d.H, d.W, d.cam0, d.n_call = B, n, H, W, cam0, n_call
    d.min_depth, d.max_depth = float(t['min_depth']), float(t['max_depth'])
    d.mask, d.K, d.invK = mask.data_ptr(), K.data_ptr(), invK.data_ptr()
    rel = cfg['data']['rel_cam_list']
    idx = []
    for c in range(n):
        srcs = ([s if s < n else -1 for s in list(rel[c]) + [c]] + [-1] * 3)[:3]
        for i in range(3):
            d.src[c][i] = srcs[i]
        idx.append([max(s, 0) for s in srcs])
    sel = _device_index([s_ for row in idx[cam0:cam0 + n_call] for s_ in row], depth.device).view(n_call, 3)
    aug_inv = torch.inverse(inputs['extrinsics_aug'][:, cam0:cam0 + n_call])         # :206
    T = (aug_inv[:, :, None] @ inputs['extrinsics'][:, sel]).contiguous()            # :224  [B,n_call,3,4,4]
    Tinv = torch.inverse(T).contiguous()                                             # :96
    views = None
    if want_views:
        views = (torch.empty((B, n_call, 3, H, W), device=depth.device), torch.empty((B, n_call, 3, H, W), device=depth.device))
    num, den = _DepthSyn.apply(depth.contiguous(), aug_depth.contiguous(), T, Tinv, d, (mask, K, invK), views)
    con = num / (den + 1e-8)
    sm = smooth_loss(None, disp_aug)
    return con, sm, ({'tform_depth': views[0], 'tform_depth_mask': views[1]} if views else {})


class _ToDepth(torch.autograd.Function):
    """vfd_to_depth_fwd / _bwd: decoder output -> metric depth for every camera of the batch, gradient to disp."""

    @staticmethod
    def forward(ctx, disp, K, min_depth, max_depth, focal, H, W):
        B, n, hs, ws = disp.shape
        _chk(disp, 'disp', (B, n, hs, ws))
        _chk(K, 'K', (B, n, 4, 4))
        depth = torch.empty((B, n, H, W), device=disp.device)
        ctx.args = (B, n, hs, ws, H, W, ctypes.c_float(min_depth), ctypes.c_float(max_depth), ctypes.c_float(focal))
        B, n, hs, ws, H, W, lo, hi, fs = ctx.args
        _lib.call('vfd_to_depth_fwd', B, n, hs, ws, H, W, _lib.ptr(disp), _lib.ptr(K), lo, hi, fs, _lib.ptr(depth),
                  _stream())
        ctx.save_for_backward(disp, K)
        return depth

    @staticmethod
    def backward(ctx, d_depth):
        disp, K = ctx.saved_tensors
        B, n, hs, ws, H, W, lo, hi, fs = ctx.args
        d_disp = torch.empty_like(disp)
        _lib.call('vfd_to_depth_bwd', B, n, hs, ws, H, W, _lib.ptr(disp), _lib.ptr(K), lo, hi, fs,
                  _lib.ptr(d_depth.contiguous()), _lib.ptr(d_disp), _stream())
        return d_disp, None, None, None, None, None, None


def to_depth(cfg, disp, K):
    """models/vfdepth.py:277-288 for all cameras at once.
    disp [B,cams,hs,ws] (decoder output of one scale), K [B,cams,4,4] = inputs['K', 0] -> depth [B,cams,H,W]."""
    t = cfg['training']
    return _ToDepth.apply(disp.contiguous(), K.contiguous(), float(t['min_depth']), float(t['max_depth']),
                          float(t['focal_length_scale']), int(t['height']), int(t['width']))


def combine_losses(cfg, num, den, smooth, spatial=True):
    """per-camera total of multi_cam_loss.py:118-137 (``spatial=False``: single_cam_loss.py:79-80) from the kernel's
    sums. num/den [n,3], smooth [n] -> [n]"""
    L = cfg['loss']
    terms = num / (den + 1e-8)
    cam = terms[:, 0] + L['disparity_smoothness'] * smooth
    if spatial:
        cam = cam + (L['spatio_coeff'] * terms[:, 1] + L['spatio_tempo_coeff'] * terms[:, 2])
    return cam / len(cfg['training']['scales']), terms


def warp_and_loss(cfg, inputs, depths, disps, cam_T_cam, noises, want_views=()):
    """The view-synthesis half of a training step for a whole rig batch in three launches
    (models/vfdepth.py:290-313: loop over cameras, mean over cameras).
    depths / disps / noises: list over cameras of [B,1,H,W] / [B,1,H,W] / [B,n_src,H,W] (the reference's per-camera
    outputs), or the already stacked [B,cams,H,W] / [B,cams,H,W] / [B,cams,n_src,H,W]; cam_T_cam {f: [B,cams,4,4]}.
    -> (total loss, dict(terms [cams,3], cam_loss [cams], views))."""
    wi = WarpInputs(cfg, inputs)
    depth = depths if torch.is_tensor(depths) else torch.cat(depths, 1)
    noise = noises if torch.is_tensor(noises) else torch.stack(noises, 1)
    T = rig_transforms(wi, inputs, cam_T_cam)
    num, den, views = warp_loss_sums(wi, depth, T, noise, 0, None, want_views)
    color0 = inputs[('color', 0, 0)]
    sm = smooth_loss(color0, disps if torch.is_tensor(disps) else torch.cat(disps, 1))
    cam_loss, terms = combine_losses(cfg, num, den, sm)
    return cam_loss.sum() / wi.num_cams, dict(terms=terms, cam_loss=cam_loss, smooth=sm, views=views, T=T)


# --------------------------------------------------------------------------------------------------
# drop-in modules with the reference's constructor / forward signatures
# --------------------------------------------------------------------------------------------------
_CTX = '_vfd_warp'      # private key in outputs[('cam', cam)] that hands pass-1 results to the loss module


class ViewRendering(torch.nn.Module):
    """Drop-in for models/geometry/view_rendering.py:8 (``ViewRendering(cfg, rank)``;
    ``forward(inputs, outputs, cam, rel_pose_dict)`` mutating ``outputs[('cam', cam)]``).

    The reference calls this once per camera (models/vfdepth.py:296-301).  When the first call of a step finds the depth
    and the ``cam_T_cam`` of EVERY camera in ``outputs`` -- the reference's own call order guarantees it
    (compute_depth_maps and predict_pose run before compute_losses, vfdepth.py:218-232) -- pass 1 (warps + intensity
    statistics) runs for the whole rig in one launch and the other five calls find their camera already done; the
    loss module then evaluates all cameras in ONE training-mode launch (the path bench.py times).  Otherwise the
    camera is processed on its own.  The reference's image keys ``('color', f, s)``, ``('color_mask', f, s)``,
    ``('overlap', f, s)``, ``('overlap_mask', f, s)`` are written only when ``materialize`` is true (the reference's
    logger reads them at checkpoints only, utils/logger.py:120-170) -- they are non-differentiable copies; gradients
    flow through the fused loss."""

    def __init__(self, cfg, rank, materialize=False):
        super().__init__()
        vcfg.flatten_into(self, cfg)
        self.cfg, self.rank, self.materialize = cfg, rank, materialize
        # aug_depth (view_rendering.py:201-241): the warped source depths are produced, fused with their loss, by
        # DepthSynLoss.forward, which also writes ('tform_depth', 0) / ('tform_depth_mask', 0)
        if list(self.scales) != [0]:
            raise NotImplementedError('only scales == [0] (all reference configs)')

    @staticmethod
    def _rig_key(wi, outputs):
        """identity of the step's per-camera tensors, or None when some camera's depth / poses are not there yet"""
        key = []
        for c in range(wi.num_cams):
            tv = outputs.get(('cam', c))
            if tv is None or ('depth', 0) not in tv or any(('cam_T_cam', 0, f) not in tv for f in wi.frame_ids[1:]):
                return None
            key.append(id(tv[('depth', 0)]))
            key += [id(tv[('cam_T_cam', 0, f)]) for f in wi.frame_ids[1:]]
        return tuple(key)

    def forward(self, inputs, outputs, cam, rel_pose_dict):
        wi = inputs.get(_CTX)
        if wi is None or wi.color[0].data_ptr() != inputs[('color', 0, 0)].data_ptr():
            wi = inputs[_CTX] = WarpInputs(self.cfg, inputs)
        tv = outputs[('cam', cam)]
        for k in _expected_rel_keys(wi, cam, getattr(self, 'spatio', True), getattr(self, 'spatio_temporal', True)):
            if k not in rel_pose_dict:
                raise KeyError(k)          # the reference reads rel_pose_dict[(frame, neighbour)] (view_rendering.py:176)
        key = self._rig_key(wi, outputs)
        rig = outputs.get(_CTX)
        if key is not None and 'extrinsics_inv' in inputs and (rig is None or rig['key'] != key):
            n = wi.num_cams
            depth = torch.cat([outputs[('cam', c)][('depth', 0)].reshape(wi.B, 1, wi.H, wi.W) for c in range(n)], 1)
            cam_T_cam = {f: torch.stack([outputs[('cam', c)][('cam_T_cam', 0, f)] for c in range(n)], 1)
                         for f in wi.frame_ids[1:]}
            T = rig_transforms(wi, inputs, cam_T_cam)          # = pose.py:66-96 for every camera
            stats = warp_stats(wi, depth.detach().contiguous(), T.detach().contiguous(), 0)
            rig = outputs[_CTX] = dict(key=key, wi=wi, depth=depth, T=T, stats=stats, loss={})
        if rig is not None and rig['key'] == key and key is not None:
            tv[_CTX] = dict(wi=wi, rig=rig)
        else:
            depth = tv[('depth', 0)].reshape(wi.B, 1, wi.H, wi.W)
            T = camera_transforms(wi, tv, rel_pose_dict, cam)
            stats = warp_stats(wi, depth.detach().contiguous(), T.detach().contiguous(), cam)
            tv[_CTX] = dict(wi=wi, depth=depth, T=T, stats=stats)
        if self.materialize:
            self.render(inputs, outputs, cam)

    def get_virtual_image(self, src_img, src_mask, tar_depth, tar_invK, src_K, T, scale=0):
        """view_rendering.py:61-82 for one (source, target) pair: (warped image, valid * warped mask); differentiable
        w.r.t. tar_depth, T and src_K (csrc/warp_image.cu)."""
        from . import projection
        return projection.get_virtual_image(src_img, src_mask, tar_depth, tar_invK, src_K, T, scale)

    def get_norm_image_single(self, src_img, src_mask, warp_img, warp_mask):
        """view_rendering.py:39-59 for one warped image per batch element (csrc/warp_image.cu)."""
        from . import projection
        return projection.get_norm_image_single(src_img, src_mask, warp_img, warp_mask)

    @torch.no_grad()
    def get_virtual_depth(self, src_depth, src_mask, src_invK, src_K, tar_depth, tar_invK, tar_K, T, min_depth,
                          max_depth, scale=0):
        """view_rendering.py:84-116 for one (source, target) pair: (source depth seen from the target view, clamped to
        [min_depth, max_depth]; its mask).  Forward-only: the differentiable form is the fused consistency loss
        (``depth_syn_loss`` / ``DepthSynLoss``), which this calls with a two-camera descriptor (camera 1 = source)."""
        B, _, H, W = src_depth.shape
        c = lambda t: t.float().contiguous()
        depth = c(torch.cat([tar_depth, src_depth], 1))
        mask = c(torch.cat([torch.zeros_like(src_mask), src_mask], 1))
        K, invK = c(torch.stack([tar_K, src_K], 1)), c(torch.stack([tar_invK, src_invK], 1))
        _chk(depth, 'depth', (B, 2, H, W))
        _chk(mask, 'mask', (B, 2, H, W))
        d = _lib.SynDesc()
        d.B, d.num_cams, d.H, d.W, d.cam0, d.n_call = B, 2, H, W, 0, 1
        d.min_depth, d.max_depth = float(min_depth), float(max_depth)
        d.mask, d.K, d.invK = mask.data_ptr(), K.data_ptr(), invK.data_ptr()
        for cam in range(2):
            for i in range(3):
                d.src[cam][i] = 1 if (cam == 0 and i == 0) else -1
        T3 = c(T[:, None, None].expand(B, 1, 3, 4, 4))
        Tinv3 = c(torch.inverse(T3))
        out_d = torch.zeros((B, 1, 3, H, W), device=depth.device)
        out_m = torch.zeros_like(out_d)
        sums = torch.empty((1, 2), device=depth.device, dtype=torch.float64)
        _lib.call('vfd_depth_syn_fwd', ctypes.byref(d), _lib.ptr(depth), _lib.ptr(c(tar_depth)), _lib.ptr(T3),
                  _lib.ptr(Tinv3), _lib.ptr(sums), _lib.ptr(out_d), _lib.ptr(out_m), _stream())
        return out_d[:, 0, 0:1].contiguous(), out_m[:, 0, 0:1].contiguous()

    @torch.no_grad()
    def render(self, inputs, outputs, cam):
        """write the reference's image keys for camera ``cam`` (on demand)."""
        tv = outputs[('cam', cam)]
        c = tv[_CTX]
        wi = c['wi']
        if 'rig' in c:          # the camera's slice of the rig pass; its statistics are recomputed (logging time only)
            depth = c['rig']['depth'][:, cam:cam + 1].detach().contiguous()
            T = c['rig']['T'][:, cam:cam + 1].detach().contiguous()
            stats = warp_stats(wi, depth, T, cam)
        else:
            depth, T, stats = c['depth'].detach(), c['T'].detach(), c['stats']
        noise = torch.zeros((wi.B, 1, wi.n_src, wi.H, wi.W), device=wi.device)
        _, _, v = warp_loss_sums(wi, depth, T, noise, cam, stats, ('color', 'color_mask', 'overlap', 'overlap_mask'))
        for k, f in enumerate(wi.frame_ids[1:]):
            tv[('color', f, 0)] = v['color'][:, 0, k]
            tv[('color_mask', f, 0)] = v['color_mask'][:, 0, k:k + 1]
        for k, f in enumerate(wi.frame_ids):
            tv[('overlap', f, 0)] = v['overlap'][:, 0, k]
            tv[('overlap_mask', f, 0)] = v['overlap_mask'][:, 0, k:k + 1]


class MultiCamLoss(torch.nn.Module):
    """Drop-in for models/losses/multi_cam_loss.py:9 (``MultiCamLoss(cfg, rank)``; ``forward(inputs, outputs, cam)``
    -> ``(cam_loss, loss_dict)``), pose_model == 'fusion' (no pose-consistency term, multi_cam_loss.py:113).

    After a rig-level ``ViewRendering`` pass the first call of a step evaluates ALL cameras in one training-mode launch
    (loss sums and gradients together) and one smoothness launch, reads the loss terms back once, and every call
    returns its camera's entry -- six calls cost what one does, and ``loss_dict`` needs one host synchronisation per
    step where the reference has four per camera (multi_cam_loss.py:127-130).

    ``materialize`` (default False): also write ``('reproj_loss', 0)`` / ``('reproj_mask', 0)`` into ``outputs`` (the
    reference's logger reads them at checkpoints; they cost an extra forward launch).  ``device_noise`` (default
    False): draw the tie-break noise of single_cam_loss.py:45-46 on the device instead of replaying the reference's
    CPU stream (``torch.randn`` per camera, in camera order) and shipping it.  ``host_loss_dict`` (default False): the
    reference fills ``loss_dict`` with Python floats (``.item()``, multi_cam_loss.py:127-130), which makes the host wait
    for the forward before it can enqueue the backward; here the entries are 0-dim device tensors (they add, divide and
    format like floats; the wait moves to whoever prints them) unless this is set."""

    def __init__(self, cfg, rank, materialize=False, device_noise=False, host_loss_dict=False):
        super().__init__()
        self.cfg, self.rank = cfg, rank
        self.materialize, self.device_noise, self.host_loss_dict = materialize, device_noise, host_loss_dict
        for k, v in cfg['loss'].items():
            setattr(self, k, v)
        self.scales, self.frame_ids = cfg['training']['scales'], cfg['training']['frame_ids']
        self.pose_model = cfg['model'].get('pose_model', 'fusion')
        if self.pose_model == 'fsm':
            raise NotImplementedError('FSM pose-consistency loss (multi_cam_loss.py:61-92) is outside the hot path')

    spatial_terms = True

    def _noise(self, wi, n_call):
        shape = (wi.B, wi.n_src, wi.H, wi.W)
        if self.device_noise:
            return torch.randn((wi.B, n_call) + shape[1:], device=wi.device)
        # one CPU draw per camera in camera order = the reference's stream (its six draws follow each other directly)
        return torch.stack([torch.randn(shape) for _ in range(n_call)], 1).to(wi.device)

    def _evaluate(self, inputs, wi, depth, T, stats, disp, color0, cam0, n_call):
        noise = self._noise(wi, n_call)
        num, den, _ = warp_loss_sums(wi, depth, T, noise, cam0, stats)
        sm = smooth_loss(color0, disp)
        cam_loss, terms = combine_losses(self.cfg, num, den, sm, spatial=self.spatial_terms)
        host = torch.cat([terms.detach(), sm.detach()[:, None]], 1)
        res = dict(cam_loss=cam_loss, host=host.cpu() if self.host_loss_dict else host, views=None)
        if self.materialize:
            with torch.no_grad():
                _, _, res['views'] = warp_loss_sums(wi, depth.detach(), T.detach(), noise, cam0, stats,
                                                    ('reproj_loss', 'reproj_mask'))
        return res

    def forward(self, inputs, outputs, cam):
        tv = outputs[('cam', cam)]
        c = tv[_CTX]
        wi = c['wi']
        if 'rig' in c:
            rig, i = c['rig'], cam
            res = rig['loss'].get(self.spatial_terms)
            if res is None:
                n = wi.num_cams
                disp = torch.cat([outputs[('cam', k)][('disp', 0)].reshape(wi.B, 1, wi.H, wi.W) for k in range(n)], 1)
                res = rig['loss'][self.spatial_terms] = self._evaluate(inputs, wi, rig['depth'], rig['T'], rig['stats'],
                                                                       disp, inputs[('color', 0, 0)], 0, n)
        else:
            i = 0
            res = self._evaluate(inputs, wi, c['depth'], c['T'], c['stats'], tv[('disp', 0)].reshape(wi.B, 1, wi.H, wi.W),
                                 inputs[('color', 0, 0)][:, cam:cam + 1], cam, 1)
        if res['views'] is not None:
            tv[('reproj_loss', 0)] = res['views']['reproj_loss'][:, i:i + 1]
            tv[('reproj_mask', 0)] = res['views']['reproj_mask'][:, i:i + 1]
        t = res['host'][i]
        num = float if self.host_loss_dict else (lambda v: v)
        loss_dict = {'reproj_loss': num(t[0]), 'smooth': num(t[3])}
        if self.spatial_terms:
            loss_dict.update(spatio_loss=num(t[1]), spatio_tempo_loss=num(t[2]))
        depth_log = tv[('depth', 0)].detach()           # base_loss.py:29-33
        loss_dict['depth/mean'], loss_dict['depth/max'], loss_dict['depth/min'] = \
            depth_log.mean(), depth_log.max(), depth_log.min()
        if cam == 0:
            pose_t = tv[('cam_T_cam', 0, -1)].detach() if ('cam_T_cam', 0, -1) in tv else None
            if pose_t is not None:
                for k, n in enumerate(('pose/tx', 'pose/ty', 'pose/tz')):
                    loss_dict[n] = pose_t[:, k, 3].abs().mean()
        return res['cam_loss'][i], loss_dict


class SingleCamLoss(MultiCamLoss):
    """Drop-in for models/losses/single_cam_loss.py:11 (``SingleCamLoss(cfg, rank)``, the loss of the configurations
    with ``training.spatio == False``, models/vfdepth.py:50-53): reprojection + disparity_smoothness * smoothness
    (:66-95).  Same kernels as MultiCamLoss -- the temporal warps, the auto-mask and the masked sums do not depend on
    the spatial warps, whose two terms are simply left out of the total and of ``loss_dict``."""
    spatial_terms = False


class DepthSynLoss(MultiCamLoss):
    """Drop-in for models/losses/depth_synthesis_loss.py:8 (``DepthSynLoss(cfg, rank)``; ``forward(inputs, outputs,
    cam)`` -> ``(cam_loss, loss_dict)``): the MultiCamLoss terms plus depth_con_coeff * consistency + depth_sm_coeff *
    smoothness of the novel-view prediction.  Reads ('depth', 0) of every camera, ('depth', 0, 'aug') and
    ('disp', 0, 'aug') of ``cam`` and inputs['extrinsics_aug'] (written by VFNet.forward); writes ('tform_depth', 0) /
    ('tform_depth_mask', 0) as the reference's ViewRendering does (view_rendering.py:240-241)."""

    def forward(self, inputs, outputs, cam):
        cam_loss, loss_dict = super().forward(inputs, outputs, cam)
        n = self.cfg['data']['num_cams']
        tv = outputs[('cam', cam)]
        depth = torch.cat([outputs[('cam', c)][('depth', 0)] for c in range(n)], 1)
        con, sm, v = depth_syn_loss(self.cfg, inputs, depth, tv[('depth', 0, 'aug')], tv[('disp', 0, 'aug')], cam,
                                    want_views=True)
        slots = [i for i, s in enumerate((list(self.cfg['data']['rel_cam_list'][cam]) + [cam])[:3]) if s < n]
        tv[('tform_depth', 0)] = [v['tform_depth'][:, 0, i:i + 1] for i in slots]
        tv[('tform_depth_mask', 0)] = [v['tform_depth_mask'][:, 0, i:i + 1] for i in slots]
        depthsyn = self.depth_con_coeff * con[0] + self.depth_sm_coeff * sm[0]
        cam_loss = cam_loss + depthsyn / len(self.scales)
        num = float if self.host_loss_dict else (lambda t: t)
        loss_dict.update(depth_loss=num(depthsyn.detach()), depth_sm_loss=num(sm[0].detach()),
                         depth_con_loss=num(con[0].detach()))
        loss_dict['cam_loss'] = num(cam_loss.detach())          # depth_synthesis_loss.py:90
        return cam_loss, loss_dict
