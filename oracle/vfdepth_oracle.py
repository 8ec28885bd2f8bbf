"""ORACLE -- test infrastructure, not product code.

A CPU (torch, fp32 or fp64) restatement of the VFDepth fusion + view-synthesis
hot path, function by function, each citing the reference lines it follows.
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this file; the product package
``vfdepth_b200`` never does.

Pinning: the reference has no golden vectors of its own, so the
oracle is pinned by fixtures produced by *running the reference's own Python*
(/root/reference imported read-only, see ``oracle/ref_import.py`` and
``tests/golden/make_golden.py``) and committed under ``tests/golden/``;
``tests/test_oracle_golden.py`` checks the oracle against them bit for bit,
and, where /root/reference is mounted, against a live run of the reference.

The oracle deliberately issues the same ATen calls in the same order as the
reference (``matmul``, ``grid_sample``, ``avg_pool2d`` ...) so that on the CPU
it is bit-identical to it; arithmetic that lives in ATen (grid_sample's
un-normalise / floor / nearbyint) is therefore the real thing,
not a re-derivation.
"""
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------------------
# A0  constant grids            (network/volumetric_fusionnet.py:22-40, 67-103)
# --------------------------------------------------------------------------------------
def voxel_end(cfg):
    m = cfg['model']
    return [m['voxel_str_p'][i] + m['voxel_unit_size'][i] * (m['voxel_size'][i] - 1) for i in range(3)]


def voxel_axes(cfg):
    """three 1-D linspace axes (x, y, z) -- volumetric_fusionnet.py:72."""
    m = cfg['model']
    end = voxel_end(cfg)
    return [torch.linspace(m['voxel_str_p'][i], end[i], m['voxel_size'][i]) for i in range(3)]


def voxel_points(cfg, batch):
    """homogeneous voxel centres [B,4,N], x fastest then y then z -- :67-80, :24-29."""
    ax = voxel_axes(cfg)
    X, Y, Z = cfg['model']['voxel_size']
    gx = ax[0].view(1, 1, X).expand(Z, Y, X)
    gy = ax[1].view(1, Y, 1).expand(Z, Y, X)
    gz = ax[2].view(Z, 1, 1).expand(Z, Y, X)
    pts = torch.stack([gx, gy, gz, torch.ones(Z, Y, X)], 0).reshape(1, 4, Z * Y * X)
    return pts.expand(batch, 4, Z * Y * X).contiguous()


def feature_hw(cfg):
    t, lvl = cfg['training'], cfg['model']['fusion_level'] + 1
    return t['height'] // (2 ** lvl), t['width'] // (2 ** lvl)


def pixel_grid(batch, h, w):
    """(u, v, 1) with u fastest, [B,3,h*w] -- :82-91."""
    v, u = torch.meshgrid(torch.arange(h), torch.arange(w), indexing='ij')
    g = torch.stack([u.reshape(-1), v.reshape(-1), torch.ones(h * w, dtype=torch.long)], 0)
    return g.unsqueeze(0).repeat(batch, 1, 1).float()


def depth_bins(cfg):
    """:39 -- torch.linspace is NOT start+i*step in fp32; always take it from torch."""
    m = cfg['model']
    return torch.linspace(m['proj_d_str'], m['proj_d_end'], m['proj_d_bins'])


# --------------------------------------------------------------------------------------
# A2/A3  voxel -> pixel coordinates and validity   (:166-195)
# --------------------------------------------------------------------------------------
def voxel_pixel_coords(K, v_local, w_dim, h_dim, eps=1e-8):
    """-> normalised sampling grid [B,N,1,2] -- :166-181."""
    B, _, N = v_local.shape
    cam = torch.matmul(K[:, :3, :3], v_local)
    pix = cam[:, :2, :] / (cam[:, 2, :].unsqueeze(1) + eps)
    if not torch.all(torch.isfinite(pix)):
        pix = torch.clamp(pix, min=-w_dim * 2, max=w_dim * 2)
    pix = pix.view(B, 2, N, 1).permute(0, 2, 3, 1)
    pix[:, :, :, 0] = pix[:, :, :, 0] / (w_dim - 1)
    pix[:, :, :, 1] = pix[:, :, :, 1] / (h_dim - 1)
    return (pix - 0.5) * 2


def voxel_valid_mask(mask_small, grid, v_local):
    """self-occlusion (nearest, >0.5) * in-front * in-bounds(strict) -> bool [B,1,N] -- :183-195."""
    selfocc = F.grid_sample(mask_small, grid, mode='nearest', padding_mode='zeros', align_corners=True) > 0.5
    front = v_local[:, 2:3, :] > 0
    g = grid.permute(0, 3, 1, 2)
    inb = ~(torch.logical_or(g > 1, g < -1).sum(dim=1, keepdim=True) > 0)
    return selfocc.squeeze(-1) * front * inb.squeeze(-1)


def unnormalize(g, size):
    """ATen grid_sampler_unnormalize, align_corners=True: ((g+1)/2)*(size-1)."""
    return ((g + 1) / 2) * (size - 1)


# --------------------------------------------------------------------------------------
# A1/A4  voxel gather            (:116-164, 197-230)
# --------------------------------------------------------------------------------------
def gather_camera(cfg, feats_img, mask_img, K, E_inv, vox_pts):
    """one camera of the loop at :124-149. -> dict(feat [B,C+1,N], valid bool [B,1,N], grid, v_local)."""
    h_dim, w_dim = feats_img.shape[-2:]
    mask_small = F.interpolate(mask_img, [h_dim, w_dim], mode='bilinear', align_corners=True)
    v_local = torch.matmul(E_inv[:, :3, :], vox_pts)
    grid = voxel_pixel_coords(K, v_local, w_dim, h_dim)
    valid = voxel_valid_mask(mask_small, grid, v_local)
    feat = F.grid_sample(feats_img, grid, mode='bilinear', padding_mode='zeros', align_corners=True)
    feat = torch.cat([feat.squeeze(-1), v_local[:, 2:3, :] / cfg['model']['voxel_size'][0]], dim=1)
    feat = feat * valid.float()
    return {'feat': feat, 'valid': valid, 'grid': grid, 'v_local': v_local, 'mask_small': mask_small}


def overlap_groups(num_cams):
    """:218-225"""
    if num_cams == 3:
        return [0], [1, 2]
    if num_cams == 6:
        return [0, 3, 4], [1, 2, 5]
    raise NotImplementedError


def backproject_into_voxel(cfg, feats_agg, input_mask, K, E_inv, model='depth', mlp=None, detail=False,
                           slopes_from=None):
    """-> voxel_feat [B,64,N] (depth; needs ``mlp`` = dict(w_no,b_no,w_ov,b_ov)) or [B,C+1,N] (pose).

    With ``detail`` also returns the intermediates the parity tests compare: per-camera grids, valid masks,
    the camera count, and the two MLP inputs ``x_no`` [B,C+1,N], ``x_ov`` [B,2(C+1),N].
    """
    B = feats_agg.shape[0]
    ncam = cfg['data']['num_cams']
    vox = voxel_points(cfg, B).to(device=feats_agg.device, dtype=feats_agg.dtype)   # the reference's type_check (:105-114)
    cams = [gather_camera(cfg, feats_agg[:, c], input_mask[:, c], K[:, c], E_inv[:, c], vox) for c in range(ncam)]
    feats = [c['feat'] for c in cams]
    count = torch.sum(torch.cat([c['valid'] for c in cams], dim=1), dim=1, keepdim=True)
    out = {}
    if model == 'depth':
        no_mask = (count == 1)
        x_no = sum(feats) * no_mask.float()
        ga, gb = overlap_groups(ncam)
        x_ov = torch.cat([sum(feats[i] for i in ga), sum(feats[i] for i in gb)], dim=1)
        ov_mask = (count == 2)
        if mlp is not None:
            z_no, z_ov = F.conv1d(x_no, mlp['w_no'], mlp['b_no']), F.conv1d(x_ov, mlp['w_ov'], mlp['b_ov'])
            if slopes_from is None:
                y_no, y_ov = F.leaky_relu(z_no, 0.1) * no_mask.float(), F.leaky_relu(z_ov, 0.1) * ov_mask.float()
            else:
                # test aid: LeakyReLU with the slope pattern PINNED to another implementation's output [B,Cv,N] (same
                # values except where a pre-activation is ~0 and two fp32 implementations pick different slopes)
                slope = torch.where(slopes_from > 0, 1.0, 0.1)
                y_no, y_ov = z_no * slope * no_mask.float(), z_ov * slope * ov_mask.float()
            voxel_feat = y_no + y_ov
        else:
            voxel_feat = None
        out.update(x_no=x_no, x_ov=x_ov)
    elif model == 'pose':
        voxel_feat = torch.sum(torch.stack(feats, dim=1), dim=1) / (count + 1e-7)
    else:
        raise ValueError(model)
    if not detail:
        return voxel_feat
    out.update(voxel_feat=voxel_feat, count=count, grids=[c['grid'] for c in cams],
               valids=[c['valid'] for c in cams], v_local=[c['v_local'] for c in cams],
               mask_small=[c['mask_small'] for c in cams])
    return out


# --------------------------------------------------------------------------------------
# A5  ray sampling               (:232-262, without the reduce_dim convs of :265)
# --------------------------------------------------------------------------------------
def ray_grid(cfg, inv_K, E, batch, dtype=torch.float32):
    # constant grids are built on the CPU and moved, as the reference's type_check (:105-114) does
    """one camera: normalised 3-D sampling grid [B,D,h,w,3] -- :245-260."""
    m = cfg['model']
    h, w = feature_hw(cfg)
    P, D = h * w, m['proj_d_bins']
    pix = pixel_grid(batch, h, w).to(device=inv_K.device, dtype=dtype)
    bins = depth_bins(cfg).to(device=inv_K.device, dtype=dtype)
    dgrid = bins.view(1, 1, D, 1).expand(batch, 3, D, P)
    ones = torch.ones(batch, 1, D, P, dtype=dtype, device=inv_K.device)
    cam = torch.matmul(inv_K[:, :3, :3], pix)
    cam = dgrid * cam.view(batch, 3, 1, P)
    cam = torch.cat([cam, ones], dim=1).view(batch, 4, -1)
    pts = torch.matmul(E[:, :3, :], cam)
    grid = pts.permute(0, 2, 1)
    end = voxel_end(cfg)
    for i in range(3):
        length = end[i] - m['voxel_str_p'][i]
        grid[:, :, i] = (grid[:, :, i] - m['voxel_str_p'][i]) / length * 2. - 1.
    return grid.view(batch, D, h, w, 3)


def project_voxel_into_image(cfg, voxel_feat, inv_K, E, detail=False):
    """voxel_feat [B,Cv,N] -> list over cameras of [B, Cv*D, h, w] (channel = c*D + d) -- :232-262."""
    B, Cv, _ = voxel_feat.shape
    X, Y, Z = cfg['model']['voxel_size']
    h, w = feature_hw(cfg)
    D = cfg['model']['proj_d_bins']
    vol = voxel_feat.view(B, Cv, Z, Y, X)
    outs, grids = [], []
    for c in range(cfg['data']['num_cams']):
        grid = ray_grid(cfg, inv_K[:, c], E[:, c], B, voxel_feat.dtype)
        s = F.grid_sample(vol, grid, mode='bilinear', padding_mode='zeros', align_corners=True)
        outs.append(s.view(B, D * Cv, h, w))
        grids.append(grid)
    return (outs, grids) if detail else outs


# --------------------------------------------------------------------------------------
# A7  Projection                 (models/geometry/geometry_util.py:33-82)
# --------------------------------------------------------------------------------------
def homo_points(batch, height, width, dtype=torch.float32):
    """(u, v, 1), u fastest, [B,3,H*W] -- geometry_util.py:44-50."""
    v, u = torch.meshgrid(torch.arange(height), torch.arange(width), indexing='ij')
    pts = torch.stack([u.reshape(-1), v.reshape(-1)], 0).to(dtype).repeat(batch, 1, 1)
    return torch.cat([pts, torch.ones(batch, 1, height * width, dtype=dtype)], 1)


def reproject_grid(depth, T, inv_K, K):
    """depth [B,1,H,W], T/inv_K/K [B,4,4] -> normalised grid [B,H,W,2] -- :52-82.
    Note the pre-multiplied (K @ T) of :67."""
    B, _, H, W = depth.shape
    homo = homo_points(B, H, W, depth.dtype).to(depth.device)
    pts = depth.view(B, 1, -1) * torch.matmul(inv_K[:, :3, :3], homo)
    pts = torch.cat([pts, homo[:, 2:3]], 1)
    p2 = (K @ T)[:, :3, :] @ pts
    n = p2[:, :2, :] / (p2[:, 2:, :] + 1e-7)
    n = n.view(B, 2, H, W).permute(0, 2, 3, 1)
    n[..., 0] /= W - 1
    n[..., 1] /= H - 1
    return (n - 0.5) * 2


# --------------------------------------------------------------------------------------
# A8/A9  warp + intensity alignment   (models/geometry/view_rendering.py:30-82)
# --------------------------------------------------------------------------------------
def virtual_image(src_img, src_mask, tar_depth, tar_inv_K, src_K, T):
    """-> (warped image [B,3,H,W], warped mask [B,1,H,W]) -- view_rendering.py:61-82. No z>0 test."""
    grid = reproject_grid(tar_depth, T, tar_inv_K, src_K)
    img = F.grid_sample(src_img, grid, mode='bilinear', padding_mode='zeros', align_corners=True)
    msk = F.grid_sample(src_mask, grid, mode='nearest', padding_mode='zeros', align_corners=True)
    img[torch.isnan(img)] = 2.0
    msk[torch.isnan(msk)] = 0
    g = grid.permute(0, 3, 1, 2)
    invalid = torch.logical_or(g > 1, g < -1).sum(dim=1, keepdim=True) > 0
    return img, (~invalid).float() * msk


def _mean_std(x, mask):
    """masked mean, UNMASKED variance over c*h*w of the mask -- :30-37."""
    _, c, h, w = mask.size()
    mean = (x * mask).sum(dim=(1, 2, 3), keepdim=True) / (mask.sum(dim=(1, 2, 3), keepdim=True) + 1e-8)
    var = ((x - mean) ** 2).sum(dim=(1, 2, 3), keepdim=True) / (c * h * w)
    return mean, torch.sqrt(var + 1e-16)


def norm_image_single(ref_img, ref_mask, warp_img, warp_mask):
    """:39-59 -- batch-wide early exit returns the un-normalised, un-masked warp."""
    warp_mask = warp_mask.detach()
    with torch.no_grad():
        mask = (ref_mask * warp_mask).bool()
        if mask.size(1) != 3:
            mask = mask.repeat(1, 3, 1, 1)
        if torch.any(mask.sum(dim=(-3, -2, -1)) == 0):
            return warp_img
        s_mean, s_std = _mean_std(ref_img, mask)
        w_mean, w_std = _mean_std(warp_img, mask)
    return ((warp_img - w_mean) / (w_std + 1e-8) * s_std + s_mean) * warp_mask.float()


def relative_cam_poses(cfg, inputs, cam_T_cam, cam):
    """{(frame, neighbour): [B,4,4]} -- models/geometry/pose.py:66-96.
    ``cam_T_cam`` = {frame: [B,4,4]} of THIS camera."""
    t, d = cfg['training'], cfg['data']
    ref_ext = inputs['extrinsics'][:, cam]
    rel = {}
    if t['spatio']:
        for nb in d['rel_cam_list'][cam]:
            if nb >= d['num_cams']:
                continue
            rel[(0, nb)] = torch.matmul(inputs['extrinsics_inv'][:, nb], ref_ext)
    if t['spatio_temporal']:
        for f in t['frame_ids'][1:]:
            for nb in d['rel_cam_list'][cam]:
                if nb >= d['num_cams']:
                    continue
                rel[(f, nb)] = torch.matmul(rel[(0, nb)], cam_T_cam[f])
    return rel


def view_rendering(cfg, inputs, depth, cam_T_cam, cam):
    """One camera of ViewRendering.forward (view_rendering.py:118-198, scale 0, no aug_depth).
    -> dict with ('color',f), ('color_mask',f) for f in frame_ids[1:] and ('overlap',f), ('overlap_mask',f)
    for every f."""
    t, d = cfg['training'], cfg['data']
    ref_color = inputs[('color', 0, 0)][:, cam]
    ref_mask = inputs['mask'][:, cam]
    ref_K, ref_iK = inputs[('K', 0)][:, cam], inputs[('inv_K', 0)][:, cam]
    view = {}
    for f in t['frame_ids'][1:]:
        img, msk = virtual_image(inputs[('color', f, 0)][:, cam], inputs['mask'][:, cam], depth, ref_iK, ref_K,
                                 cam_T_cam[f])
        if t['intensity_align']:
            img = norm_image_single(ref_color, ref_mask, img, msk)
        view[('color', f)] = img
        view[('color_mask', f)] = msk
    if t['spatio'] or t['spatio_temporal']:
        rel = relative_cam_poses(cfg, inputs, cam_T_cam, cam)
        for f in t['frame_ids']:
            o_img = torch.zeros_like(ref_color)
            o_msk = torch.zeros_like(ref_mask)
            for nb in d['rel_cam_list'][cam]:
                if nb >= d['num_cams']:
                    continue
                img, msk = virtual_image(inputs[('color', f, 0)][:, nb], inputs['mask'][:, nb], depth, ref_iK,
                                         inputs[('K', 0)][:, nb], rel[(f, nb)])
                if t['intensity_align']:
                    img = norm_image_single(ref_color, ref_mask, img, msk)
                o_img = o_img + img
                o_msk = o_msk + msk
            view[('overlap', f)] = o_img
            view[('overlap_mask', f)] = o_msk
    return view


# --------------------------------------------------------------------------------------
# A11-A13  losses                 (models/losses/loss_util.py, single_cam_loss.py, multi_cam_loss.py)
# --------------------------------------------------------------------------------------
def ssim_loss(pred, target):
    """loss_util.py:43-67"""
    pred = F.pad(pred, (1, 1, 1, 1), mode='reflect')
    target = F.pad(target, (1, 1, 1, 1), mode='reflect')
    mu_p = F.avg_pool2d(pred, kernel_size=3, stride=1)
    mu_t = F.avg_pool2d(target, kernel_size=3, stride=1)
    musq_p, musq_t, mu_pt = mu_p.pow(2), mu_t.pow(2), mu_p * mu_t
    sig_p = F.avg_pool2d(pred.pow(2), kernel_size=3, stride=1) - musq_p
    sig_t = F.avg_pool2d(target.pow(2), kernel_size=3, stride=1) - musq_t
    sig_pt = F.avg_pool2d(pred * target, kernel_size=3, stride=1) - mu_pt
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim = ((2 * mu_pt + C1) * (2 * sig_pt + C2)) / ((musq_p + musq_t + C1) * (sig_p + sig_t + C2) + 1e-8)
    return torch.clamp((1 - ssim) / 2, 0, 1)


def photometric_loss(pred, target):
    """0.85 * SSIM + 0.15 * L1, channel mean -> [B,1,H,W] -- loss_util.py:70-78."""
    l1 = torch.abs(target - pred).mean(1, True)
    return 0.85 * ssim_loss(pred, target).mean(1, True) + 0.15 * l1


def auto_masks(reprojection_loss, identity_reprojection_loss):
    """loss_util.py:6-18"""
    if identity_reprojection_loss is None:
        return torch.ones_like(reprojection_loss)
    idxs = torch.argmin(torch.cat([reprojection_loss, identity_reprojection_loss], dim=1), dim=1, keepdim=True)
    return (idxs == 0).float()


def edg_smooth_loss_raw(rgb, disp_map):
    """loss_util.py:28-40 as the free function is written (no mean normalisation)."""
    grx = (rgb[:, :, :, :-1] - rgb[:, :, :, 1:]).abs().mean(1, True)
    gry = (rgb[:, :, :-1, :] - rgb[:, :, 1:, :]).abs().mean(1, True)
    gdx = (disp_map[:, :, :, :-1] - disp_map[:, :, :, 1:]).abs() * (-1.0 * grx).exp()
    gdy = (disp_map[:, :, :-1, :] - disp_map[:, :, 1:, :]).abs() * (-1.0 * gry).exp()
    return gdx.mean() + gdy.mean()


def masked_loss(loss, mask):
    """loss_util.py:21-25 -- sums run over the whole local batch."""
    return (loss * mask).sum() / (mask.sum() + 1e-8)


def edge_smooth_loss(rgb, disp):
    """loss_util.py:28-40 applied to the mean-normalised disparity of single_cam_loss.py:57-65."""
    mean_disp = disp.mean(2, True).mean(3, True)
    nd = disp / (mean_disp + 1e-8)
    grx = (rgb[:, :, :, :-1] - rgb[:, :, :, 1:]).abs().mean(1, True)
    gry = (rgb[:, :, :-1, :] - rgb[:, :, 1:, :]).abs().mean(1, True)
    gdx = (nd[:, :, :, :-1] - nd[:, :, :, 1:]).abs() * (-1.0 * grx).exp()
    gdy = (nd[:, :, :-1, :] - nd[:, :, 1:, :]).abs() * (-1.0 * gry).exp()
    return gdx.mean() + gdy.mean()


def multi_cam_loss(cfg, inputs, view, disp, cam, noise, decisions=None):
    """One camera of MultiCamLoss.forward (multi_cam_loss.py:94-138, scale 0, pose_model='fusion').
    ``noise`` [B,2,H,W] is the CPU-drawn ``torch.randn`` of single_cam_loss.py:45-46, passed in so
    that every implementation sees the same draw.  -> (cam_loss, detail dict).

    ``decisions`` (tests only) pins the three per-pixel hard decisions to given values instead of taking them from
    this run's own fp32 comparisons: dict(reproj_mask [B,1,H,W] float = auto-mask * ref_mask, sel_reproj / sel_st
    [B,1,H,W] long = the frame each torch.min picked).  Everything else is computed exactly as without it."""
    t, L = cfg['training'], cfg['loss']
    target = inputs[('color', 0, 0)][:, cam]
    ref_mask = inputs['mask'][:, cam]
    fr = t['frame_ids'][1:]
    # single_cam_loss.py:17-55
    reproj_all = torch.cat([photometric_loss(view[('color', f)], target) for f in fr], 1)
    reproj, _ = torch.min(reproj_all, dim=1, keepdim=True)
    if decisions is not None:
        reproj = torch.gather(reproj_all, 1, decisions['sel_reproj'])
    ident_all = torch.cat([photometric_loss(inputs[('color', f, 0)][:, cam], target) for f in fr], 1)
    ident_all = ident_all + 0.00001 * noise
    ident, _ = torch.min(ident_all, dim=1, keepdim=True)
    idx = torch.argmin(torch.cat([reproj, ident], dim=1), dim=1, keepdim=True)
    auto = (idx == 0).float() * ref_mask
    if decisions is not None:
        auto = decisions['reproj_mask'].to(reproj.dtype)
    reproj_loss = masked_loss(reproj, auto)
    smooth = edge_smooth_loss(target, disp)
    # multi_cam_loss.py:16-29
    sp_mask = ref_mask * view[('overlap_mask', 0)]
    sp_map = photometric_loss(view[('overlap', 0)], target)
    spatio = masked_loss(sp_map, sp_mask)
    # multi_cam_loss.py:31-59
    st_maps = torch.cat([photometric_loss(view[('overlap', f)], target) for f in fr], 1)
    st_masks = torch.cat([ref_mask * view[('overlap_mask', f)] * auto for f in fr], 1)
    st_map, _ = torch.min(st_maps, dim=1, keepdim=True)
    if decisions is not None:
        st_map = torch.gather(st_maps, 1, decisions['sel_st'])
    st_mask, _ = torch.max(st_masks.float(), dim=1, keepdim=True)
    spatio_tempo = masked_loss(st_map, st_mask)
    # accumulation order of multi_cam_loss.py:118-121
    cam_loss = 0.
    cam_loss += reproj_loss
    cam_loss += L['disparity_smoothness'] * smooth / (2 ** 0)
    cam_loss += L['spatio_coeff'] * spatio + L['spatio_tempo_coeff'] * spatio_tempo
    cam_loss += L['pose_loss_coeff'] * 0
    cam_loss /= len(t['scales'])
    detail = dict(reproj_loss=reproj_loss, smooth=smooth, spatio_loss=spatio, spatio_tempo_loss=spatio_tempo,
                  reproj_mask=auto, reproj_map=auto * reproj, spatio_mask=sp_mask, spatio_tempo_mask=st_mask)
    # decisions and their margins (how far each comparison is from flipping), for the parity tests
    with torch.no_grad():
        big = torch.full_like(reproj, 1e30)
        gap = lambda x: (x[:, 0:1] - x[:, 1:2]).abs() if x.shape[1] > 1 else big
        detail.update(sel_reproj=torch.argmin(reproj_all, 1, keepdim=True), sel_st=torch.argmin(st_maps, 1, keepdim=True),
                      margin_auto=(reproj - ident).abs(), margin_reproj=gap(reproj_all), margin_st=gap(st_maps))
    return cam_loss, detail


def warp_and_loss(cfg, inputs, depths, disps, cam_T_cam, noises, decisions=None):
    """The view-synthesis half of a training step for a whole rig batch:
    models/vfdepth.py:290-313 (loop over cameras, mean over cameras).
    depths/disps/noises: list over cameras; cam_T_cam {f: [B,6,4,4]}."""
    total = 0
    details = []
    for cam in range(cfg['data']['num_cams']):
        T = {f: cam_T_cam[f][:, cam] for f in cfg['training']['frame_ids'][1:]}
        view = view_rendering(cfg, inputs, depths[cam], T, cam)
        loss, det = multi_cam_loss(cfg, inputs, view, disps[cam], cam, noises[cam],
                                   None if decisions is None else decisions[cam])
        det['view'] = view
        details.append(det)
        total = total + loss
    return total / cfg['data']['num_cams'], details


def to_depth(cfg, disp_in, K_in):
    """models/vfdepth.py:277-288 (VFDepthAlgo.to_depth): disp [B,1,hs,ws] of one camera, K [B,4,4] -> depth [B,1,H,W]."""
    t = cfg['training']
    min_disp = 1 / t['max_depth']
    max_disp = 1 / t['min_depth']
    disp_range = max_disp - min_disp
    disp_in = F.interpolate(disp_in, [t['height'], t['width']], mode='bilinear', align_corners=False)
    disp = min_disp + disp_range * disp_in
    depth = 1 / disp
    return depth * K_in[:, 0:1, 0:1].unsqueeze(2) / t['focal_length_scale']


def aggregate_features(packed_feats, lev, weight, bias, slope=0.1):
    """network/fusion_depthnet.py:57-66 (same lines in fusion_posenet.py:62-70): resize the encoder levels above
    ``lev`` to level ``lev`` (bilinear, align_corners=True), concatenate, 1x1 conv + LeakyReLU (blocks.py:41-61 conv2d
    with kernel_size=1).  packed_feats: list of [B*cams,C_l,h_l,w_l]; weight [Cout,sum C_l,1,1] -> [B*cams,Cout,h,w]."""
    _, _, up_h, up_w = packed_feats[lev].size()
    feats = packed_feats[lev:lev + 1] + [F.interpolate(feat, [up_h, up_w], mode='bilinear', align_corners=True)
                                         for feat in packed_feats[lev + 1:]]
    return F.leaky_relu(F.conv2d(torch.cat(feats, dim=1), weight, bias), slope)


# --------------------------------------------------------------------------------------
# 8f-3  depth synthesis (aug_depth)   view_rendering.py:84-116, :201-241; depth_synthesis_loss.py:15-43
# --------------------------------------------------------------------------------------
def virtual_depth(src_depth, src_mask, src_inv_K, src_K, tar_depth, tar_inv_K, tar_K, T, min_depth, max_depth):
    """-> (source depth warped into the target view [B,1,H,W], its mask) -- view_rendering.py:84-116."""
    b, _, h, w = src_depth.size()
    homo = homo_points(b, h, w, src_depth.dtype).to(src_depth.device)
    src_points = src_depth.view(b, 1, -1) * torch.matmul(src_inv_K[:, :3, :3], homo)
    src_points = torch.cat([src_points, homo[:, 2:3]], 1)
    src_points_warped = torch.matmul(T[:, :3, :], src_points)
    src_depth_warped = src_points_warped.reshape(b, 3, h, w)[:, 2:3, :, :]
    pix_coords = reproject_grid(tar_depth, torch.inverse(T), tar_inv_K, src_K)
    depth_warped = F.grid_sample(src_depth_warped, pix_coords, mode='bilinear', padding_mode='zeros',
                                 align_corners=True)
    mask_warped = F.grid_sample(src_mask, pix_coords, mode='nearest', padding_mode='zeros', align_corners=True)
    depth_warped[torch.isnan(depth_warped)] = 2.0
    mask_warped[torch.isnan(mask_warped)] = 0
    g = pix_coords.permute(0, 3, 1, 2)
    invalid = torch.logical_or(g > 1, g < -1).sum(dim=1, keepdim=True) > 0
    valid_min = depth_warped > min_depth
    depth_warped[~valid_min] = min_depth
    valid_max = depth_warped < max_depth
    depth_warped[~valid_max] = max_depth
    return depth_warped, (~invalid).float() * mask_warped * valid_min * valid_max


def depth_synthesis(cfg, inputs, depths, aug_depth, cam):
    """the aug_depth branch of ViewRendering.forward (:201-241) for one camera: depths = list over cameras of
    ('depth', 0), aug_depth = ('depth', 0, 'aug') of `cam` -> (tform_depth list, tform_mask list)."""
    t = cfg['training']
    n = cfg['data']['num_cams']
    aug_ext_inv = torch.inverse(inputs['extrinsics_aug'][:, cam])
    K, inv_K = inputs[('K', 0)], inputs[('inv_K', 0)]
    tform_depth, tform_mask = [], []
    for src in cfg['data']['rel_cam_list'][cam] + [cam]:
        if src >= n:
            continue
        rel_pose = torch.matmul(aug_ext_inv, inputs['extrinsics'][:, src])
        d, m = virtual_depth(depths[src], inputs['mask'][:, src], inv_K[:, src], K[:, src], aug_depth, inv_K[:, cam],
                             K[:, cam], rel_pose, t['min_depth'], t['max_depth'])
        tform_depth.append(d)
        tform_mask.append(m)
    return tform_depth, tform_mask


def aug_losses(aug_depth, tform_depth, tform_mask, disp_aug):
    """DepthSynLoss.compute_aug_losses (depth_synthesis_loss.py:15-43) -> (depth_con_loss, depth_sm_loss)."""
    losses, masks = [], []
    for d, m in zip(tform_depth, tform_mask):
        pl = (aug_depth - d).abs() / (aug_depth + d + 1e-8)
        losses.append(torch.clamp(pl, 0., 1.))
        masks.append(m)
    con = masked_loss(torch.cat(losses, 0), torch.cat(masks, 0))
    mean_disp = disp_aug.mean(2, True).mean(3, True)
    nd = disp_aug / (mean_disp + 1e-8)
    gx = torch.abs(nd[:, :, :, :-1] - nd[:, :, :, 1:])
    gy = torch.abs(nd[:, :, :-1, :] - nd[:, :, 1:, :])
    return con, gx.mean() + gy.mean()
