"""Seeded synthetic surround rigs.

There is no dataset on the box, so every test and benchmark runs on rigs drawn
here: 6 jittered cameras on a ring (ego frame x forward / y left / z up, camera
frame z forward / x right / y down, as the reference's DDAD convention,
/root/reference/network/volumetric_fusionnet.py:20), intrinsics per scale as
/root/reference/dataset/data_util.py:63-68 builds them, N(0,1) feature maps,
a binary self-occlusion mask, smooth random images for frames 0/-1/+1, a
smooth disparity turned into depth as /root/reference/models/vfdepth.py:277-288
does, and a small canonical ego-motion distributed to every camera as
/root/reference/models/geometry/pose.py:44-64 does.

Everything is drawn on the CPU with an explicit ``torch.Generator`` so the same
tensors can be handed to the oracle and to the CUDA kernels.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

_YAWS_DEG = [0.0, 60.0, -60.0, 120.0, -120.0, 180.0]


def _rot_z(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, -s, 0.0], [s, c, 0.0], [0.0, 0.0, 1.0]])


def axis_angle_to_matrix(v):
    """Rodrigues formula, [...,3] -> [...,3,3] (stands in for pytorch3d's function,
    which only produces *inputs* of the hot path)."""
    theta = v.norm(dim=-1, keepdim=True).clamp_min(1e-12)
    k = v / theta
    kx, ky, kz = k.unbind(-1)
    zero = torch.zeros_like(kx)
    Kx = torch.stack([zero, -kz, ky, kz, zero, -kx, -ky, kx, zero], -1).reshape(v.shape[:-1] + (3, 3))
    th = theta[..., None]
    eye = torch.eye(3, dtype=v.dtype, device=v.device).expand(v.shape[:-1] + (3, 3))
    return eye + torch.sin(th) * Kx + (1 - torch.cos(th)) * (Kx @ Kx)


def _smooth_field(gen, shape, coarse, fine_amp=0.25):
    """Low-pass random field in [0,1]: coarse uniform noise upsampled + a finer octave."""
    b, c, h, w = shape
    lo = torch.rand((b, c, max(h // coarse, 2), max(w // coarse, 2)), generator=gen)
    hi = torch.rand((b, c, max(h // (coarse // 4 or 1), 2), max(w // (coarse // 4 or 1), 2)), generator=gen)
    lo = F.interpolate(lo, size=(h, w), mode='bicubic', align_corners=False)
    hi = F.interpolate(hi, size=(h, w), mode='bilinear', align_corners=False)
    return ((1 - fine_amp) * lo + fine_amp * hi).clamp(0.0, 1.0)


def make_calibration(gen, batch, height, width, dataset='ddad'):
    """-> K0 [B,6,4,4] (scale 0, fp64), extrinsics camera->ego [B,6,4,4] (fp64)."""
    f_front, f_side = (721.0, 350.0) if dataset == 'ddad' else (506.0, 324.0)
    f_front *= width / 640.0
    f_side *= width / 640.0
    r_base = np.array([[0.0, 0.0, 1.0], [-1.0, 0.0, 0.0], [0.0, -1.0, 0.0]])  # cam xyz -> ego
    K = np.zeros((batch, 6, 4, 4))
    E = np.zeros((batch, 6, 4, 4))
    u = lambda n: torch.rand(n, generator=gen, dtype=torch.float64).numpy()
    for b in range(batch):
        for c in range(6):
            yaw = math.radians(_YAWS_DEG[c] + 3.0 * float(u(1)[0]))
            # small pitch/roll so that no projection is axis-aligned
            tilt = axis_angle_to_matrix(torch.from_numpy((u(3) - 0.5) * math.radians(2.0))).numpy()
            R = _rot_z(yaw) @ tilt @ r_base
            t = np.array([1.5 * math.cos(yaw), 1.5 * math.sin(yaw), 1.5]) + 0.1 * u(3)
            E[b, c, :3, :3] = R
            E[b, c, :3, 3] = t
            E[b, c, 3, 3] = 1.0
            f = (f_front if c == 0 else f_side) * (1.0 + 0.02 * (float(u(1)[0]) - 0.5))
            K[b, c] = np.eye(4)
            K[b, c, 0, 0] = f
            K[b, c, 1, 1] = f * (1.0 + 0.004 * (float(u(1)[0]) - 0.5))
            K[b, c, 0, 2] = width / 2.0 + float(u(1)[0])
            K[b, c, 1, 2] = height / 2.0 + float(u(1)[0])
    return K, E


def scaled_intrinsics(K0, scale):
    """rows 0-1 divided by 2**scale, inverse by fp64 pinv (data_util.py:63-68)."""
    Ks = K0.copy()
    Ks[..., :2, :] /= (2 ** scale)
    return Ks, np.linalg.pinv(Ks)


def make_mask(gen, batch, height, width):
    """binary self-occlusion mask [B,6,1,H,W]: bottom 15 % rows and 3 rectangles are 0."""
    m = torch.ones(batch, 6, 1, height, width)
    m[..., int(height * 0.85):, :] = 0
    for b in range(batch):
        for c in range(6):
            for _ in range(3):
                r = torch.rand(4, generator=gen)
                y0 = int(r[0] * height * 0.7)
                x0 = int(r[1] * width * 0.9)
                hh = 2 + int(r[2] * height * 0.12)
                ww = 2 + int(r[3] * width * 0.12)
                m[b, c, 0, y0:y0 + hh, x0:x0 + ww] = 0
    return m


def distribute_pose(ref_T, E, E_inv):
    """canonical (front camera) motion -> every camera (pose.py:44-64). [B,4,4],[B,6,4,4] -> [B,6,4,4]."""
    out = []
    for c in range(E.shape[1]):
        out.append(E_inv[:, c] @ E[:, 0] @ ref_T @ E_inv[:, 0] @ E[:, c])
    return torch.stack(out, 1)


def to_depth(disp, K0, min_depth, max_depth, focal_length_scale, height, width):
    """disparity [B,1,h',w'] in [0,1] -> metric depth [B,1,H,W] (vfdepth.py:277-288)."""
    min_disp, max_disp = 1 / max_depth, 1 / min_depth
    disp = F.interpolate(disp, [height, width], mode='bilinear', align_corners=False)
    depth = 1 / (min_disp + (max_disp - min_disp) * disp)
    return depth * K0[:, 0:1, 0:1].unsqueeze(2) / focal_length_scale


def make_rig(cfg, seed=1234, feat_channels=None, with_feats=True, with_images=True):
    """Draw one batch of ``cfg['training']['batch_size']`` rigs.  Returns (inputs, extras).

    ``inputs`` carries exactly the dict keys the reference consumes;
    ``extras`` carries ``feats_agg`` [B,6,C,h,w], ``depth`` [6][B,1,H,W], ``disp``,
    ``cam_T_cam`` {f: [B,6,4,4]} and ``noise`` [6][B,2,H,W] (the CPU-drawn tie-break noise of
    /root/reference/models/losses/single_cam_loss.py:45-46).
    """
    tr, md = cfg['training'], cfg['model']
    B, H, W = tr['batch_size'], tr['height'], tr['width']
    lvl = md['fusion_level'] + 1
    h, w = H // (2 ** lvl), W // (2 ** lvl)
    C = feat_channels if feat_channels is not None else md['fusion_feat_in_dim']
    gen = torch.Generator().manual_seed(int(seed))

    K0, E = make_calibration(gen, B, H, W, cfg['data'].get('dataset', 'ddad'))
    inputs = {}
    for s in sorted({0, lvl}):
        Ks, iKs = scaled_intrinsics(K0, s)
        inputs[('K', s)] = torch.from_numpy(Ks).float()
        inputs[('inv_K', s)] = torch.from_numpy(iKs).float()
    inputs['extrinsics'] = torch.from_numpy(E).float()
    inputs['extrinsics_inv'] = torch.inverse(inputs['extrinsics'])
    inputs['mask'] = make_mask(gen, B, H, W)

    extras = {}
    if with_feats:
        extras['feats_agg'] = torch.randn((B, 6, C, h, w), generator=gen)
    if with_images:
        base = _smooth_field(gen, (B * 6, 3, H, W), 32).view(B, 6, 3, H, W)
        for f in tr['frame_ids']:
            if f == 0:
                img = base
            else:
                pert = _smooth_field(gen, (B * 6, 3, H, W), 16).view(B, 6, 3, H, W)
                img = (torch.roll(base, shifts=(f, 3 * f), dims=(-2, -1)) * 0.9 + 0.1 * pert).clamp(0, 1)
            inputs[('color', f, 0)] = img.contiguous()
            inputs[('color_aug', f, 0)] = img.contiguous()
        disp = torch.sigmoid(4.0 * (_smooth_field(gen, (B * 6, 1, H, W), 32, 0.1) - 0.5)).view(B, 6, 1, H, W)
        extras['disp'] = [disp[:, c].contiguous() for c in range(6)]
        extras['depth'] = [to_depth(extras['disp'][c], inputs[('K', 0)][:, c], tr['min_depth'], tr['max_depth'],
                                    tr['focal_length_scale'], H, W) for c in range(6)]
        extras['cam_T_cam'] = {}
        for f in tr['frame_ids'][1:]:
            aa = (torch.rand((B, 3), generator=gen) - 0.5) * math.radians(2.0)
            T = torch.eye(4).repeat(B, 1, 1)
            T[:, :3, :3] = axis_angle_to_matrix(aa)
            T[:, :3, 3] = torch.tensor([0.0, 0.0, 0.8 * f]) + 0.05 * (torch.rand((B, 3), generator=gen) - 0.5)
            extras['cam_T_cam'][f] = distribute_pose(T, inputs['extrinsics'], inputs['extrinsics_inv'])
        extras['noise'] = [torch.randn((B, 2, H, W), generator=gen) for _ in range(6)]
    return inputs, extras
