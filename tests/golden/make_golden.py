"""Generate the golden fixtures by RUNNING THE REFERENCE ITSELF (CPU, fp32).

    python tests/golden/make_golden.py            # needs /root/reference (or $VFDEPTH_REF)

Writes ``golden_fusion.npz``, ``golden_warp.npz`` and ``golden_to_depth.npz``, ``golden_depth_syn.npz``,
``golden_feat_agg.npz``, ``golden_loss_free.npz`` next to this file (``--only-new`` regenerates all but the first two,
``--only-agg`` / ``--only-loss`` / ``--only-sweep`` just that one).  The
fixtures hold the seeded inputs and what the reference's own classes
(``VFNet``, ``ViewRendering``, ``Pose``, ``MultiCamLoss``) returned for them, at
sizes small enough to commit.  ``tests/test_oracle_golden.py`` then pins
``oracle/vfdepth_oracle.py`` to these bit for bit, and the GPU parity tests
check the CUDA path against the same files.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402
from vfdepth_b200 import config as vcfg  # noqa: E402
from vfdepth_b200.rig import make_rig  # noqa: E402

FUSION_CFG = dict(model__voxel_size=[14, 12, 4], model__voxel_str_p=[-17.5, -15.0, -3.0],
                  model__voxel_unit_size=[2.5, 2.5, 1.5], model__proj_d_bins=6, model__proj_d_str=2,
                  model__proj_d_end=16, model__fusion_feat_in_dim=8, model__voxel_pre_dim=[4],
                  training__height=64, training__width=96)
WARP_CFG = dict(training__height=48, training__width=64, training__focal_length_scale=30)


def fusion_cfg():
    return vcfg.default_config('ddad', batch_size=2, **FUSION_CFG)


def warp_cfg():
    return vcfg.default_config('ddad', batch_size=2, **WARP_CFG)


def _np(t):
    return t.detach().cpu().numpy()


def make_fusion(ref, path):
    cfg = fusion_cfg()
    inputs, extras = make_rig(cfg, seed=4242, with_images=False)
    C = cfg['model']['fusion_feat_in_dim']
    lvl = cfg['model']['fusion_level'] + 1
    out = {'feats_agg': _np(extras['feats_agg']), 'mask': _np(inputs['mask']),
           'K': _np(inputs[('K', lvl)]), 'inv_K': _np(inputs[('inv_K', lvl)]),
           'extrinsics': _np(inputs['extrinsics']), 'extrinsics_inv': _np(inputs['extrinsics_inv'])}
    torch.manual_seed(7)
    net = ref.VFNet(cfg, C, 16, model='depth')
    net.reduce_dim = torch.nn.Identity()        # keep the raw ray samples (:262), skip the convs of :265
    net.type_check(extras['feats_agg'][0, 0])
    feats = extras['feats_agg'].clone().requires_grad_(True)
    vox = net.backproject_into_voxel(feats, inputs['mask'], inputs[('K', lvl)], inputs['extrinsics_inv'])
    proj = net.project_voxel_into_image(vox, inputs[('inv_K', lvl)], inputs['extrinsics'])
    gen = torch.Generator().manual_seed(99)
    g_up = [torch.randn(p.shape, generator=gen) for p in proj]
    loss = sum((p * g).sum() for p, g in zip(proj, g_up))
    params = [net.conv_non_overlap[0].weight, net.conv_non_overlap[0].bias,
              net.conv_overlap[0].weight, net.conv_overlap[0].bias]
    grads = torch.autograd.grad(loss, [feats] + params)
    out.update(w_no=_np(params[0]), b_no=_np(params[1]), w_ov=_np(params[2]), b_ov=_np(params[3]),
               voxel_feat=_np(vox), proj=_np(torch.stack(proj, 1)), g_up=_np(torch.stack(g_up, 1)),
               d_feats=_np(grads[0]), d_w_no=_np(grads[1]), d_b_no=_np(grads[2]), d_w_ov=_np(grads[3]),
               d_b_ov=_np(grads[4]))
    # per-camera coordinates / masks straight from the reference's own helpers
    grids, valids = [], []
    import torch.nn.functional as F
    h, w = extras['feats_agg'].shape[-2:]
    for c in range(6):
        v_local = torch.matmul(inputs['extrinsics_inv'][:, c, :3, :], net.voxel_pts)
        g = net.calculate_sample_pixel_coords(inputs[('K', lvl)][:, c], v_local, w, h)
        m = F.interpolate(inputs['mask'][:, c], [h, w], mode='bilinear', align_corners=True)
        grids.append(_np(g))
        valids.append(_np(net.calculate_valid_mask(m, g, v_local)))
    out['grids'] = np.stack(grids, 1)
    out['valids'] = np.stack(valids, 1)
    # pose model
    netp = ref.VFNet(cfg, C, 16, model='pose')
    netp.type_check(extras['feats_agg'][0, 0])
    out['voxel_feat_pose'] = _np(netp.backproject_into_voxel(extras['feats_agg'], inputs['mask'],
                                                             inputs[('K', lvl)], inputs['extrinsics_inv']))
    np.savez_compressed(path, **out)
    print('wrote', path, {k: v.shape for k, v in out.items()})


def make_warp(ref, path):
    cfg = warp_cfg()
    inputs, extras = make_rig(cfg, seed=777, with_feats=False)
    B = cfg['training']['batch_size']
    out = {'mask': _np(inputs['mask']), 'K': _np(inputs[('K', 0)]), 'inv_K': _np(inputs[('inv_K', 0)]),
           'extrinsics': _np(inputs['extrinsics']), 'extrinsics_inv': _np(inputs['extrinsics_inv'])}
    for f in (0, -1, 1):
        out['color_%d' % f] = _np(inputs[('color', f, 0)])
    out['disp'] = _np(torch.stack(extras['disp'], 1))
    out['depth'] = _np(torch.stack(extras['depth'], 1))
    out['T_m1'] = _np(extras['cam_T_cam'][-1])
    out['T_p1'] = _np(extras['cam_T_cam'][1])
    vr = ref.ViewRendering(cfg, 'cpu')
    pose = ref.Pose(cfg)
    lossm = ref.MultiCamLoss(cfg, 'cpu')
    depths = [d.clone().requires_grad_(True) for d in extras['depth']]
    Ts = {f: extras['cam_T_cam'][f].clone().requires_grad_(True) for f in (-1, 1)}
    outputs = {}
    total = 0
    noises, scal = [], []
    for cam in range(6):
        outputs[('cam', cam)] = {('depth', 0): depths[cam], ('disp', 0): extras['disp'][cam],
                                 ('cam_T_cam', 0, -1): Ts[-1][:, cam], ('cam_T_cam', 0, 1): Ts[1][:, cam]}
    for cam in range(6):
        rel = pose.compute_relative_cam_poses(inputs, outputs, cam)
        vr(inputs, outputs, cam, rel)
        torch.manual_seed(1000 + cam)
        noises.append(torch.randn(B, 2, cfg['training']['height'], cfg['training']['width']))
        torch.manual_seed(1000 + cam)          # the loss module draws the same tensor from the global stream
        cam_loss, ld = lossm(inputs, outputs, cam)
        total = total + cam_loss
        scal.append([ld['reproj_loss'], ld['spatio_loss'], ld['spatio_tempo_loss'], ld['smooth'], float(cam_loss)])
    total = total / 6
    grads = torch.autograd.grad(total, depths + [Ts[-1], Ts[1]])
    out['noise'] = _np(torch.stack(noises, 1))
    out['cam_scalars'] = np.asarray(scal, dtype=np.float64)
    out['total'] = _np(total)
    out['d_depth'] = _np(torch.stack(grads[:6], 1))
    out['d_T_m1'] = _np(grads[6])
    out['d_T_p1'] = _np(grads[7])
    for cam in (0, 3):
        tv = outputs[('cam', cam)]
        for f in (-1, 1):
            out['cam%d_color_%d' % (cam, f)] = _np(tv[('color', f, 0)])
            out['cam%d_color_mask_%d' % (cam, f)] = _np(tv[('color_mask', f, 0)])
        for f in (0, -1, 1):
            out['cam%d_overlap_%d' % (cam, f)] = _np(tv[('overlap', f, 0)])
            out['cam%d_overlap_mask_%d' % (cam, f)] = _np(tv[('overlap_mask', f, 0)])
        out['cam%d_reproj_loss' % cam] = _np(tv[('reproj_loss', 0)])
        out['cam%d_reproj_mask' % cam] = _np(tv[('reproj_mask', 0)])
    np.savez_compressed(path, **out)
    print('wrote', path, 'total loss', float(total))


def make_to_depth(path):
    """VFDepthAlgo.to_depth (models/vfdepth.py:277-288) called unbound on a namespace with the attributes it reads,
    per camera as vfdepth.py:266-275 does, for decoder outputs of scale 0, 1 and 2; gradients for a seeded upstream."""
    import types
    algo = ref_import.load_algo()
    cfg = warp_cfg()
    t = cfg['training']
    inputs, _ = make_rig(cfg, seed=777, with_feats=False)
    K = inputs[('K', 0)]
    B, H, W = t['batch_size'], t['height'], t['width']
    ns = types.SimpleNamespace(max_depth=t['max_depth'], min_depth=t['min_depth'], height=H, width=W,
                               focal_length_scale=t['focal_length_scale'])
    gen = torch.Generator().manual_seed(31)
    out = {'K': _np(K)}
    for s in (0, 1, 2):
        disp = torch.rand((B, 6, H >> s, W >> s), generator=gen).requires_grad_(True)
        depth = torch.cat([algo.to_depth(ns, disp[:, c:c + 1], K[:, c]) for c in range(6)], 1)
        g_up = torch.randn(depth.shape, generator=gen)
        (d_disp,) = torch.autograd.grad((depth * g_up).sum(), disp)
        out.update({'disp_%d' % s: _np(disp), 'depth_%d' % s: _np(depth), 'g_up_%d' % s: _np(g_up),
                    'd_disp_%d' % s: _np(d_disp)})
    np.savez_compressed(path, **out)
    print('wrote', path)


SYN_CFG = dict(training__aug_depth=True, loss__depth_con_coeff=0.03, loss__depth_sm_coeff=0.05)


def syn_cfg():
    return vcfg.default_config('ddad', batch_size=2, **dict(WARP_CFG, **SYN_CFG))


def syn_case(cfg, seed):
    """seeded inputs of the depth-synthesis path: a rig, augmented extrinsics (small rotations so that the novel views
    overlap their sources), the novel-view depth / disparity predictions."""
    from vfdepth_b200.rig import axis_angle_to_matrix
    inputs, extras = make_rig(cfg, seed=seed, with_feats=False)
    B = cfg['training']['batch_size']
    gen = torch.Generator().manual_seed(seed + 1)
    angle = (torch.rand((B, 6, 3), generator=gen) - 0.5) * torch.tensor([0.2, 0.2, 0.5])
    tform = torch.eye(4).repeat(B, 6, 1, 1)
    tform[:, :, :3, :3] = axis_angle_to_matrix(angle)
    inputs['extrinsics_aug'] = tform @ inputs['extrinsics']
    depth = torch.cat(extras['depth'], 1)
    aug = depth * (1.0 + 0.1 * (torch.rand(depth.shape, generator=gen) - 0.5))
    disp_aug = torch.rand(depth.shape, generator=gen) * 0.5 + 0.05
    w = torch.rand((6, 2), generator=gen) + 0.5
    return inputs, extras, depth, aug, disp_aug, w


def make_depth_syn(ref, path):
    """ViewRendering.forward with aug_depth (view_rendering.py:201-241 -> get_virtual_depth :84-116) and
    DepthSynLoss.compute_aug_losses (depth_synthesis_loss.py:15-43), the reference's own classes."""
    ref_import.load_algo()
    import sys as _sys
    DepthSynLoss = _sys.modules['models.losses'].DepthSynLoss
    cfg = syn_cfg()
    inputs, extras, depth, aug, disp_aug, w = syn_case(cfg, 779)
    vr = ref.ViewRendering(cfg, 'cpu')
    pose = ref.Pose(cfg)
    lossm = DepthSynLoss(cfg, 'cpu')
    depths = [depth[:, c:c + 1].clone().requires_grad_(True) for c in range(6)]
    augs = [aug[:, c:c + 1].clone().requires_grad_(True) for c in range(6)]
    disps = [disp_aug[:, c:c + 1].clone().requires_grad_(True) for c in range(6)]
    outputs = {}
    for cam in range(6):
        outputs[('cam', cam)] = {('depth', 0): depths[cam], ('disp', 0): extras['disp'][cam],
                                 ('depth', 0, 'aug'): augs[cam], ('disp', 0, 'aug'): disps[cam],
                                 ('cam_T_cam', 0, -1): extras['cam_T_cam'][-1][:, cam],
                                 ('cam_T_cam', 0, 1): extras['cam_T_cam'][1][:, cam]}
    total, con, sm = 0, [], []
    out = {k: _np(inputs[k]) for k in ('mask', 'extrinsics', 'extrinsics_aug')}
    out.update(K=_np(inputs[('K', 0)]), inv_K=_np(inputs[('inv_K', 0)]), depth=_np(depth), aug=_np(aug),
               disp_aug=_np(disp_aug), w=_np(w))
    for cam in range(6):
        rel = pose.compute_relative_cam_poses(inputs, outputs, cam)
        vr(inputs, outputs, cam, rel)
        c, s_ = lossm.compute_aug_losses(outputs[('cam', cam)], 0)
        con.append(c)
        sm.append(s_)
        total = total + w[cam, 0] * c + w[cam, 1] * s_
        if cam in (0, 3):
            out['cam%d_tform_depth' % cam] = _np(torch.cat(outputs[('cam', cam)][('tform_depth', 0)], 1))
            out['cam%d_tform_mask' % cam] = _np(torch.cat(outputs[('cam', cam)][('tform_depth_mask', 0)], 1))
    grads = torch.autograd.grad(total, depths + augs + disps)
    out.update(con=_np(torch.stack(con)), sm=_np(torch.stack(sm)), d_depth=_np(torch.cat(grads[:6], 1)),
               d_aug=_np(torch.cat(grads[6:12], 1)), d_disp_aug=_np(torch.cat(grads[12:], 1)))
    np.savez_compressed(path, **out)
    print('wrote', path, 'con', [float(c) for c in con])


AGG_CFG = dict(training__height=64, training__width=96, model__fusion_feat_in_dim=64, model__voxel_size=[20, 20, 4],
               model__voxel_str_p=[-10.0, -10.0, -2.0], model__voxel_unit_size=[1.0, 1.0, 1.0], model__proj_d_bins=8)


def make_feat_agg(path):
    """The reference's FusedDepthNet (network/fusion_depthnet.py:13-92, randomly initialised ResNet-18 encoder) run on
    a seeded rig with forward hooks on its encoder and its conv1x1 block: the fixture holds the encoder levels
    lev.. (the inputs of :57-64), conv1x1's parameters, its output (= feats_agg before unpacking) and the gradients of a
    seeded upstream w.r.t. the levels and the parameters."""
    import importlib
    ref_import.load_algo()
    fdn = importlib.import_module('network.fusion_depthnet')
    cfg = vcfg.default_config('ddad', batch_size=1, **AGG_CFG)
    cfg['model']['weights_init'] = False
    torch.manual_seed(0)
    net = fdn.FusedDepthNet(cfg)
    with torch.no_grad():
        net.conv1x1[0].bias.normal_(0, 0.05)
    inputs, _ = make_rig(cfg, seed=5, with_feats=False)
    inputs[('color_aug', 0, 0)] = inputs[('color', 0, 0)]
    cap = {}
    net.encoder.register_forward_hook(lambda m, i, o: cap.update(feats=list(o)))
    net.conv1x1.register_forward_hook(lambda m, i, o: cap.update(agg=o))
    net.conv1x1.register_forward_pre_hook(lambda m, i: cap.update(cat=i[0]))
    net.train()
    net(inputs)
    lev = cfg['model']['fusion_level']
    levels = cap['feats'][lev:]
    gen = torch.Generator().manual_seed(77)
    g_up = torch.randn(cap['agg'].shape, generator=gen)
    params = [net.conv1x1[0].weight, net.conv1x1[0].bias]
    # gradients through the reference's graph w.r.t. the concatenated maps and the parameters.  (Inside the encoder
    # level l+1 is a function of level l, so the graph's gradient w.r.t. a level would include that path; the partial
    # derivative of lines :57-64 alone is the slice of d_cat pulled back through the same F.interpolate call.)
    d_cat, d_w, d_b = torch.autograd.grad((cap['agg'] * g_up).sum(), [cap['cat']] + params)
    out = {'lev': np.int64(lev), 'agg': _np(cap['agg']), 'g_up': _np(g_up), 'weight': _np(params[0]), 'bias': _np(params[1]),
           'd_weight': _np(d_w), 'd_bias': _np(d_b)}
    up_h, up_w = levels[0].shape[-2:]
    off = 0
    for i, x in enumerate(levels):
        c = x.shape[1]
        leaf = x.detach().clone().requires_grad_(True)
        y = leaf if i == 0 else torch.nn.functional.interpolate(leaf, [up_h, up_w], mode='bilinear', align_corners=True)
        (g,) = torch.autograd.grad((y * d_cat[:, off:off + c]).sum(), leaf)
        off += c
        out['feat_%d' % i] = _np(x)
        out['d_feat_%d' % i] = _np(g)
    np.savez_compressed(path, **out)
    print('wrote', path, [tuple(x.shape) for x in levels], tuple(cap['agg'].shape))


def make_loss_free(ref, path):
    """The reference's loss_util free functions (models/losses/loss_util.py:6-78) on seeded smooth images."""
    gen = torch.Generator().manual_seed(515)
    B, Hh, Ww = 2, 24, 40
    base = torch.rand(B, 3, Hh // 4, Ww // 4, generator=gen)
    target = torch.nn.functional.interpolate(base, size=(Hh, Ww), mode='bilinear', align_corners=True)
    pred = (target + 0.05 * torch.randn(B, 3, Hh, Ww, generator=gen)).clamp(0, 1)
    disp = torch.rand(B, 1, Hh, Ww, generator=gen)
    reproj = torch.rand(B, 1, Hh, Ww, generator=gen)
    ident = torch.rand(B, 2, Hh, Ww, generator=gen)
    ident[:, 0, ::3] = reproj[:, 0, ::3]                      # ties: the first minimum wins
    mask = (torch.rand(B, 1, Hh, Ww, generator=gen) > 0.4).float()
    lu = ref.loss_util
    pr = pred.clone().requires_grad_(True)
    ssim = lu.compute_ssim_loss(pr, target)
    photo = lu.compute_photometric_loss(pr, target)
    dd = disp.clone().requires_grad_(True)
    sm = lu.compute_edg_smooth_loss(target, dd)
    ll = photo.detach().clone().requires_grad_(True)
    ml = lu.compute_masked_loss(ll, mask)
    out = {'pred': pred, 'target': target, 'disp': disp, 'reproj': reproj, 'ident': ident, 'mask': mask,
           'ssim': ssim, 'photo': photo, 'auto_mask': lu.compute_auto_masks(reproj, ident),
           'auto_mask_none': lu.compute_auto_masks(reproj, None), 'smooth': sm, 'masked': ml,
           'd_disp': torch.autograd.grad(sm, dd)[0], 'd_loss': torch.autograd.grad(ml, ll)[0],
           'd_pred': torch.autograd.grad(photo.sum(), pr)[0]}
    np.savez_compressed(path, **{k: _np(v) for k, v in out.items()})
    print('wrote', path)


SWEEP_PICK = list(range(0, 767, 37)) + [151, 152, 254, 255, 355, 356, 406, 407, 766]


def make_syn_sweep(ref, path):
    """The synthesis-visualisation sweep of VFNet.forward (volumetric_fusionnet.py:320-336, utils/visualize.py:9-57) run
    by the reference on the rig of golden_fusion.npz: every view's angles, and for the views of SWEEP_PICK the
    augmented inverse intrinsics and camera 0's raw ray samples (reduce_dim replaced by the identity)."""
    import importlib
    cfg = fusion_cfg()
    cfg['eval'] = dict(cfg.get('eval', {}), syn_visualize=True)
    inputs, extras = make_rig(cfg, seed=4242, with_images=False)
    C = cfg['model']['fusion_feat_in_dim']
    lvl = cfg['model']['fusion_level'] + 1
    torch.manual_seed(7)
    net = ref.VFNet(cfg, C, 16, model='depth')
    net.reduce_dim = torch.nn.Identity()
    assert net.syn_visualize
    with torch.no_grad():
        fd = net(inputs, extras['feats_agg'])
    params = importlib.import_module('utils.visualize').aug_depth_params(inputs[('K', lvl)])
    assert len(params) == len(fd['syn_feat']) == 767
    out = {'angles': np.array([[float(a) for a in p[1:]] for p in params], dtype=np.float64),
           'pick': np.array(SWEEP_PICK), 'inv_K_aug': np.stack([_np(params[i][0]) for i in SWEEP_PICK]),
           'syn_feat': np.stack([_np(fd['syn_feat'][i]) for i in SWEEP_PICK])}
    np.savez_compressed(path, **out)
    print('wrote', path, {k: v.shape for k, v in out.items()})


if __name__ == '__main__':
    if '--only-sweep' in sys.argv:
        make_syn_sweep(ref_import.load(), os.path.join(HERE, 'golden_syn_sweep.npz'))
        sys.exit(0)
    if '--only-loss' in sys.argv:
        make_loss_free(ref_import.load(), os.path.join(HERE, 'golden_loss_free.npz'))
        sys.exit(0)
    if '--only-agg' in sys.argv:
        make_feat_agg(os.path.join(HERE, 'golden_feat_agg.npz'))
        sys.exit(0)
    ref = ref_import.load()
    if '--only-new' not in sys.argv:
        make_fusion(ref, os.path.join(HERE, 'golden_fusion.npz'))
        make_warp(ref, os.path.join(HERE, 'golden_warp.npz'))
    make_to_depth(os.path.join(HERE, 'golden_to_depth.npz'))
    make_depth_syn(ref, os.path.join(HERE, 'golden_depth_syn.npz'))
    make_feat_agg(os.path.join(HERE, 'golden_feat_agg.npz'))
    make_loss_free(ref, os.path.join(HERE, 'golden_loss_free.npz'))
    make_syn_sweep(ref, os.path.join(HERE, 'golden_syn_sweep.npz'))
