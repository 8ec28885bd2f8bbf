"""CPU: pin oracle/vfdepth_oracle.py to the fixtures produced by running the reference itself
(tests/golden/make_golden.py), bit for bit."""
import torch

from oracle import vfdepth_oracle as orc
from tests import helpers as H


def _mlp(g):
    return dict(w_no=g['w_no'], b_no=g['b_no'], w_ov=g['w_ov'], b_ov=g['b_ov'])


def test_fusion_forward_bit_exact():
    g = H.load_npz('golden_fusion.npz')
    cfg = H.fusion_cfg()
    d = orc.backproject_into_voxel(cfg, g['feats_agg'], g['mask'], g['K'], g['extrinsics_inv'], 'depth',
                                   _mlp(g), detail=True)
    assert torch.equal(torch.stack(d['grids'], 1), g['grids'])
    assert torch.equal(torch.stack(d['valids'], 1), g['valids'])
    assert torch.equal(d['voxel_feat'], g['voxel_feat'])
    proj = orc.project_voxel_into_image(cfg, d['voxel_feat'], g['inv_K'], g['extrinsics'])
    assert torch.equal(torch.stack(proj, 1), g['proj'])
    pose = orc.backproject_into_voxel(cfg, g['feats_agg'], g['mask'], g['K'], g['extrinsics_inv'], 'pose')
    assert torch.equal(pose, g['voxel_feat_pose'])


def test_fusion_backward_bit_exact():
    g = H.load_npz('golden_fusion.npz')
    cfg = H.fusion_cfg()
    feats = g['feats_agg'].clone().requires_grad_(True)
    mlp = {k: v.clone().requires_grad_(True) for k, v in _mlp(g).items()}
    vox = orc.backproject_into_voxel(cfg, feats, g['mask'], g['K'], g['extrinsics_inv'], 'depth', mlp)
    proj = orc.project_voxel_into_image(cfg, vox, g['inv_K'], g['extrinsics'])
    loss = sum((p * g['g_up'][:, c]).sum() for c, p in enumerate(proj))
    grads = torch.autograd.grad(loss, [feats, mlp['w_no'], mlp['b_no'], mlp['w_ov'], mlp['b_ov']])
    for got, key in zip(grads, ['d_feats', 'd_w_no', 'd_b_no', 'd_w_ov', 'd_b_ov']):
        assert torch.allclose(got, g[key], rtol=1e-6, atol=1e-7), key
    assert torch.equal(grads[0], g['d_feats'])


def test_warp_loss_bit_exact():
    g = H.load_npz('golden_warp.npz')
    cfg = H.warp_cfg()
    inp = H.warp_inputs(g)
    depths = [g['depth'][:, c].clone().requires_grad_(True) for c in range(6)]
    Ts = {-1: g['T_m1'].clone().requires_grad_(True), 1: g['T_p1'].clone().requires_grad_(True)}
    disps = [g['disp'][:, c] for c in range(6)]
    noises = [g['noise'][:, c] for c in range(6)]
    total, det = orc.warp_and_loss(cfg, inp, depths, disps, Ts, noises)
    assert torch.equal(total.detach(), g['total'])
    for cam in range(6):
        got = [det[cam][k].item() for k in ('reproj_loss', 'spatio_loss', 'spatio_tempo_loss', 'smooth')]
        assert got == g['cam_scalars'][cam, :4].tolist()
    for cam in (0, 3):
        v = det[cam]['view']
        for f in (-1, 1):
            assert torch.equal(v[('color', f)].detach(), g['cam%d_color_%d' % (cam, f)])
            assert torch.equal(v[('color_mask', f)].detach(), g['cam%d_color_mask_%d' % (cam, f)])
            assert torch.equal(v[('overlap_mask', f)].detach(), g['cam%d_overlap_mask_%d' % (cam, f)])
        for f in (0, -1, 1):
            assert torch.equal(v[('overlap', f)].detach(), g['cam%d_overlap_%d' % (cam, f)])
        # compute_spatio_loss overwrites ('overlap_mask', 0) with ref_mask * overlap_mask (multi_cam_loss.py:28)
        assert torch.equal(det[cam]['spatio_mask'], g['cam%d_overlap_mask_0' % cam])
        assert torch.equal(det[cam]['reproj_mask'], g['cam%d_reproj_mask' % cam])
        assert torch.equal(det[cam]['reproj_map'].detach(), g['cam%d_reproj_loss' % cam])
    grads = torch.autograd.grad(total, depths + [Ts[-1], Ts[1]])
    assert torch.equal(torch.stack(grads[:6], 1), g['d_depth'])
    assert torch.allclose(grads[6], g['d_T_m1'], rtol=1e-6, atol=1e-9)
    assert torch.allclose(grads[7], g['d_T_p1'], rtol=1e-6, atol=1e-9)


def test_to_depth_bit_exact():
    """oracle.to_depth against VFDepthAlgo.to_depth itself (models/vfdepth.py:277-288) at scales 0, 1, 2."""
    g = H.load_npz('golden_to_depth.npz')
    cfg = H.warp_cfg()
    for s in (0, 1, 2):
        disp = g['disp_%d' % s].clone().requires_grad_(True)
        depth = torch.cat([orc.to_depth(cfg, disp[:, c:c + 1], g['K'][:, c]) for c in range(6)], 1)
        assert torch.equal(depth, g['depth_%d' % s]), s
        d_disp, = torch.autograd.grad((depth * g['g_up_%d' % s]).sum(), disp)
        assert torch.equal(d_disp, g['d_disp_%d' % s]), s


def test_depth_synthesis_bit_exact():
    """oracle.depth_synthesis / aug_losses against the reference's ViewRendering (aug_depth branch, :201-241 ->
    get_virtual_depth :84-116) and DepthSynLoss.compute_aug_losses (depth_synthesis_loss.py:15-43)."""
    g = H.load_npz('golden_depth_syn.npz')
    cfg = H.syn_cfg()
    inputs = H.syn_inputs(g)
    depths = [g['depth'][:, c:c + 1].clone().requires_grad_(True) for c in range(6)]
    augs = [g['aug'][:, c:c + 1].clone().requires_grad_(True) for c in range(6)]
    disps = [g['disp_aug'][:, c:c + 1].clone().requires_grad_(True) for c in range(6)]
    total = 0
    for cam in range(6):
        td, tm = orc.depth_synthesis(cfg, inputs, depths, augs[cam], cam)
        if cam in (0, 3):
            assert torch.equal(torch.cat(td, 1), g['cam%d_tform_depth' % cam])
            assert torch.equal(torch.cat(tm, 1), g['cam%d_tform_mask' % cam])
        con, sm = orc.aug_losses(augs[cam], td, tm, disps[cam])
        assert torch.equal(con, g['con'][cam]) and torch.equal(sm, g['sm'][cam])
        total = total + g['w'][cam, 0] * con + g['w'][cam, 1] * sm
    grads = torch.autograd.grad(total, depths + augs + disps)
    assert torch.equal(torch.cat(grads[:6], 1), g['d_depth'])
    assert torch.equal(torch.cat(grads[6:12], 1), g['d_aug'])
    assert torch.equal(torch.cat(grads[12:], 1), g['d_disp_aug'])


def test_aggregate_features_bit_exact():
    """oracle.aggregate_features against what the reference's FusedDepthNet computed between its encoder and its fusion
    net (network/fusion_depthnet.py:57-66; fixture from forward hooks on the real module), forward and gradients."""
    g = H.load_npz('golden_feat_agg.npz')
    levels = [g['feat_%d' % i].clone().requires_grad_(True) for i in range(3)]
    w, b = g['weight'].clone().requires_grad_(True), g['bias'].clone().requires_grad_(True)
    agg = orc.aggregate_features(levels, 0, w, b)
    assert torch.equal(agg, g['agg'])
    grads = torch.autograd.grad((agg * g['g_up']).sum(), levels + [w, b])
    for i in range(3):
        assert torch.equal(grads[i], g['d_feat_%d' % i]), i
    assert torch.equal(grads[3], g['d_weight']) and torch.equal(grads[4], g['d_bias'])


def test_loss_free_functions_bit_exact():
    """oracle ssim_loss / photometric_loss / auto_masks / masked_loss / edg_smooth_loss_raw against the reference's
    loss_util free functions (models/losses/loss_util.py:6-78; fixture: make_golden.py --only-loss), values and
    gradients bit for bit."""
    g = H.load_npz('golden_loss_free.npz')
    pr = g['pred'].clone().requires_grad_(True)
    photo = orc.photometric_loss(pr, g['target'])
    assert torch.equal(orc.ssim_loss(pr, g['target']).detach(), g['ssim'])
    assert torch.equal(photo.detach(), g['photo'])
    assert torch.equal(torch.autograd.grad(photo.sum(), pr)[0], g['d_pred'])
    assert torch.equal(orc.auto_masks(g['reproj'], g['ident']), g['auto_mask'])
    assert torch.equal(orc.auto_masks(g['reproj'], None), g['auto_mask_none'])
    assert 0.2 < float(g['auto_mask'].mean()) < 0.9
    dd = g['disp'].clone().requires_grad_(True)
    sm = orc.edg_smooth_loss_raw(g['target'], dd)
    assert torch.equal(sm.detach(), g['smooth'])
    assert torch.equal(torch.autograd.grad(sm, dd)[0], g['d_disp'])
    ll = g['photo'].clone().requires_grad_(True)
    ml = orc.masked_loss(ll, g['mask'])
    assert torch.equal(ml.detach(), g['masked'])
    assert torch.equal(torch.autograd.grad(ml, ll)[0], g['d_loss'])


def test_syn_sweep_params_and_samples_bit_exact():
    """The synthesis-visualisation sweep (volumetric_fusionnet.py:320-336): vfdepth_b200.vfnet.aug_depth_params against
    the reference's utils/visualize.py:9-57 (all 767 angle triples, the augmented inverse intrinsics of the picked
    views), and the oracle's ray sampling of those views against the reference's 'syn_feat' -- bit for bit."""
    from vfdepth_b200.vfnet import VFNet, aug_depth_params
    g, s = H.load_npz('golden_fusion.npz'), H.load_npz('golden_syn_sweep.npz')
    cfg = H.fusion_cfg()
    params = aug_depth_params(g['K'])
    assert len(params) == 767
    assert torch.equal(torch.tensor([[float(a) for a in p[1:]] for p in params], dtype=torch.float64), s['angles'])
    for k, i in enumerate(s['pick'].tolist()):
        inv_K_aug, *ang = params[i]
        assert torch.equal(inv_K_aug, s['inv_K_aug'][k])
        E = VFNet.rotate_extrinsics(g['extrinsics'], ang)
        cam0 = orc.project_voxel_into_image(cfg, g['voxel_feat'], inv_K_aug, E)[0]
        assert torch.equal(cam0, s['syn_feat'][k]), i
