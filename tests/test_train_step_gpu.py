"""Step-level parity of the chained fusion-depth training step (BASELINE.json configs[3]): the
drop-in modules chained the way VFDepthAlgo.process_batch chains them (/root/reference/models/vfdepth.py:191-313 --
two pose-network forwards, the depth network, to_depth, per-rig warp + losses, ONE backward) against the same chain
with every in-scope piece replaced by the oracle on the CPU, on a reduced nuScenes-shaped rig: total loss and the
gradient of EVERY parameter (encoders, 1x1 aggregation conv, VFNet MLPs + reduce_dim of both networks, decoders)."""
import copy

import pytest
import torch

from oracle import vfdepth_oracle as orc
from tests import helpers as H
from vfdepth_b200 import config as vcfg
from vfdepth_b200 import train_step as ts
from vfdepth_b200.rig import make_rig

pytestmark = pytest.mark.gpu


def _cfg(batch):
    return vcfg.default_config('nuscenes', batch_size=batch, model__voxel_size=[24, 20, 6],
                               model__voxel_str_p=[-24.0, -20.0, -3.0], model__voxel_unit_size=[2.0, 2.0, 1.5],
                               model__proj_d_bins=10, model__proj_d_end=30, model__fusion_feat_in_dim=32,
                               model__voxel_pre_dim=[16], training__height=88, training__width=160,
                               training__focal_length_scale=75)


class _OracleFusion:
    """mixin: the in-scope pieces of _FusionNet on the oracle (CPU)"""
    slopes = None      # LeakyReLU pattern of the kernel's voxel volume (depth model), pinned as in tests/test_fusion_gpu.py

    def aggregate(self, packed_feats):
        conv = self.conv1x1[0]
        y = orc.aggregate_features(packed_feats, self.lev, conv.weight, conv.bias)
        n = self.num_cams
        return y.view(y.shape[0] // n, n, *y.shape[1:])

    def fuse(self, inputs, feats_agg):
        vf, cfg = self.fusion_net, self.cfg
        lvl = cfg['model']['fusion_level'] + 1
        args = (cfg, feats_agg, inputs['mask'], inputs[('K', lvl)], inputs['extrinsics_inv'])
        if vf.model == 'pose':
            vox = orc.backproject_into_voxel(*args, 'pose')
            B = vox.shape[0]
            X, Y, Z = cfg['model']['voxel_size']
            return vf.reduce_dim(vox.view(B, -1, Y, X))                       # volumetric_fusionnet.py:339-342
        mlp = dict(w_no=vf.conv_non_overlap[0].weight, b_no=vf.conv_non_overlap[0].bias,
                   w_ov=vf.conv_overlap[0].weight, b_ov=vf.conv_overlap[0].bias)
        vox = orc.backproject_into_voxel(*args, 'depth', mlp, slopes_from=self.slopes)
        proj = orc.project_voxel_into_image(cfg, vox, inputs[('inv_K', lvl)], inputs['extrinsics'])
        return {'proj_feat': torch.stack([vf.reduce_dim(p) for p in proj], 1).flatten(0, 1)}


class _OracleDepthNet(_OracleFusion, ts.FusionDepthNet):
    pass


class _OraclePoseNet(_OracleFusion, ts.FusionPoseNet):
    pass


def _oracle_loss(cfg, depth_net, pose_net, inputs, noise, decisions):
    n = cfg['data']['num_cams']
    inputs = dict(inputs)
    inputs['extrinsics_inv'] = torch.inverse(inputs['extrinsics'])
    cam_T_cam = ts.predict_poses(pose_net, inputs, cfg['training']['frame_ids'], n)
    disp = depth_net(inputs)
    disps = [disp[:, c:c + 1] for c in range(n)]
    depths = [orc.to_depth(cfg, d, inputs[('K', 0)][:, c]) for c, d in enumerate(disps)]
    total, _ = orc.warp_and_loss(cfg, inputs, depths, disps, cam_T_cam, [noise[:, c] for c in range(n)], decisions)
    return total


@pytest.fixture
def fp32_convs():
    """cuDNN may run fp32 convolutions on TF32 tensor cores (torch default, which the reference keeps: train.py:23-25 only
    switches matmul TF32 off) -- a 1e-3 error in every conv of the stand-in encoders / decoders that would drown the
    comparison of the pieces under test"""
    prev = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32 = prev


def test_chained_step_matches_oracle_chain(fp32_convs):
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    dev = torch.device('cuda:0')
    B = 2
    cfg = _cfg(B)
    inputs, ex = make_rig(cfg, seed=4242, with_feats=False)
    noise = torch.stack(ex['noise'], 1)                                   # [B,cams,2,H,W]
    step = ts.TrainStep(cfg, dev, seed=3)
    # a random-init network predicts a nearly constant disparity and no ego-motion at all -- a degenerate geometry (with
    # T = identity the depth gradient of the warp vanishes and fp32 autograd returns cancellation noise): give the last
    # layers some signal and the pose decoder a motion of the size real rigs have (~0.6 m, ~0.3 degrees per frame)
    with torch.no_grad():
        step.depth_net.decoder.dispconv[0].weight.mul_(20.0)
        step.pose_net.pose_decoder.out.bias.copy_(torch.tensor([0.5, -0.3, 0.2, 10.0, -5.0, 60.0]))
    inp_d = {k: v.to(dev) for k, v in inputs.items()}

    seen = {}
    orig = step.depth_net.fusion_net.backproject_into_voxel

    def spy(*a, **k):
        out = orig(*a, **k)
        seen['vox'] = out.detach()
        return out

    step.depth_net.fusion_net.backproject_into_voxel = spy
    total, det = ts.forward_loss(cfg, step.depth_net, step.pose_net, inp_d, noise.to(dev),
                                 want_views=('reproj_mask', 'selection'))
    total.backward()
    v = det['views']
    dec = []
    for cam in range(6):
        sel = v['selection'][:, cam:cam + 1].cpu().long()
        dec.append(dict(reproj_mask=v['reproj_mask'][:, cam:cam + 1].cpu().float(), sel_reproj=sel % 2, sel_st=sel // 2))

    # the same networks (same weights) on the CPU with the oracle in place of every kernel, in fp32 and in fp64: the
    # fp64 run arbitrates (tests/test_warp_gpu.py does the same for the loss gradients, whose fp32 reference is itself up
    # to a few 1e-4 away from fp64 -- the warp's depth gradient is a difference of nearly equal terms)
    def oracle_grads(dtype):
        o_depth, o_pose = _OracleDepthNet(cfg), _OraclePoseNet(cfg)
        o_depth.load_state_dict(copy.deepcopy(step.depth_net.state_dict()))
        o_pose.load_state_dict(copy.deepcopy(step.pose_net.state_dict()))
        o_depth, o_pose = o_depth.to(dtype), o_pose.to(dtype)
        o_depth.slopes = seen['vox'].cpu()
        cv = lambda t: t.to(dtype) if t.is_floating_point() else t
        t_ref = _oracle_loss(cfg, o_depth, o_pose, {k: cv(v) for k, v in inputs.items()}, cv(noise), dec)
        t_ref.backward()
        return float(t_ref), [q.grad for q in list(o_depth.parameters()) + list(o_pose.parameters())]

    t32, g32 = oracle_grads(torch.float32)
    t64, g64 = oracle_grads(torch.float64)
    names = [n for n, _ in list(step.depth_net.named_parameters()) + list(step.pose_net.named_parameters())]
    got = [p.grad for p in list(step.depth_net.parameters()) + list(step.pose_net.parameters())]
    rec = {'total': float(total), 'total_oracle_fp64': t64, 'total_rel_err': abs(float(total) - t64) / abs(t64),
           'total_rel_err_oracle_fp32': abs(t32 - t64) / abs(t64), 'params': {}}
    failed = []
    for name, a, b32, b64 in zip(names, got, g32, g64):
        assert a is not None and b64 is not None, name
        e, e32 = H.rel_err(a.cpu(), b64), H.rel_err(b32, b64)
        bar = max(1e-3, 2 * e32)
        rec['params'][name] = {'err': e, 'oracle_fp32_err': e32, 'bar': bar, 'checksum': float(a.double().sum()),
                               'checksum_oracle_fp64': float(b64.sum())}
        if e > bar:
            failed.append((name, e, bar))
    rec['worst_param_grad_err'] = max(v['err'] for v in rec['params'].values())
    from tests.test_fusion_gpu import _report
    _report('chained_step_vs_oracle_chain', rec)
    assert rec['total_rel_err'] < 1e-4
    assert not failed, failed


def test_train_step_runs_and_descends():
    """one TrainStep.step at the reduced size: finite loss, every parameter gets a gradient and moves"""
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    dev = torch.device('cuda:0')
    cfg = _cfg(2)
    inputs, _ = make_rig(cfg, seed=99, with_feats=False)
    step = ts.TrainStep(cfg, dev, seed=5)
    before = [p.detach().clone() for p in step.params]
    inp_d = {k: v.to(dev) for k, v in inputs.items()}
    losses = [float(step.step(inp_d)) for _ in range(3)]
    assert all(torch.isfinite(torch.tensor(losses)))
    moved = sum(int(not torch.equal(a, b.detach())) for a, b in zip(before, step.params))
    assert moved == len(step.params), (moved, len(step.params))
