"""The chained fusion-depth training step (BASELINE.json configs[3] and [4]) on the drop-in modules.

What the reference runs per step (models/vfdepth.py:191-313): two pose-network forwards (frames [-1,0] and [0,1],
models/geometry/pose.py:32-41; network/fusion_posenet.py:43-75: encoder -> aggregation -> pose VFNet -> BEV -> pose
decoder), the canonical pose distributed to every camera (pose.py:44-64), one depth-network forward
(network/fusion_depthnet.py:41-77: encoder -> aggregation -> depth VFNet -> decoder), ``to_depth`` (vfdepth.py:266-288),
then per camera ViewRendering + MultiCamLoss (vfdepth.py:290-313), one backward and an Adam step; under DDP
(vfdepth.py:62-70) the gradients of both networks are all-reduced during the backward.

In scope here is everything between the encoders and the decoders -- ``aggregate_features``, ``VFNet`` (depth and
pose), ``to_depth``, the rig-level ``warp_and_loss`` -- on the CUDA kernels.  The ResNet encoders and the depth / pose
decoders are OUT of scope: ``StandInEncoder`` and the two ``*DecoderLite`` classes are random-init
stand-ins with the reference's tensor shapes at every interface (five encoder levels at 1/2 .. 1/32 with 64, 64, 128,
256, 512 channels; disparity at full resolution; 6 pose numbers from the 25 x 25 BEV map), so that the step has the
reference's data flow, parameter-gradient traffic and kernel shapes without claiming its backbone.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import aggregate, view
from .rig import axis_angle_to_matrix
from .vfnet import VFNet

ENC_CH = [64, 64, 128, 256, 512]        # ResnetEncoder.num_ch_enc (external/layers.py), levels 1/2 .. 1/32


class StandInEncoder(nn.Module):
    """five stride-2 conv3x3 + ReLU stages with the ResNet-18/34 level widths (stand-in: the backbone is out of scope)"""

    def __init__(self, n_images):
        super().__init__()
        chans = [3 * n_images] + ENC_CH
        self.stages = nn.ModuleList(nn.Conv2d(chans[i], chans[i + 1], 3, stride=2, padding=1) for i in range(5))
        self.num_ch_enc = list(ENC_CH)

    def forward(self, x):
        feats = []
        for conv in self.stages:
            x = F.relu(conv(x))
            feats.append(x)
        return feats


def _conv(cin, cout, k=3, act='ELU'):
    """the reference's conv2d factory (network/blocks.py:41-61): (conv, identity, activation)"""
    a = {'ELU': nn.ELU(inplace=True), 'LRU': nn.LeakyReLU(0.1, inplace=True), None: nn.Identity()}[act]
    return nn.Sequential(nn.Conv2d(cin, cout, k, padding=(k - 1) // 2, padding_mode='reflect'), nn.Identity(), a)


class DepthDecoderLite(nn.Module):
    """DepthDecoder of network/fusion_depthnet.py:80-129 for use_skips=False, scales=[0]: from the fused 1/8-resolution
    features up to a full-resolution sigmoid disparity (three conv / nearest-upsample / conv stages)."""

    def __init__(self, level_in, ch_in):
        super().__init__()
        dec = [16, 32, 64, 128, 256]
        layers = []
        for i in range(level_in, -1, -1):
            cin = ch_in if i == level_in else dec[i + 1]
            layers += [_conv(cin, dec[i]), _conv(dec[i], dec[i])]
        self.convs = nn.ModuleList(layers)
        self.dispconv = _conv(dec[0], 1, act=None)

    def forward(self, x):
        for i in range(0, len(self.convs), 2):
            x = self.convs[i](x)
            x = F.interpolate(x, scale_factor=2, mode='nearest')
            x = self.convs[i + 1](x)
        return torch.sigmoid(self.dispconv(x))


class PoseDecoderLite(nn.Module):
    """stand-in for external/layers.py PoseDecoder: BEV features -> (axis-angle, translation), each [B,1,1,3]"""

    def __init__(self, ch_in):
        super().__init__()
        self.squeeze = nn.Conv2d(ch_in, 256, 1)
        self.conv = nn.Conv2d(256, 256, 3, stride=2, padding=1)
        self.out = nn.Conv2d(256, 6, 1)

    def forward(self, bev):
        x = F.relu(self.squeeze(bev))
        x = F.relu(self.conv(x))
        x = 0.01 * self.out(x).mean((2, 3)).view(-1, 1, 1, 6)
        return x[..., :3], x[..., 3:]


class _FusionNet(nn.Module):
    """shared front half of FusedDepthNet / FusedPoseNet: encoder -> multi-level aggregation -> VFNet"""

    def __init__(self, cfg, n_images, model):
        super().__init__()
        m = cfg['model']
        self.cfg, self.lev = cfg, m['fusion_level']
        self.batch_size, self.num_cams = cfg['training']['batch_size'], cfg['data']['num_cams']
        self.encoder = StandInEncoder(n_images)
        self.conv1x1 = _conv(sum(ENC_CH[self.lev:]), m['fusion_feat_in_dim'], k=1, act='LRU')
        self.fusion_net = VFNet(cfg, m['fusion_feat_in_dim'], ENC_CH[self.lev], model=model)

    def aggregate(self, packed_feats):
        """fusion_depthnet.py:57-66 -> feats_agg [B,cams,C,h,w] (channel-last storage)"""
        return aggregate.aggregate_features(packed_feats, self.lev, self.conv1x1, packed_feats[0].shape[0] // self.num_cams,
                                            self.num_cams)

    def fuse(self, inputs, feats_agg):
        return self.fusion_net(inputs, feats_agg)


class FusionDepthNet(_FusionNet):
    """FusedDepthNet.forward (network/fusion_depthnet.py:41-77) -> ('disp', 0) of all cameras, [B,cams,H,W]"""

    def __init__(self, cfg):
        super().__init__(cfg, 1, 'depth')
        self.decoder = DepthDecoderLite(self.lev, ENC_CH[self.lev])

    def forward(self, inputs):
        img = inputs[('color_aug', 0, 0)]
        B, n = img.shape[:2]
        feats = self.encoder(img.flatten(0, 1))
        proj = self.fuse(inputs, self.aggregate(feats))['proj_feat']        # [B*cams, 128, h, w]
        disp = self.decoder(proj)
        return disp.view(B, n, *disp.shape[-2:])


class FusionPoseNet(_FusionNet):
    """FusedPoseNet.forward (network/fusion_posenet.py:43-75) -> canonical (axis-angle, translation)"""

    def __init__(self, cfg):
        super().__init__(cfg, 2, 'pose')
        self.pose_decoder = PoseDecoderLite(ENC_CH[self.lev])

    def forward(self, inputs, frame_ids):
        img = torch.cat([inputs[('color_aug', frame_ids[0], 0)], inputs[('color_aug', frame_ids[1], 0)]], 2)
        feats = self.encoder(img.flatten(0, 1))
        bev = self.fuse(inputs, self.aggregate(feats))
        axis_angle, translation = self.pose_decoder(bev)
        return axis_angle, torch.clamp(translation, -4.0, 4.0)


def vec_to_matrix(rot_angle, trans_vec, invert=False):
    """models/geometry/geometry_util.py:8-30: [B,1,3] axis-angle + [B,1,3] translation -> [B,4,4]"""
    b = rot_angle.shape[0]
    R = torch.eye(4, device=rot_angle.device, dtype=rot_angle.dtype).repeat(b, 1, 1)
    T = torch.eye(4, device=rot_angle.device, dtype=rot_angle.dtype).repeat(b, 1, 1)
    R[:, :3, :3] = axis_angle_to_matrix(rot_angle).squeeze(1)
    t = trans_vec.clone().contiguous().view(-1, 3, 1)
    if invert:
        R = R.transpose(1, 2)
        t = -1 * t
    T[:, :3, 3:] = t
    return torch.matmul(R, T) if invert else torch.matmul(T, R)


def predict_poses(pose_net, inputs, frame_ids, num_cams):
    """Pose.compute_pose for pose_model='fusion' (pose.py:22-64): one network call per source frame, the canonical
    motion distributed to every camera.  -> {f: [B,cams,4,4]}"""
    E, Einv = inputs['extrinsics'], inputs['extrinsics_inv']
    out = {}
    for f in frame_ids[1:]:
        pair = [-1, 0] if f < 0 else [0, 1]
        axis_angle, translation = pose_net(inputs, pair)
        ref_T = vec_to_matrix(axis_angle[:, 0], translation[:, 0], invert=(f < 0)).to(E.dtype)     # .float() at pose.py:57
        out[f] = torch.stack([Einv[:, c] @ E[:, 0] @ ref_T @ Einv[:, 0] @ E[:, c] for c in range(num_cams)], 1)
    return out


def forward_loss(cfg, depth_net, pose_net, inputs, noise=None, want_views=()):
    """VFDepthAlgo.process_batch (models/vfdepth.py:191-204) -> (total loss, detail dict)."""
    t = cfg['training']
    n = cfg['data']['num_cams']
    inputs = dict(inputs)
    inputs['extrinsics_inv'] = torch.inverse(inputs['extrinsics'])                     # vfdepth.py:211
    cam_T_cam = predict_poses(pose_net, inputs, t['frame_ids'], n)
    disp = depth_net(inputs)                                                           # [B,cams,H,W]
    depth = view.to_depth(cfg, disp, inputs[('K', 0)])
    if noise is None:
        B = disp.shape[0]
        noise = torch.randn((B, n, len(t['frame_ids']) - 1, t['height'], t['width']), device=disp.device)
    total, det = view.warp_and_loss(cfg, inputs, depth, disp, cam_T_cam, noise, want_views=want_views)
    det.update(disp=disp, depth=depth, cam_T_cam=cam_T_cam)
    return total, det


class TrainStep:
    """both networks, Adam, and one optimisation step; ``ddp=True`` wraps the networks in DistributedDataParallel
    (the process group must exist) as models/vfdepth.py:62-70 does, so the gradient all-reduce of every parameter --
    the VFNet MLPs and reduce_dim convs included -- runs inside ``step`` overlapped with the backward."""

    def __init__(self, cfg, device, ddp=False, lr=1e-4, seed=0, depth_cls=None, pose_cls=None, loss_fn=None):
        """``depth_cls`` / ``pose_cls`` / ``loss_fn`` default to the CUDA-backed networks and ``forward_loss``; the CPU
        tests of the multi-rank logic pass oracle-backed ones (tests/test_train_ddp_cpu.py)."""
        torch.manual_seed(seed)
        self.cfg, self.device = cfg, device
        self.loss_fn = loss_fn or forward_loss
        self.depth_net = (depth_cls or FusionDepthNet)(cfg).to(device)
        self.pose_net = (pose_cls or FusionPoseNet)(cfg).to(device)
        self.nets = [self.depth_net, self.pose_net]
        if ddp:
            from torch.nn.parallel import DistributedDataParallel as DDP
            idx = device.index if device.type == 'cuda' else None
            self.nets = [DDP(m, device_ids=None if idx is None else [idx], broadcast_buffers=True) for m in self.nets]
        self.params = [p for m in self.nets for p in m.parameters()]
        self.opt = torch.optim.Adam(self.params, lr=lr)

    def grad_bytes(self):
        return sum(p.numel() * p.element_size() for p in self.params)

    def step(self, inputs, noise=None):
        self.opt.zero_grad(set_to_none=True)
        total, _ = self.loss_fn(self.cfg, self.nets[0], self.nets[1], inputs, noise)
        total.backward()
        self.opt.step()
        return total.detach()
