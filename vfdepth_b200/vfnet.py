"""VFNet -- drop-in for /root/reference/network/volumetric_fusionnet.py:11-343 on top of the CUDA kernels.

Same constructor, same ``forward(inputs, feats_agg)``, same sub-methods and the same state-dict keys
(``conv_overlap.0.*``, ``conv_non_overlap.0.*``, ``reduce_dim.0.*``, ``reduce_dim.3.*``) so published
checkpoints load.  What changes is underneath:

* the per-camera loops of ``backproject_into_voxel`` (:124-149) and ``project_voxel_into_image`` (:243-262)
  are one kernel launch each (vfdepth_b200/csrc/voxel_gather.cu, ray_sample.cu);
* everything is channel-last: the gather writes the two MLP inputs as rows ``[B,N,257(+pad)]`` so the 1x1
  ``Conv1d`` of :197-230 is a plain GEMM (``F.linear`` with the Conv1d's own weight) whose output IS the
  channel-last voxel volume the ray kernel reads, and the ray kernel emits samples ``[B*6,h,w,D,Cv]`` (one
  contiguous 256 B run per sample) which are handed to ``reduce_dim`` as a channels_last NCHW tensor with channel
  = d*Cv + c; the reference's order c*D + d (:262) is restored by permuting the INPUT channels of
  ``reduce_dim.0.weight`` on the fly (the parameter itself keeps the reference's shape and meaning).

The learnable parts (the two MLPs and ``reduce_dim``) stay in PyTorch / cuBLAS / cuDNN.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import config as vcfg
from . import ops
from .rig import axis_angle_to_matrix


def _act():
    return nn.LeakyReLU(0.1, inplace=True)


def _conv1x1_block(cin, cout):
    # the reference's conv1d() factory (network/blocks.py:64-83) with kernel_size=1: (conv, identity, act)
    return nn.Sequential(nn.Conv1d(cin, cout, kernel_size=1, padding=0, bias=True, padding_mode='reflect'),
                         nn.Identity(), _act())


def _conv3x3_layers(cin, cout, stride):
    # the reference's conv2d() factory (network/blocks.py:41-61): (conv, identity, act)
    return [nn.Conv2d(cin, cout, kernel_size=3, stride=stride, padding=1, bias=True, padding_mode='reflect'),
            nn.Identity(), _act()]


_DEG = 0.0174533      # the reference's rounded pi / 180 (utils/visualize.py:6)


def _sweep(top, step=2):
    """0 -> top -> -top -> 0 in steps of ``step`` (both turning points visited twice, as the reference's lists do)."""
    return list(range(0, top + 1, step)) + list(range(top, -top - 1, -step)) + list(range(-top, 1, step))


def aug_depth_params(K, n_steps=75):
    """Virtual-view sweep of the synthesis visualisation (reference: utils/visualize.py:9-57): a list of
    ``[inv_K_aug [B,cams,4,4], ang_x, ang_y, ang_z]`` -- roll +-10 deg, pitch +-5 deg, two focal-length ramps (camera
    1 / camera 0 focal ratio, divided by 1.5 and back) and a full yaw turn at the last ramp's intrinsics."""
    inv_K = torch.inverse(K)
    params = [[inv_K, _DEG * (a / n_steps * 10.), 0.0, 0.0] for a in _sweep(n_steps)]
    params += [[inv_K, 0.0, _DEG * (a / 10.), 0.0] for a in _sweep(50)]
    ratio = K[:, 1, 0, 0] / K[:, 0, 0, 0]
    ratio_aug = ratio / 1.5
    ramps = [(i / 100. * ratio_aug + (1 - i / 100.)) for i in range(101)]
    ramps += [(i / 50. * ratio + (1 - i / 50.) * ratio_aug) for i in range(51)]
    for scale in ramps:
        K_aug = K.clone()
        K_aug[:, :, 0, 0] *= scale[:, None]
        K_aug[:, :, 1, 1] *= scale[:, None]
        params.append([torch.inverse(K_aug), 0.0, 0.0, 0.0])
    inv_last = torch.inverse(K_aug)
    params += [[inv_last, 0.0, 0.0, _DEG * 360 * (i / 360.)] for i in range(360)]
    return params


class VFNet(nn.Module):
    """Surround-view volumetric fusion (reference: VFNet, volumetric_fusionnet.py:11)."""

    def __init__(self, cfg, feat_in_dim, feat_out_dim, model='depth'):
        super().__init__()
        vcfg.flatten_into(self, cfg)          # the reference's read_config (:62-65)
        self.eps = 1e-8
        self.model = model
        self.voxel_end_p = [self.voxel_str_p[i] + self.voxel_unit_size[i] * (self.voxel_size[i] - 1)
                            for i in range(3)]
        self.x_dim, self.y_dim, self.z_dim = self.voxel_size
        self.n_voxels = self.x_dim * self.y_dim * self.z_dim
        self.img_h = self.height // (2 ** (self.fusion_level + 1))
        self.img_w = self.width // (2 ** (self.fusion_level + 1))
        self.num_pix = self.img_h * self.img_w
        self.feat_in_dim = feat_in_dim

        if model == 'depth':
            self.v_dim_o = [(feat_in_dim + 1) * 2] + list(self.voxel_pre_dim)
            self.v_dim_no = [feat_in_dim + 1] + list(self.voxel_pre_dim)
            self.conv_overlap = _conv1x1_block(self.v_dim_o[0], self.v_dim_o[1])
            self.conv_non_overlap = _conv1x1_block(self.v_dim_no[0], self.v_dim_no[1])
            encoder_dims = self.proj_d_bins * self.v_dim_o[-1]
            stride = 1
            cv = self.v_dim_o[-1]
        else:
            encoder_dims = (feat_in_dim + 1) * self.z_dim
            stride = 2
            cv = feat_in_dim + 1
        self.reduce_dim = nn.Sequential(*_conv3x3_layers(encoder_dims, 256, stride),
                                        *_conv3x3_layers(256, feat_out_dim, stride))
        self.geom = ops.FusionGeometry(self.batch_size, self.num_cams, feat_in_dim, self.img_h, self.img_w,
                                       self.voxel_size, self.voxel_str_p, self.voxel_unit_size, self.proj_d_bins,
                                       self.proj_d_str, self.proj_d_end, cv)
        self.last_count = None      # [B,N] uint8 camera count of the last gather (:152), for inspection
        self.fused_mlp = True       # gather + MLPs in one kernel (else: compact gather rows + cuBLAS GEMMs)

    # ------------------------------------------------------------------------------------------
    @staticmethod
    def _channel_last(feats_agg):
        """[B,cams,C,h,w] -> contiguous [B,cams,h,w,C]; free when the producer already emits channels_last."""
        return feats_agg.permute(0, 1, 3, 4, 2).contiguous()

    def _mask_small(self, input_mask):
        # the resize of :129 stays in torch so the thresholded values are the reference's
        B, ncam = input_mask.shape[:2]
        m = F.interpolate(input_mask.reshape(B * ncam, 1, *input_mask.shape[-2:]).float(),
                          [self.img_h, self.img_w], mode='bilinear', align_corners=True)
        return m.view(B, ncam, self.img_h, self.img_w).contiguous()

    def _mlp(self, rows, block, out_mask):
        """1x1 Conv1d + LeakyReLU(0.1) of :197-230 on channel-last rows [B,N,ld] (zero-padded columns)."""
        conv = block[0]
        wgt = conv.weight.squeeze(-1)
        return F.leaky_relu(F.linear(rows, wgt, conv.bias), 0.1) * out_mask

    def gather_rows(self, feats_agg, input_mask, intrinsics, extrinsics_inv, detail=False):
        """The geometry core of ``backproject_into_voxel``: the kernel's raw channel-last outputs."""
        feats_cl = self._channel_last(feats_agg)
        mask_small = self._mask_small(input_mask)
        K = intrinsics.float().contiguous()
        Einv = extrinsics_inv.float().contiguous()
        if self.model == 'depth':
            return ops.voxel_gather_depth(feats_cl, mask_small, K, Einv, self.geom, detail)
        return ops.voxel_gather_pose(feats_cl, mask_small, K, Einv, self.geom)

    def backproject_into_voxel(self, feats_agg, input_mask, intrinsics, extrinsics_inv):
        """-> voxel_feat [B,C',N] (a permuted view of channel-last rows [B,N,C']).  Reference :116-164."""
        g = self.geom
        if self.model == 'depth':
            # only voxels seen by exactly 1 / exactly 2 cameras reach an MLP (:200-206, :227-229): classify first,
            # gather compact GEMM rows for those, scatter the MLP outputs into the (otherwise zero) volume
            feats_cl = self._channel_last(feats_agg)
            mask_small = self._mask_small(input_mask)
            K = intrinsics.float().contiguous()
            Einv = extrinsics_inv.float().contiguous()
            if self.fused_mlp and g.C <= 256 and g.Cv <= 64 and g.Cv % 4 == 0:
                # the MLPs commute with the bilinear gather -- project every feature pixel once (dense
                # tcgen05 + TMA product), gather Cv-float rows per voxel; no host round trip, no GEMM row in HBM
                count, table = ops.voxel_classify(mask_small, K, Einv, g, table=True)
                self.last_count = count
                vol = ops.voxel_fuse(feats_cl, self.conv_non_overlap[0].weight.squeeze(-1),
                                     self.conv_non_overlap[0].bias, self.conv_overlap[0].weight.squeeze(-1),
                                     self.conv_overlap[0].bias, count, table, g)
                return vol.permute(0, 2, 1)
            count, slot, idx_no, idx_ov = ops.voxel_classify(mask_small, K, Einv, g)
            self.last_count = count
            x_no, x_ov = ops.voxel_gather_depth_compact(feats_cl, mask_small, K, Einv, slot, idx_no.numel(),
                                                        idx_ov.numel(), g)
            c1 = self.feat_in_dim + 1
            w_no = F.pad(self.conv_non_overlap[0].weight.squeeze(-1), (0, g.ld_no - c1))
            w_ov = self.conv_overlap[0].weight.squeeze(-1)
            w_ov = torch.cat([F.pad(w_ov[:, :c1], (0, g.half_ov - c1)),
                              F.pad(w_ov[:, c1:], (0, g.ld_ov - g.half_ov - c1))], 1)
            y_no = F.leaky_relu(F.linear(x_no, w_no, self.conv_non_overlap[0].bias), 0.1)
            y_ov = F.leaky_relu(F.linear(x_ov, w_ov, self.conv_overlap[0].bias), 0.1)
            B = feats_cl.shape[0]
            rows = torch.zeros((B * g.N, y_no.shape[1]), device=feats_cl.device, dtype=torch.float32)
            rows = rows.index_copy(0, idx_no, y_no).index_copy(0, idx_ov, y_ov)
            return rows.view(B, g.N, -1).permute(0, 2, 1)     # [B,N,Cv] == channel-last voxel volume
        avg, count = self.gather_rows(feats_agg, input_mask, intrinsics, extrinsics_inv)
        self.last_count = count
        return avg[:, :, :self.feat_in_dim + 1].permute(0, 2, 1)

    def ray_rows(self, voxel_feat, inv_K, extrinsics, want_grid=False):
        """The geometry core of ``project_voxel_into_image``: ray samples [B,cams,h,w,D,Cv], c fastest."""
        B, cv, _ = voxel_feat.shape
        vol = voxel_feat.permute(0, 2, 1).contiguous().view(B, self.z_dim, self.y_dim, self.x_dim, cv)
        return ops.ray_sample(vol, inv_K.float().contiguous(), extrinsics.float().contiguous(), self.geom, want_grid)

    def project_voxel_into_image(self, voxel_feat, inv_K, extrinsics):
        """-> list over cameras of [B,feat_out,h,w].  Reference :232-267."""
        rows = self.ray_rows(voxel_feat, inv_K, extrinsics)
        B, ncam, _, _, D, cv = rows.shape
        x = rows.view(B * ncam, self.img_h, self.img_w, D * cv).permute(0, 3, 1, 2)   # channels_last NCHW
        conv0 = self.reduce_dim[0]
        # input channel c*D+d of the reference (:262) is channel d*cv+c here: permute the weight, not the samples
        w0 = conv0.weight.view(conv0.out_channels, cv, D, 3, 3).permute(0, 2, 1, 3, 4)
        w0 = w0.reshape(conv0.out_channels, D * cv, 3, 3)
        y = F.conv2d(F.pad(x, (1, 1, 1, 1), mode='reflect'), w0, conv0.bias, stride=conv0.stride)
        for layer in list(self.reduce_dim)[1:]:
            y = layer(y)
        y = y.view(B, ncam, *y.shape[1:])
        return [y[:, c] for c in range(ncam)]

    def augment_extrinsics(self, ext):
        """random rotation of every camera's extrinsics (:269-287), aug_depth only."""
        with torch.no_grad():
            b, cam = ext.shape[:2]
            angle = torch.rand(b, cam, 3)
            for i in range(3):
                angle[:, :, i] = (angle[:, :, i] - 0.5) * self.aug_angle[i]
            tform = torch.eye(4).repeat(b, cam, 1, 1)
            tform[:, :, :3, :3] = axis_angle_to_matrix(angle)
            return tform.to(device=ext.device, dtype=ext.dtype) @ ext

    @staticmethod
    def rotate_extrinsics(ext, angles):
        """(rotation by the axis-angle vector ``angles``) @ extrinsics for every camera (:321-327)."""
        tform = torch.eye(4)
        tform[:3, :3] = axis_angle_to_matrix(torch.tensor([list(angles)], dtype=torch.float32))[0]
        return tform.to(device=ext.device, dtype=ext.dtype) @ ext

    def forward(self, inputs, feats_agg):
        """Reference :289-343."""
        lvl = self.fusion_level + 1
        mask, K, inv_K = inputs['mask'], inputs['K', lvl], inputs['inv_K', lvl]
        extrinsics, extrinsics_inv = inputs['extrinsics'], inputs['extrinsics_inv']
        fusion_dict = {('cam', cam): {} for cam in range(self.num_cams)}
        if self.model == 'pose':
            # no [B,C',N] -> [B,C'*Z,Y,X] reshape copy: the gather writes the BEV map channels-last
            return self.reduce_bev(self.bev_rows(feats_agg, mask, K, extrinsics_inv))
        voxel_feat = self.backproject_into_voxel(feats_agg, mask, K, extrinsics_inv)
        if self.model == 'depth':
            feats = self.project_voxel_into_image(voxel_feat, inv_K, extrinsics)
            fusion_dict['proj_feat'] = torch.stack(feats, 1).flatten(0, 1)
            if self.aug_depth:
                inputs['extrinsics_aug'] = self.augment_extrinsics(extrinsics)
                feats = self.project_voxel_into_image(voxel_feat, inv_K, inputs['extrinsics_aug'])
                fusion_dict['proj_feat_aug'] = torch.stack(feats, 1).flatten(0, 1)
            if getattr(self, 'syn_visualize', False):
                # synthesis sweep (:320-336): camera 0's projected features for every virtual view of the sweep
                with torch.no_grad():
                    fusion_dict['syn_feat'] = [
                        self.project_voxel_into_image(voxel_feat, inv_K_aug, self.rotate_extrinsics(extrinsics, ang))[0]
                        for inv_K_aug, *ang in aug_depth_params(K)]
            return fusion_dict
        raise ValueError(self.model)

    def bev_rows(self, feats_agg, input_mask, intrinsics, extrinsics_inv):
        """pose gather with z innermost: [B,Y,X,Z*ld] = the BEV map of :339-340 in channels-last storage."""
        g = self.geom
        avg, count = ops.voxel_gather_pose(self._channel_last(feats_agg), self._mask_small(input_mask),
                                           intrinsics.float().contiguous(), extrinsics_inv.float().contiguous(), g,
                                           bev=True)
        self.last_count = count
        return avg.view(avg.shape[0], self.y_dim, self.x_dim, self.z_dim * g.ld_no)

    def reduce_bev(self, bev_rows):
        """reduce_dim (:341) on the channels-last BEV rows.  Channel z*ld + c here is channel c*Z + z of the reference's
        reshape (:340); the first conv's INPUT channels are permuted / zero-padded on the fly, the parameter keeps
        the reference's shape."""
        g = self.geom
        c1 = self.feat_in_dim + 1
        conv0 = self.reduce_dim[0]
        w0 = conv0.weight.view(conv0.out_channels, c1, self.z_dim, 3, 3).permute(0, 2, 1, 3, 4)
        w0 = F.pad(w0, (0, 0, 0, 0, 0, g.ld_no - c1)).reshape(conv0.out_channels, self.z_dim * g.ld_no, 3, 3)
        x = bev_rows.permute(0, 3, 1, 2)                      # NCHW view of channels_last storage
        y = F.conv2d(F.pad(x, (1, 1, 1, 1), mode='reflect'), w0, conv0.bias, stride=conv0.stride)
        for layer in list(self.reduce_dim)[1:]:
            y = layer(y)
        return y
