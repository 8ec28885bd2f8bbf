"""Shared helpers for the parity tests: load golden fixtures into the dict layout the oracle and
the product modules consume."""
import os

import numpy as np
import torch

from vfdepth_b200 import config as vcfg

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')

FUSION_CFG = dict(model__voxel_size=[14, 12, 4], model__voxel_str_p=[-17.5, -15.0, -3.0],
                  model__voxel_unit_size=[2.5, 2.5, 1.5], model__proj_d_bins=6, model__proj_d_str=2,
                  model__proj_d_end=16, model__fusion_feat_in_dim=8, model__voxel_pre_dim=[4],
                  training__height=64, training__width=96)
WARP_CFG = dict(training__height=48, training__width=64, training__focal_length_scale=30)


def fusion_cfg():
    return vcfg.default_config('ddad', batch_size=2, **FUSION_CFG)


def warp_cfg():
    return vcfg.default_config('ddad', batch_size=2, **WARP_CFG)


def load_npz(name):
    with np.load(os.path.join(GOLDEN, name)) as z:
        return {k: torch.from_numpy(z[k]) for k in z.files}


def fusion_inputs(g, cfg):
    lvl = cfg['model']['fusion_level'] + 1
    return {'mask': g['mask'], ('K', lvl): g['K'], ('inv_K', lvl): g['inv_K'],
            'extrinsics': g['extrinsics'], 'extrinsics_inv': g['extrinsics_inv']}


def warp_inputs(g):
    inp = {'mask': g['mask'], ('K', 0): g['K'], ('inv_K', 0): g['inv_K'],
           'extrinsics': g['extrinsics'], 'extrinsics_inv': g['extrinsics_inv']}
    for f in (0, -1, 1):
        inp[('color', f, 0)] = g['color_%d' % f]
    return inp


def rel_err(a, b):
    """max |a-b| / max(|b|) -- the 'relative to the tensor's scale' error used for fp32 parity."""
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


SYN_CFG = dict(training__aug_depth=True, loss__depth_con_coeff=0.03, loss__depth_sm_coeff=0.05)


def syn_cfg():
    return vcfg.default_config('ddad', batch_size=2, **dict(WARP_CFG, **SYN_CFG))


def syn_inputs(g):
    return {'mask': g['mask'], ('K', 0): g['K'], ('inv_K', 0): g['inv_K'], 'extrinsics': g['extrinsics'],
            'extrinsics_aug': g['extrinsics_aug']}
