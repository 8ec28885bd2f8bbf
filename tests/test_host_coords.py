"""CPU: the coordinate header the CUDA kernels include (vfdepth_b200/csrc/vfd_coords.h), compiled as plain C,
must reproduce the oracle's (= the reference's) sampling coordinates BIT FOR BIT at full DDAD / nuScenes size.
This is what makes "bit-exact voxel-to-pixel indices and validity masks" checkable without a GPU."""
import ctypes
import os
import subprocess
import tempfile

import numpy as np
import pytest
import torch

from oracle import vfdepth_oracle as orc
from vfdepth_b200 import config as vcfg
from vfdepth_b200.rig import make_rig

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope='module')
def hc():
    out = os.path.join(tempfile.mkdtemp(prefix='vfd_hc_'), 'libhost_coords.so')
    subprocess.check_call(['gcc', '-O2', '-ffp-contract=off', '-mfma', '-shared', '-fPIC', '-o', out,
                           os.path.join(HERE, 'host_coords.c'), '-lm'])
    return ctypes.CDLL(out)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _f32(t):
    return np.ascontiguousarray(t.detach().cpu().numpy().astype(np.float32))


@pytest.mark.parametrize('dataset,seed', [('ddad', 11), ('nuscenes', 12)])
def test_voxel_to_pixel_bits(hc, dataset, seed):
    cfg = vcfg.default_config(dataset, batch_size=1)
    inputs, _ = make_rig(cfg, seed=seed, with_feats=False, with_images=False)
    lvl = cfg['model']['fusion_level'] + 1
    h, w = orc.feature_hw(cfg)
    X, Y, Z = cfg['model']['voxel_size']
    N = X * Y * Z
    ax = [_f32(a) for a in orc.voxel_axes(cfg)]
    vox = orc.voxel_points(cfg, 1)
    for c in range(6):
        Einv, K = inputs['extrinsics_inv'][:, c], inputs[('K', lvl)][:, c]
        v_ref = torch.matmul(Einv[:, :3, :], vox)
        g_ref = orc.voxel_pixel_coords(K, v_ref, w, h)
        v = np.empty((3, N), np.float32)
        g = np.empty((N, 2), np.float32)
        hc.hc_voxel_grid(_p(_f32(Einv[0, :3])), _p(_f32(K[0])), _p(ax[0]), _p(ax[1]), _p(ax[2]), X, Y, Z, w, h,
                         _p(v), _p(g))
        assert np.array_equal(v, v_ref[0].numpy())
        assert np.array_equal(g, g_ref[0, :, 0, :].numpy())


@pytest.mark.parametrize('dataset,seed', [('ddad', 21), ('nuscenes', 22)])
def test_ray_grid_bits(hc, dataset, seed):
    cfg = vcfg.default_config(dataset, batch_size=1)
    inputs, _ = make_rig(cfg, seed=seed, with_feats=False, with_images=False)
    lvl = cfg['model']['fusion_level'] + 1
    h, w = orc.feature_hw(cfg)
    m = cfg['model']
    D = m['proj_d_bins']
    end = orc.voxel_end(cfg)
    start = np.asarray(m['voxel_str_p'], np.float32)
    length = np.asarray([end[i] - m['voxel_str_p'][i] for i in range(3)], np.float32)
    bins = _f32(orc.depth_bins(cfg))
    for c in range(6):
        g_ref = orc.ray_grid(cfg, inputs[('inv_K', lvl)][:, c], inputs['extrinsics'][:, c], 1)
        g = np.empty((D, h, w, 3), np.float32)
        hc.hc_ray_grid(_p(_f32(inputs[('inv_K', lvl)][0, c])), _p(_f32(inputs['extrinsics'][0, c, :3])), _p(bins),
                       D, h, w, _p(start), _p(length), _p(g))
        assert np.array_equal(g, g_ref[0].numpy())


def test_warp_grid_bits(hc):
    cfg = vcfg.default_config('ddad', batch_size=1)
    inputs, extras = make_rig(cfg, seed=31, with_feats=False)
    H, W = cfg['training']['height'], cfg['training']['width']
    for c in range(6):
        for f in (-1, 1):
            T = extras['cam_T_cam'][f][:, c]
            g_ref = orc.reproject_grid(extras['depth'][c], T, inputs[('inv_K', 0)][:, c], inputs[('K', 0)][:, c])
            g = np.empty((H, W, 2), np.float32)
            hc.hc_warp_grid(_p(_f32(inputs[('inv_K', 0)][0, c])), _p(_f32(inputs[('K', 0)][0, c])), _p(_f32(T[0])),
                            _p(_f32(extras['depth'][c][0, 0])), H, W, _p(g))
            assert np.array_equal(g, g_ref[0].numpy())
