"""CPU: the C-ABI library builds, loads, exports every symbol include/vfdepth_b200.h declares, agrees with the ctypes
structs on layout, and rejects bad arguments with the documented status codes BEFORE touching the device (no compute
calls are made here -- there is no GPU in the build container)."""
import ctypes
import os
import re
import subprocess
import tempfile

import pytest

from vfdepth_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, 'include', 'vfdepth_b200.h')


def _declared():
    with open(HEADER) as f:
        return sorted(set(re.findall(r'\b(vfd_[a-z0-9_]+)\s*\(', f.read())))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), n
    assert lib.vfd_version() == 200


def test_header_is_plain_c_and_structs_match_ctypes():
    """the header must compile as C (no C++/torch types in the boundary) and the struct layouts must be the ones the
    ctypes mirror assumes."""
    src = r'''
#include <stdio.h>
#include <stddef.h>
#include "vfdepth_b200.h"
int main(void) {
  printf("%zu %zu %zu ", sizeof(vfd_fusion_desc), sizeof(vfd_warp_desc), sizeof(vfd_warp_views));
  printf("%zu %zu %zu ", offsetof(vfd_fusion_desc, group), offsetof(vfd_fusion_desc, depth_div), offsetof(vfd_fusion_desc, axis_x));
  printf("%zu %zu %zu ", offsetof(vfd_warp_desc, nbr), offsetof(vfd_warp_desc, color), offsetof(vfd_warp_desc, invK));
  printf("%zu %zu %zu %zu\n", sizeof(vfd_syn_desc), sizeof(vfd_agg_desc), offsetof(vfd_agg_desc, cw), offsetof(vfd_agg_desc, slope));
  return 0; }
'''
    d = tempfile.mkdtemp(prefix='vfd_abi_')
    c, exe = os.path.join(d, 't.c'), os.path.join(d, 't')
    with open(c, 'w') as f:
        f.write(src)
    subprocess.check_call(['gcc', '-std=c99', '-Wall', '-Werror', '-I', os.path.join(ROOT, 'include'), c, '-o', exe])
    got = [int(x) for x in subprocess.check_output([exe]).split()]
    F, W, V = _lib.FusionDesc, _lib.WarpDesc, _lib.WarpViews
    want = [ctypes.sizeof(F), ctypes.sizeof(W), ctypes.sizeof(V), F.group.offset, F.depth_div.offset, F.axis_x.offset,
            W.nbr.offset, W.color.offset, W.invK.offset,
            ctypes.sizeof(_lib.SynDesc), ctypes.sizeof(_lib.AggDesc), _lib.AggDesc.cw.offset, _lib.AggDesc.slope.offset]
    assert got == want


def test_error_strings():
    lib = _lib.load()
    assert lib.vfd_last_error(0) == b'ok'
    for code in range(-7, 0):
        assert lib.vfd_last_error(code) not in (b'ok', b'unknown vfd_status')
    assert lib.vfd_last_error(-99) == b'unknown vfd_status'


def test_bad_arguments_are_rejected_before_any_launch():
    lib = _lib.load()
    null = ctypes.c_void_p(0)
    # NULL descriptor
    assert lib.vfd_voxel_gather_depth_fwd(null, null, null, null, null, null, null, null, null, null, null) == -1
    assert lib.vfd_ray_sample_fwd(null, null, null, null, null, null, null) == -1
    assert lib.vfd_voxel_classify(null, null, null, null, null, null, null, null, null, null, null, ctypes.c_size_t(0), null) == -1
    assert lib.vfd_voxel_gather_depth_compact_fwd(null, null, null, null, null, null, null, null, null) == -1
    assert lib.vfd_voxel_gather_depth_compact_bwd(null, null, null, null, null, null, null, null, null) == -1
    assert lib.vfd_voxel_classify_workspace_bytes(null) == 0
    assert lib.vfd_voxel_fuse_fwd(*([null] * 10), ctypes.c_size_t(0), null) == -1
    assert lib.vfd_voxel_fuse_bwd(*([null] * 14), ctypes.c_size_t(0), null) == -1
    assert lib.vfd_voxel_table_bytes(null) == 0
    assert lib.vfd_voxel_fuse_bwd_workspace_bytes(null) == 0
    assert lib.vfd_voxel_fuse_fwd_workspace_bytes(null) == 0
    fd = _lib.FusionDesc()
    fd.B, fd.num_cams, fd.C, fd.h, fd.w, fd.X, fd.Y, fd.Z, fd.D, fd.Cv = 1, 6, 256, 48, 80, 100, 100, 20, 50, 64
    # P [6*3840][128] + three [2][128][256] weight planes, each block 256-byte aligned
    assert lib.vfd_voxel_fuse_fwd_workspace_bytes(ctypes.byref(fd)) == 6 * 3840 * 128 * 4 + 3 * 2 * 128 * 256 * 4
    fd.Cv = 66                                                                             # Cv % 4
    assert lib.vfd_voxel_fuse_fwd(ctypes.byref(fd), *([null] * 9), ctypes.c_size_t(0), null) == -6
    assert lib.vfd_smooth_loss_fwd(1, 6, 8, 8, null, null, null, null, null) == -1      # colour may be NULL, disp not
    assert lib.vfd_smooth_loss_fwd(1, 6, 1, 8, null, null, null, null, null) == -2
    assert lib.vfd_smooth_loss_bwd(1, 6, 8, 8, null, null, null, null, null, null, null) == -1
    assert lib.vfd_smooth_loss_train_fwd(1, 6, 8, 8, null, null, null, null, null, null, null) == -1
    assert lib.vfd_smooth_loss_train_bwd(1, 6, 8, 8, null, null, null, null, null, null) == -1
    assert lib.vfd_warp_loss_train_fwd(null, null, null, null, null, null, null, null, null, ctypes.c_size_t(0), null) == -1
    assert lib.vfd_warp_loss_train_bwd(null, null, null, null, ctypes.c_size_t(0), null, null, null) == -1
    f = ctypes.c_float
    assert lib.vfd_to_depth_fwd(1, 6, 8, 8, 8, 8, null, null, f(1.5), f(200.0), f(300.0), null, null) == -1
    assert lib.vfd_to_depth_fwd(1, 6, 8, 8, 4, 8, null, null, f(1.5), f(200.0), f(300.0), null, null) == -2
    assert lib.vfd_to_depth_fwd(1, 6, 8, 8, 8, 8, null, null, f(0.0), f(200.0), f(300.0), null, null) == -2
    assert lib.vfd_to_depth_bwd(1, 6, 8, 8, 8, 8, null, null, f(1.5), f(200.0), f(300.0), null, null, null) == -1
    assert lib.vfd_depth_syn_fwd(null, null, null, null, null, null, null, null, null) == -1
    assert lib.vfd_depth_syn_bwd(null, null, null, null, null, null, null, null, null) == -1
    sd = _lib.SynDesc()
    assert lib.vfd_depth_syn_fwd(ctypes.byref(sd), null, null, null, null, null, null, null, null) == -2
    assert lib.vfd_warp_stats(null, null, null, null, null, null, null) == -1
    assert lib.vfd_warp_raw_bytes(null) == 0
    assert lib.vfd_voxel_gather_pose_bev_fwd(null, null, null, null, null, null, null, null) == -1
    assert lib.vfd_voxel_gather_pose_bev_bwd(null, null, null, null, null, null, null) == -1
    assert lib.vfd_warp_image_blocks() == 148
    assert lib.vfd_warp_image_fwd(2, 8, 8, null, null, null, null, null, null, null, null, null, null) == -1
    assert lib.vfd_warp_image_fwd(2, 1, 8, null, null, null, null, null, null, null, null, null, null) == -2
    assert lib.vfd_warp_image_bwd(2, 8, 8, null, null, null, null, null, null, null, null, null, null, null) == -1
    assert lib.vfd_projection_backproject(2, 8, 8, null, null, null, null) == -1
    assert lib.vfd_photometric_loss_fwd(2, 8, 8, null, null, null, null, null) == -1
    assert lib.vfd_photometric_loss_bwd(2, 8, 8, null, null, null, null, null, null) == -1
    assert lib.vfd_photometric_loss_fwd(2, 8, 1, null, null, null, null, null) == -2
    assert lib.vfd_auto_masks(2, 2, 2, 64, null, null, null, null) == -1
    assert lib.vfd_auto_masks(2, 0, 2, 64, null, null, null, null) == -2
    assert lib.vfd_masked_loss_fwd(ctypes.c_size_t(64), null, null, null, null) == -1
    assert lib.vfd_masked_loss_fwd(ctypes.c_size_t(0), null, null, null, null) == -2
    assert lib.vfd_masked_loss_bwd(ctypes.c_size_t(64), null, null, null, null, null) == -1
    assert lib.vfd_edg_smooth_loss_fwd(2, 3, 8, 8, null, null, null, null) == -1
    assert lib.vfd_edg_smooth_loss_fwd(2, 3, 1, 8, null, null, null, null) == -2
    assert lib.vfd_edg_smooth_loss_bwd(2, 3, 8, 8, null, null, null, null, null) == -1
    assert lib.vfd_norm_image_fwd(2, 8, 8, null, null, null, null, null, null, null) == -1
    assert lib.vfd_norm_image_bwd(2, 8, 1, null, null, null, null, null) == -2
    assert lib.vfd_projection_reproject(0, 8, 8, null, null, null, null, null) == -2
    assert lib.vfd_feat_agg_fwd(null, null, null, null, null, null) == -1
    assert lib.vfd_feat_agg_bwd(null, null, null, null, null, null) == -1
    ad = _lib.AggDesc()
    assert lib.vfd_feat_agg_fwd(ctypes.byref(ad), null, null, null, null, null) == -2      # zero sizes
    ad.n_img, ad.h, ad.w, ad.C, ad.n_coarse = 6, 8, 12, 64, 1
    ad.ch[0], ad.cw[0] = 9, 6                                                              # coarser level larger than h
    assert lib.vfd_feat_agg_fwd(ctypes.byref(ad), null, null, null, null, null) == -2
    ad.ch[0] = 4
    assert lib.vfd_feat_agg_fwd(ctypes.byref(ad), null, null, null, null, null) == -1      # valid descriptor, NULL data
    ad.C = 66
    assert lib.vfd_feat_agg_bwd(ctypes.byref(ad), null, null, null, null, null) == -2      # C % 4
    # descriptor out of range
    d = _lib.FusionDesc()
    assert lib.vfd_voxel_gather_depth_fwd(ctypes.byref(d), null, null, null, null, null, null, null, null, null, null) == -2
    w = _lib.WarpDesc()
    assert lib.vfd_warp_loss_fwd(ctypes.byref(w), null, null, null, null, null, null, null, null) == -2
    # unsupported channel count (not a multiple of 4)
    d.B, d.num_cams, d.C, d.h, d.w, d.X, d.Y, d.Z, d.D, d.Cv = 1, 6, 6, 4, 4, 2, 2, 2, 2, 4
    assert lib.vfd_voxel_gather_depth_fwd(ctypes.byref(d), null, null, null, null, null, null, null, null, null, null) == -6
    # valid sizes but NULL data pointers
    w.B, w.num_cams, w.H, w.W, w.n_src, w.cam0, w.n_call = 1, 6, 8, 8, 2, 0, 6
    assert lib.vfd_warp_stats(ctypes.byref(w), null, null, null, null, null, null) == -1
    w.color[0] = w.color[1] = w.color[2] = w.mask = w.K = w.invK = 16        # non-NULL: only the sizes are read
    assert lib.vfd_warp_raw_bytes(ctypes.byref(w)) == 1 * 6 * 8 * 64 * 16 + 1 * 6 * 64
    with pytest.raises(RuntimeError, match='NULL'):
        _lib.check(-1, 'vfd_test')


def test_host_layer_refuses_cpu_tensors():
    """no CPU fallback: the host layer raises instead of computing anything off the device."""
    import torch
    from vfdepth_b200 import config as vcfg
    from vfdepth_b200.vfnet import VFNet
    cfg = vcfg.default_config('ddad', batch_size=1, model__voxel_size=[4, 4, 2], model__fusion_feat_in_dim=8,
                              model__voxel_pre_dim=[4], training__height=32, training__width=32)
    net = VFNet(cfg, 8, 8, 'depth')
    feats = torch.zeros(1, 6, 8, 4, 4)
    mask = torch.ones(1, 6, 1, 32, 32)
    eye = torch.eye(4).expand(1, 6, 4, 4)
    with pytest.raises(ValueError, match='CUDA'):
        net.backproject_into_voxel(feats, mask, eye, eye)


def test_state_dict_keys_match_reference():
    """published checkpoints must load: SURVEY.md section 5 lists the VFNet keys of the reference."""
    from vfdepth_b200 import config as vcfg
    from vfdepth_b200.vfnet import VFNet
    cfg = vcfg.default_config('ddad', batch_size=1, model__voxel_size=[4, 4, 2], model__fusion_feat_in_dim=8,
                              model__voxel_pre_dim=[4], training__height=32, training__width=32)
    keys = set(VFNet(cfg, 8, 8, 'depth').state_dict().keys())
    assert keys == {'conv_overlap.0.weight', 'conv_overlap.0.bias', 'conv_non_overlap.0.weight',
                    'conv_non_overlap.0.bias', 'reduce_dim.0.weight', 'reduce_dim.0.bias', 'reduce_dim.3.weight',
                    'reduce_dim.3.bias'}
    assert set(VFNet(cfg, 8, 8, 'pose').state_dict().keys()) == {'reduce_dim.0.weight', 'reduce_dim.0.bias',
                                                                 'reduce_dim.3.weight', 'reduce_dim.3.bias'}


def test_warp_backward_workspace_layout():
    """vfd_warp_bwd_workspace_bytes covers what the training-mode launches put behind each other: the d_T partial sums
    of every 28 x 8 tile ([B*n_call][tiles][8 warps * 12] floats), one flag byte per tile (rounded up to 16 bytes) and one
    auto-mask decision byte per pixel (csrc/warp_loss_bwd.cuh); a short workspace is refused before any launch."""
    lib = _lib.load()
    wd = _lib.WarpDesc()
    wd.B, wd.num_cams, wd.H, wd.W, wd.n_src, wd.cam0, wd.n_call, wd.intensity_align = 8, 6, 384, 640, 2, 0, 6, 1
    dummy = ctypes.c_void_p(256)          # non-NULL, never dereferenced: the checks below fail first
    for f in range(3):
        wd.color[f] = 256
    wd.mask, wd.K, wd.invK = 256, 256, 256
    tiles = ((640 + 27) // 28) * (384 // 8)
    expect = 8 * 6 * tiles * 8 * 12 * 4 + (8 * 6 * tiles + 15) // 16 * 16 + 8 * 6 * 384 * 640
    assert lib.vfd_warp_bwd_workspace_bytes(ctypes.byref(wd)) == expect
    wd.H, wd.W = 75, 130                  # ragged: 5 x 10 tiles
    tiles = 5 * 10
    expect = 8 * 6 * tiles * 96 * 4 + (8 * 6 * tiles + 15) // 16 * 16 + 8 * 6 * 75 * 130
    assert lib.vfd_warp_bwd_workspace_bytes(ctypes.byref(wd)) == expect
    short = ctypes.c_size_t(expect - 1)
    args = [ctypes.byref(wd)] + [dummy] * 8
    assert lib.vfd_warp_loss_train_fwd(*args, short, None) == -4           # VFD_ERR_WORKSPACE
    assert lib.vfd_warp_loss_train_bwd(ctypes.byref(wd), dummy, dummy, dummy, short, dummy, dummy, None) == -4
    wd.n_src = 3
    assert lib.vfd_warp_bwd_workspace_bytes(ctypes.byref(wd)) == 0         # unsupported descriptor
