"""Per-kernel timing of the C-ABI entry points on synthetic DDAD-shaped rigs (CUDA events, inputs resident).

    python tools/kernel_bench.py [--batch 8] [--iters 10] [--only gather_fwd,ray_bwd,...] [--dataset ddad]

Prints one line per kernel: ms per launch, algorithmic GB per launch and GB/s against the measured
HBM peak.  Also the command ncu is pointed at (one kernel via --only, --iters 1).
"""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from vfdepth_b200 import _lib, config as vcfg, ops  # noqa: E402
from vfdepth_b200.rig import make_rig  # noqa: E402


def peak_gbs():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.isfile(p):
        return json.load(open(p))['hbm_gbs'], 'measured'
    return 6650.0, 'fallback'


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=8)
    ap.add_argument('--iters', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--dataset', default='ddad')
    ap.add_argument('--only', default='')
    args = ap.parse_args()
    dev = torch.device('cuda:0')
    B = args.batch
    cfg = vcfg.default_config(args.dataset, batch_size=1)
    inputs, extras = make_rig(cfg, seed=1, with_images=False)
    lvl = cfg['model']['fusion_level'] + 1
    g = ops.FusionGeometry.from_cfg(cfg)
    g.B = B
    rep = lambda t: t.to(dev).repeat(B, *([1] * (t.dim() - 1))).contiguous()
    feats = torch.randn(B, 6, g.h, g.w, g.C, device=dev)
    mask = rep(inputs['mask'])
    mask_small = torch.nn.functional.interpolate(mask.view(B * 6, 1, *mask.shape[-2:]), [g.h, g.w], mode='bilinear',
                                                 align_corners=True).view(B, 6, g.h, g.w).contiguous()
    K, iK = rep(inputs[('K', lvl)]), rep(inputs[('inv_K', lvl)])
    E, Einv = rep(inputs['extrinsics']), rep(inputs['extrinsics_inv'])
    d = g.desc(dev, B)
    N = g.N
    x_no = torch.empty(B, N, g.ld_no, device=dev)
    x_ov = torch.empty(B, N, g.ld_ov, device=dev)
    count = torch.empty(B, N, dtype=torch.uint8, device=dev)
    d_feats = torch.empty_like(feats)
    vol = torch.randn(B, g.Z, g.Y, g.X, g.Cv, device=dev)
    out = torch.empty(B, 6, g.h, g.w, g.D, g.Cv, device=dev)
    d_vol = torch.empty_like(vol)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = _lib.ptr
    MB = 1e6
    feat_b = B * 6 * g.h * g.w * g.C * 4
    c1 = g.C + 1
    kernels = {
        'gather_fwd': (lambda: _lib.call('vfd_voxel_gather_depth_fwd', ctypes.byref(d), p(feats), p(mask_small), p(K),
                                         p(Einv), p(x_no), p(x_ov), p(count), p(None), p(None), st),
                       feat_b + B * N * 3 * c1 * 4),
        'gather_bwd': (lambda: _lib.call('vfd_voxel_gather_depth_bwd', ctypes.byref(d), p(mask_small), p(K), p(Einv),
                                         p(x_no), p(x_ov), p(d_feats), st),
                       feat_b + B * N * 3 * c1 * 4),
        'pose_fwd': (lambda: _lib.call('vfd_voxel_gather_pose_fwd', ctypes.byref(d), p(feats), p(mask_small), p(K),
                                       p(Einv), p(x_no), p(count), st),
                     feat_b + B * N * c1 * 4),
        'pose_bwd': (lambda: _lib.call('vfd_voxel_gather_pose_bwd', ctypes.byref(d), p(mask_small), p(K), p(Einv),
                                       p(x_no), p(d_feats), st),
                     feat_b + B * N * c1 * 4),
        'ray_fwd': (lambda: _lib.call('vfd_ray_sample_fwd', ctypes.byref(d), p(vol), p(iK), p(E), p(out), p(None), st),
                    vol.numel() * 4 + out.numel() * 4),
        'ray_bwd': (lambda: _lib.call('vfd_ray_sample_bwd', ctypes.byref(d), p(iK), p(E), p(out), p(d_vol), st),
                    vol.numel() * 4 + out.numel() * 4),
    }
    # compact lists + fused gather/MLP kernels
    cnt8, slot, idx_no, idx_ov, totals = ops.voxel_classify(mask_small, K, Einv, g, sizes=False)
    m1, m2 = totals.tolist()
    cnt8, table = ops.voxel_classify(mask_small, K, Einv, g, table=True)
    w_no, b_no = torch.randn(g.Cv, c1, device=dev) * 0.05, torch.randn(g.Cv, device=dev) * 0.1
    w_ov, b_ov = torch.randn(g.Cv, 2 * c1, device=dev) * 0.05, torch.randn(g.Cv, device=dev) * 0.1
    fvol = torch.empty(B, N, g.Cv, device=dev)
    gvol = torch.randn(B, N, g.Cv, device=dev)
    dw_no, db_no, dw_ov, db_ov = (torch.empty_like(t) for t in (w_no, b_no, w_ov, b_ov))
    nws = _lib.load().vfd_voxel_fuse_bwd_workspace_bytes(ctypes.byref(d))
    ws = torch.empty(nws, device=dev, dtype=torch.uint8)
    nws_f = _lib.load().vfd_voxel_fuse_fwd_workspace_bytes(ctypes.byref(d))
    ws_f = torch.empty(nws_f, device=dev, dtype=torch.uint8)
    nws_c = _lib.load().vfd_voxel_classify_workspace_bytes(ctypes.byref(d))
    ws_c = torch.empty(max(nws_c, 8), device=dev, dtype=torch.uint8)
    cx_no = torch.empty(max(m1, 1), g.ld_no, device=dev)
    cx_ov = torch.empty(max(m2, 1), g.ld_ov, device=dev)
    rows_b = (m1 + 2 * m2) * c1 * 4
    vol_b = B * N * g.Cv * 4
    kernels.update({
        'compact_fwd': (lambda: _lib.call('vfd_voxel_gather_depth_compact_fwd', ctypes.byref(d), p(feats), p(mask_small),
                                          p(K), p(Einv), p(slot), p(cx_no), p(cx_ov), st), feat_b + rows_b),
        'compact_bwd': (lambda: _lib.call('vfd_voxel_gather_depth_compact_bwd', ctypes.byref(d), p(mask_small), p(K),
                                          p(Einv), p(slot), p(cx_no), p(cx_ov), p(d_feats), st), feat_b + rows_b),
        'classify': (lambda: _lib.call('vfd_voxel_classify', ctypes.byref(d), p(mask_small), p(K), p(Einv), p(cnt8), None,
                                       None, None, None, p(table), p(ws_c), ctypes.c_size_t(ws_c.numel()), st),
                     B * 6 * g.h * g.w * 4 + B * N + (m1 + 2 * m2) * 40),
        'fuse_fwd': (lambda: _lib.call('vfd_voxel_fuse_fwd', ctypes.byref(d), p(feats), p(cnt8), p(table), p(w_no), p(b_no),
                                       p(w_ov), p(b_ov), p(fvol), p(ws_f), ctypes.c_size_t(nws_f), st),
                     feat_b + vol_b),
        'fuse_bwd': (lambda: _lib.call('vfd_voxel_fuse_bwd', ctypes.byref(d), p(feats), p(cnt8), p(table), p(w_no), p(w_ov),
                                       p(fvol), p(gvol), p(d_feats), p(dw_no), p(db_no), p(dw_ov), p(db_ov), p(ws),
                                       ctypes.c_size_t(nws), st),
                     2 * feat_b + 2 * vol_b),
    })
    print('count-1 voxels %d, count-2 voxels %d of %d' % (m1, m2, B * N), flush=True)
    only = [k for k in args.only.split(',') if k]
    peak, how = peak_gbs()
    # make x_no / x_ov / out hold realistic upstream gradients for the backward kernels
    kernels['gather_fwd'][0]()
    kernels['ray_fwd'][0]()
    kernels['compact_fwd'][0]()
    kernels['fuse_fwd'][0]()
    torch.cuda.synchronize()
    for name, (fn, nbytes) in kernels.items():
        if only and name not in only:
            continue
        for _ in range(args.warmup):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.iters
        gbs = nbytes / ms / 1e6
        print('%-11s B=%d  %8.3f ms  %8.1f MB alg  %7.1f GB/s  %.3f of %s HBM peak (%.0f GB/s)' %
              (name, B, ms, nbytes / MB, gbs, gbs / peak, how, peak), flush=True)


if __name__ == '__main__':
    main()
