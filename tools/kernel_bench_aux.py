"""Per-kernel timing of the small full-resolution kernels either side of the warp path at the
bench batch: to_depth, edge-aware smoothness, depth synthesis.  CUDA events on the launch stream, inputs larger than
L2 in total (B=8: 47 MB per [B,6,H,W] map, 8 maps live).

    python tools/kernel_bench_aux.py [--batch 8] [--iters 20]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tools'))

from kernel_bench import peak_gbs  # noqa: E402
from vfdepth_b200 import config as vcfg  # noqa: E402
from vfdepth_b200 import view  # noqa: E402
from vfdepth_b200.rig import axis_angle_to_matrix, make_rig  # noqa: E402


def timed(fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def bench_loss_free(B, cfg, iters, report, dev):
    """stand-alone loss_util free functions (csrc/loss_free.cu, photometric.cu) on [B*6,.,H,W] maps."""
    from vfdepth_b200 import loss_util
    H, W = cfg['training']['height'], cfg['training']['width']
    n = B * 6
    gen = torch.Generator().manual_seed(2)
    rgb = torch.rand((n, 3, H, W), generator=gen).to(dev)
    pred = torch.rand((n, 3, H, W), generator=gen).to(dev).requires_grad_(True)
    disp = torch.rand((n, 1, H, W), generator=gen).to(dev).requires_grad_(True)
    rp = torch.rand((n, 1, H, W), generator=gen).to(dev)
    idl = torch.rand((n, 2, H, W), generator=gen).to(dev)
    px = n * H * W * 4
    one = torch.ones((), device=dev)
    with torch.no_grad():
        report('auto_masks', timed(lambda: loss_util.compute_auto_masks(rp, idl), iters), 4 * px)
        mask = loss_util.compute_auto_masks(rp, idl)
        report('masked_loss_fwd', timed(lambda: loss_util.compute_masked_loss(disp, mask), iters), 2 * px)
        report('edg_smooth_fwd', timed(lambda: loss_util.compute_edg_smooth_loss(rgb, disp), iters), 4 * px)
        report('photometric_fwd', timed(lambda: loss_util.compute_photometric_loss(pred, rgb), iters), 7 * px)
    ml = loss_util.compute_masked_loss(disp, mask)
    report('masked_loss_bwd', timed(lambda: torch.autograd.grad(ml, disp, one, retain_graph=True), iters), 2 * px)
    sm = loss_util.compute_edg_smooth_loss(rgb, disp)
    report('edg_smooth_bwd', timed(lambda: torch.autograd.grad(sm, disp, one, retain_graph=True), iters), 5 * px)
    ph = loss_util.compute_photometric_loss(pred, rgb)
    upp = torch.ones_like(ph)
    report('photometric_bwd', timed(lambda: torch.autograd.grad(ph, pred, upp, retain_graph=True), iters), 10 * px)


def bench_aggregation(B, iters, report, dev):
    """feature aggregation (fusion_depthnet.py:57-66) at the DDAD sizes: ResNet levels 128@48x80, 256@24x40, 512@12x20
    -> 256 channels.  Ours (per-level GEMMs + vfd_feat_agg) against the reference's formulation in torch on the same
    GPU (interpolate, cat, conv2d, LeakyReLU, then the channel-last copy the voxel kernels need)."""
    import torch.nn as nn
    import torch.nn.functional as F
    from vfdepth_b200 import _lib, aggregate
    n = B * 6
    gen = torch.Generator().manual_seed(1)
    lv = [torch.randn((n, c, h, w), generator=gen).to(dev).requires_grad_(True)
          for c, h, w in ((128, 48, 80), (256, 24, 40), (512, 12, 20))]
    blk = nn.Sequential(nn.Conv2d(896, 256, 1, bias=True), nn.Identity(), nn.LeakyReLU(0.1, inplace=True)).to(dev)
    params = [blk[0].weight, blk[0].bias]
    up = torch.randn((B, 6, 256, 48, 80), device=dev)
    out_bytes = n * 48 * 80 * 256 * 4
    in_bytes = sum(x.numel() for x in lv) * 4

    def ours():
        o = aggregate.aggregate_features(lv, 0, blk, B, 6)
        return torch.autograd.grad(o, lv + params, up)

    def ref():
        cat = torch.cat([lv[0]] + [F.interpolate(x, [48, 80], mode='bilinear', align_corners=True) for x in lv[1:]], 1)
        o = blk(cat).view(B, 6, 256, 48, 80)
        cl = o.permute(0, 1, 3, 4, 2).contiguous()
        return torch.autograd.grad(cl, lv + params, up.permute(0, 1, 3, 4, 2))

    with torch.no_grad():
        report('feat_agg ours fwd', timed(lambda: aggregate.aggregate_features(lv, 0, blk, B, 6), iters), in_bytes + out_bytes)
        report('feat_agg ref  fwd', timed(lambda: blk(torch.cat([lv[0]] + [F.interpolate(x, [48, 80], mode='bilinear', align_corners=True) for x in lv[1:]], 1)).permute(0, 2, 3, 1).contiguous(), iters), in_bytes + out_bytes)
    report('feat_agg ours f+b', timed(ours, iters), 3 * (in_bytes + out_bytes))
    report('feat_agg ref  f+b', timed(ref, iters), 3 * (in_bytes + out_bytes))
    # the kernels alone (projections resident): fwd reads y0 + coarse, writes out; bwd reads out, d_out, writes d_y*
    _lib.start_event_trace()
    for _ in range(iters):
        ours()
    tr = _lib.stop_event_trace()
    coarse = n * 256 * 4 * (24 * 40 + 12 * 20)
    for name, nbytes in (('vfd_feat_agg_fwd', 2 * out_bytes + coarse), ('vfd_feat_agg_bwd', 3 * out_bytes + coarse)):
        if name in tr:
            report(name, tr[name][1] / tr[name][0], nbytes)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=8)
    ap.add_argument('--iters', type=int, default=20)
    args = ap.parse_args()
    dev = torch.device('cuda:0')
    B = args.batch
    cfg = vcfg.default_config('ddad', batch_size=1)
    inputs, ex = make_rig(cfg, seed=3, with_feats=False)
    rep = lambda t: t.to(dev).repeat(B, *([1] * (t.dim() - 1))).contiguous()
    inp = {k: rep(v) for k, v in inputs.items() if k in ('mask', 'extrinsics', ('K', 0), ('inv_K', 0), ('color', 0, 0))}
    gen = torch.Generator().manual_seed(9)
    angle = (torch.rand((1, 6, 3), generator=gen) - 0.5) * torch.tensor([0.2, 0.2, 0.5])
    tform = torch.eye(4).repeat(1, 6, 1, 1)
    tform[:, :, :3, :3] = axis_angle_to_matrix(angle)
    inp['extrinsics_aug'] = rep(tform @ inputs['extrinsics'])
    disp = rep(torch.cat(ex['disp'], 1))
    H, W = disp.shape[-2:]
    px = B * 6 * H * W * 4
    peak, how = peak_gbs()

    def report(name, ms, nbytes):
        gbs = nbytes / ms / 1e6
        print('%-16s B=%d  %7.3f ms  %7.1f MB alg  %7.1f GB/s  %.3f of %s HBM peak (%.0f GB/s)' %
              (name, B, ms, nbytes / 1e6, gbs, gbs / peak, how, peak), flush=True)

    # to_depth
    d_leaf = disp.clone().requires_grad_(True)
    depth = view.to_depth(cfg, d_leaf, inp[('K', 0)])
    up = torch.randn_like(depth)
    report('to_depth_fwd', timed(lambda: view.to_depth(cfg, disp, inp[('K', 0)]), args.iters), 2 * px)
    report('to_depth_bwd', timed(lambda: torch.autograd.grad(depth, d_leaf, up, retain_graph=True), args.iters), 3 * px)
    # smoothness
    sm = view.smooth_loss(inp[('color', 0, 0)], d_leaf)
    ups = torch.ones_like(sm)
    report('smooth_fwd', timed(lambda: view.smooth_loss(inp[('color', 0, 0)], disp), args.iters), 4 * px)
    report('smooth_bwd', timed(lambda: torch.autograd.grad(sm, d_leaf, ups, retain_graph=True), args.iters), 5 * px)
    # depth synthesis: novel-view depth = the camera's own depth perturbed by 5 %
    dep = depth.detach().clone().requires_grad_(True)
    aug = (depth.detach() * (1.0 + 0.1 * (torch.rand_like(depth) - 0.5))).requires_grad_(True)
    con, sm2, _ = view.depth_syn_loss(cfg, inp, dep, aug, disp)
    upc = torch.ones_like(con)
    with torch.no_grad():
        report('depth_syn_fwd', timed(lambda: view.depth_syn_loss(cfg, inp, dep, aug, disp), args.iters), 4 * px)
    report('depth_syn_bwd', timed(lambda: torch.autograd.grad(con, [dep, aug], upc, retain_graph=True), args.iters), 5 * px)
    print('depth_con_loss per camera:', [round(float(c), 4) for c in con])
    bench_loss_free(B, cfg, args.iters, report, dev)
    bench_aggregation(B, args.iters, report, dev)
    print('(depth_syn_fwd includes the colour-less smoothness of disp_aug and two batched 4x4 inversions in torch)')


if __name__ == '__main__':
    main()
