"""Diagnostic: warp + photometric loss kernels against the golden fixture (reference run) and the oracle.

    python tools/warp_check.py [--size H W] [--batch B]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import vfdepth_oracle as orc  # noqa: E402
from tests import helpers as H  # noqa: E402
from vfdepth_b200 import config as vcfg, view  # noqa: E402
from vfdepth_b200.rig import make_rig  # noqa: E402


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def frac_bad(a, b, tol=1e-4):
    a, b = a.double().cpu(), b.double().cpu()
    e = (a - b).abs() / b.abs().max().clamp_min(1e-30)
    return float((e > tol).double().mean()), int((e > tol).sum())


def run(cfg, inp, depths, disps, Ts, noises, dev):
    to = lambda t: t.to(dev)
    inp_d = {k: to(v) for k, v in inp.items()}
    depths_d = [to(d).requires_grad_(True) for d in depths]
    Ts_d = {f: to(t).requires_grad_(True) for f, t in Ts.items()}
    total, det = view.warp_and_loss(cfg, inp_d, depths_d, [to(d) for d in disps], Ts_d, [to(n) for n in noises],
                                    want_views=view.VIEW_KEYS)
    grads = torch.autograd.grad(total, depths_d + [Ts_d[f] for f in sorted(Ts_d)])
    return total, det, grads


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--size', type=int, nargs=2, default=None)
    ap.add_argument('--batch', type=int, default=2)
    args = ap.parse_args()
    dev = torch.device('cuda:0')
    if args.size is None:
        g = H.load_npz('golden_warp.npz')
        cfg = H.warp_cfg()
        inp = H.warp_inputs(g)
        depths = [g['depth'][:, c] for c in range(6)]
        disps = [g['disp'][:, c] for c in range(6)]
        noises = [g['noise'][:, c] for c in range(6)]
        Ts = {-1: g['T_m1'], 1: g['T_p1']}
    else:
        cfg = vcfg.default_config('ddad', batch_size=args.batch, training__height=args.size[0],
                                  training__width=args.size[1],
                                  training__focal_length_scale=300 * args.size[1] / 640)
        inp, ex = make_rig(cfg, seed=99, with_feats=False)
        depths, disps, noises, Ts = ex['depth'], ex['disp'], ex['noise'], ex['cam_T_cam']
    # oracle
    d_ref = [d.clone().requires_grad_(True) for d in depths]
    T_ref = {f: t.clone().requires_grad_(True) for f, t in Ts.items()}
    tot_ref, det_ref = orc.warp_and_loss(cfg, inp, d_ref, disps, T_ref, noises)
    g_ref = torch.autograd.grad(tot_ref, d_ref + [T_ref[f] for f in sorted(T_ref)])
    total, det, grads = run(cfg, inp, depths, disps, Ts, noises, dev)
    from vfdepth_b200 import _lib
    _lib.start_event_trace()
    for _ in range(3):
        run(cfg, inp, depths, disps, Ts, noises, dev)
    print('kernel ms (3 runs):', _lib.stop_event_trace())
    print('total', float(total), float(tot_ref), 'rel', abs(float(total) - float(tot_ref)) / abs(float(tot_ref)))
    for cam in range(6):
        r = det_ref[cam]
        ref_terms = [float(r['reproj_loss']), float(r['spatio_loss']), float(r['spatio_tempo_loss'])]
        got = det['terms'][cam].tolist()
        print('cam', cam, ['%.3e' % (abs(a - b) / max(abs(b), 1e-30)) for a, b in zip(got, ref_terms)], got)
    v = det['views']
    fr = cfg['training']['frame_ids']
    for cam in range(6):
        rv = det_ref[cam]['view']
        msg = []
        for k, f in enumerate(fr[1:]):
            msg.append('color%+d %.1e' % (f, rel(v['color'][:, cam, k], rv[('color', f)])))
            msg.append('cmask%+d %d' % (f, int((v['color_mask'][:, cam, k:k + 1].cpu() != rv[('color_mask', f)]).sum())))
        for k, f in enumerate(fr):
            msg.append('ov%+d %.1e' % (f, rel(v['overlap'][:, cam, k], rv[('overlap', f)])))
            msg.append('om%+d %d' % (f, int((v['overlap_mask'][:, cam, k:k + 1].cpu() != rv[('overlap_mask', f)]).sum())))
        msg.append('automask_mismatch %d' % int((v['reproj_mask'][:, cam:cam + 1].cpu() != det_ref[cam]['reproj_mask']).sum()))
        msg.append('reproj_map %.1e' % rel(v['reproj_loss'][:, cam:cam + 1], det_ref[cam]['reproj_map']))
        print('cam', cam, ' '.join(msg))
    for cam in range(6):
        print('d_depth cam', cam, 'rel %.2e' % rel(grads[cam], g_ref[cam]), 'bad', frac_bad(grads[cam], g_ref[cam]))
    for i, f in enumerate(sorted(Ts)):
        print('d_T', f, 'rel %.2e' % rel(grads[6 + i], g_ref[6 + i]))
        if args.size is None:
            print(grads[6 + i][0, 0].cpu(), g_ref[6 + i][0, 0])


if __name__ == '__main__':
    main()
