"""diagnostic: d total / d disp and d total / d cam_T_cam of the rig-level warp_and_loss against the fp64 oracle, term by
term (loss coefficients switched on one at a time), on the chained test's geometry (random decoder disparity, small motion)"""
import copy
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import vfdepth_oracle as orc
from tests.test_train_step_gpu import _cfg
from vfdepth_b200 import view, train_step as ts
from vfdepth_b200.rig import make_rig

dev = torch.device('cuda:0')
cfg0 = _cfg(2)
inputs, ex = make_rig(cfg0, seed=4242, with_feats=False)
noise = torch.stack(ex['noise'], 1)
step = ts.TrainStep(cfg0, dev, seed=3)
with torch.no_grad():
    step.depth_net.decoder.dispconv[0].weight.mul_(20.0)
    step.pose_net.pose_decoder.out.bias.copy_(torch.tensor([0.5, -0.3, 0.2, 10.0, -5.0, 60.0]))
inp_d = {k: v.to(dev) for k, v in inputs.items()}
with torch.no_grad():
    _, det = ts.forward_loss(cfg0, step.depth_net, step.pose_net, inp_d, noise.to(dev))
disp0 = det['disp'].detach().cpu()
T0 = {f: t.detach().cpu() for f, t in det['cam_T_cam'].items()}
inputs['extrinsics_inv'] = torch.inverse(inputs['extrinsics'])
inp_d['extrinsics_inv'] = inputs['extrinsics_inv'].to(dev)
print('disp range', float(disp0.min()), float(disp0.max()))

def rel(a, b):
    return float((a.double().cpu() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))

for name, coeffs in (('all', None), ('reproj', (0, 0, 0)), ('smooth', (1, 0, 0)), ('spatial', (0, 1, 0)), ('st', (0, 0, 1))):
    cfg = copy.deepcopy(cfg0)
    if coeffs is not None:
        cfg['loss']['disparity_smoothness'] = 0.001 * coeffs[0]
        cfg['loss']['spatio_coeff'] = 0.03 * coeffs[1]
        cfg['loss']['spatio_tempo_coeff'] = 0.1 * coeffs[2]
    # GPU
    d = disp0.to(dev).requires_grad_(True)
    Ts = {f: t.to(dev).requires_grad_(True) for f, t in T0.items()}
    depth = view.to_depth(cfg, d, inp_d[('K', 0)])
    total, dd = view.warp_and_loss(cfg, inp_d, depth, d, Ts, noise.to(dev), want_views=('reproj_mask', 'selection'))
    g = torch.autograd.grad(total, [d] + list(Ts.values()))
    v = dd['views']
    dec = []
    for cam in range(6):
        sel = v['selection'][:, cam:cam + 1].cpu().long()
        dec.append(dict(reproj_mask=v['reproj_mask'][:, cam:cam + 1].cpu().float(), sel_reproj=sel % 2, sel_st=sel // 2))
    # fused path too
    d2 = disp0.to(dev).requires_grad_(True)
    Ts2 = {f: t.to(dev).requires_grad_(True) for f, t in T0.items()}
    total2, _ = view.warp_and_loss(cfg, inp_d, view.to_depth(cfg, d2, inp_d[('K', 0)]), d2, Ts2, noise.to(dev))
    g2 = torch.autograd.grad(total2, [d2] + list(Ts2.values()))
    res = {}
    for dt in (torch.float32, torch.float64):
        cv = lambda t: t.to(dt) if t.is_floating_point() else t
        inp = {k: cv(v_) for k, v_ in inputs.items()}
        dr = cv(disp0).clone().requires_grad_(True)
        Tr = {f: cv(t).clone().requires_grad_(True) for f, t in T0.items()}
        disps = [dr[:, c:c + 1] for c in range(6)]
        depths = [orc.to_depth(cfg, x, inp[('K', 0)][:, c]) for c, x in enumerate(disps)]
        t_ref, _ = orc.warp_and_loss(cfg, inp, depths, disps, Tr, [cv(noise)[:, c] for c in range(6)], dec)
        res[dt] = (float(t_ref), torch.autograd.grad(t_ref, [dr] + list(Tr.values())))
    t64, g64 = res[torch.float64]
    t32, g32 = res[torch.float32]
    print('%-8s total gpu %.8f fused %.8f fp64 %.8f | d_disp err: gpu %.2e fused %.2e oracle32 %.2e | d_T err: gpu %.2e %.2e oracle32 %.2e %.2e | sum d_disp gpu %.6e fp64 %.6e'
          % (name, float(total), float(total2), t64, rel(g[0], g64[0]), rel(g2[0], g64[0]), rel(g32[0], g64[0]),
             rel(g[1], g64[1]), rel(g[2], g64[2]), rel(g32[1], g64[1]), rel(g32[2], g64[2]),
             float(g[0].double().sum()), float(g64[0].sum())))
