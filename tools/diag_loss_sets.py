"""GPU diagnostic: loss sums / g3 / d_T of the training-mode launch against the separate forward + backward launches
on a nuScenes-shaped rig (352 x 640), per camera and column.   python tools/diag_loss_sets.py [H W dataset]"""
import sys
import torch
sys.path.insert(0, '.')
from vfdepth_b200 import config as vcfg, view
from vfdepth_b200.rig import make_rig

H_, W_ = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (352, 640)
ds = sys.argv[3] if len(sys.argv) > 3 else 'nuscenes'
dev = torch.device('cuda:0')
cfg = vcfg.default_config(ds, batch_size=1, training__height=H_, training__width=W_, training__focal_length_scale=300 * W_ / 640)
inp, ex = make_rig(cfg, seed=8, with_feats=False)
to = lambda t: t.to(dev)
inp_d = {k: to(v) for k, v in inp.items()}
wi = view.WarpInputs(cfg, inp_d) if hasattr(view, 'WarpInputs') else None
res = {}
for fused in (True, False):
    view._WarpLoss.fuse_training = fused
    depths = [to(d).requires_grad_(True) for d in ex['depth']]
    Ts = {f: to(t).requires_grad_(True) for f, t in ex['cam_T_cam'].items()}
    total, det = view.warp_and_loss(cfg, inp_d, depths, [to(d) for d in ex['disp']], Ts, [to(n) for n in ex['noise']])
    fr = cfg['training']['frame_ids'][1:]
    grads = torch.autograd.grad(total, depths + [Ts[f] for f in fr])
    res[fused] = (total.detach().cpu(), det, [g.cpu() for g in grads])
view._WarpLoss.fuse_training = True
a, b = res[True], res[False]
print('total', float(a[0]), float(b[0]), 'rel', abs(float(a[0]) - float(b[0])) / abs(float(b[0])))
for k in a[1]:
    va, vb = a[1][k], b[1][k]
    if torch.is_tensor(va):
        va, vb = va.detach().cpu().double(), vb.detach().cpu().double()
        print(k, 'max abs diff', float((va - vb).abs().max()), '\n  fused', va.flatten()[:24].tolist(), '\n  sep  ', vb.flatten()[:24].tolist())
for i, (ga, gb) in enumerate(zip(a[2], b[2])):
    dlt = (ga - gb).abs()
    print('grad', i, tuple(ga.shape), 'max abs', float(dlt.max()), 'scale', float(gb.abs().max()), 'n>1e-6*scale', int((dlt > 1e-6 * gb.abs().max()).sum()))
    if ga.dim() == 4 and float(dlt.max()) > 1e-5 * float(gb.abs().max()):
        idx = (dlt > 1e-5 * gb.abs().max()).nonzero()
        print('   first bad pixels', idx[:12].tolist(), ' y range', int(idx[:, 2].min()), int(idx[:, 2].max()), 'x range', int(idx[:, 3].min()), int(idx[:, 3].max()))
