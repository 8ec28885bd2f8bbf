"""diagnostic: fused vs cuBLAS-compact vs dense path gradients at several channel counts (run on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from vfdepth_b200 import config as vcfg, ops
from vfdepth_b200.rig import make_rig
from vfdepth_b200.vfnet import VFNet

dev = torch.device('cuda:0')


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


for C, Cv, B in [(64, 64, 1), (128, 64, 1), (132, 64, 1), (256, 64, 1), (256, 8, 1), (256, 64, 2)]:
    cfg = vcfg.default_config('ddad', batch_size=B, model__fusion_feat_in_dim=C, model__voxel_pre_dim=[Cv])
    inputs, _ = make_rig(cfg, seed=41, with_images=False, with_feats=False)
    lvl = cfg['model']['fusion_level'] + 1
    torch.manual_seed(3)
    net = VFNet(cfg, C, 16, 'depth').to(dev)
    g = net.geom
    feats0 = torch.randn(B, 6, C, g.h, g.w, device=dev)
    up = torch.randn(B, Cv, g.N, device=dev)
    to = lambda t: t.to(dev)
    args = (to(inputs['mask']), to(inputs[('K', lvl)]), to(inputs['extrinsics_inv']))
    params = [net.conv_non_overlap[0].weight, net.conv_non_overlap[0].bias, net.conv_overlap[0].weight, net.conv_overlap[0].bias]
    out = {}
    for name in ('fused', 'cublas', 'dense'):
        f = feats0.clone().requires_grad_(True)
        if name == 'dense':
            x_no, x_ov, count = net.gather_rows(f, *args)[:3]
            c1 = C + 1
            w_no = F.pad(params[0].squeeze(-1), (0, g.ld_no - c1))
            w_ov = params[2].squeeze(-1)
            w_ov = torch.cat([F.pad(w_ov[:, :c1], (0, g.half_ov - c1)), F.pad(w_ov[:, c1:], (0, g.ld_ov - g.half_ov - c1))], 1)
            cnt = count.unsqueeze(-1)
            y = F.leaky_relu(F.linear(x_no, w_no, params[1]), 0.1) * (cnt == 1).float() + \
                F.leaky_relu(F.linear(x_ov, w_ov, params[3]), 0.1) * (cnt == 2).float()
            vox = y.permute(0, 2, 1)
        else:
            net.fused_mlp = name == 'fused'
            vox = net.backproject_into_voxel(f, *args)
        grads = torch.autograd.grad((vox * up).sum(), [f] + params)
        out[name] = [vox.detach()] + [x.detach() for x in grads]
    names = ['vox', 'd_feats', 'd_w_no', 'd_b_no', 'd_w_ov', 'd_b_ov']
    print('C=%d Cv=%d B=%d' % (C, Cv, B))
    for i, n in enumerate(names):
        print('   %-8s fused-vs-dense %.2e  cublas-vs-dense %.2e  fused-vs-cublas %.2e' % (
            n, rel(out['fused'][i], out['dense'][i]), rel(out['cublas'][i], out['dense'][i]), rel(out['fused'][i], out['cublas'][i])))
    df = (out['fused'][1] - out['dense'][1]).abs()       # [B,6,C,h,w]
    perch = df.amax(dim=(0, 1, 3, 4))
    bad = (perch > 1e-4 * out['dense'][1].abs().max()).nonzero().flatten().tolist()
    print('   fused bad channels:', bad[:20], '... n=%d' % len(bad))
    dc = (out['cublas'][1] - out['dense'][1]).abs().amax(dim=(0, 1, 3, 4))
    badc = (dc > 1e-4 * out['dense'][1].abs().max()).nonzero().flatten().tolist()
    print('   cublas bad channels:', badc[:20], '... n=%d' % len(badc))
