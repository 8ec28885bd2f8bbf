"""GPU diagnostic: per-pixel comparison of the training-mode launch (sums, auto-mask bytes, g3) with the forward kernel's
views (reproj_loss, reproj_mask).   python tools/diag_loss_dump.py H W dataset [out.pt]"""
import ctypes
import sys
import torch
sys.path.insert(0, '.')
from vfdepth_b200 import config as vcfg, view, _lib
from vfdepth_b200.rig import make_rig

H_, W_, ds = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
dev = torch.device('cuda:0')
cfg = vcfg.default_config(ds, batch_size=1, training__height=H_, training__width=W_, training__focal_length_scale=300 * W_ / 640)
inp, ex = make_rig(cfg, seed=8, with_feats=False)
to = lambda t: t.to(dev)
inp_d = {k: to(v) for k, v in inp.items()}
wi = view.WarpInputs(cfg, inp_d)
depth = torch.cat([to(d) for d in ex['depth']], 1).contiguous()
noise = torch.stack([to(n) for n in ex['noise']], 1).contiguous()
T = view.rig_transforms(wi, inp_d, {f: to(t) for f, t in ex['cam_T_cam'].items()}).contiguous()
params, raw = view.warp_stats(wi, depth, T, 0)
n_call = depth.shape[1]
d = wi.desc(0, n_call)
B, Hh, Ww = wi.B, wi.H, wi.W
# training-mode launch
sums = torch.empty((n_call, 8), device=dev, dtype=torch.float64)
nbytes = _lib.load().vfd_warp_bwd_workspace_bytes(ctypes.byref(d))
ws = torch.zeros((nbytes,), device=dev, dtype=torch.uint8)
g3 = torch.full((B, n_call, 3, Hh, Ww), float('nan'), device=dev)
_lib.call('vfd_warp_loss_train_fwd', ctypes.byref(d), _lib.ptr(depth), _lib.ptr(T), _lib.ptr(params), _lib.ptr(raw), _lib.ptr(noise),
          _lib.ptr(sums), _lib.ptr(g3), _lib.ptr(ws), ctypes.c_size_t(nbytes), view._stream())
# forward kernel with views
vstruct, views = view._alloc_views(wi, n_call, ('reproj_loss', 'reproj_mask', 'selection'))
sums2 = torch.empty_like(sums)
_lib.call('vfd_warp_loss_fwd', ctypes.byref(d), _lib.ptr(depth), _lib.ptr(T), _lib.ptr(params), _lib.ptr(raw), _lib.ptr(noise),
          _lib.ptr(sums2), ctypes.byref(vstruct), view._stream())
torch.cuda.synchronize()
print('sums train', sums[:, :6].cpu().tolist()[0])
print('sums fwd  ', sums2[:, :6].cpu().tolist()[0])
print('sum diffs ', (sums - sums2)[:, :6].abs().max(0).values.cpu().tolist())
tiles = ((Ww + 27) // 28) * ((Hh + 7) // 8)      # kSetW = 28 (csrc/warp_loss_bwd.cuh)
pbytes = B * n_call * tiles * 96 * 4
fbytes = (B * n_call * tiles + 15) // 16 * 16
abit = ws[pbytes + fbytes:pbytes + fbytes + B * n_call * Hh * Ww].view(B, n_call, Hh, Ww)
flags = ws[pbytes:pbytes + B * n_call * tiles].view(B * n_call, tiles)
print('tile flags set', int(flags.sum()), 'of', flags.numel(), ' g3 NaN left', int(torch.isnan(g3).sum()))
mask = inp_d['mask'].float().view(B, n_call, Hh, Ww)
auto_fwd = views['reproj_mask'].view(B, n_call, Hh, Ww)
auto_train = abit.float() * mask
bad = (auto_fwd != auto_train).nonzero()
print('auto-mask mismatches', bad.shape[0], bad[:20].tolist())
# what the training sums should be from the forward kernel's per-pixel maps
rl = views['reproj_loss'].view(B, n_call, Hh, Ww).double()
print('fwd maps: sum reproj*auto', rl.sum((0, 2, 3)).cpu().tolist()[:2], ' sum auto', auto_fwd.double().sum((0, 2, 3)).cpu().tolist()[:2])
if len(sys.argv) > 4:
    torch.save({'sums': sums.cpu(), 'g3': g3.cpu(), 'abit': abit.cpu(), 'flags': flags.cpu(), 'partial': ws[:pbytes].cpu()}, sys.argv[4])
