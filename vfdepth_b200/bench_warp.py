"""Workload of bench.py's view-synthesis half: decoder disparity -> depth (to_depth), 6 cameras x 3 frames warp +
SSIM/L1 photometric loss + edge-aware smoothness, forward and backward to the 6 disparity maps (through the depth maps)
and the 12 cam_T_cam (BASELINE.json configs[2]), on synthetic DDAD-shaped rigs."""
import torch

from . import config as vcfg
from . import view
from .rig import make_rig


class WarpWorkload:
    def __init__(self, batch, dev, seed, dataset='ddad'):
        self.cfg = vcfg.default_config(dataset, batch_size=batch)
        self.dev, self.B = dev, batch
        inputs, ex = make_rig(self.cfg, seed=seed, with_feats=False)
        t = self.cfg['training']
        self.H, self.W, self.frames = t['height'], t['width'], t['frame_ids']
        pin = lambda x: x.contiguous().pin_memory()
        keys = [('color', f, 0) for f in self.frames] + ['mask', ('K', 0), ('inv_K', 0), 'extrinsics', 'extrinsics_inv']
        self.h_inputs = {k: pin(inputs[k]) for k in keys}
        # end to end the images travel as uint8 -- what every dataset stores and a loader hands over -- and become the
        # float tensors the reference consumes on the device (the reference's loaders do ToTensor on the host)
        self.color_keys = [('color', f, 0) for f in self.frames]
        # (the self-occlusion masks too: the datasets keep them as 0 / 255 PNGs)
        self.h_u8 = {k: pin((inputs[k] * 255.0).round().clamp(0, 255).to(torch.uint8)) for k in self.color_keys + ['mask']}
        self.h_disp = pin(torch.cat(ex['disp'], 1))                   # [B,6,H,W] decoder output of scale 0
        self.h_noise = pin(torch.stack(ex['noise'], 1))               # [B,6,2,H,W]
        self.h_T = {f: pin(ex['cam_T_cam'][f]) for f in self.frames[1:]}
        self.h_out = torch.empty((batch, 6, self.H, self.W)).pin_memory()
        self.h_loss = torch.zeros(1).pin_memory()
        self.inputs = {k: v.to(dev) for k, v in self.h_inputs.items()}
        self.disp, self.noise = self.h_disp.to(dev), self.h_noise.to(dev)
        self.T = {f: v.to(dev) for f, v in self.h_T.items()}
        self._stage = {}

    def step(self, inputs, disp, noise, T):
        n = self.cfg['data']['num_cams']
        if noise is None:
            # tie-break noise of the identity-reprojection loss: drawn on the device, inside the step (the reference
            # draws it on the CPU and ships it, single_cam_loss.py:45-46; same distribution, no PCIe traffic)
            noise = torch.randn((self.B, n, len(self.frames) - 1, self.H, self.W), device=self.dev)
        disp = disp.detach().requires_grad_(True)            # all cameras stacked: [B,6,H,W]
        Ts = {f: t.detach().requires_grad_(True) for f, t in T.items()}
        depth = view.to_depth(self.cfg, disp, inputs[('K', 0)])          # models/vfdepth.py:266-275
        total, _ = view.warp_and_loss(self.cfg, inputs, depth, disp, Ts, noise)
        grads = torch.autograd.grad(total, [disp] + list(Ts.values()))
        return total, grads

    def mask_density(self):
        """fractions of the pixels whose loss masks are non-zero on this workload's rigs (outside the timed region): the
        set kernels skip work where a mask is zero, so a bench line says how sparse its synthetic data is"""
        wi = view.WarpInputs(self.cfg, self.inputs)
        depth = view.to_depth(self.cfg, self.disp, self.inputs[('K', 0)])
        T = view.rig_transforms(wi, self.inputs, self.T)
        _, den, _ = view.warp_loss_sums(wi, depth, T, self.noise)
        px = float(self.B * self.H * self.W)
        f = (den.double().sum(0) / (px * wi.num_cams)).tolist()
        return {'auto_mask_kept': f[0], 'spatial_overlap': f[1], 'spatio_temporal': f[2],
                'self_occlusion_mask_zero': float((self.inputs['mask'] == 0).double().mean())}

    def step_resident(self):
        return self.step(self.inputs, self.disp, self.noise, self.T)

    def upload(self, slot=0):
        """host (pinned) -> device copies of one step's inputs into the staging buffers of `slot`, on the current
        stream (preallocated: no allocator traffic inside the timed region)."""
        if slot not in self._stage:
            mk = lambda t: torch.empty(t.shape, dtype=t.dtype, device=self.dev)
            self._stage[slot] = ({k: mk(v) for k, v in self.h_inputs.items()}, mk(self.h_disp),
                                 {f: mk(v) for f, v in self.h_T.items()}, {k: mk(v) for k, v in self.h_u8.items()})
        inputs, disp, T, u8 = self._stage[slot]
        for k, v in self.h_inputs.items():
            if k in self.h_u8:
                u8[k].copy_(self.h_u8[k], non_blocking=True)
            else:
                inputs[k].copy_(v, non_blocking=True)
        disp.copy_(self.h_disp, non_blocking=True)
        for f, v in self.h_T.items():
            T[f].copy_(v, non_blocking=True)
        return inputs, disp, None, T, u8

    def compute(self, staged):
        inputs, disp, noise, T, u8 = staged
        for k, v in u8.items():          # uint8 -> float in [0, 1] on the compute stream (ToTensor's x / 255), preallocated
            torch.div(v, 255.0, out=inputs[k])
        return self.step(inputs, disp, noise, T)

    def download(self, result, grads_too=False):
        """the step's result: the loss.  The gradients stay on the device, where the decoder's backward consumes them in
        training (``grads_too``: also d loss / d disparity, 47 MB per 8 rigs)"""
        total, grads = result
        if grads_too:
            self.h_out.copy_(grads[0], non_blocking=True)
        self.h_loss.copy_(total.detach().reshape(1), non_blocking=True)

    def h2d_bytes(self):
        ts = [self.h_u8.get(k, v) for k, v in self.h_inputs.items()] + [self.h_disp] + list(self.h_T.values())
        return sum(t.numel() * t.element_size() for t in ts)

    def d2h_bytes(self, grads_too=False):
        return (self.h_out.numel() * 4 if grads_too else 0) + 4

    def algorithmic_bytes(self):
        """per launch: unique inputs read once + required outputs written once."""
        px = self.B * 6 * self.H * self.W * 4
        imgs, depth, mask, noise = 3 * 3 * px, px, px, 2 * px
        return {'vfd_warp_stats': imgs + depth + mask, 'vfd_warp_loss_fwd': imgs + depth + mask + noise,
                'vfd_warp_loss_bwd': imgs + depth + mask + noise + px,
                # training mode: sums and per-set gradients in one launch (3 gradient maps out), then the combine
                'vfd_warp_loss_train_fwd': imgs + depth + mask + noise + 3 * px, 'vfd_warp_loss_train_bwd': 4 * px,
                'vfd_smooth_loss_fwd': 3 * px + px, 'vfd_smooth_loss_bwd': 3 * px + 2 * px,
                'vfd_smooth_loss_train_fwd': 3 * px + 2 * px, 'vfd_smooth_loss_train_bwd': 2 * px,
                'vfd_to_depth_fwd': 2 * px, 'vfd_to_depth_bwd': 3 * px}
