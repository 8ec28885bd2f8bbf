"""Stand-alone forms of the reference's photometric free functions (/root/reference/models/losses/loss_util.py:43-78)
on csrc/photometric.cu: ``compute_ssim_loss(pred, target)`` -> [B,3,H,W] and ``compute_photometric_loss(pred, target)``
-> [B,1,H,W], differentiable w.r.t. ``pred``.  The training path does not call these -- ``view.warp_and_loss`` /
``MultiCamLoss`` evaluate the same arithmetic for all warps of a rig inside the fused loss kernels, together with
``compute_auto_masks`` / ``compute_masked_loss`` (loss_util.py:6-25) and the smoothness term (:28-40), which therefore
exist only in fused form (``vfd_warp_loss_*``, ``vfd_smooth_loss_*``)."""
import torch

from . import _lib
from .ops import _chk, _stream


class _Photometric(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target, want_ssim):
        B, _, H, W = pred.shape
        _chk(pred, 'pred', (B, 3, H, W))
        _chk(target, 'target', (B, 3, H, W))
        out = torch.empty((B, 3 if want_ssim else 1, H, W), device=pred.device)
        _lib.call('vfd_photometric_loss_fwd', B, H, W, _lib.ptr(pred), _lib.ptr(target),
                  _lib.ptr(out if want_ssim else None), _lib.ptr(None if want_ssim else out), _stream())
        ctx.want_ssim = want_ssim
        ctx.save_for_backward(pred, target)
        return out

    @staticmethod
    def backward(ctx, d_out):
        pred, target = ctx.saved_tensors
        B, _, H, W = pred.shape
        d_out = d_out.contiguous()
        d_pred = torch.empty_like(pred)
        _lib.call('vfd_photometric_loss_bwd', B, H, W, _lib.ptr(pred), _lib.ptr(target),
                  _lib.ptr(d_out if ctx.want_ssim else None), _lib.ptr(None if ctx.want_ssim else d_out),
                  _lib.ptr(d_pred), _stream())
        return d_pred, None, None


def compute_ssim_loss(pred, target):
    """loss_util.py:43-67 -> clamp((1 - SSIM) / 2, 0, 1), [B,3,H,W]."""
    return _Photometric.apply(pred.float().contiguous(), target.float().contiguous(), True)


def compute_photometric_loss(pred=None, target=None):
    """loss_util.py:70-78 -> 0.85 * mean_c SSIM loss + 0.15 * mean_c |target - pred|, [B,1,H,W]."""
    return _Photometric.apply(pred.float().contiguous(), target.float().contiguous(), False)
