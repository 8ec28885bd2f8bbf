"""Stand-alone forms of the reference's loss free functions (/root/reference/models/losses/loss_util.py), same names and
arguments: ``compute_ssim_loss(pred, target)`` -> [B,3,H,W] and ``compute_photometric_loss(pred, target)`` -> [B,1,H,W]
(:43-78, csrc/photometric.cu, differentiable w.r.t. ``pred``); ``compute_auto_masks`` (:6-18, no gradient),
``compute_masked_loss`` (:21-25, gradient to ``loss``) and ``compute_edg_smooth_loss`` (:28-40, gradient to ``disp_map``)
on csrc/loss_free.cu.  The training path does not call these -- ``view.warp_and_loss`` / ``MultiCamLoss`` evaluate the
same arithmetic for all warps of a rig inside the fused kernels (``vfd_warp_loss_*``, ``vfd_smooth_loss_*``); they exist
so that code calling the reference's free functions keeps working.  CUDA tensors only; there is no CPU path."""
import ctypes

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


def compute_auto_masks(reprojection_loss, identity_reprojection_loss):
    """loss_util.py:6-18 -> float mask [B,1,H,W]: 1 where the arg-min over cat[reprojection, identity] is channel 0
    (all ones, shaped like ``reprojection_loss``, when ``identity_reprojection_loss`` is None)."""
    rp = reprojection_loss.detach().float().contiguous()
    _chk(rp, 'reprojection_loss')
    if rp.dim() != 4:
        raise ValueError('reprojection_loss must be [B,F,H,W]')
    if identity_reprojection_loss is None:
        return torch.ones_like(rp)
    B, F1, H, W = rp.shape
    idl = identity_reprojection_loss.detach().float().contiguous()
    if idl.dim() != 4:
        raise ValueError('identity_reprojection_loss must be [B,F,H,W]')
    F2 = idl.shape[1]
    _chk(idl, 'identity_reprojection_loss', (B, F2, H, W))
    mask = torch.empty((B, 1, H, W), device=rp.device)
    _lib.call('vfd_auto_masks', B, F1, F2, H * W, _lib.ptr(rp), _lib.ptr(idl), _lib.ptr(mask), _stream())
    return mask


class _MaskedLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, loss, mask):
        _chk(loss, 'loss')
        _chk(mask, 'mask', loss.shape)
        sums = torch.empty(2, device=loss.device, dtype=torch.float64)
        _lib.call('vfd_masked_loss_fwd', ctypes.c_size_t(loss.numel()), _lib.ptr(loss), _lib.ptr(mask), _lib.ptr(sums),
                  _stream())
        ctx.save_for_backward(mask, sums)
        return sums[0].float() / (sums[1].float() + 1e-8)

    @staticmethod
    def backward(ctx, d_out):
        mask, sums = ctx.saved_tensors
        d_out = d_out.float().contiguous()
        d_loss = torch.empty_like(mask)
        _lib.call('vfd_masked_loss_bwd', ctypes.c_size_t(mask.numel()), _lib.ptr(mask), _lib.ptr(sums), _lib.ptr(d_out),
                  _lib.ptr(d_loss), _stream())
        return d_loss, None


def compute_masked_loss(loss, mask):
    """loss_util.py:21-25 -> scalar (loss * mask).sum() / (mask.sum() + 1e-8); the mask is a constant for autograd."""
    shape = torch.broadcast_shapes(loss.shape, mask.shape)
    return _MaskedLoss.apply(loss.float().expand(shape).contiguous(), mask.detach().float().expand(shape).contiguous())


class _EdgeSmooth(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rgb, disp):
        B, C, H, W = rgb.shape
        _chk(rgb, 'rgb')
        _chk(disp, 'disp_map', (B, 1, H, W))
        sums = torch.empty(2, device=rgb.device, dtype=torch.float64)
        _lib.call('vfd_edg_smooth_loss_fwd', B, C, H, W, _lib.ptr(rgb), _lib.ptr(disp), _lib.ptr(sums), _stream())
        ctx.save_for_backward(rgb, disp)
        return (sums[0] / (B * H * (W - 1)) + sums[1] / (B * (H - 1) * W)).float()

    @staticmethod
    def backward(ctx, d_out):
        rgb, disp = ctx.saved_tensors
        B, C, H, W = rgb.shape
        d_out = d_out.float().contiguous()
        d_disp = torch.empty_like(disp)
        _lib.call('vfd_edg_smooth_loss_bwd', B, C, H, W, _lib.ptr(rgb), _lib.ptr(disp), _lib.ptr(d_out), _lib.ptr(d_disp),
                  _stream())
        return None, d_disp


def compute_edg_smooth_loss(rgb, disp_map):
    """loss_util.py:28-40 -> scalar edge-aware smoothness of ``disp_map`` [B,1,H,W] under ``rgb`` [B,C,H,W] (no mean
    normalisation: that is single_cam_loss.py:57-65, fused in ``view.smooth_loss``); gradient to ``disp_map``."""
    return _EdgeSmooth.apply(rgb.detach().float().contiguous(), disp_map.float().contiguous())
