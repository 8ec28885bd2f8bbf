"""The sub-method surface of the reference's view synthesis on the CUDA kernels of csrc/warp_image.cu:

* ``Projection(batch_size, height, width, device)`` with ``backproject(invK, depth)``, ``reproject(K, points3D, T)``
  and ``forward(depth, T, bp_invK, rp_K)`` -- /root/reference/models/geometry/geometry_util.py:33-82;
* ``get_virtual_image(src_img, src_mask, tar_depth, tar_invK, src_K, T)`` -- models/geometry/view_rendering.py:61-82.

``Projection.forward`` and ``get_virtual_image`` are differentiable w.r.t. the target depth, ``T`` and ``K`` (what the
reference's training step differentiates; the source image is an input).  ``backproject`` / ``reproject`` on their own
are forward-only.  The training path does not go through these: ``view.warp_and_loss`` / ``ViewRendering.forward`` +
``MultiCamLoss.forward`` fuse all warps of a rig with their loss (csrc/warp_loss_*.cu); the coordinates and samples
here come from the same device helpers, so both paths agree bit for bit.
"""
import torch

from . import _lib
from .ops import _chk, _stream


def _mats(B, *ms):
    out = []
    for m, name in ms:
        _chk(m, name, (B, 4, 4))
        out.append(m)
    return out


def _grads_from_dP(dP_part, K, T):
    """dP_part [B,blocks,12] -> (d_K, d_T) for P = (K @ T)[:, :3, :]."""
    B = K.shape[0]
    dM = torch.zeros((B, 4, 4), device=K.device, dtype=torch.float32)
    dM[:, :3, :] = dP_part.sum(1).view(B, 3, 4)
    return dM @ T.transpose(1, 2), K.transpose(1, 2) @ dM


class _WarpImage(torch.autograd.Function):
    """vfd_warp_image_fwd / _bwd with a source image: (img, valid * mask)."""

    @staticmethod
    def forward(ctx, src_img, src_mask, depth, invK, K, T):
        B, _, H, W = depth.shape
        _chk(src_img, 'src_img', (B, 3, H, W))
        _chk(src_mask, 'src_mask', (B, 1, H, W))
        _chk(depth, 'tar_depth', (B, 1, H, W))
        _mats(B, (invK, 'tar_invK'), (K, 'src_K'), (T, 'T'))
        img = torch.empty_like(src_img)
        mask = torch.empty_like(src_mask)
        _lib.call('vfd_warp_image_fwd', B, H, W, _lib.ptr(src_img), _lib.ptr(src_mask), _lib.ptr(depth), _lib.ptr(invK),
                  _lib.ptr(K), _lib.ptr(T), _lib.ptr(img), _lib.ptr(mask), _lib.ptr(None), _stream())
        ctx.save_for_backward(src_img, src_mask, depth, invK, K, T)
        ctx.mark_non_differentiable(mask)
        return img, mask

    @staticmethod
    def backward(ctx, d_img, _):
        src_img, src_mask, depth, invK, K, T = ctx.saved_tensors
        B, _, H, W = depth.shape
        d_depth = torch.empty_like(depth)
        part = torch.empty((B, _lib.load().vfd_warp_image_blocks(), 12), device=depth.device)
        _lib.call('vfd_warp_image_bwd', B, H, W, _lib.ptr(src_img), _lib.ptr(src_mask), _lib.ptr(depth), _lib.ptr(invK),
                  _lib.ptr(K), _lib.ptr(T), _lib.ptr(d_img.contiguous()), _lib.ptr(None), _lib.ptr(d_depth),
                  _lib.ptr(part), _stream())
        d_K, d_T = _grads_from_dP(part, K, T)
        return None, None, d_depth, None, d_K, d_T


class _ProjectionGrid(torch.autograd.Function):
    """vfd_warp_image_fwd / _bwd without an image: the normalised sampling grid [B,H,W,2]."""

    @staticmethod
    def forward(ctx, depth, invK, K, T):
        B, _, H, W = depth.shape
        _chk(depth, 'depth', (B, 1, H, W))
        _mats(B, (invK, 'bp_invK'), (K, 'rp_K'), (T, 'T'))
        grid = torch.empty((B, H, W, 2), device=depth.device)
        _lib.call('vfd_warp_image_fwd', B, H, W, _lib.ptr(None), _lib.ptr(None), _lib.ptr(depth), _lib.ptr(invK),
                  _lib.ptr(K), _lib.ptr(T), _lib.ptr(None), _lib.ptr(None), _lib.ptr(grid), _stream())
        ctx.save_for_backward(depth, invK, K, T)
        return grid

    @staticmethod
    def backward(ctx, d_grid):
        depth, invK, K, T = ctx.saved_tensors
        B, _, H, W = depth.shape
        d_depth = torch.empty_like(depth)
        part = torch.empty((B, _lib.load().vfd_warp_image_blocks(), 12), device=depth.device)
        _lib.call('vfd_warp_image_bwd', B, H, W, _lib.ptr(None), _lib.ptr(None), _lib.ptr(depth), _lib.ptr(invK),
                  _lib.ptr(K), _lib.ptr(T), _lib.ptr(None), _lib.ptr(d_grid.contiguous()), _lib.ptr(d_depth),
                  _lib.ptr(part), _stream())
        d_K, d_T = _grads_from_dP(part, K, T)
        return d_depth, None, d_K, d_T


def get_virtual_image(src_img, src_mask, tar_depth, tar_invK, src_K, T, scale=0):
    """view_rendering.py:61-82: the source image seen from the target view and its validity mask
    (in-bounds * nearest(source self-occlusion mask); NaN samples -> 2.0 / 0).  ``scale`` is accepted for signature
    parity (the reference indexes its per-scale Projection with it; sizes are taken from the tensors here)."""
    c = lambda t: t.float().contiguous()
    return _WarpImage.apply(c(src_img), c(src_mask), c(tar_depth), c(tar_invK), c(src_K), c(T))


class Projection(torch.nn.Module):
    """Drop-in for geometry_util.py:33 ``Projection(batch_size, height, width, device)``."""

    def __init__(self, batch_size, height, width, device):
        super().__init__()
        self.batch_size, self.width, self.height = batch_size, width, height

    def _check(self, B, H, W):
        if (H, W) != (self.height, self.width):
            raise ValueError('Projection built for %dx%d, got %dx%d' % (self.height, self.width, H, W))

    @torch.no_grad()
    def backproject(self, invK, depth):
        """-> points3D [B,4,H*W] (:52-60); forward-only."""
        B, _, H, W = depth.shape
        self._check(B, H, W)
        depth, invK = depth.float().contiguous(), invK.float().contiguous()
        _chk(depth, 'depth', (B, 1, H, W))
        _chk(invK, 'invK', (B, 4, 4))
        pts = torch.empty((B, 4, H * W), device=depth.device)
        _lib.call('vfd_projection_backproject', B, H, W, _lib.ptr(depth), _lib.ptr(invK), _lib.ptr(pts), _stream())
        return pts

    @torch.no_grad()
    def reproject(self, K, points3D, T):
        """-> pix_coords [B,H,W,2] in [-1,1] (:62-77); forward-only, row 3 of points3D is taken as 1."""
        B = points3D.shape[0]
        H, W = self.height, self.width
        pts, K, T = points3D.float().contiguous(), K.float().contiguous(), T.float().contiguous()
        _chk(pts, 'points3D', (B, 4, H * W))
        _chk(K, 'K', (B, 4, 4))
        _chk(T, 'T', (B, 4, 4))
        grid = torch.empty((B, H, W, 2), device=pts.device)
        _lib.call('vfd_projection_reproject', B, H, W, _lib.ptr(pts), _lib.ptr(K), _lib.ptr(T), _lib.ptr(grid), _stream())
        return grid

    def forward(self, depth, T, bp_invK, rp_K):
        """:79-82, differentiable w.r.t. depth, T and rp_K."""
        B, _, H, W = depth.shape
        self._check(B, H, W)
        c = lambda t: t.float().contiguous()
        return _ProjectionGrid.apply(c(depth), c(bp_invK), c(rp_K), c(T))


class _NormImage(torch.autograd.Function):
    """vfd_norm_image_fwd / _bwd: intensity alignment of one warped image per batch element."""

    @staticmethod
    def forward(ctx, ref_img, ref_mask, warp_img, warp_mask):
        B, _, H, W = warp_img.shape
        _chk(ref_img, 'src_img', (B, 3, H, W))
        _chk(ref_mask, 'src_mask', (B, 1, H, W))
        _chk(warp_img, 'warp_img', (B, 3, H, W))
        _chk(warp_mask, 'warp_mask', (B, 1, H, W))
        stats = torch.empty((B, 8), device=warp_img.device, dtype=torch.float64)
        out = torch.empty_like(warp_img)
        _lib.call('vfd_norm_image_fwd', B, H, W, _lib.ptr(ref_img), _lib.ptr(ref_mask), _lib.ptr(warp_img),
                  _lib.ptr(warp_mask), _lib.ptr(stats), _lib.ptr(out), _stream())
        ctx.save_for_backward(warp_mask, stats)
        return out

    @staticmethod
    def backward(ctx, d_out):
        warp_mask, stats = ctx.saved_tensors
        B, _, H, W = warp_mask.shape
        d_warp = torch.empty((B, 3, H, W), device=d_out.device)
        _lib.call('vfd_norm_image_bwd', B, H, W, _lib.ptr(warp_mask), _lib.ptr(stats), _lib.ptr(d_out.contiguous()),
                  _lib.ptr(d_warp), _stream())
        return None, None, d_warp, None


def get_norm_image_single(src_img, src_mask, warp_img, warp_mask):
    """view_rendering.py:39-59 (``src_img`` / ``src_mask`` are the TARGET frame's image and mask there): the warped
    image with its masked mean / standard deviation aligned to the target's over their overlap, times ``warp_mask``;
    returned unchanged when any batch element has no overlap.  Gradient flows to ``warp_img`` only."""
    c = lambda t: t.float().contiguous()
    return _NormImage.apply(c(src_img), c(src_mask), c(warp_img), c(warp_mask.detach()))
