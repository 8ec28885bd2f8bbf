"""Multi-level feature aggregation in front of the volumetric fusion -- the lines
/root/reference/network/fusion_depthnet.py:57-66 (identical in network/fusion_posenet.py:62-70):

    packed_feats_list = packed_feats[lev:lev+1] + [F.interpolate(feat, [up_h, up_w], 'bilinear', align_corners=True) ...]
    packed_feats_agg = self.conv1x1(torch.cat(packed_feats_list, dim=1))
    feats_agg = unpack_cam_feat(packed_feats_agg, self.batch_size, self.num_cams)

``conv1x1`` (network/blocks.py:41-61 with kernel_size=1: conv, identity, LeakyReLU(0.1)) and the bilinear resize are
both linear and act on different axes, so the 1x1 convolution is applied to every level at its own resolution (plain
cuBLAS GEMMs on channel-last rows with the matching column block of the conv's weight; 128 + 256/4 + 512/16 = 224
instead of 896 multiply-adds per output element at the shipped sizes) and ``vfd_feat_agg_fwd`` resizes the coarse
projections, adds them, applies bias and LeakyReLU and writes ``feats_agg`` channel-last -- the layout the voxel gather
reads, so ``VFNet._channel_last`` becomes a no-op.  Neither the resized maps (566 MB at B=8) nor their concatenation
(755 MB) exist.  The parameter keeps the reference's shape and state-dict key (``conv1x1.0.weight`` [Cout,sum C_l,1,1]).
"""
import ctypes

import torch
import torch.nn.functional as F

from . import _lib
from .ops import _chk, _stream


def _desc(y0, ycs, slope):
    n, h, w, C = y0.shape
    d = _lib.AggDesc()
    d.n_img, d.h, d.w, d.C, d.n_coarse, d.slope = n, h, w, C, len(ycs), slope
    for i, y in enumerate(ycs):
        d.ch[i], d.cw[i] = y.shape[1], y.shape[2]
    return d


def _ptr_array(ts):
    return (ctypes.c_void_p * _lib.VFD_AGG_LEVELS)(*[t.data_ptr() for t in ts])


class _FeatAgg(torch.autograd.Function):
    """vfd_feat_agg_fwd / _bwd: channel-last projections of the levels + bias -> channel-last feats_agg."""

    @staticmethod
    def forward(ctx, bias, slope, y0, *ycs):
        n, h, w, C = y0.shape
        if len(ycs) > _lib.VFD_AGG_LEVELS:
            raise ValueError('at most %d coarse levels' % _lib.VFD_AGG_LEVELS)
        _chk(y0, 'y0', (n, h, w, C))
        _chk(bias, 'bias', (C,))
        for i, y in enumerate(ycs):
            _chk(y, 'y%d' % (i + 1), (n, y.shape[1], y.shape[2], C))
        out = torch.empty_like(y0)
        d = _desc(y0, ycs, slope)
        _lib.call('vfd_feat_agg_fwd', ctypes.byref(d), _lib.ptr(y0), _ptr_array(ycs), _lib.ptr(bias), _lib.ptr(out),
                  _stream())
        ctx.desc = d
        ctx.shapes = [tuple(y.shape) for y in ycs]
        ctx.save_for_backward(out)
        return out

    @staticmethod
    def backward(ctx, d_out):
        (out,) = ctx.saved_tensors
        d_out = d_out.contiguous()
        d_y0 = torch.empty_like(out)
        d_ycs = [torch.empty(s, device=out.device) for s in ctx.shapes]
        _lib.call('vfd_feat_agg_bwd', ctypes.byref(ctx.desc), _lib.ptr(out), _lib.ptr(d_out), _lib.ptr(d_y0),
                  _ptr_array(d_ycs), _stream())
        d_bias = d_y0.sum((0, 1, 2)) if ctx.needs_input_grad[0] else None
        return (d_bias, None, d_y0) + tuple(d_ycs)


def aggregate_rows(levels, weight, bias, slope=0.1):
    """levels: list of [n_img,C_l,h_l,w_l] (finest first); weight [Cout,sum C_l] or [Cout,sum C_l,1,1]; bias [Cout]
    -> feats_agg channel-last [n_img,h,w,Cout]."""
    wgt = weight.reshape(weight.shape[0], -1)
    if wgt.shape[1] != sum(int(f.shape[1]) for f in levels):
        raise ValueError('conv1x1 expects %d input channels, the levels have %d'
                         % (wgt.shape[1], sum(int(f.shape[1]) for f in levels)))
    ys, off = [], 0
    for f in levels:
        c = f.shape[1]
        # channel-last rows of the level (a free view when the encoder runs channels_last) x its column block
        ys.append(F.linear(f.permute(0, 2, 3, 1), wgt[:, off:off + c]).contiguous())
        off += c
    return _FeatAgg.apply(bias.contiguous(), float(slope), ys[0], *ys[1:])


def aggregate_features(packed_feats, lev, conv1x1, batch_size, num_cams):
    """Drop-in for fusion_depthnet.py:57-66: ``packed_feats`` the encoder's list of [B*cams,C_l,h_l,w_l], ``conv1x1`` the
    reference's block (conv, identity, LeakyReLU) -> feats_agg [B,cams,Cout,h,w], a permuted view of channel-last
    storage."""
    conv = conv1x1[0]
    act = conv1x1[-1]
    slope = getattr(act, 'negative_slope', None)
    if slope is None:
        raise ValueError('conv1x1 must end in LeakyReLU (network/blocks.py:46-47)')
    if conv.kernel_size != (1, 1) or conv.bias is None:
        raise ValueError('conv1x1 must be a biased 1x1 convolution')
    rows = aggregate_rows(list(packed_feats[lev:]), conv.weight, conv.bias, slope)
    n, h, w, C = rows.shape
    if n != batch_size * num_cams:
        raise ValueError('packed batch %d != batch_size * num_cams = %d' % (n, batch_size * num_cams))
    return rows.view(batch_size, num_cams, h, w, C).permute(0, 1, 4, 2, 3)
