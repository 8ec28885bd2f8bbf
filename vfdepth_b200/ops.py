"""torch.autograd bindings of the fusion kernels: tensors in, raw pointers + the current stream out.

PyTorch is plumbing here (device memory, streams, autograd graph); all arithmetic of the path happens in
libvfdepth_b200.so.  Layouts are channel-last (include/vfdepth_b200.h).
"""
import ctypes

import torch

from . import _lib


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _chk(t, name, shape=None):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise ValueError('%s must be a CUDA tensor (vfdepth_b200 has no CPU path)' % name)
    if t.dtype != torch.float32:
        raise ValueError('%s must be float32, got %s' % (name, t.dtype))
    if not t.is_contiguous():
        raise ValueError('%s must be contiguous' % name)
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise ValueError('%s has shape %s, expected %s' % (name, tuple(t.shape), tuple(shape)))
    return t


def _round4(n):
    return (n + 3) // 4 * 4


class FusionGeometry:
    """Sizes + constant axes of VFNet (network/volumetric_fusionnet.py:22-40) as a vfd_fusion_desc.

    The voxel axes and depth bins are taken from ``torch.linspace`` on the CPU exactly as the reference builds
    them (:72, :39) and uploaded once per device: a device-side ``start + i*step`` does not reproduce
    ``linspace`` bit for bit.
    """

    def __init__(self, batch, num_cams, C, h, w, voxel_size, voxel_str_p, voxel_unit_size, D, d_str, d_end, Cv):
        if num_cams == 6:
            group = [0, 1, 1, 0, 0, 1]          # :222-223
        elif num_cams == 3:
            group = [0, 1, 1]                   # :219-220
        else:
            raise NotImplementedError('overlap grouping is defined for 3 or 6 cameras only')
        self.B, self.num_cams, self.C, self.h, self.w = batch, num_cams, C, h, w
        self.X, self.Y, self.Z = voxel_size
        self.N = self.X * self.Y * self.Z
        self.D, self.Cv = D, Cv
        self.ld_no = _round4(C + 1)
        self.half_ov = _round4(C + 1)
        self.ld_ov = 2 * self.half_ov
        end = [voxel_str_p[i] + voxel_unit_size[i] * (voxel_size[i] - 1) for i in range(3)]
        self._axes_cpu = [torch.linspace(voxel_str_p[i], end[i], voxel_size[i]) for i in range(3)]
        self._bins_cpu = torch.linspace(d_str, d_end, D)
        self._start = [float(voxel_str_p[i]) for i in range(3)]
        self._length = [float(end[i] - voxel_str_p[i]) for i in range(3)]
        self._group = group
        self._depth_div = float(voxel_size[0])
        self._dev = {}

    @classmethod
    def from_cfg(cls, cfg, C=None, Cv=None):
        m, t = cfg['model'], cfg['training']
        lvl = m['fusion_level'] + 1
        return cls(t['batch_size'], cfg['data']['num_cams'], C if C is not None else m['fusion_feat_in_dim'],
                   t['height'] // 2 ** lvl, t['width'] // 2 ** lvl, m['voxel_size'], m['voxel_str_p'],
                   m['voxel_unit_size'], m['proj_d_bins'], m['proj_d_str'], m['proj_d_end'],
                   Cv if Cv is not None else m['voxel_pre_dim'][-1])

    def desc(self, device, batch=None):
        """-> (FusionDesc, keep-alive tensors) for ``device``."""
        key = (str(device), batch or self.B)
        if key not in self._dev:
            axes = [a.to(device) for a in self._axes_cpu]
            bins = self._bins_cpu.to(device)
            d = _lib.FusionDesc()
            d.B, d.num_cams, d.C, d.h, d.w = batch or self.B, self.num_cams, self.C, self.h, self.w
            d.X, d.Y, d.Z, d.D, d.Cv = self.X, self.Y, self.Z, self.D, self.Cv
            d.ld_no, d.ld_ov, d.half_ov = self.ld_no, self.ld_ov, self.half_ov
            for i, g in enumerate(self._group):
                d.group[i] = g
            d.depth_div = self._depth_div
            for i in range(3):
                d.vox_start[i] = self._start[i]
                d.vox_length[i] = self._length[i]
            d.axis_x, d.axis_y, d.axis_z = axes[0].data_ptr(), axes[1].data_ptr(), axes[2].data_ptr()
            d.bins = bins.data_ptr()
            self._dev[key] = (d, axes + [bins])
        return self._dev[key][0]


class _VoxelGatherDepth(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, mask_small, K, Einv, geom, want_detail):
        B = feats.shape[0]
        g = geom
        _chk(feats, 'feats', (B, g.num_cams, g.h, g.w, g.C))
        _chk(mask_small, 'mask_small', (B, g.num_cams, g.h, g.w))
        _chk(K, 'K', (B, g.num_cams, 4, 4))
        _chk(Einv, 'Einv', (B, g.num_cams, 4, 4))
        dev = feats.device
        d = g.desc(dev, B)
        x_no = torch.empty((B, g.N, g.ld_no), device=dev, dtype=torch.float32)
        x_ov = torch.empty((B, g.N, g.ld_ov), device=dev, dtype=torch.float32)
        count = torch.empty((B, g.N), device=dev, dtype=torch.uint8)
        grid = valid = None
        if want_detail:
            grid = torch.empty((B, g.num_cams, g.N, 2), device=dev, dtype=torch.float32)
            valid = torch.empty((B, g.num_cams, g.N), device=dev, dtype=torch.uint8)
        _lib.call('vfd_voxel_gather_depth_fwd', ctypes.byref(d), _lib.ptr(feats), _lib.ptr(mask_small), _lib.ptr(K),
                  _lib.ptr(Einv), _lib.ptr(x_no), _lib.ptr(x_ov), _lib.ptr(count), _lib.ptr(grid), _lib.ptr(valid),
                  _stream())
        ctx.geom = g
        ctx.save_for_backward(mask_small, K, Einv)
        ctx.mark_non_differentiable(count)
        if want_detail:
            ctx.mark_non_differentiable(grid, valid)
            return x_no, x_ov, count, grid, valid
        return x_no, x_ov, count

    @staticmethod
    def backward(ctx, d_no, d_ov, *unused):
        mask_small, K, Einv = ctx.saved_tensors
        g = ctx.geom
        B = mask_small.shape[0]
        dev = mask_small.device
        d = g.desc(dev, B)
        d_no = torch.zeros((B, g.N, g.ld_no), device=dev) if d_no is None else d_no.contiguous()
        d_ov = torch.zeros((B, g.N, g.ld_ov), device=dev) if d_ov is None else d_ov.contiguous()
        d_feats = torch.empty((B, g.num_cams, g.h, g.w, g.C), device=dev, dtype=torch.float32)
        _lib.call('vfd_voxel_gather_depth_bwd', ctypes.byref(d), _lib.ptr(mask_small), _lib.ptr(K), _lib.ptr(Einv),
                  _lib.ptr(d_no), _lib.ptr(d_ov), _lib.ptr(d_feats), _stream())
        return d_feats, None, None, None, None, None


class _VoxelGatherPose(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, mask_small, K, Einv, geom, bev=False):
        B = feats.shape[0]
        g = geom
        _chk(feats, 'feats', (B, g.num_cams, g.h, g.w, g.C))
        _chk(mask_small, 'mask_small', (B, g.num_cams, g.h, g.w))
        _chk(K, 'K', (B, g.num_cams, 4, 4))
        _chk(Einv, 'Einv', (B, g.num_cams, 4, 4))
        dev = feats.device
        d = g.desc(dev, B)
        avg = torch.empty((B, g.N, g.ld_no), device=dev, dtype=torch.float32)
        count = torch.empty((B, g.N), device=dev, dtype=torch.uint8)
        _lib.call('vfd_voxel_gather_pose_bev_fwd' if bev else 'vfd_voxel_gather_pose_fwd', ctypes.byref(d),
                  _lib.ptr(feats), _lib.ptr(mask_small), _lib.ptr(K), _lib.ptr(Einv), _lib.ptr(avg), _lib.ptr(count),
                  _stream())
        ctx.geom = g
        ctx.bev = bev
        ctx.save_for_backward(mask_small, K, Einv)
        ctx.mark_non_differentiable(count)
        return avg, count

    @staticmethod
    def backward(ctx, d_avg, _):
        mask_small, K, Einv = ctx.saved_tensors
        g = ctx.geom
        B = mask_small.shape[0]
        dev = mask_small.device
        d = g.desc(dev, B)
        d_avg = d_avg.contiguous()
        d_feats = torch.empty((B, g.num_cams, g.h, g.w, g.C), device=dev, dtype=torch.float32)
        _lib.call('vfd_voxel_gather_pose_bev_bwd' if ctx.bev else 'vfd_voxel_gather_pose_bwd', ctypes.byref(d),
                  _lib.ptr(mask_small), _lib.ptr(K), _lib.ptr(Einv), _lib.ptr(d_avg), _lib.ptr(d_feats), _stream())
        return d_feats, None, None, None, None, None


class _RaySample(torch.autograd.Function):
    @staticmethod
    def forward(ctx, vol, invK, E, geom, want_grid):
        g = geom
        B = vol.shape[0]
        _chk(vol, 'vol', (B, g.Z, g.Y, g.X, g.Cv))
        _chk(invK, 'inv_K', (B, g.num_cams, 4, 4))
        _chk(E, 'extrinsics', (B, g.num_cams, 4, 4))
        dev = vol.device
        d = g.desc(dev, B)
        out = torch.empty((B, g.num_cams, g.h, g.w, g.D, g.Cv), device=dev, dtype=torch.float32)
        grid = torch.empty((B, g.num_cams, g.D, g.h, g.w, 3), device=dev) if want_grid else None
        _lib.call('vfd_ray_sample_fwd', ctypes.byref(d), _lib.ptr(vol), _lib.ptr(invK), _lib.ptr(E), _lib.ptr(out),
                  _lib.ptr(grid), _stream())
        ctx.geom = g
        ctx.save_for_backward(invK, E)
        if want_grid:
            ctx.mark_non_differentiable(grid)
            return out, grid
        return out

    @staticmethod
    def backward(ctx, d_out, *unused):
        invK, E = ctx.saved_tensors
        g = ctx.geom
        B = invK.shape[0]
        dev = invK.device
        d = g.desc(dev, B)
        d_out = d_out.contiguous()
        d_vol = torch.empty((B, g.Z, g.Y, g.X, g.Cv), device=dev, dtype=torch.float32)
        _lib.call('vfd_ray_sample_bwd', ctypes.byref(d), _lib.ptr(invK), _lib.ptr(E), _lib.ptr(d_out),
                  _lib.ptr(d_vol), _stream())
        return d_vol, None, None, None, None


def voxel_classify(mask_small, K, Einv, geom, sizes=True, table=False):
    """-> count [B,N] u8, slot [B,N] i32, idx_no [M1] i64, idx_ov [M2] i64 (only count-1 / count-2
    voxels reach an MLP).  With ``sizes`` the two totals are read back to the host (one 8-byte copy) to slice the
    lists; without, the full-capacity lists and the device-side totals [2] are returned (no synchronisation).
    ``table=True`` (the fused path): -> (count, table) only -- one launch, no lists."""
    g = geom
    B = mask_small.shape[0]
    _chk(mask_small, 'mask_small', (B, g.num_cams, g.h, g.w))
    _chk(K, 'K', (B, g.num_cams, 4, 4))
    _chk(Einv, 'Einv', (B, g.num_cams, 4, 4))
    dev = mask_small.device
    d = g.desc(dev, B)
    count = torch.empty((B, g.N), device=dev, dtype=torch.uint8)
    nbytes = _lib.load().vfd_voxel_classify_workspace_bytes(ctypes.byref(d))
    ws = torch.empty((max(nbytes, 8) // 8,), device=dev, dtype=torch.int64)
    if table:   # per-voxel corner table for the fused kernels (80 bytes per voxel of the grid)
        tab = torch.empty((_lib.load().vfd_voxel_table_bytes(ctypes.byref(d)) // 16,), device=dev,
                          dtype=torch.complex128)
        _lib.call('vfd_voxel_classify', ctypes.byref(d), _lib.ptr(mask_small), _lib.ptr(K), _lib.ptr(Einv),
                  _lib.ptr(count), None, None, None, None, _lib.ptr(tab), _lib.ptr(ws),
                  ctypes.c_size_t(ws.numel() * 8), _stream(), kernels=1)
        return count, tab
    slot = torch.empty((B, g.N), device=dev, dtype=torch.int32)
    idx_no = torch.empty((B * g.N,), device=dev, dtype=torch.int64)
    idx_ov = torch.empty((B * g.N,), device=dev, dtype=torch.int64)
    totals = torch.empty((2,), device=dev, dtype=torch.int32)
    _lib.call('vfd_voxel_classify', ctypes.byref(d), _lib.ptr(mask_small), _lib.ptr(K), _lib.ptr(Einv),
              _lib.ptr(count), _lib.ptr(slot), _lib.ptr(idx_no), _lib.ptr(idx_ov), _lib.ptr(totals), None,
              _lib.ptr(ws), ctypes.c_size_t(ws.numel() * 8), _stream())
    if not sizes:
        return count, slot, idx_no, idx_ov, totals
    m1, m2 = totals.tolist()
    return count, slot, idx_no[:m1], idx_ov[:m2]


class _VoxelGatherDepthCompact(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, mask_small, K, Einv, slot, m1, m2, geom):
        B = feats.shape[0]
        g = geom
        _chk(feats, 'feats', (B, g.num_cams, g.h, g.w, g.C))
        _chk(mask_small, 'mask_small', (B, g.num_cams, g.h, g.w))
        _chk(K, 'K', (B, g.num_cams, 4, 4))
        _chk(Einv, 'Einv', (B, g.num_cams, 4, 4))
        if slot.dtype != torch.int32 or tuple(slot.shape) != (B, g.N) or not slot.is_cuda or not slot.is_contiguous():
            raise ValueError('slot must be a contiguous CUDA int32 tensor [B,N]')
        dev = feats.device
        d = g.desc(dev, B)
        # at least one (never read) row so the pointers stay valid when a class is empty
        x_no = torch.empty((max(m1, 1), g.ld_no), device=dev, dtype=torch.float32)
        x_ov = torch.empty((max(m2, 1), g.ld_ov), device=dev, dtype=torch.float32)
        _lib.call('vfd_voxel_gather_depth_compact_fwd', ctypes.byref(d), _lib.ptr(feats), _lib.ptr(mask_small),
                  _lib.ptr(K), _lib.ptr(Einv), _lib.ptr(slot), _lib.ptr(x_no), _lib.ptr(x_ov), _stream())
        ctx.geom, ctx.m = g, (m1, m2)
        ctx.save_for_backward(mask_small, K, Einv, slot)
        return x_no[:m1], x_ov[:m2]

    @staticmethod
    def backward(ctx, d_no, d_ov):
        mask_small, K, Einv, slot = ctx.saved_tensors
        g = ctx.geom
        m1, m2 = ctx.m
        B = mask_small.shape[0]
        dev = mask_small.device
        d = g.desc(dev, B)
        d_no = torch.zeros((max(m1, 1), g.ld_no), device=dev) if d_no is None or m1 == 0 else d_no.contiguous()
        d_ov = torch.zeros((max(m2, 1), g.ld_ov), device=dev) if d_ov is None or m2 == 0 else d_ov.contiguous()
        d_feats = torch.empty((B, g.num_cams, g.h, g.w, g.C), device=dev, dtype=torch.float32)
        _lib.call('vfd_voxel_gather_depth_compact_bwd', ctypes.byref(d), _lib.ptr(mask_small), _lib.ptr(K),
                  _lib.ptr(Einv), _lib.ptr(slot), _lib.ptr(d_no), _lib.ptr(d_ov), _lib.ptr(d_feats), _stream())
        return d_feats, None, None, None, None, None, None, None


def voxel_gather_depth_compact(feats_cl, mask_small, K, Einv, slot, m1, m2, geom):
    """feats_cl [B,cams,h,w,C] -> x_no [m1,ld_no] (count-1 voxels), x_ov [m2,ld_ov] (count-2 voxels)."""
    return _VoxelGatherDepthCompact.apply(feats_cl, mask_small, K, Einv, slot, m1, m2, geom)


def _workspace(nbytes, dev):
    """256-byte aligned scratch (the caching allocator hands out 512-byte aligned blocks)"""
    return torch.empty((max(nbytes, 256),), device=dev, dtype=torch.uint8)


class _VoxelFuse(torch.autograd.Function):
    """gather + both 1x1 MLPs (vfd_voxel_fuse_fwd: dense projection of the feature pixels, then the voxel gather);
    backward in vfd_voxel_fuse_bwd."""

    @staticmethod
    def forward(ctx, feats, w_no, b_no, w_ov, b_ov, count, table, geom):
        B = feats.shape[0]
        g = geom
        c1 = g.C + 1
        _chk(feats, 'feats', (B, g.num_cams, g.h, g.w, g.C))
        w_no, b_no, w_ov, b_ov = (t.detach().contiguous() for t in (w_no, b_no, w_ov, b_ov))
        _chk(w_no, 'w_no', (g.Cv, c1))
        _chk(b_no, 'b_no', (g.Cv,))
        _chk(w_ov, 'w_ov', (g.Cv, 2 * c1))
        _chk(b_ov, 'b_ov', (g.Cv,))
        if count.dtype != torch.uint8 or tuple(count.shape) != (B, g.N) or not count.is_cuda or not count.is_contiguous():
            raise ValueError('count must be a contiguous CUDA uint8 tensor [B,N]')
        dev = feats.device
        d = g.desc(dev, B)
        vol = torch.empty((B, g.N, g.Cv), device=dev, dtype=torch.float32)
        nbytes = _lib.load().vfd_voxel_fuse_fwd_workspace_bytes(ctypes.byref(d))
        if nbytes == 0:
            raise ValueError('vfd_voxel_fuse_fwd does not take C=%d, Cv=%d' % (g.C, g.Cv))
        ws = _workspace(nbytes, dev)
        _lib.call('vfd_voxel_fuse_fwd', ctypes.byref(d), _lib.ptr(feats), _lib.ptr(count), _lib.ptr(table),
                  _lib.ptr(w_no), _lib.ptr(b_no), _lib.ptr(w_ov), _lib.ptr(b_ov), _lib.ptr(vol), _lib.ptr(ws),
                  ctypes.c_size_t(ws.numel()), _stream())
        ctx.geom = g
        ctx.save_for_backward(feats, w_no, w_ov, count, table, vol)
        return vol

    @staticmethod
    def backward(ctx, d_vol):
        feats, w_no, w_ov, count, table, vol = ctx.saved_tensors
        g = ctx.geom
        B = feats.shape[0]
        dev = feats.device
        d = g.desc(dev, B)
        d_vol = d_vol.contiguous()
        d_feats = torch.empty_like(feats)
        d_w_no, d_b_no = torch.empty_like(w_no), torch.empty((g.Cv,), device=dev)
        d_w_ov, d_b_ov = torch.empty_like(w_ov), torch.empty((g.Cv,), device=dev)
        ws = _workspace(_lib.load().vfd_voxel_fuse_bwd_workspace_bytes(ctypes.byref(d)), dev)
        _lib.call('vfd_voxel_fuse_bwd', ctypes.byref(d), _lib.ptr(feats), _lib.ptr(count), _lib.ptr(table),
                  _lib.ptr(w_no), _lib.ptr(w_ov), _lib.ptr(vol), _lib.ptr(d_vol), _lib.ptr(d_feats), _lib.ptr(d_w_no),
                  _lib.ptr(d_b_no), _lib.ptr(d_w_ov), _lib.ptr(d_b_ov), _lib.ptr(ws), ctypes.c_size_t(ws.numel()),
                  _stream())
        return d_feats, d_w_no, d_b_no, d_w_ov, d_b_ov, None, None, None


def voxel_fuse(feats_cl, w_no, b_no, w_ov, b_ov, count, table, geom):
    """feats_cl [B,cams,h,w,C] + MLP parameters -> channel-last voxel volume [B,N,Cv]."""
    return _VoxelFuse.apply(feats_cl, w_no, b_no, w_ov, b_ov, count, table, geom)


def voxel_gather_depth(feats_cl, mask_small, K, Einv, geom, detail=False):
    """feats_cl [B,cams,h,w,C] -> x_no [B,N,ld_no], x_ov [B,N,ld_ov], count [B,N] (+ grid, valid with ``detail``)."""
    return _VoxelGatherDepth.apply(feats_cl, mask_small, K, Einv, geom, detail)


def voxel_gather_pose(feats_cl, mask_small, K, Einv, geom, bev=False):
    """-> avg [B,N,ld_no], count [B,N]; bev=True: rows ordered (y,x,z) = the channels-last BEV map [B,Y,X,Z*ld_no]"""
    return _VoxelGatherPose.apply(feats_cl, mask_small, K, Einv, geom, bev)


def ray_sample(vol_cl, invK, E, geom, want_grid=False):
    """vol_cl [B,Z,Y,X,Cv] -> samples [B,cams,h,w,D,Cv] (c fastest), (+ grid [B,cams,D,h,w,3])."""
    return _RaySample.apply(vol_cl, invK, E, geom, want_grid)
