"""ctypes binding of libvfdepth_b200.so (the C ABI declared in include/vfdepth_b200.h).

There is no fallback: if the library is missing it is built with nvcc, and if that fails, or a call returns a
non-zero status, a RuntimeError is raised.  Nothing in this package computes the hot path any other way.
"""
import ctypes
import os
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, 'lib', 'libvfdepth_b200.so')
VFD_MAX_CAMS = 8

c_float_p = ctypes.c_void_p


class FusionDesc(ctypes.Structure):
    """struct vfd_fusion_desc"""
    _fields_ = [
        ('B', ctypes.c_int32), ('num_cams', ctypes.c_int32), ('C', ctypes.c_int32),
        ('h', ctypes.c_int32), ('w', ctypes.c_int32),
        ('X', ctypes.c_int32), ('Y', ctypes.c_int32), ('Z', ctypes.c_int32),
        ('D', ctypes.c_int32), ('Cv', ctypes.c_int32),
        ('ld_no', ctypes.c_int32), ('ld_ov', ctypes.c_int32), ('half_ov', ctypes.c_int32),
        ('group', ctypes.c_int32 * VFD_MAX_CAMS),
        ('depth_div', ctypes.c_float),
        ('vox_start', ctypes.c_float * 3), ('vox_length', ctypes.c_float * 3),
        ('axis_x', ctypes.c_void_p), ('axis_y', ctypes.c_void_p), ('axis_z', ctypes.c_void_p),
        ('bins', ctypes.c_void_p),
    ]


VFD_WARP_SRC = 2
VFD_WARP_NBR = 2


class WarpDesc(ctypes.Structure):
    """struct vfd_warp_desc"""
    _fields_ = [
        ('B', ctypes.c_int32), ('num_cams', ctypes.c_int32), ('H', ctypes.c_int32), ('W', ctypes.c_int32),
        ('n_src', ctypes.c_int32), ('cam0', ctypes.c_int32), ('n_call', ctypes.c_int32),
        ('intensity_align', ctypes.c_int32),
        ('nbr', (ctypes.c_int32 * VFD_WARP_NBR) * VFD_MAX_CAMS),
        ('color', ctypes.c_void_p * (1 + VFD_WARP_SRC)),
        ('mask', ctypes.c_void_p), ('K', ctypes.c_void_p), ('invK', ctypes.c_void_p),
    ]


class SynDesc(ctypes.Structure):
    """struct vfd_syn_desc"""
    _fields_ = [
        ('B', ctypes.c_int32), ('num_cams', ctypes.c_int32), ('H', ctypes.c_int32), ('W', ctypes.c_int32),
        ('cam0', ctypes.c_int32), ('n_call', ctypes.c_int32),
        ('src', (ctypes.c_int32 * 3) * VFD_MAX_CAMS),
        ('min_depth', ctypes.c_float), ('max_depth', ctypes.c_float),
        ('mask', ctypes.c_void_p), ('K', ctypes.c_void_p), ('invK', ctypes.c_void_p),
    ]


VFD_AGG_LEVELS = 3


class AggDesc(ctypes.Structure):
    """struct vfd_agg_desc"""
    _fields_ = [
        ('n_img', ctypes.c_int32), ('h', ctypes.c_int32), ('w', ctypes.c_int32), ('C', ctypes.c_int32),
        ('n_coarse', ctypes.c_int32),
        ('ch', ctypes.c_int32 * VFD_AGG_LEVELS), ('cw', ctypes.c_int32 * VFD_AGG_LEVELS),
        ('slope', ctypes.c_float),
    ]


class WarpViews(ctypes.Structure):
    """struct vfd_warp_views"""
    _fields_ = [(n, ctypes.c_void_p) for n in ('color', 'color_mask', 'overlap', 'overlap_mask', 'reproj_loss',
                                               'reproj_mask', 'grid', 'selection')]


_lock = threading.Lock()
_lib = None
launches = 0   # number of kernels launched through the ABI by this process (bench.py reports it)
# kernels one call of each entry point launches (memsets not counted); anything not listed launches one
KERNELS_PER_CALL = {'vfd_voxel_classify': 3, 'vfd_voxel_fuse_fwd': 3, 'vfd_voxel_fuse_bwd': 5, 'vfd_warp_stats': 2, 'vfd_warp_loss_bwd': 2, 'vfd_warp_loss_train_fwd': 2, 'vfd_warp_loss_train_bwd': 2,
                    'vfd_smooth_loss_fwd': 2, 'vfd_smooth_loss_fwd_static': 2, 'vfd_smooth_loss_bwd': 2, 'vfd_smooth_loss_train_fwd': 2,
                    'vfd_feat_agg_bwd': 3, 'vfd_norm_image_fwd': 2}     # fine + one launch per coarse level (two at the shipped configs)


def load():
    """-> ctypes.CDLL; builds the library first if it is not there."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        from . import build as _build
        if _build._stale():          # missing, or older than any source / header it is built from
            _build.build()
        lib = ctypes.CDLL(LIB_PATH)
        lib.vfd_version.restype = ctypes.c_int
        lib.vfd_last_error.restype = ctypes.c_char_p
        lib.vfd_last_error.argtypes = [ctypes.c_int]
        lib.vfd_warp_bwd_workspace_bytes.restype = ctypes.c_size_t
        lib.vfd_warp_raw_bytes.restype = ctypes.c_size_t
        lib.vfd_voxel_classify_workspace_bytes.restype = ctypes.c_size_t
        lib.vfd_voxel_fuse_bwd_workspace_bytes.restype = ctypes.c_size_t
        lib.vfd_voxel_fuse_fwd_workspace_bytes.restype = ctypes.c_size_t
        lib.vfd_voxel_table_bytes.restype = ctypes.c_size_t
        _lib = lib
    return _lib


def check(code, what):
    if code != 0:
        raise RuntimeError('%s failed: %s (status %d)' % (what, load().vfd_last_error(code).decode(), code))


def ptr(t):
    """device pointer of a torch tensor (None -> NULL)."""
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


_events = None   # when a dict: name -> [(start, end) CUDA events] recorded around every ABI call (bench.py)


def start_event_trace():
    """Record a CUDA-event pair on the current stream around every ABI call until ``stop_event_trace``."""
    global _events
    _events = {}


def stop_event_trace():
    """-> {name: (launch count, total ms)}; synchronises the device."""
    global _events
    import torch
    ev, _events = _events, None
    torch.cuda.synchronize()
    return {k: (len(v), sum(a.elapsed_time(b) for a, b in v)) for k, v in (ev or {}).items()}


def call(name, *args, kernels=None):
    global launches
    fn = getattr(load(), name)
    launches += KERNELS_PER_CALL.get(name, 1) if kernels is None else kernels
    if _events is None:
        check(fn(*args), name)
        return
    import torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    rc = fn(*args)
    e1.record()
    _events.setdefault(name, []).append((e0, e1))
    check(rc, name)
