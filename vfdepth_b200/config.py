"""Configuration for the VFDepth hot path.

The reference flattens *every* section of its YAML into instance attributes
(``read_config``: /root/reference/network/volumetric_fusionnet.py:62-65,
/root/reference/models/geometry/view_rendering.py:19-22), so the key names
below are the drop-in contract.  ``get_config`` accepts the reference's own
YAML files unchanged and adds the same derived keys as
/root/reference/utils/misc.py:44-72 (``num_cams``, ``rel_cam_list``, ``mode``).
"""
import copy
import os
from collections import defaultdict

# camera adjacency of a 6-camera surround rig (/root/reference/utils/misc.py:10):
#        3 1
#  rear  5   0  front
#        4 2
REL_CAM = {0: [1, 2], 1: [0, 3], 2: [0, 4], 3: [1, 5], 4: [2, 5], 5: [3, 4]}

_DDAD_CAMS = ['camera_01', 'camera_05', 'camera_06', 'camera_07', 'camera_08', 'camera_09']
_NUSC_CAMS = ['CAM_FRONT', 'CAM_FRONT_LEFT', 'CAM_FRONT_RIGHT', 'CAM_BACK_LEFT', 'CAM_BACK_RIGHT', 'CAM_BACK']


def rel_cam_list(cameras):
    """camera-name list -> {cam index: [neighbour indices present in the list]}."""
    idx = []
    for name in cameras:
        if name in _DDAD_CAMS:
            idx.append(_DDAD_CAMS.index(name))
        elif name in _NUSC_CAMS:
            idx.append(_NUSC_CAMS.index(name))
        else:
            idx.append(None)
    out = defaultdict(list)
    for i in idx:
        out[i] = [j for j in REL_CAM[i] if j in idx]
    return out


def _base(dataset):
    ddad = dataset == 'ddad'
    cams = _DDAD_CAMS if ddad else _NUSC_CAMS
    return {
        'ddp': {'ddp_enable': False, 'world_size': 1, 'gpus': [0]},
        'model': {
            'num_layers': 18, 'weights_init': False,
            'depth_model': 'fusion', 'pose_model': 'fusion',
            'fusion_level': 2, 'fusion_feat_in_dim': 256, 'use_skips': False,
            'voxel_unit_size': [1.0, 1.0, 1.5], 'voxel_size': [100, 100, 20],
            'voxel_str_p': [-50.0, -50.0, -15.0], 'voxel_pre_dim': [64],
            'proj_d_bins': 50, 'proj_d_str': 2, 'proj_d_end': 50,
            'mode': 'train',
        },
        'data': {'dataset': 'ddad' if ddad else 'nuscenes', 'cameras': list(cams),
                 'num_cams': 6, 'rel_cam_list': rel_cam_list(cams)},
        'training': {
            'height': 384 if ddad else 352, 'width': 640, 'scales': [0], 'frame_ids': [0, -1, 1],
            'batch_size': 2, 'min_depth': 1.5, 'max_depth': 200.0 if ddad else 80.0,
            'spatio': True, 'spatio_temporal': True, 'intensity_align': True,
            'focal_length_scale': 300, 'aug_depth': False, 'aug_angle': [15, 15, 40],
        },
        'loss': {'disparity_smoothness': 0.001, 'spatio_coeff': 0.03,
                 'spatio_tempo_coeff': 0.1, 'pose_loss_coeff': 0.0},
        'eval': {'syn_visualize': False},
    }


def default_config(dataset='ddad', batch_size=None, **overrides):
    """The fusion config of ``configs/{ddad,nuscenes}/*_surround_fusion.yaml`` as a dict.

    ``overrides`` are ``section__key=value`` pairs, e.g. ``model__voxel_size=[10, 10, 4]``.
    """
    cfg = _base(dataset)
    if batch_size is not None:
        cfg['training']['batch_size'] = int(batch_size)
    for k, v in overrides.items():
        sec, key = k.split('__', 1)
        cfg[sec][key] = v
    return cfg


def get_config(path, mode='train', weight_path=None):
    """Read a reference YAML and add the derived keys (utils/misc.py:44-72)."""
    import yaml
    with open(path, 'r') as f:
        cfg = yaml.load(f, Loader=yaml.FullLoader)
    name = os.path.splitext(os.path.basename(path))[0]
    log_path = os.path.join(cfg['data'].get('log_dir', './results/'), name)
    cfg['data']['log_path'] = log_path
    cfg['data']['save_weights_root'] = os.path.join(log_path, 'models')
    if weight_path is None:
        weight_path = os.path.join(log_path, 'models', cfg.get('load', {}).get('weights', ''))
    cfg['data']['load_weights_dir'] = weight_path
    cfg['data']['num_cams'] = len(cfg['data']['cameras'])
    cfg['model']['mode'] = mode
    cfg['data']['rel_cam_list'] = rel_cam_list(cfg['data']['cameras'])
    if mode == 'train':
        cfg.setdefault('eval', {})['syn_visualize'] = False
    elif mode == 'eval':
        cfg['ddp']['world_size'] = 1
        cfg['ddp']['gpus'] = [0]
        cfg['training']['batch_size'] = cfg['eval']['eval_batch_size']
        cfg['training']['depth_flip'] = False
    return cfg


def flatten_into(obj, cfg):
    """Set every ``cfg[section][key]`` as ``obj.key`` (the reference's ``read_config``)."""
    for sec in cfg.keys():
        for k, v in cfg[sec].items():
            setattr(obj, k, v)


def clone(cfg):
    return copy.deepcopy(cfg)
