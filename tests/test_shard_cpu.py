"""CPU: the N>1 host logic (rig partition, per-rank seeds, max-over-ranks timing) under a real 2-process gloo group,
and the config contract (the reference's YAML keys, rel_cam_list, get_config on the reference's own files)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vfdepth_b200 import config as vcfg
from vfdepth_b200 import shard


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        assert shard.dist_env() == (rank, world, rank)
        rigs = shard.rank_rigs(17, rank, world)
        # every rank contributes its rig list; together they must be disjoint and cover 2 * (17 // 2) rigs
        gathered = [None] * world
        dist.all_gather_object(gathered, rigs)
        # rank r pretends its timed region took 10 + 5 r ms for 8 rigs
        rps, ms = shard.job_throughput(8, 10.0 + 5.0 * rank)
        # the synthetic rigs of two ranks must differ (per-rank seeds), and be reproducible
        from vfdepth_b200.rig import make_rig
        cfg = vcfg.default_config('ddad', batch_size=1, training__height=32, training__width=64)
        a, _ = make_rig(cfg, seed=shard.rank_seed(1234, rank), with_feats=False, with_images=False)
        k = [None] * world
        dist.all_gather_object(k, a[('K', 0)].sum().item())
        if rank == 0:
            torch.save(dict(gathered=gathered, rps=rps, ms=ms, k=k), out)
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo(tmp_path):
    out = str(tmp_path / 'r.pt')
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = torch.load(out)
    a, b = r['gathered']
    assert len(a) == len(b) == 8 and not set(a) & set(b) and set(a) | set(b) == set(range(16))
    assert r['ms'] == 15.0                                  # slowest rank
    assert abs(r['rps'] - 16 / 15e-3) < 1e-6                # all ranks' rigs / slowest rank's time
    assert r['k'][0] != r['k'][1]


def test_rank_rigs_partition():
    for n, w in ((16, 4), (17, 4), (3, 8), (8, 1)):
        parts = [shard.rank_rigs(n, r, w) for r in range(w)]
        flat = [i for p in parts for i in p]
        assert len(set(flat)) == len(flat) and len({len(p) for p in parts}) == 1
        keep_all = [i for r in range(w) for i in shard.rank_rigs(n, r, w, drop_last=False)]
        assert sorted(keep_all) == list(range(n))
    with pytest.raises(ValueError):
        shard.rank_rigs(4, 4, 4)


def test_single_process_reductions_are_identity():
    assert shard.max_over_ranks(3.5) == 3.5 and shard.sum_over_ranks(2) == 2.0


def test_config_contract():
    cfg = vcfg.default_config('ddad')
    assert cfg['data']['rel_cam_list'][0] == [1, 2] and cfg['data']['rel_cam_list'][5] == [3, 4]
    assert cfg['model']['voxel_size'] == [100, 100, 20] and cfg['training']['frame_ids'] == [0, -1, 1]
    nu = vcfg.default_config('nuscenes')
    assert (nu['training']['height'], nu['training']['width'], nu['training']['max_depth']) == (352, 640, 80.0)
    o = vcfg.default_config('ddad', batch_size=3, model__proj_d_bins=7)
    assert o['training']['batch_size'] == 3 and o['model']['proj_d_bins'] == 7

    class Obj:
        pass
    x = Obj()
    vcfg.flatten_into(x, cfg)          # the reference's read_config: every key of every section becomes an attribute
    assert x.voxel_str_p == [-50.0, -50.0, -15.0] and x.spatio_coeff == 0.03 and x.num_cams == 6


@pytest.mark.parametrize('name', ['ddad/ddad_surround_fusion.yaml', 'nuscenes/nusc_surround_fusion.yaml'])
def test_get_config_reads_reference_yaml(name):
    path = os.path.join(os.environ.get('VFDEPTH_REF', '/root/reference'), 'configs', name)
    if not os.path.isfile(path):
        pytest.skip('reference tree not mounted')
    cfg = vcfg.get_config(path)
    base = vcfg.default_config('ddad' if 'ddad' in name else 'nuscenes')
    for sec, key in (('model', 'voxel_size'), ('model', 'voxel_str_p'), ('model', 'voxel_unit_size'),
                     ('model', 'proj_d_bins'), ('model', 'fusion_feat_in_dim'), ('training', 'height'),
                     ('training', 'width'), ('training', 'frame_ids'), ('training', 'intensity_align'),
                     ('loss', 'spatio_coeff'), ('loss', 'spatio_tempo_coeff'), ('loss', 'disparity_smoothness')):
        assert cfg[sec][key] == base[sec][key], (sec, key)
    assert cfg['data']['num_cams'] == 6 and dict(cfg['data']['rel_cam_list']) == dict(base['data']['rel_cam_list'])
