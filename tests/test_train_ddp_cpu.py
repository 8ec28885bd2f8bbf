"""CPU, 2-process gloo: the multi-rank logic of the chained training step (BASELINE.json configs[4];
/root/reference/models/vfdepth.py:62-70).  ``TrainStep(ddp=True)`` wraps both networks in DistributedDataParallel; the
pose network is called twice per step before the single backward (models/geometry/pose.py:32-41), every rank owns its
own rigs.  With the kernels replaced by the oracle (there is no GPU here) the test checks what the collective must
deliver: after one step every rank holds the SAME parameters, and they equal a single-process step on the mean of the
two ranks' gradients."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vfdepth_b200 import config as vcfg
from vfdepth_b200 import train_step as ts
from vfdepth_b200.rig import make_rig


def _cfg():
    return vcfg.default_config('nuscenes', batch_size=1, model__voxel_size=[10, 8, 4],
                               model__voxel_str_p=[-10.0, -8.0, -3.0], model__voxel_unit_size=[2.0, 2.0, 1.5],
                               model__proj_d_bins=4, model__proj_d_end=12, model__fusion_feat_in_dim=8,
                               model__voxel_pre_dim=[4], training__height=64, training__width=96,
                               training__focal_length_scale=45)


def _classes():
    from tests.test_train_step_gpu import _OracleDepthNet, _OraclePoseNet, _oracle_loss

    def loss_fn(cfg, depth_net, pose_net, inputs, noise):
        return _oracle_loss(cfg, depth_net, pose_net, inputs, noise, None), {}

    return _OracleDepthNet, _OraclePoseNet, loss_fn


def _rig(rank):
    cfg = _cfg()
    inputs, ex = make_rig(cfg, seed=700 + rank, with_feats=False)
    return cfg, inputs, torch.stack(ex['noise'], 1)


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(2)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        cfg, inputs, noise = _rig(rank)
        d_cls, p_cls, loss_fn = _classes()
        step = ts.TrainStep(cfg, torch.device('cpu'), ddp=True, seed=21, depth_cls=d_cls, pose_cls=p_cls, loss_fn=loss_fn)
        loss = step.step(inputs, noise)
        flat = torch.cat([p.detach().flatten() for p in step.params])
        gathered = [torch.zeros_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        if rank == 0:
            torch.save(dict(params=gathered, loss=float(loss), grad_bytes=step.grad_bytes()), out)
    finally:
        dist.destroy_process_group()


def test_two_rank_ddp_step_equals_mean_gradient_step(tmp_path):
    out = str(tmp_path / 'ddp.pt')
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = torch.load(out)
    assert torch.equal(r['params'][0], r['params'][1])          # the all-reduce left both ranks with the same model
    # single process: gradients of the two rigs averaged by hand, then the same Adam step
    d_cls, p_cls, loss_fn = _classes()
    cfg = _cfg()
    ref = ts.TrainStep(cfg, torch.device('cpu'), ddp=False, seed=21, depth_cls=d_cls, pose_cls=p_cls, loss_fn=loss_fn)
    grads = None
    for rank in range(2):
        _, inputs, noise = _rig(rank)
        ref.opt.zero_grad(set_to_none=True)
        total, _ = loss_fn(cfg, ref.depth_net, ref.pose_net, inputs, noise)
        total.backward()
        g = [p.grad.clone() for p in ref.params]
        grads = g if grads is None else [a + b for a, b in zip(grads, g)]
    ref.opt.zero_grad(set_to_none=True)
    for p, g in zip(ref.params, grads):
        p.grad = g / 2
    ref.opt.step()
    flat = torch.cat([p.detach().flatten() for p in ref.params])
    assert r['grad_bytes'] == flat.numel() * 4
    # Adam's first step moves every weight by ~lr * sign(g): compare the updates, not the weights
    torch.manual_seed(21)
    init = ts.TrainStep(cfg, torch.device('cpu'), ddp=False, seed=21, depth_cls=d_cls, pose_cls=p_cls, loss_fn=loss_fn)
    w0 = torch.cat([p.detach().flatten() for p in init.params])
    du, dv = r['params'][0] - w0, flat - w0
    agree = ((du - dv).abs() <= 1e-3 * dv.abs().max()).float().mean()
    assert float(agree) > 0.999, float(agree)
