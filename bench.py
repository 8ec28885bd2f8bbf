"""bench.py -- rigs/second of the VFDepth fusion + view-synthesis hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 ... bench.py --gpus N ...

Headline (`value`, BASELINE.json configs[1] + configs[2]).  One step = one pass of the hot path, forward AND backward,
over one batch of synthetic DDAD-shaped rigs (8 rigs of 6 cameras, 384x640 images -> 48x80x256 feature maps,
100x100x20 voxels, 50 depth bins): VFNet.backproject_into_voxel (voxel classification, the two 1x1 MLPs as a dense
tcgen05 projection of the feature pixels, voxel gather) -> project_voxel_into_image without reduce_dim (ray sampling) ->
backward of both with a resident upstream gradient; then to_depth + the 6-camera x 3-frame view-synthesis warp + SSIM/L1
+ smoothness loss forward+backward.  Every rank owns whole rigs (weak scaling, no data-path collective); value = rigs all
ranks processed / max-over-ranks device time.

The same JSON line also carries
  roofline / rooflines   the two calls with the most time per step against the measured HBM peak
  e2e                    the same step fed from pinned host buffers (uint8 images, as datasets store them)
  full_step              + what a reference step runs around the headline: the two pose-network gathers with the BEV
                         hand-off, and the reduce_dim convolutions (cuDNN) of the depth and pose VFNets, timed apart
  train_step             BASELINE.json configs[3]/[4]: the chained nuScenes-shaped fusion-depth training step (two pose
                         forwards, depth forward, to_depth, warp + losses, one backward, Adam) under
                         DistributedDataParallel when N > 1 -- the one collective of the path, inside the timed step
  torch_cuda_baseline    the reference's own torch ops (the oracle) on the same B200 with the reference's flags
  cpu_baseline           the same ops on the host cores: all cores, and 1 thread (the reference launcher's setting)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = 'vfnet_fusion_warp_fwd_bwd_rigs_per_sec'
UNIT = 'rigs/s'
BATCH = 8                 # rigs per GPU per step (configs[1])
WORKLOAD = ('configs[1]: VFNet fusion fwd+bwd (voxel classification + 1x1 MLPs + voxel gather + ray sampling, no reduce_dim)%s, '
            '%d DDAD-shaped rigs (6x384x640) per GPU per step')
WARP_PART = ' + configs[2]: to_depth + 6-cam x 3-frame view-synthesis warp + SSIM/L1 photometric + smoothness loss fwd+bwd (to disparity and poses)'
DATASET = 'ddad'


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--batch', type=int, default=BATCH)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--e2e-grads', action='store_true',
                    help='e2e also copies the gradients (d_feats, d_disp: 236 MB per 8 rigs) to the host every step')
    ap.add_argument('--no-graph', action='store_true', help='launch the resident step eagerly instead of replaying a CUDA graph')
    ap.add_argument('--no-warp', action='store_true', help='time the fusion half only')
    ap.add_argument('--no-extras', action='store_true', help='headline only: no full_step / train_step / torch_cuda_baseline')
    return ap.parse_args()


def dist_env():
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    return rank, world, local


def measured_peak():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.isfile(p):
        with open(p) as f:
            return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    return 6650.0, 'fallback (B200_PROFILING.md)'


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '200'], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(',')]
            if len(f) < 6 or not f[0].isdigit():
                continue
            sm.append(int(f[0]))
            mx = int(f[1])
            for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), f[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': mx, 'reasons': sorted(reasons),
                'samples': len(sm)}


# --------------------------------------------------------------------------------------------------
# workload
# --------------------------------------------------------------------------------------------------
class FusionWorkload:
    """BATCH synthetic DDAD-shaped rigs on one device + the drop-in VFNet (random-init MLP weights)."""

    def __init__(self, batch, dev, seed):
        from vfdepth_b200 import config as vcfg
        from vfdepth_b200.rig import make_rig
        from vfdepth_b200.vfnet import VFNet
        self.cfg = vcfg.default_config(DATASET, batch_size=batch)
        self.dev, self.B = dev, batch
        inputs, _ = make_rig(self.cfg, seed=seed, with_feats=False, with_images=False)
        m = self.cfg['model']
        self.lvl = m['fusion_level'] + 1
        torch.manual_seed(seed)
        self.net = VFNet(self.cfg, m['fusion_feat_in_dim'], 128, 'depth').to(dev)
        g = self.net.geom
        self.g = g
        gen = torch.Generator().manual_seed(seed)
        # host (pinned) copies: what a data loader / the encoder hand-off would hold
        pin = lambda t: t.contiguous().pin_memory()
        # channel-last feature maps, as a channels_last 1x1 conv emits them; logical shape [B,6,C,h,w]
        self.h_feats = pin(torch.randn((batch, g.num_cams, g.h, g.w, g.C), generator=gen))
        self.h_small = {k: pin(inputs[k]) for k in ('mask', 'extrinsics', 'extrinsics_inv')}
        self.h_small['K'] = pin(inputs[('K', self.lvl)])
        self.h_small['inv_K'] = pin(inputs[('inv_K', self.lvl)])
        self.d_feats_host = torch.empty_like(self.h_feats).pin_memory()
        self.h_scalar = torch.zeros(1).pin_memory()
        # resident copies
        self.feats = self.h_feats.to(dev)
        self.small = {k: v.to(dev) for k, v in self.h_small.items()}
        # upstream gradient of the ray samples: produced on the device by reduce_dim's backward in training
        self.g_up = torch.randn((batch, g.num_cams, g.h, g.w, g.D, g.Cv), device=dev)
        self._stage = {}

    def step(self, feats, small):
        """fusion forward + backward through the drop-in API; returns (ray rows, d_feats)."""
        f = feats.permute(0, 1, 4, 2, 3).requires_grad_(True)        # [B,6,C,h,w] view, channels_last strides
        vox = self.net.backproject_into_voxel(f, small['mask'], small['K'], small['extrinsics_inv'])
        rows = self.net.ray_rows(vox, small['inv_K'], small['extrinsics'])
        for p in self.net.parameters():
            p.grad = None
        torch.autograd.backward(rows, self.g_up)
        return rows, f.grad

    def step_resident(self):
        return self.step(self.feats, self.small)

    def upload(self, slot=0):
        """host (pinned) -> device copies of one step's inputs into the staging buffers of `slot`, on the current
        stream (preallocated: no allocator traffic inside the timed region)."""
        if slot not in self._stage:
            mk = lambda t: torch.empty(t.shape, dtype=t.dtype, device=self.dev)
            self._stage[slot] = (mk(self.h_feats), {k: mk(v) for k, v in self.h_small.items()})
        feats, small = self._stage[slot]
        feats.copy_(self.h_feats, non_blocking=True)
        for k, v in self.h_small.items():
            small[k].copy_(v, non_blocking=True)
        return feats, small

    def compute(self, staged):
        return self.step(*staged)

    def download(self, result, grads_too=False):
        """the step's result read back by the host: one element of the ray samples (this half has no loss of its own).
        ``grads_too``: also d_feats, 189 MB per 8 rigs -- in training it never leaves the device (the encoder's backward
        consumes it there)"""
        rows, d_feats = result
        if grads_too:
            self.d_feats_host.copy_(d_feats.permute(0, 1, 3, 4, 2), non_blocking=True)
        self.h_scalar.copy_(rows[0, 0, 0, 0, 0, :1], non_blocking=True)

    def h2d_bytes(self):
        return self.h_feats.numel() * 4 + sum(v.numel() * 4 for v in self.h_small.values())

    def d2h_bytes(self, grads_too=False):
        return (self.d_feats_host.numel() * 4 if grads_too else 0) + 4

    def algorithmic_bytes(self):
        """per launch, SURVEY.md 8d: unique inputs read once + required outputs written once (fp32)."""
        g, B = self.g, self.B
        feats = B * g.num_cams * g.h * g.w * g.C * 4
        msk = B * g.num_cams * g.h * g.w * 4
        vol = B * g.N * g.Cv * 4
        ray = B * g.num_cams * g.h * g.w * g.Cv * g.D * 4
        # corner table: 40 bytes per (voxel, camera) entry of the voxels seen by exactly one / exactly two cameras
        # (data dependent; read from the last classification), written once, read by the forward and the backward
        cnt = self.net.last_count
        m1 = int((cnt == 1).sum()) if cnt is not None else B * g.N
        m2 = int((cnt == 2).sum()) if cnt is not None else B * g.N
        table = (m1 + 2 * m2) * 40
        return {'vfd_voxel_classify': msk + B * g.N + table,
                # fused gather + MLPs: neither the GEMM rows nor the projected pixels reach HBM as
                # required traffic.  fwd: features + table + count in, volume out; bwd: features + volume + its gradient
                # + table in, feature gradient out
                'vfd_voxel_fuse_fwd': feats + table + B * g.N + vol, 'vfd_voxel_fuse_bwd': 2 * feats + table + B * g.N + 2 * vol,
                'vfd_ray_sample_fwd': vol + ray, 'vfd_ray_sample_bwd': vol + ray}


def timed(fn, steps, world):
    """K steps bracketed by barrier + synchronize on both sides, CUDA events on the launching stream."""
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


# --------------------------------------------------------------------------------------------------
# baselines: the reference's own torch path (the oracle, bit-identical to the reference on the CPU --
# tests/test_oracle_golden.py) on the host cores or on the GPU
# --------------------------------------------------------------------------------------------------
def reference_step(batch, device, seed=1234, with_warp=True):
    """-> a closure running one step (fusion fwd+bwd + warp/loss fwd+bwd of `batch` rigs) through the oracle on
    `device`; the rigs are drawn once, outside the timed region."""
    from oracle import vfdepth_oracle as orc
    from vfdepth_b200 import config as vcfg
    from vfdepth_b200.rig import make_rig
    cfg = vcfg.default_config(DATASET, batch_size=batch)
    inputs, extras = make_rig(cfg, seed=seed, with_images=with_warp)
    lvl = cfg['model']['fusion_level'] + 1
    C, Cv = cfg['model']['fusion_feat_in_dim'], cfg['model']['voxel_pre_dim'][-1]
    gen = torch.Generator().manual_seed(seed)
    mlp = dict(w_no=torch.randn((Cv, C + 1, 1), generator=gen) * 0.05, b_no=torch.zeros(Cv),
               w_ov=torch.randn((Cv, 2 * (C + 1), 1), generator=gen) * 0.05, b_ov=torch.zeros(Cv))
    h, w = orc.feature_hw(cfg)
    g_up = torch.randn((batch, Cv * cfg['model']['proj_d_bins'], h, w), generator=gen)
    to = lambda t: t.to(device)
    inputs = {k: to(v) for k, v in inputs.items()}
    feats = to(extras['feats_agg'])
    mlp = {k: to(v) for k, v in mlp.items()}
    g_up = to(g_up)
    if with_warp:
        disp0 = [to(d) for d in extras['disp']]
        T0 = {k: to(v) for k, v in extras['cam_T_cam'].items()}
        noise = [to(n) for n in extras['noise']]

    def one():
        f = feats.clone().requires_grad_(True)
        p = {k: v.clone().requires_grad_(True) for k, v in mlp.items()}
        vox = orc.backproject_into_voxel(cfg, f, inputs['mask'], inputs[('K', lvl)], inputs['extrinsics_inv'], 'depth', p)
        proj = orc.project_voxel_into_image(cfg, vox, inputs[('inv_K', lvl)], inputs['extrinsics'])
        torch.autograd.backward(proj, [g_up] * len(proj))
        if with_warp:
            disps = [d.clone().requires_grad_(True) for d in disp0]
            depths = [orc.to_depth(cfg, d, inputs[('K', 0)][:, c]) for c, d in enumerate(disps)]
            Ts = {k: v.clone().requires_grad_(True) for k, v in T0.items()}
            total, _ = orc.warp_and_loss(cfg, inputs, depths, disps, Ts, noise)
            torch.autograd.grad(total, disps + list(Ts.values()))
        return f.grad

    return one


def cpu_step_pass(steps, warmup, threads=None, with_warp=True):
    """One rig per step through the oracle on the host cores.  -> (rigs/s, threads, seconds per rig)."""
    cores = os.cpu_count() or 1
    threads = cores if threads is None else threads
    prev = torch.get_num_threads()
    torch.set_num_threads(threads)
    try:
        one = reference_step(1, torch.device('cpu'), with_warp=with_warp)
        for _ in range(warmup):
            one()
        t0 = time.perf_counter()
        for _ in range(steps):
            one()
        dt = time.perf_counter() - t0
    finally:
        torch.set_num_threads(prev)
    return steps / dt, threads, dt / steps


def torch_cuda_pass(batch, dev, with_warp=True):
    """The 'reference torch path on the same B200' of BASELINE.json configs[1]: the reference's ATen calls (the oracle) on
    cuda with /root/reference/train.py:23-25's flags, CUDA-event timed.  -> dict (or an 'unavailable' note)."""
    flags = (torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark = True, False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        for b in (batch, 2, 1):
            try:
                one = reference_step(b, dev, with_warp=with_warp)
                one()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                n = 3
                e0.record()
                for _ in range(n):
                    one()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / n
                return {'value': b / (ms * 1e-3), 'unit': UNIT, 'ms_per_step': ms, 'rigs_per_step': b,
                        'how': 'the reference\'s torch ops (oracle/vfdepth_oracle.py = network/volumetric_fusionnet.py, '
                               'models/geometry, models/losses) on cuda:0, cudnn.deterministic=True, benchmark=False, '
                               'matmul.allow_tf32=False (train.py:23-25), 1 warm-up + %d timed steps, CUDA events' % n}
            except torch.cuda.OutOfMemoryError:
                one = None
                torch.cuda.empty_cache()
        return {'unavailable': 'out of memory at every batch size tried'}
    finally:
        torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark = flags[0], flags[1]
        torch.backends.cuda.matmul.allow_tf32 = flags[2]
        torch.cuda.empty_cache()


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    rps, cores, sec = cpu_step_pass(steps, warmup, with_warp=not args.no_warp)
    sample = ('fusion fwd+bwd (gather + 1x1 MLPs + ray sampling, no reduce_dim)%s of 1 DDAD rig per step, %d warm-up + %d '
              'timed steps, reference torch ops on %d host threads' % ('' if args.no_warp else ' + 6-cam x 3-frame warp + SSIM/L1 loss fwd+bwd', warmup, steps, cores))
    line = {'impl': 'reference', 'metric': METRIC, 'value': rps, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': steps,
            'warmup': warmup, 'ms_per_step': sec * 1e3, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': WORKLOAD % ('' if args.no_warp else WARP_PART, 1) + '; reference torch path on the host cores',
                       'rigs_per_step_per_gpu': 1},
            'cpu_baseline': {'value': rps, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
            'e2e': {'value': rps, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# extras: what a reference step runs around the headline, and the chained training step
# --------------------------------------------------------------------------------------------------
def _event_ms(fn, iters):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def full_step_pass(wl, headline_ms, iters):
    """SURVEY.md 8d config 2 asks for the geometry core and reduce_dim apart, and a reference step does three voxel
    gathers (one depth + two pose, network/fusion_posenet.py:70 via models/geometry/pose.py:32-41): time the two pose
    gathers with the BEV hand-off (fwd+bwd), the pose reduce_dim (fwd+bwd, cuDNN) and the depth reduce_dim over the ray
    samples (fwd+bwd, cuDNN) on the same rigs."""
    from vfdepth_b200.vfnet import VFNet
    dev, B, g = wl.dev, wl.B, wl.g
    torch.manual_seed(7)
    pose = VFNet(wl.cfg, g.C, 128, 'pose').to(dev)
    feats = [torch.randn_like(wl.feats) for _ in range(2)]                      # the two pose encoders' feature maps
    s = wl.small
    out = {}
    with torch.no_grad():
        bev0 = pose.bev_rows(feats[0].permute(0, 1, 4, 2, 3), s['mask'], s['K'], s['extrinsics_inv'])
    g_bev = torch.randn_like(bev0)

    def pose_gathers():
        for f in feats:
            x = f.permute(0, 1, 4, 2, 3).requires_grad_(True)
            rows = pose.bev_rows(x, s['mask'], s['K'], s['extrinsics_inv'])
            torch.autograd.backward(rows, g_bev)

    def pose_reduce():
        for _ in feats:
            x = bev0.detach().requires_grad_(True)
            y = pose.reduce_bev(x)
            y.sum().backward()

    rows0, _ = wl.step_resident()
    rows0 = rows0.detach()

    def depth_reduce():
        x = rows0.requires_grad_(True)
        Bc, n, h, w, D, cv = x.shape
        xi = x.view(Bc * n, h, w, D * cv).permute(0, 3, 1, 2)
        conv0 = wl.net.reduce_dim[0]
        w0 = conv0.weight.view(conv0.out_channels, cv, D, 3, 3).permute(0, 2, 1, 3, 4).reshape(conv0.out_channels, D * cv, 3, 3)
        y = torch.nn.functional.conv2d(torch.nn.functional.pad(xi, (1, 1, 1, 1), mode='reflect'), w0, conv0.bias)
        for layer in list(wl.net.reduce_dim)[1:]:
            y = layer(y)
        y.sum().backward()
        x.grad = None

    for name, fn in (('pose_gathers_2x_fwd_bwd_ms', pose_gathers), ('reduce_dim_pose_2x_fwd_bwd_ms', pose_reduce),
                     ('reduce_dim_depth_fwd_bwd_ms', depth_reduce)):
        fn()
        out[name] = _event_ms(fn, iters)
    out['hot_path_ms'] = headline_ms
    out['geometry_core_ms'] = headline_ms + out['pose_gathers_2x_fwd_bwd_ms']
    out['rigs_per_s_geometry_core'] = B / (out['geometry_core_ms'] * 1e-3)
    total = out['geometry_core_ms'] + out['reduce_dim_pose_2x_fwd_bwd_ms'] + out['reduce_dim_depth_fwd_bwd_ms']
    out['ms_per_step'] = total
    out['rigs_per_s'] = B / (total * 1e-3)
    pose_bytes = 2 * 2 * (B * g.num_cams * g.h * g.w * g.C * 4 + B * g.N * (g.C + 1) * 4)
    out['pose_gathers_frac_of_hbm_peak'] = pose_bytes / (out['pose_gathers_2x_fwd_bwd_ms'] * 1e-3) / 1e9 / measured_peak()[0]
    out['what'] = ('headline hot path + the two pose-network voxel gathers with the BEV hand-off (geometry core), + reduce_dim of '
                   'the pose VFNet (2x) and of the depth VFNet over the ray samples -- cuDNN convolutions that stay torch '
                   '(SURVEY.md 8a A6), timed separately, per GPU, %d rigs' % B)
    del pose, feats, bev0, g_bev, rows0
    torch.cuda.empty_cache()
    return out


def dense_mask_pass(warp, iters):
    """the view-synthesis half with the auto-mask ON at every pixel (a huge tie-break noise makes the identity
    reprojection lose everywhere, as tests/test_warp_gpu.py does): the loss kernels skip work where a mask is zero and the
    synthetic rigs' auto-mask keeps under 1 % of the pixels, so this is the other end of the range -- driving data, whose
    auto-mask typically keeps half of the pixels or more, lies in between"""
    from vfdepth_b200 import _lib
    noise = torch.full_like(warp.noise, 1e9)
    fn = lambda: warp.step(warp.inputs, warp.disp, noise, warp.T)
    fn()
    _lib.start_event_trace()
    ms = _event_ms(fn, iters)
    tr = _lib.stop_event_trace()
    per = {name: tot / max(1, n) for name, (n, tot) in tr.items()}      # ms per call
    return {'ms_per_step': ms, 'vfd_warp_loss_train_fwd_ms': per.get('vfd_warp_loss_train_fwd'),
            'sparse_ms_per_step': _event_ms(lambda: warp.step_resident(), iters),
            'what': 'to_depth + warp + losses fwd+bwd of the same rigs with the auto-mask on at every pixel (eager launches), '
                    'next to the same call on the bench rigs as they are (auto-mask sparse)'}


def dropin_pass(warp, iters):
    """the view-synthesis half through the reference's own call pattern (models/vfdepth.py:296-313: per camera
    ViewRendering.forward then MultiCamLoss.forward on the outputs dict) instead of the rig-level view.warp_and_loss --
    the drop-in modules share one rig-level pass, so this should cost what the rig-level call costs"""
    from vfdepth_b200 import view
    cfg, n = warp.cfg, warp.cfg['data']['num_cams']
    vr, lossm = view.ViewRendering(cfg, 0), view.MultiCamLoss(cfg, 0, device_noise=True)
    inputs = warp.inputs
    E, Einv = inputs['extrinsics'], inputs['extrinsics_inv']
    rel_list = cfg['data']['rel_cam_list']

    def one():
        disp = warp.disp.detach().requires_grad_(True)
        Ts = {f: t.detach().requires_grad_(True) for f, t in warp.T.items()}
        depth = view.to_depth(cfg, disp, inputs[('K', 0)])
        outputs = {('cam', c): {('depth', 0): depth[:, c:c + 1], ('disp', 0): disp[:, c:c + 1],
                                **{('cam_T_cam', 0, f): Ts[f][:, c] for f in Ts}} for c in range(n)}
        total = 0
        for cam in range(n):
            rel = {}
            for nb in rel_list[cam]:                       # Pose.compute_relative_cam_poses (pose.py:66-96)
                rel[(0, nb)] = torch.matmul(Einv[:, nb], E[:, cam])
                for f in Ts:
                    rel[(f, nb)] = torch.matmul(rel[(0, nb)], outputs[('cam', cam)][('cam_T_cam', 0, f)])
            vr(inputs, outputs, cam, rel)
            cam_loss, _ = lossm(inputs, outputs, cam)
            total = total + cam_loss
        (total / n).backward()

    one()
    ms = _event_ms(one, iters)
    rig = _event_ms(lambda: warp.step_resident(), iters)
    # the same per-camera call pattern replayed from a CUDA graph (the modules never synchronise or read back: the
    # ~150 small torch launches of the pose products and the dict plumbing then cost no host time)
    graph_ms = None
    try:
        from vfdepth_b200 import graph as vgraph
        replay, _ = vgraph.capture(one, warmup=1)
        replay()
        graph_ms = _event_ms(replay, iters)
        del replay
    except Exception as e:  # noqa: BLE001 -- reported, not fatal: the eager number stands
        print('bench: CUDA graph capture of the per-camera modules failed (%s)' % e, file=sys.stderr, flush=True)
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
    return {'per_camera_modules_ms': ms, 'rig_level_call_ms': rig, 'per_camera_modules_cuda_graph_ms': graph_ms,
            'what': 'to_depth + view synthesis + losses fwd+bwd of %d rigs: six ViewRendering / MultiCamLoss calls on the outputs '
                    'dict (the reference\'s call pattern, incl. the per-camera pose products; loss_dict entries stay on the '
                    'device) against one view.warp_and_loss call, both as eager launches; and the per-camera pattern replayed from a '
                    'CUDA graph' % warp.B}


def train_step_pass(dev, rank, world, steps):
    """BASELINE.json configs[3] (nuScenes-shaped full fusion-depth training step, random init, B=2 per GPU as
    configs/nuscenes/nusc_surround_fusion.yaml:47) and configs[4] (the same under DDP: the gradient all-reduce of every
    parameter over NCCL/NVLink inside the timed step, no per-step barrier).  Encoders / decoders are stand-ins
    (vfdepth_b200/train_step.py) -- they are outside the hot path."""
    from vfdepth_b200 import config as vcfg
    from vfdepth_b200 import train_step as ts
    from vfdepth_b200.rig import make_rig
    B = 2
    cfg = vcfg.default_config('nuscenes', batch_size=B)
    inputs, _ = make_rig(cfg, seed=4321 + 1000 * rank, with_feats=False)
    inputs = {k: v.to(dev) for k, v in inputs.items()}
    step = ts.TrainStep(cfg, dev, ddp=world > 1, seed=11)
    for _ in range(3):
        step.step(inputs)
    n = max(3, min(steps, 10))
    ms = timed(lambda: step.step(inputs), n, world) / n
    out = {'value': B * world / (ms * 1e-3), 'unit': UNIT, 'ms_per_step': ms, 'rigs_per_step_per_gpu': B, 'steps': n,
           'n_gpus': world, 'dataset_shape': 'nuScenes 6 x 352 x 640, 44 x 80 x 256 features, 100 x 100 x 20 voxels',
           'param_grad_bytes': step.grad_bytes(),
           'collective': ('DistributedDataParallel over NCCL: %d bytes of parameter gradients (encoders, 1x1 aggregation, VFNet '
                          'MLPs + reduce_dim of both networks, decoders) all-reduced per step in 25 MB buckets overlapped with '
                          'the backward; no per-step barrier' % step.grad_bytes()) if world > 1 else 'none (single GPU)',
           'what': 'two pose-network forwards + depth-network forward + to_depth + rig-level warp/loss + one backward + Adam; '
                   'stand-in encoders / decoders (out of scope), everything between them on the CUDA kernels'}
    del step
    torch.cuda.empty_cache()
    return out


def run_ours(args):
    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device: vfdepth_b200 has no CPU path')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    numa = 'unbound (single process)'
    if world > 1:
        # before any pinned allocation: first-touch places the staging buffers on the GPU's own NUMA node
        from vfdepth_b200 import shard
        numa = shard.bind_to_gpu_numa(local)
        dist.init_process_group('nccl', device_id=dev)
    torch.backends.cuda.matmul.allow_tf32 = False      # the reference's setting (train.py:25)
    from vfdepth_b200 import _lib
    wl = FusionWorkload(args.batch, dev, seed=1234 + 1000 * rank)
    warp = None
    if not args.no_warp:
        from vfdepth_b200.bench_warp import WarpWorkload
        warp = WarpWorkload(args.batch, dev, seed=1234 + 1000 * rank)

    def step():
        wl.step_resident()
        if warp is not None:
            warp.step_resident()

    parts = [wl] + ([warp] if warp is not None else [])
    copy_stream, out_stream = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def _tensors(x):
        if isinstance(x, torch.Tensor):
            yield x
        elif isinstance(x, dict):
            for v in x.values():
                yield from _tensors(v)
        elif isinstance(x, (list, tuple)):
            for v in x:
                yield from _tensors(v)

    free_ev = {0: None, 1: None}

    def stage(slot):
        """issue the H2D copies of ONE step's inputs on the copy stream, into staging slot `slot` (two slots: the copy of
        step i+1 overlaps the compute of step i; a slot is rewritten only after the step that read it has finished)"""
        with torch.cuda.stream(copy_stream):
            if free_ev[slot] is not None:
                copy_stream.wait_event(free_ev[slot])
            staged = [p.upload(slot) for p in parts]
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        return staged, ev

    e2e_graphs, dl_ev = {}, {0: None, 1: None}

    def compute_slot(slot, staged):
        """the public-API calls of one step on the staged inputs of `slot`; replayed from a CUDA graph captured per
        staging slot (the staging tensors are fixed per slot) unless --no-graph"""
        if args.no_graph:
            return [p.compute(x) for p, x in zip(parts, staged)]
        if slot not in e2e_graphs:
            from vfdepth_b200 import graph as vgraph
            e2e_graphs[slot] = vgraph.capture(lambda: [p.compute(x) for p, x in zip(parts, staged)], warmup=1)
        replay, outs = e2e_graphs[slot]
        replay()
        return outs

    def run_e2e(steps):
        """the public API fed from pinned host buffers: every step's inputs are copied host->device and its results
        device->host inside the timed region; the copy of step i+1 is issued before the compute of step i (two extra
        streams), so copies overlap compute the way a training input pipeline does.  Exactly `steps` uploads."""
        main = torch.cuda.current_stream()
        copy_stream.wait_stream(main)
        nxt = stage(0)
        for i in range(steps):
            staged, ev = nxt
            if i + 1 < steps:
                nxt = stage((i + 1) % 2)
            main.wait_event(ev)
            if dl_ev[i % 2] is not None:       # the graph rewrites this slot's result tensors in place
                main.wait_event(dl_ev[i % 2])
            results = compute_slot(i % 2, staged)
            done = torch.cuda.Event()
            done.record(main)
            free_ev[i % 2] = done
            with torch.cuda.stream(out_stream):
                out_stream.wait_event(done)
                for p, r in zip(parts, results):
                    p.download(r, args.e2e_grads)
                dl_ev[i % 2] = torch.cuda.Event()
                dl_ev[i % 2].record(out_stream)
            if args.no_graph:                  # eager results come from the caching allocator: keep them alive for the D2H
                for t in _tensors(results):
                    t.record_stream(out_stream)
        main.wait_stream(out_stream)
        main.wait_stream(copy_stream)

    for _ in range(max(3, args.warmup)):
        step()
    torch.cuda.synchronize()
    # the resident step is replayed from a CUDA graph (vfdepth_b200/graph.py): same kernels, no launch gaps
    run_step, graphed, per_step = step, False, None
    if not args.no_graph:
        try:
            from vfdepth_b200 import graph as vgraph
            n0 = _lib.launches
            run_step, _ = vgraph.capture(step, warmup=1)
            per_step = (_lib.launches - n0) // 2          # one warm-up call + the captured call
            graphed = True
            for _ in range(2):
                run_step()
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001 -- fall back to eager launches, and say so in the line
            print('bench: CUDA graph capture failed (%s); eager launches' % e, file=sys.stderr, flush=True)
            run_step, graphed = step, False
            torch.cuda.synchronize()
            torch.cuda.empty_cache()         # drop the aborted capture's private pool, then re-warm the eager path
            for _ in range(3):
                step()
            torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = _lib.launches
    ms = timed(run_step, args.steps, world)
    launches = per_step * args.steps if graphed else _lib.launches - n0
    rigs = args.batch * world * args.steps
    value = rigs / (ms * 1e-3)

    # per-kernel device time inside a second timed pass (events around every ABI call on the launch stream)
    _lib.start_event_trace()
    for _ in range(args.steps):
        step()
    trace = _lib.stop_event_trace()

    e2e = None
    if not args.no_e2e:
        run_e2e(4)          # allocator steady state for the per-step results
        ms_e2e = timed(lambda: run_e2e(args.steps), 1, world)
        h2d = wl.h2d_bytes() + (warp.h2d_bytes() if warp else 0)
        d2h = wl.d2h_bytes(args.e2e_grads) + (warp.d2h_bytes(args.e2e_grads) if warp else 0)
        e2e = {'value': rigs / (ms_e2e * 1e-3), 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
               'ms_per_step': ms_e2e / args.steps,
               'cuda_graph': not args.no_graph,
               'how': 'pinned host buffers; H2D of step i+1 on a copy stream while step i computes, D2H on a third stream; the '
                      'public-API calls of a step are replayed from a CUDA graph captured per staging slot. D2H = the step\'s result '
                      '(loss + one ray sample)%s. '
                      'Images and masks travel as uint8 (what every dataset stores) and are converted on the device inside the timed '
                      'region; %d of the H2D bytes are the fp32 feature maps, which stand in for an encoder that lives on '
                      'the device in training (they never exist on the host there)'
                      % (' and the gradients d_feats / d_disp' if args.e2e_grads else '; the gradients stay on the device, where '
                         'the encoder / decoder backward consumes them in training (--e2e-grads copies them too)',
                         wl.h_feats.numel() * 4)}
    clocks = sampler.stop() if rank == 0 else None

    extras = {}
    if not args.no_extras:
        del run_step
        torch.cuda.empty_cache()
        full = full_step_pass(wl, ms / args.steps, max(2, min(args.steps, 5)))
        if rank == 0:
            extras['full_step'] = full
            if warp is not None:
                extras['dropin_modules'] = dropin_pass(warp, max(2, min(args.steps, 5)))
                extras['warp_dense_auto_mask'] = dense_mask_pass(warp, max(2, min(args.steps, 5)))
        wl.g_up = None
        torch.cuda.empty_cache()
        tr = train_step_pass(dev, rank, world, args.steps)
        if rank == 0:
            extras['train_step'] = tr

    if rank == 0:
        peak, how = measured_peak()
        alg = wl.algorithmic_bytes()
        if warp is not None:
            alg.update(warp.algorithmic_bytes())
        kernels = {}
        for name, (n, tot) in trace.items():
            per = tot / max(n, 1)
            k = {'launches_per_step': n / args.steps, 'ms_per_launch': per, 'ms_per_step': tot / args.steps}
            if name in alg:
                k['algorithmic_mb'] = alg[name] / 1e6
                k['gbs'] = alg[name] / per / 1e6
                k['frac_of_hbm_peak'] = k['gbs'] / peak
            kernels[name] = k
        ours_ms = sum(k['ms_per_step'] for k in kernels.values())
        traffic = {}
        tpath = os.path.join(ROOT, 'profiles', 'traffic.json')
        if os.path.isfile(tpath):
            with open(tpath) as f:
                traffic = json.load(f)

        def roof(name):
            return {'bound': 'hbm', 'kernel': name, 'achieved': kernels[name].get('gbs'), 'peak': peak, 'unit': 'GB/s',
                    'frac': kernels[name].get('frac_of_hbm_peak'), 'traffic': traffic.get(name), 'peak_source': how,
                    'algorithmic_bytes_per_launch': alg.get(name), 'ms_per_launch': kernels[name]['ms_per_launch'],
                    'share_of_step': kernels[name]['ms_per_step'] / ours_ms}

        ranked = sorted(kernels, key=lambda n: -kernels[n]['ms_per_step'])
        step_alg = sum(alg.get(n, 0) * kernels[n]['launches_per_step'] for n in kernels)
        line = {'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
                'warmup': max(3, args.warmup), 'ms_per_step': ms / args.steps, 'higher_is_better': True,
                'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
                'config': {'workload': WORKLOAD % (WARP_PART if warp else '', args.batch),
                           'rigs_per_step_per_gpu': args.batch, 'l2': 'inputs and outputs larger than L2 '
                           '(189 MB feature maps, 410 MB voxel volume, 2.4 GB ray samples, 0.6 GB of images per step)',
                           'sharding': 'whole rigs per rank, no data-path collective', 'host_affinity': numa, 'cuda_graph': graphed,
                           'kernel_ms_per_step': ours_ms, 'torch_glue_ms_per_step': ms / args.steps - ours_ms,
                           'loss_mask_density': warp.mask_density() if warp else None},
                'clocks': clocks, 'e2e': e2e, 'gpu_launches': launches, 'roofline': roof(ranked[0]),
                'rooflines': [roof(n) for n in ranked[:2]],
                'step_roofline': {'algorithmic_bytes_per_step': step_alg, 'frac_of_hbm_peak': step_alg / (ms / args.steps) / 1e6 / peak},
                'kernels': kernels}
        line.update(extras)
        if world == 1 and not args.no_extras:
            tc = torch_cuda_pass(args.batch, dev, with_warp=warp is not None)
            line['torch_cuda_baseline'] = tc
            if 'value' in tc:
                line['vs_torch_cuda'] = value / tc['value']
        if world == 1 and not args.no_cpu_baseline:
            rps, cores, sec = cpu_step_pass(2, 1, with_warp=warp is not None)
            line['cpu_baseline'] = {'value': rps, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                                    'sample': 'the same step (fusion%s fwd+bwd) of 1 DDAD rig through the oracle (reference '
                                              'torch ops), 1 warm-up + 2 timed, %.1f s per rig' % (' + warp/loss' if warp is not None else '', sec)}
            # the reference launcher pins OMP / MKL / NUMEXPR to ONE thread (train.py:17-20): that number too
            rps1, _, sec1 = cpu_step_pass(1, 0, threads=1, with_warp=warp is not None)
            line['cpu_baseline']['one_thread'] = {'value': rps1, 'unit': UNIT, 'cores': 1,
                                                  'sample': 'the same step, 1 rig, 1 timed step without warm-up, %.1f s' % sec1}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
