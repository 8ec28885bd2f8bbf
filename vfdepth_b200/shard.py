"""Multi-GPU plumbing of the hot path: whole rigs per rank, no data-path collective.

Every function on the path is independent per batch element, while the 6 cameras of a rig interact (overlap counts,
neighbour warps), so the only sensible partition is by rig -- the reference's DistributedSampler
(/root/reference/models/vfdepth.py:119-125).  The one collective the reference's training step has (the DDP gradient
all-reduce, models/vfdepth.py:70) follows the encoder/decoder backward, not these kernels; what lives here is the
index partition, the per-rank seeding (utils/ddp.py:24-29) and the max-over-ranks timing reduction bench.py uses.
Works with any torch.distributed backend (NCCL on the GPU box, gloo in the CPU tests).
"""
import os

import torch
import torch.distributed as dist


def dist_env():
    """(rank, world_size, local_rank) from the torchrun environment; (0, 1, 0) when launched plainly."""
    return int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))


def bind_to_gpu_numa(local_rank):
    """Pin this process to the host cores NVML reports as local to GPU ``local_rank`` (its NUMA node), so that the
    pinned staging buffers allocated afterwards are first-touched on the node the GPU's PCIe root hangs off.  torchrun
    does not bind its workers; with 8 ranks feeding 8 GPUs from one node's memory the host->device copies of the
    end-to-end step share one socket's memory controllers and the inter-socket link.  Returns a short description
    (recorded in the bench line) and never raises: without NVML, or when the local cores are not in this process's
    allowed set, nothing changes."""
    try:
        import pynvml
        import torch.cuda as tc
        pynvml.nvmlInit()
        try:
            bus = getattr(tc.get_device_properties(local_rank), 'pci_bus_id', None)
            dom = getattr(tc.get_device_properties(local_rank), 'pci_domain_id', 0)
            dev_id = getattr(tc.get_device_properties(local_rank), 'pci_device_id', 0)
            if bus is not None:
                handle = pynvml.nvmlDeviceGetHandleByPciBusId(('%08x:%02x:%02x.0' % (dom, bus, dev_id)).encode())
            else:
                handle = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
            words = (os.cpu_count() + 63) // 64
            mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
        finally:
            pynvml.nvmlShutdown()
        local = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        cpus = sorted(local & allowed)
        if not cpus:
            return 'unbound (no GPU-local core in the allowed set of %d)' % len(allowed)
        if len(cpus) == len(allowed):
            return 'unbound (all %d allowed cores are GPU-local)' % len(allowed)
        os.sched_setaffinity(0, cpus)
        return 'bound to %d GPU-local cores (%d-%d) of %d' % (len(cpus), cpus[0], cpus[-1], len(allowed))
    except Exception as e:      # noqa: BLE001 -- affinity is an optimisation, never a failure
        return 'unbound (%s: %s)' % (type(e).__name__, str(e)[:60])


def rank_rigs(n_rigs, rank, world, drop_last=True):
    """indices of the rigs rank ``rank`` owns: r, r+world, r+2*world, ... (DistributedSampler without shuffling).
    With ``drop_last`` every rank gets the same count (the reference's loaders use drop_last=True, vfdepth.py:114)."""
    if not 0 <= rank < world:
        raise ValueError('rank %d outside world of %d' % (rank, world))
    per = n_rigs // world if drop_last else -(-n_rigs // world)
    return [i for i in range(rank, n_rigs, world)][:per]


def rank_seed(base, rank, rig=0):
    """seed of rig ``rig`` on rank ``rank`` (1234 + 1000*rank + rig index)."""
    return int(base) + 1000 * int(rank) + int(rig)


def max_over_ranks(value, device=None):
    """device-timed milliseconds -> the slowest rank's (what bench.py reports); identity without a process group."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value, device=None):
    """e.g. rigs processed by all ranks."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def job_throughput(rigs_this_rank, ms_this_rank, device=None):
    """whole-job rigs/s = rigs of all ranks / slowest rank's device time."""
    ms = max_over_ranks(ms_this_rank, device)
    return sum_over_ranks(rigs_this_rank, device) / (ms * 1e-3), ms
