"""Host<->device copy bandwidth of pinned buffers on this box, per NUMA node the buffer was first touched on, alone and
with a concurrent copy in the other direction -- the bound of bench.py's e2e number (755 MB in / 236 MB out per step).

    python tools/pcie_probe.py
"""
import glob
import os
import subprocess

import torch


def cpus_of(node):
    with open('/sys/devices/system/node/node%d/cpulist' % node) as f:
        out = []
        for part in f.read().strip().split(','):
            if '-' in part:
                a, b = part.split('-')
                out.extend(range(int(a), int(b) + 1))
            elif part:
                out.append(int(part))
        return out


def bw(fn, nbytes, iters=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return nbytes * iters / e0.elapsed_time(e1) / 1e6


def main():
    dev = torch.device('cuda:0')
    torch.zeros(1, device=dev)
    try:
        print(subprocess.run(['nvidia-smi', 'topo', '-m'], capture_output=True, text=True, timeout=20).stdout)
        q = subprocess.run(['nvidia-smi', '--query-gpu=pci.bus_id,pcie.link.gen.current,pcie.link.width.current',
                            '--format=csv,noheader'], capture_output=True, text=True, timeout=20).stdout.strip()
        print('gpu pci:', q)
        bus = q.split(',')[0].strip().lower()
        if bus.startswith('0000'):
            bus = bus[4:]
        for p in glob.glob('/sys/bus/pci/devices/*%s/numa_node' % bus):
            print(p, open(p).read().strip())
    except Exception as e:  # noqa: BLE001
        print('nvidia-smi query failed:', e)
    nodes = sorted(int(p.rsplit('node', 1)[1]) for p in glob.glob('/sys/devices/system/node/node[0-9]*'))
    print('numa nodes:', nodes, 'cpus:', os.cpu_count(), 'affinity:', len(os.sched_getaffinity(0)))
    n = 512 << 20
    d_buf = torch.empty(n, dtype=torch.uint8, device=dev)
    d_buf2 = torch.empty(n, dtype=torch.uint8, device=dev)
    s2 = torch.cuda.Stream()
    all_cpus = os.sched_getaffinity(0)
    for node in [None] + nodes:
        if node is not None:
            cp = [c for c in cpus_of(node) if c in all_cpus]
            if not cp:
                continue
            os.sched_setaffinity(0, cp)
        h = torch.empty(n, dtype=torch.uint8).pin_memory()
        h.fill_(1)
        h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
        h2.fill_(1)
        h2d = bw(lambda: d_buf.copy_(h, non_blocking=True), n)
        d2h = bw(lambda: h2.copy_(d_buf2, non_blocking=True), n)

        def both():
            d_buf.copy_(h, non_blocking=True)
            with torch.cuda.stream(s2):
                h2.copy_(d_buf2, non_blocking=True)
            torch.cuda.current_stream().wait_stream(s2)
        bi = bw(both, n)
        print('first touch on node %s: H2D %.1f GB/s, D2H %.1f GB/s, concurrent %.1f GB/s each way' %
              (node, h2d, d2h, bi), flush=True)
        os.sched_setaffinity(0, all_cpus)


if __name__ == '__main__':
    main()
