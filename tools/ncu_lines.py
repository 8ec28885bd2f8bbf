"""Stall samples / executed instructions of one kernel per SOURCE LINE: joins the SASS page of an .ncu-rep (--import-source
on) with nvdisasm -g line annotations of the shipped cubin (same instruction order).
   python tools/ncu_lines.py gpurun_out/prof_step.ncu-rep warp_loss_bwd warp_loss_bwd_kernelILb1 [top [kernel-name regex]]"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def sass_lines(obj, mangled_frag):
    """[(file, line)] per instruction of the function whose mangled name contains mangled_frag"""
    d = tempfile.mkdtemp()
    subprocess.run(['cuobjdump', '-xelf', obj + '.sm_100a.cubin', os.path.join(ROOT, 'vfdepth_b200', 'lib', 'libvfdepth_b200.so')],
                   cwd=d, capture_output=True)
    out = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(d, obj + '.sm_100a.cubin')], capture_output=True, text=True).stdout
    res, on, cur = [], False, ('?', 0)
    for ln in out.splitlines():
        if ln.startswith('//---') and '.text.' in ln:
            on = mangled_frag in ln
            continue
        if not on:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        if re.match(r'\s+/\*[0-9a-f]{4,5}\*/', ln):
            res.append(cur)
    return res


def main(rep, obj, frag, top=40, kernel=None):
    lines = sass_lines(obj, frag)
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + (kernel or obj.split('.')[0])],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, data, nk = None, [], 0
    for r in rows:
        if r and r[0] == 'Kernel Name':
            nk += 1
            if nk > 1:
                break
            continue
        if r and r[0] == 'Address':
            hdr = {h: i for i, h in enumerate(r)}
            continue
        if hdr and len(r) >= len(hdr):
            data.append(r)
    print('instructions: ncu %d, nvdisasm %d' % (len(data), len(lines)))
    smp, ins = defaultdict(float), defaultdict(float)
    for r, loc in zip(data, lines):
        smp[loc] += float(r[hdr['# Samples']] or 0)
        ins[loc] += float(r[hdr['Instructions Executed']] or 0)
    ts, ti = sum(smp.values()), sum(ins.values())
    srcs = {}
    for loc in sorted(smp, key=lambda k: -smp[k])[:int(top)]:
        f = loc[0]
        if f not in srcs:
            p = os.path.join(ROOT, 'vfdepth_b200', 'csrc', f)
            srcs[f] = open(p).read().splitlines() if os.path.isfile(p) else []
        text = srcs[f][loc[1] - 1].strip()[:95] if 0 < loc[1] <= len(srcs[f]) else ''
        print('%5.2f%% smp %5.2f%% ins  %s:%d  %s' % (100 * smp[loc] / ts, 100 * ins[loc] / ti, f, loc[1], text))


if __name__ == '__main__':
    main(*sys.argv[1:6])
