"""Per-source-line executed instructions, stall samples and opcode mix of one kernel: joins the SASS page of an .ncu-rep
captured with --import-source on with nvdisasm -g line annotations of the shipped cubin (same instruction order).  What
the round-2 kernel work was steered by (loss set kernels, ray backward, statistics pass).

    python tools/ncu_hot_lines.py <rep.ncu-rep> <cubin stem, e.g. ray_sample> <mangled-name fragment> <source file> [top [file:line,...]]
    e.g. python tools/ncu_hot_lines.py gpurun_out/prof_step.ncu-rep ray_sample bwd_vec_kernelILi64 vfdepth_b200/csrc/ray_sample.cu 30
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import Counter, defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COLS = ['# Samples', 'Instructions Executed', 'stall_long_sb', 'stall_short_sb', 'stall_wait', 'stall_math', 'stall_lg',
        'stall_not_selected', 'stall_mio', 'stall_no_inst', 'stall_branch_resolving', 'stall_barrier']


def sass_lines(stem, frag):
    d = tempfile.mkdtemp()
    lib = os.path.join(ROOT, 'vfdepth_b200', 'lib', 'libvfdepth_b200.so')
    subprocess.run(['cuobjdump', '-xelf', stem, lib], cwd=d, capture_output=True)
    cubin = [f for f in os.listdir(d) if f.endswith('.cubin')][0]
    out = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(d, cubin)], capture_output=True, text=True).stdout
    on, cur, lines, sass = False, ('?', 0), [], []
    for ln in out.splitlines():
        if ln.startswith('//---') and '.text.' in ln:
            on = frag in ln
            continue
        if not on:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r'\s+/\*[0-9a-f]{4,5}\*/\s+(.*?);', ln)
        if m:
            lines.append(cur)
            sass.append(m.group(1))
    return lines, sass


def main(rep, stem, frag, srcfile, top=28, detail=''):
    lines, sass = sass_lines(stem, frag)
    kname = re.sub(r'IL[ib]\d+.*', '', frag)      # mangled template suffix off: ncu matches demangled names
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kname],
                         capture_output=True, text=True).stdout
    hdr, data = None, []
    for r in csv.reader(io.StringIO(out)):
        if r and r[0] == 'Address':
            hdr = {h: i for i, h in enumerate(r)}
            continue
        if hdr and len(r) >= len(hdr):
            data.append(r)
    data = data[:len(lines)]        # first kernel of the report that matches
    agg, tot, opc = defaultdict(Counter), Counter(), Counter()
    for i, r in enumerate(data):
        for c in COLS:
            v = float(r[hdr[c]] or 0)
            agg[lines[i]][c] += v
            tot[c] += v
        t = sass[i].split()
        op = t[1] if t[0].startswith('@') else t[0]
        opc[op.split('.')[0]] += float(r[hdr['Instructions Executed']] or 0)
    print(len(lines), 'SASS instructions;', {c: int(tot[c]) for c in COLS})
    print('opcode mix (M warp instructions):', [(k, round(v / 1e6, 1)) for k, v in opc.most_common(24)])
    src = open(srcfile).read().split('\n')
    base = os.path.basename(srcfile)
    for k, c in sorted(agg.items(), key=lambda x: -x[1]['Instructions Executed'])[:int(top)]:
        print(k, round(c['Instructions Executed'] / 1e6, 1), 'M', 'samples', int(c['# Samples']),
              {cc[6:]: int(c[cc]) for cc in COLS[2:] if c[cc] > 200}, src[k[1] - 1].strip()[:70] if k[0] == base and k[1] <= len(src) else '')
    for want in [w for w in detail.split(',') if w]:
        f, l = want.split(':')
        oc = Counter()
        for i, r in enumerate(data):
            if lines[i] == (f, int(l)):
                t = sass[i].split()
                oc[t[1] if t[0].startswith('@') else t[0]] += float(r[hdr['Instructions Executed']] or 0)
        print(want, [(k, round(v / 1e6, 1)) for k, v in oc.most_common(14)])


if __name__ == '__main__':
    main(*sys.argv[1:7])
