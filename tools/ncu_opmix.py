"""Instruction mix of one kernel from an .ncu-rep captured with --import-source on: executed warp instructions and stall
samples per SASS mnemonic.   python tools/ncu_opmix.py gpurun_out/prof_step.ncu-rep warp_stats_kernel [top]"""
import csv
import io
import subprocess
import sys
from collections import defaultdict


def main(rep, kernel, top=25):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', kernel], capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = None
    ins, smp = defaultdict(float), defaultdict(float)
    n_k = 0
    for r in rows:
        if r and r[0] == 'Kernel Name':
            n_k += 1
            if n_k > 1:
                break
            continue
        if r and r[0] == 'Address':
            hdr = {h: i for i, h in enumerate(r)}
            continue
        if hdr is None or len(r) < len(hdr):
            continue
        sass = r[hdr['Source']].strip()
        parts = sass.split()
        if not parts:
            continue
        op = parts[1] if parts[0].startswith('@') and len(parts) > 1 else parts[0]
        op = op.split('.')[0] + ('.' + op.split('.')[1] if op.startswith(('MUFU', 'LDG', 'STG', 'LDS', 'STS', 'RED', 'ATOM')) and '.' in op else '')
        ins[op] += float(r[hdr['Instructions Executed']] or 0)
        smp[op] += float(r[hdr['# Samples']] or 0)
    ti, ts = sum(ins.values()), sum(smp.values())
    print('%s: %.3g warp instructions, %d samples' % (kernel, ti, ts))
    for op in sorted(ins, key=lambda k: -ins[k])[:int(top)]:
        print('  %-14s %6.2f%% of instructions  %6.2f%% of samples' % (op, 100 * ins[op] / ti, 100 * smp[op] / max(ts, 1)))


if __name__ == '__main__':
    main(*sys.argv[1:4])
