"""Per-kernel counts of the SASS mnemonics that prove what a kernel is built from (B200_PROFILING.md):
UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG/UBLKCP = TMA, HMMA = legacy mma.sync, REDG/RED = global
reductions, LDGSTS = cp.async, SYNCS = mbarrier.   python tools/sass_opcodes.py [lib.so] > profiles/r02_sass_opcodes.txt"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, 'vfdepth_b200', 'lib', 'libvfdepth_b200.so')
PATTERNS = [('UTC*MMA', r'\bUTC[A-Z]*MMA'), ('LDTM', r'\bLDTM'), ('STTM', r'\bSTTM'), ('UTMALDG', r'\bUTMALDG'),
            ('UTMASTG', r'\bUTMASTG'), ('UBLKCP', r'\bUBLKCP'), ('UTMAPF', r'\bUTMAPF|UTMACCTL'), ('HMMA', r'\bHMMA'),
            ('RED', r'\bRED[G]?\b|\bRED\.'), ('ATOM', r'\bATOM[GS]?\b|\bATOM\.'), ('LDGSTS', r'\bLDGSTS'),
            ('SYNCS', r'\bSYNCS'), ('FFMA2', r'\bFFMA2|\bFMUL2|\bFADD2')]


def main():
    out = subprocess.run(['cuobjdump', '-sass', LIB], capture_output=True, text=True).stdout
    counts, name, total = {}, None, {}
    for line in out.splitlines():
        m = re.search(r'Function : (\S+)', line)
        if m:
            name = subprocess.run(['c++filt', m.group(1)], capture_output=True, text=True).stdout.strip().split('(')[0]
            counts[name] = {k: 0 for k, _ in PATTERNS}
            total[name] = 0
            continue
        if name is None or not re.match(r'\s+/\*[0-9a-f]{4,5}\*/', line):
            continue
        total[name] += 1
        for k, pat in PATTERNS:
            if re.search(pat, line):
                counts[name][k] += 1
    keys = [k for k, _ in PATTERNS]
    print('%-52s %7s ' % ('kernel (sm_100a SASS of %s)' % os.path.basename(LIB), 'instrs') + ' '.join('%7s' % k for k in keys))
    for n in sorted(counts):
        print('%-52s %7d ' % (n[:52], total[n]) + ' '.join('%7d' % counts[n][k] for k in keys))
    print('%-52s %7d ' % ('TOTAL', sum(total.values())) + ' '.join('%7d' % sum(c[k] for c in counts.values()) for k in keys))


if __name__ == '__main__':
    main()
