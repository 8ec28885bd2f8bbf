"""Build libvfdepth_b200.so in-tree with nvcc for sm_100a (no torch headers: the library is a plain C ABI).

    python -m vfdepth_b200.build [--force]

The .so is git-ignored but travels to the GPU box with the repo snapshot.
"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, 'lib', 'libvfdepth_b200.so')
NVCC = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17', '--fmad=true',
         '-Xcompiler', '-fPIC', '-Xptxas', '-v']
# kernel-tuning experiments (tools/gpu_*.sh): extra -D defines for one build, e.g. VFD_NVCC_DEFINES="-DVFD_LOSS_THREADS=256";
# part of the source hash, so a library built with them is rebuilt when they change.  The shipped library uses none.
FLAGS += os.environ.get('VFD_NVCC_DEFINES', '').split()


def sources():
    return sorted(glob.glob(os.path.join(CSRC, '*.cu')))


STAMP = LIB + '.srchash'


def _deps():
    return sources() + sorted(glob.glob(os.path.join(CSRC, '*.h')) + glob.glob(os.path.join(CSRC, '*.cuh'))) + \
        [os.path.join(os.path.dirname(HERE), 'include', 'vfdepth_b200.h')]


def source_hash():
    """sha256 over every file the library is built from + the compiler flags (file times do not survive a snapshot
    copy to another box, contents do)"""
    import hashlib
    h = hashlib.sha256(' '.join(FLAGS).encode())
    for p in _deps():
        h.update(os.path.basename(p).encode())
        with open(p, 'rb') as f:
            h.update(f.read())
    return h.hexdigest()


def _stale():
    """True when the library is missing or was built from other sources than the ones in the tree now."""
    if not os.path.isfile(LIB) or not os.path.isfile(STAMP):
        return True
    with open(STAMP) as f:
        return f.read().strip() != source_hash()


def build(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    objs = []
    procs = []
    for src in sources():
        obj = os.path.join(os.path.dirname(LIB), os.path.basename(src)[:-3] + '.o')
        objs.append(obj)
        procs.append((src, subprocess.Popen([NVCC] + FLAGS + ['-c', src, '-o', obj], stdout=subprocess.PIPE,
                                            stderr=subprocess.STDOUT, text=True)))
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError('nvcc failed on %s' % src)
    # (the arch on the link line too: without it nvcc emits an empty sm_52 image and a deprecation warning)
    subprocess.check_call([NVCC, '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', LIB] + objs + ['-lcudart'])
    with open(STAMP, 'w') as f:
        f.write(source_hash() + '\n')
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose=True))
