"""
Builds percivaltts_b200/libpctts.so (the C-ABI declared in include/pctts.h) from csrc/*.cu with nvcc for sm_100a.

In-tree build: the shared object lands next to this file so that it travels with the repository snapshot to the
GPU box; objects go to percivaltts_b200/build/ (git-ignored).  Run as `python -m percivaltts_b200.build [--force]`.
"""
import glob
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, 'csrc')
OBJDIR = os.path.join(HERE, 'build')
LIBPATH = os.path.join(HERE, 'libpctts.so')

NVCC_FLAGS = ['-O3', '-std=c++17', '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo',
              '-Xcompiler', '-fPIC', '-I', os.path.join(ROOT, 'include'), '-I', CSRC]


def _nvcc():
    exe = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    if not os.path.exists(exe):
        raise RuntimeError('nvcc not found: libpctts.so cannot be built')
    return exe


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile every csrc/*.cu and link libpctts.so; returns the library path."""
    nvcc = _nvcc()
    os.makedirs(OBJDIR, exist_ok=True)
    sources = sorted(glob.glob(os.path.join(CSRC, '*.cu')))
    headers = sorted(glob.glob(os.path.join(CSRC, '*.cuh'))) + sorted(glob.glob(os.path.join(ROOT, 'include', '*.h')))
    jobs = []
    for src in sources:
        obj = os.path.join(OBJDIR, os.path.splitext(os.path.basename(src))[0] + '.o')
        if force or _stale(obj, [src] + headers):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        cmd = [nvcc] + NVCC_FLAGS + ['-c', src, '-o', obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        return src, r.returncode, r.stdout + r.stderr

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for src, rc, log in ex.map(compile_one, jobs):
            if verbose or rc != 0:
                sys.stderr.write(log)
            if rc != 0:
                raise RuntimeError('nvcc failed on %s' % src)
    objs = [os.path.join(OBJDIR, os.path.splitext(os.path.basename(s))[0] + '.o') for s in sources]
    if force or jobs or _stale(LIBPATH, objs):
        cmd = [nvcc, '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', LIBPATH] + objs
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError('linking libpctts.so failed')
    return LIBPATH


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose=True))
