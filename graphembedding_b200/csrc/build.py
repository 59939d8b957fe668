"""Build libsgnn.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).  Every .cu/.cpp
in this directory is one translation unit; units compile in parallel and are linked into one
shared library."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
# SGNN_DEFINES="-DSGNN_TILE_TRACE" SGNN_LIB_SUFFIX=_trace builds an instrumented copy next to the product library
SUFFIX = os.environ.get('SGNN_LIB_SUFFIX', '')
LIB = os.path.join(HERE, 'libsgnn%s.so' % SUFFIX)
OBJ = os.path.join(HERE, 'build' + SUFFIX)
DEP_SUFFIX = ('.cuh', '.cu', '.cpp', '.h')
FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
         '--expt-relaxed-constexpr', '--extended-lambda', '-Xcompiler', '-fPIC'] + os.environ.get('SGNN_DEFINES', '').split()


def _deps():
    deps = [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith(DEP_SUFFIX)]
    deps.append(os.path.join(HERE, '..', '..', 'include', 'sgnn.h'))
    return [d for d in deps if os.path.exists(d)]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    headers = [d for d in _deps() if d.endswith(('.cuh', '.h'))]
    srcs = sorted(f for f in os.listdir(HERE) if f.endswith(('.cu', '.cpp')))
    if not force and not _stale(LIB, _deps()):
        return LIB
    nvcc = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
    os.makedirs(OBJ, exist_ok=True)

    def compile_one(src):
        obj = os.path.join(OBJ, src + '.o')
        if not force and not _stale(obj, headers + [os.path.join(HERE, src)]):
            return obj, ''
        cmd = [nvcc] + FLAGS + (['-Xptxas', '-v'] if verbose else []) + ['-c', src, '-o', obj]
        r = subprocess.run(cmd, cwd=HERE, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError('nvcc failed on ' + src)
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(compile_one, srcs))
    if verbose:
        for _, log in results:
            sys.stderr.write(log)
    cmd = [nvcc, '-shared', '-o', LIB] + [o for o, _ in results] + ['-lcuda']
    r = subprocess.run(cmd, cwd=HERE, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError('nvcc failed linking libsgnn.so')
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
