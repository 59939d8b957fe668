"""Build libsgnn.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, 'libsgnn.so')
SOURCES = ['sgnn_api.cu', 'sgnn_pack.cpp']
HEADERS = ['sgnn_common.cuh', 'sgnn_small.cuh', 'sgnn_allpairs.cuh', 'sgnn_rank.cuh',
           os.path.join('..', '..', 'include', 'sgnn.h')]
EXTRA_GLOB_SUFFIX = ('.cuh', '.cu', '.cpp', '.h')


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith(EXTRA_GLOB_SUFFIX)]
    deps.append(os.path.join(HERE, '..', '..', 'include', 'sgnn.h'))
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
    srcs = [f for f in os.listdir(HERE) if f.endswith('.cu') or f.endswith('.cpp')]
    cmd = [nvcc, '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
           '--expt-relaxed-constexpr', '--extended-lambda', '-Xcompiler', '-fPIC', '-shared',
           '-o', LIB] + sorted(srcs) + ['-lcuda']
    if verbose:
        cmd.insert(1, '-Xptxas')
        cmd.insert(2, '-v')
    r = subprocess.run(cmd, cwd=HERE, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError('nvcc failed building libsgnn.so')
    if verbose:
        sys.stderr.write(r.stderr)
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
