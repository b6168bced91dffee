"""Builds libcoslam_b200.so (the C-ABI library, include/coslam_b200.h) in-tree with nvcc for sm_100a.

    python -m coslam_b200.build [--force]

The .so is git-ignored but travels to the GPU box with the gpurun snapshot.  nvcc cross-compiles without a GPU.
"""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
INCLUDE = os.path.join(os.path.dirname(HERE), 'include')
OBJ_DIR = os.path.join(HERE, 'build')
LIB_PATH = os.path.join(HERE, 'libcoslam_b200.so')
SOURCES = ['api.cu', 'render.cu', 'points.cu', 'encoding.cu', 'tc_selftest.cu', 'points_tc.cu', 'optim.cu', 'assemble.cu']
NVCC = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
              '-Xcompiler', '-fPIC', '--expt-relaxed-constexpr',
              '-I', INCLUDE]


def _stamp():
    h = hashlib.sha256()
    for root in (CSRC, INCLUDE):
        for name in sorted(os.listdir(root)):
            with open(os.path.join(root, name), 'rb') as f:
                h.update(name.encode())
                h.update(f.read())
    h.update(' '.join(NVCC_FLAGS).encode())
    return h.hexdigest()


def needs_build():
    stamp_file = os.path.join(OBJ_DIR, 'stamp')
    if not os.path.isfile(LIB_PATH) or not os.path.isfile(stamp_file):
        return True
    with open(stamp_file) as f:
        return f.read().strip() != _stamp()


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB_PATH
    if not os.path.isfile(NVCC):
        raise RuntimeError('nvcc not found at %s: libcoslam_b200.so cannot be built' % NVCC)
    os.makedirs(OBJ_DIR, exist_ok=True)

    def compile_one(src):
        obj = os.path.join(OBJ_DIR, src.replace('.cu', '.o'))
        cmd = [NVCC] + NVCC_FLAGS + (['-Xptxas', '-v'] if verbose else []) + ['-c', os.path.join(CSRC, src), '-o', obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError('nvcc failed on %s:\n%s\n%s' % (src, r.stdout, r.stderr))
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [NVCC, '-shared', '-o', LIB_PATH] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a', '-lcudart']
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError('link failed:\n%s\n%s' % (r.stdout, r.stderr))
    with open(os.path.join(OBJ_DIR, 'stamp'), 'w') as f:
        f.write(_stamp())
    return LIB_PATH


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
