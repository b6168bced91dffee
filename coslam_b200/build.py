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
SOURCES = ['api.cu', 'render.cu', 'points.cu', 'encoding.cu', 'tc_selftest.cu', 'points_tc.cu', 'points_ws.cu', 'cells.cu', 'optim.cu', 'assemble.cu', 'track.cu', 'runtime.cu', 'microbench.cu', 'noalloc.cu', 'track_frame.cu', 'host_rng.cu']
NVCC = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
              '-Xcompiler', '-fPIC', '--expt-relaxed-constexpr',
              '-I', INCLUDE] + os.environ.get('CB200_EXTRA_NVCC_FLAGS', '').split()


def _stamp():
    h = hashlib.sha256()
    for root in (CSRC, INCLUDE):
        for name in sorted(os.listdir(root)):
            with open(os.path.join(root, name), 'rb') as f:
                h.update(name.encode())
                h.update(f.read())
    # flags without the absolute include path: the tree is built in one place and run from another (gpurun snapshot)
    h.update(' '.join(f for f in NVCC_FLAGS if f != INCLUDE).encode())
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
    # one builder at a time (several ranks may import the package at once); the others wait, then find a fresh stamp
    import fcntl
    with open(os.path.join(OBJ_DIR, 'lock'), 'w') as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not needs_build():
                return LIB_PATH
            return _build_locked(verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(verbose):
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
    tmp = LIB_PATH + '.tmp.%d' % os.getpid()
    cmd = [NVCC, '-shared', '-o', tmp] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a', '-lcudart']
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError('link failed:\n%s\n%s' % (r.stdout, r.stderr))
    os.replace(tmp, LIB_PATH)                  # atomic: a concurrent dlopen sees the old or the new library, never a partial one
    with open(os.path.join(OBJ_DIR, 'stamp'), 'w') as f:
        f.write(_stamp())
    return LIB_PATH


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
