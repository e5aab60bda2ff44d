"""Builds ``libelephas_b200.so`` (all hand-written sm_100a kernels) in-tree with nvcc.

``python -m elephas_b200.ops.build`` or ``elephas_b200.ops.build.build()``. The library has no
torch / pybind dependency - it is a plain C ABI (``include/elephas_b200.h``) loaded with ctypes -
so a full rebuild takes well under a minute and the resulting ``.so`` travels with the repository
snapshot to the GPU box. Objects are cached per source file and rebuilt when the source, a header
or the flags change.
"""
import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CSRC = os.path.join(ROOT, 'csrc')
INCLUDE = os.path.join(ROOT, 'include')
LIB_DIR = os.path.join(ROOT, 'elephas_b200', '_lib')
LIB_PATH = os.path.join(LIB_DIR, 'libelephas_b200.so')
OBJ_DIR = os.path.join(ROOT, 'build', 'obj')

SOURCES = ['core.cu', 'gemm_sm100.cu', 'elementwise.cu', 'loss_kernels.cu', 'optim_kernels.cu', 'comm_kernels.cu', 'ps_kernels.cu', 'conv_sm100.cu', 'conv_aux.cu', 'pool_kernels.cu', 'vmm.cu',
           'host_stage.cu', 'host_cast.cpp', 'narrow_head.cu', 'mlp_fused.cu']
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17', '-Xcompiler', '-fPIC',
              '-DNDEBUG']


def _nvcc():
    exe = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found (needed to build libelephas_b200.so)")
    return exe


def _sources():
    return [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def _digest(path):
    h = hashlib.sha256()
    h.update(' '.join(NVCC_FLAGS).encode())
    with open(path, 'rb') as fh:
        h.update(fh.read())
    for d in (CSRC, INCLUDE):
        for name in sorted(os.listdir(d)):
            if name.endswith(('.cuh', '.h')):
                with open(os.path.join(d, name), 'rb') as fh:
                    h.update(fh.read())
    return h.hexdigest()


def _compile_one(src, verbose):
    path = os.path.join(CSRC, src)
    obj = os.path.join(OBJ_DIR, src.replace('.cu', '.o').replace('.cpp', '.o'))
    stamp = obj + '.sha'
    digest = _digest(path)
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == digest:
        return obj, False
    if src.endswith('.cpp'):
        # plain host code (explicit SIMD with per-function target attributes): the host compiler alone
        cmd = [shutil.which('g++') or 'g++', '-O3', '-std=c++17', '-fPIC', '-DNDEBUG', '-I', INCLUDE, '-c', path, '-o', obj]
    else:
        cmd = [_nvcc()] + NVCC_FLAGS + ['-I', INCLUDE, '-c', path, '-o', obj]
    if verbose:
        print(' '.join(cmd), flush=True)
    subprocess.run(cmd, check=True)
    with open(stamp, 'w') as fh:
        fh.write(digest)
    return obj, True


def build(verbose=False, force=False):
    """Compile every CUDA source for sm_100a and link ``libelephas_b200.so``. Returns its path."""
    os.makedirs(OBJ_DIR, exist_ok=True)
    os.makedirs(LIB_DIR, exist_ok=True)
    if force:
        for f in os.listdir(OBJ_DIR):
            os.remove(os.path.join(OBJ_DIR, f))
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as pool:
        results = list(pool.map(lambda s: _compile_one(s, verbose), _sources()))
    objs = [o for o, _ in results]
    if any(changed for _, changed in results) or not os.path.exists(LIB_PATH):
        cmd = [_nvcc(), '-shared', '-o', LIB_PATH] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a', '-lcudart']
        if verbose:
            print(' '.join(cmd), flush=True)
        subprocess.run(cmd, check=True)
    return LIB_PATH


def is_built():
    return os.path.exists(LIB_PATH)


if __name__ == '__main__':
    print(build(verbose='-q' not in sys.argv, force='--force' in sys.argv))
