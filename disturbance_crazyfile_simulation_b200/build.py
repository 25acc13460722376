"""Build ``csrc/libcf_b200.so`` (the C-ABI CUDA library) in-tree for sm_100a with nvcc."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
OUT = os.path.join(CSRC, 'libcf_b200.so')
SOURCES = ['cf_abi.cu', 'cf_kernels.cu', 'cf_policy.cu', 'cf_ppo.cu', 'cf_hj_solver.cu', 'cf_raster.cu']
DEPS = SOURCES + ['cf_device.cuh', 'cf_types.h', 'cf_kernels.h', os.path.join('..', '..', 'include', 'cf_abi.h')]
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-Xcompiler', '-fPIC', '-shared', '-cudart', 'static']


def nvcc_path() -> str:
    for cand in (os.environ.get('NVCC'), shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError('nvcc not found')


def is_stale() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in DEPS)


def build(force: bool = False, verbose: bool = False, extra_flags=(), out: str = OUT) -> str:
    """``extra_flags`` / ``out`` build an experimental variant (e.g. ``-DCF_TILE=64``) next to the product library;
    ``CF_LIB_PATH`` makes :mod:`abi` load it (tools/ only)."""
    if not force and not extra_flags and out == OUT and not is_stale():
        return OUT
    cmd = [nvcc_path()] + NVCC_FLAGS + list(extra_flags) + (['-Xptxas', '-v'] if verbose else []) + \
          [os.path.join(CSRC, s) for s in SOURCES] + ['-o', out]
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError('nvcc failed:\n' + ' '.join(cmd) + '\n' + p.stdout + p.stderr)
    if verbose:
        print(p.stderr)
    return out


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
