"""Build oracle/libcf_oracle.so (test infrastructure).  Tries the system gcc with OpenMP first,
then without OpenMP (single-threaded oracle) if libgomp is not usable."""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, 'cf_oracle.c')
OUT = os.path.join(HERE, 'libcf_oracle.so')
BASE = ['-O2', '-ffp-contract=off', '-fno-fast-math', '-fPIC', '-shared', '-Wall',
        '-Wno-unused-variable', '-Wno-unused-const-variable']


def build(force: bool = False) -> str:
    hdr = os.path.join(HERE, '..', 'include', 'cf_abi.h')
    if (not force and os.path.exists(OUT)
            and os.path.getmtime(OUT) >= max(os.path.getmtime(SRC), os.path.getmtime(hdr))):
        return OUT
    errors = []
    for cc in ('/usr/bin/gcc', shutil.which('gcc') or 'gcc', 'cc'):
        for omp in (['-fopenmp'], []):
            cmd = [cc] + BASE + omp + ['-o', OUT, SRC, '-lm']
            p = subprocess.run(cmd, capture_output=True, text=True)
            if p.returncode == 0:
                return OUT
            errors.append(' '.join(cmd) + '\n' + p.stderr)
    raise RuntimeError('could not build the CPU oracle:\n' + '\n'.join(errors))


if __name__ == '__main__':
    print(build(force='--force' in sys.argv))
