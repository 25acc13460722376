"""build_variants.py - experimental builds of the C-ABI library with extra nvcc flags, next to the product .so.

    python tools/build_variants.py name1="-DCF_TILE=64" name2="-DCF_TILE=32 -DFOO" ...

writes csrc/variants/libcf_<name>.so; run a tool against one with CF_LIB_PATH=<path>.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from disturbance_crazyfile_simulation_b200 import build as B  # noqa: E402

out_dir = os.path.join(B.CSRC, 'variants')
os.makedirs(out_dir, exist_ok=True)
for spec in sys.argv[1:]:
    name, flags = spec.split('=', 1)
    path = os.path.join(out_dir, f'libcf_{name}.so')
    B.build(force=True, extra_flags=flags.split(), out=path, verbose='-v' in sys.argv)
    print(path)
