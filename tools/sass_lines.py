"""sass_lines.py - where a kernel's SASS bytes come from: instruction counts per source function / line.

    python tools/sass_lines.py [--so PATH] [--kernel SUBSTR] [--by func|line|outer] [--top N]

Reads the library with cuobjdump / nvdisasm (-g: line info from -lineinfo builds) here on the CPU box.  `--by outer`
attributes every instruction to the line of the KERNEL body it was inlined into (the outermost frame), which is the
view that shows what an inlined device function costs per call site.
"""
import argparse
import collections
import os
import re
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, '..', 'disturbance_crazyfile_simulation_b200', 'csrc', 'libcf_b200.so')


def disasm(so, cubin_substr='cf_kernels'):
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run(['cuobjdump', '-xelf', 'all', os.path.abspath(so)], cwd=tmp, capture_output=True, text=True)
        cubins = [f for f in os.listdir(tmp) if f.endswith('.cubin') and f.startswith(cubin_substr + '.')]
        if not cubins:
            cubins = [f for f in os.listdir(tmp) if f.endswith('.cubin')]
        out = subprocess.run(['nvdisasm', '-gi', os.path.join(tmp, cubins[0])], capture_output=True, text=True).stdout
    return out


LINE = re.compile(r'//## File "([^"]+)", line (\d+)(.*)')
INL = re.compile(r'inlined at "([^"]+)", line (\d+)')
INSTR = re.compile(r'^\s+/\*[0-9a-f]+\*/\s+(\S+)')


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--so', default=SO)
    ap.add_argument('--kernel', default='k_stepILb1ELi1')
    ap.add_argument('--cubin', default='cf_kernels')
    ap.add_argument('--by', default='outer', choices=['line', 'outer', 'file'])
    ap.add_argument('--top', type=int, default=60)
    a = ap.parse_args()
    text = disasm(a.so, a.cubin)
    cur_fn, loc, chain = None, None, ()
    counts = collections.Counter()
    total = 0
    for ln in text.splitlines():
        if ln.startswith('.text.'):
            cur_fn = ln[6:].rstrip(':')
            continue
        if cur_fn is None or a.kernel not in cur_fn:
            continue
        m = LINE.search(ln)
        if m:
            loc = (os.path.basename(m.group(1)), int(m.group(2)))
            chain = tuple((os.path.basename(f), int(l)) for f, l in INL.findall(m.group(3)))
            continue
        m = INSTR.match(ln)
        if m and loc:
            total += 1
            if a.by == 'line':
                key = loc
            elif a.by == 'file':
                key = (loc[0], 0)
            else:
                key = chain[-1] if chain else loc
            counts[key] += 1
    print(f'{a.kernel}: {total} instructions = {total * 16 / 1024:.1f} KB')
    for (f, l), c in counts.most_common(a.top):
        print(f'{c:6d} {c * 16 / 1024:6.1f} KB  {f}:{l}')


if __name__ == '__main__':
    main()
