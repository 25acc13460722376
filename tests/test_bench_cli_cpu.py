"""CPU: bench.py's command line parses and prints its help (an unescaped '%' in a help string once broke --help), and the
options the driver's contract names exist."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bench_help_renders_and_names_the_contract_options():
    p = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--help'], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    for opt in ('--gpus', '--steps', '--warmup', '--impl', '--workload', '--e2e-depth', '--e2e-separate'):
        assert opt in p.stdout, opt
