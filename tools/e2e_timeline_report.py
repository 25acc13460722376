"""Reads the CSV of CF_PROBE_E2E_TIMELINE (csrc/cf_abi.cu) and prints, for the steady part of the run, the mean duration of each
phase of a host-pointer step and the idle time of the download engine between consecutive steps."""
import sys
import numpy as np

a = np.loadtxt(sys.argv[1], delimiter=',', skiprows=1)
a = a[len(a) // 4:]                                   # steady state
a = a[np.argsort(a[:, 4])]                            # by the end of the observation download
h2d, kern, obs, small = a[:, 2] - a[:, 1], a[:, 3] - a[:, 2], a[:, 4] - a[:, 3], a[:, 5] - a[:, 4]
period = np.diff(a[:, 5])
gap = a[1:, 3] - a[:-1, 5]                            # this step's kernels done -> previous step's last download done (negative: it waited)
print(f'steps {len(a)}  period {1e3 * period.mean():.1f} us (median {1e3 * np.median(period):.1f})')
for name, v in (('h2d event span', h2d), ('kernels event span', kern), ('obs d2h event span', obs), ('small d2h event span', small)):
    print(f'  {name:22s} mean {1e3 * v.mean():7.1f} us  median {1e3 * np.median(v):7.1f}  p90 {1e3 * np.percentile(v, 90):7.1f}')
print(f'  kernels-done minus previous-download-done: mean {1e3 * gap.mean():.1f} us (negative = the download engine never waits for compute)')
print(f'  obs bytes / obs span: {8912896 / (np.median(obs) * 1e-3) / 1e9:.1f} GB/s')
