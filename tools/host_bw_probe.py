"""Host-side cost of assembling [n][34] observation rows from compact halves ("send fewer bytes").

Sending only the new half of every row (68 B instead of 136 B per env-step over PCIe) means the HOST must build the row the
caller reads: row[0:17] = previous half, row[17:34] = new half.  This probe measures how fast this box's cores do exactly
that (tools/host_bw_probe.c: one memcpy per half, k pthreads, buffers rotated like the bench's shards) next to the
pinned-memory D2H rate of full rows and of halves, to decide whether the compact path can beat the plain 136 B/env DMA.
Usage: python tools/host_bw_probe.py   (prints one JSON object)"""
import json
import os
import subprocess
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    exe = os.path.join(tempfile.gettempdir(), 'host_bw_probe')
    subprocess.run(['gcc', '-O3', '-pthread', os.path.join(HERE, 'host_bw_probe.c'), '-o', exe], check=True)
    res = {'host_assembly': json.loads(subprocess.run([exe], capture_output=True, text=True, check=True).stdout),
           'cpus': os.cpu_count()}
    try:
        import torch
        if torch.cuda.is_available():
            n = 65536
            for name, w in (('full_rows_136B', 34), ('new_halves_68B', 17)):
                d = torch.empty((n, w), dtype=torch.float32, device='cuda')
                h = torch.empty((n, w), dtype=torch.float32).pin_memory()
                h.copy_(d); torch.cuda.synchronize()
                t0 = time.perf_counter()
                for _ in range(100):
                    h.copy_(d, non_blocking=True)
                torch.cuda.synchronize()
                dt = (time.perf_counter() - t0) / 100
                res[f'd2h_{name}'] = {'ms': 1e3 * dt, 'GB_per_s': n * w * 4 / dt / 1e9, 'env_steps_per_s': n / dt}
    except Exception as e:  # noqa: BLE001
        res['d2h_error'] = str(e)
    print(json.dumps(res))


if __name__ == '__main__':
    main()
