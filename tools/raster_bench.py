"""Device time of the frame renderer: F capture_image() frames (1920x1080x3 uint8) per launch; HBM-write roofline."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from disturbance_crazyfile_simulation_b200 import raster  # noqa: E402


def measure(frames=64, reps=20, mesh=None):
    r = raster.Rasteriser(0, mesh)
    rng = np.random.default_rng(0)
    poses = np.zeros((frames, 7), np.float32)
    poses[:, :3] = rng.uniform([-1, -1, 0.3], [1, 1, 1.5], size=(frames, 3))
    q = rng.standard_normal((frames, 4)) * [0.2, 0.2, 1, 1]
    poses[:, 3:] = q / np.linalg.norm(q, axis=1, keepdims=True)
    p = torch.from_numpy(poses).cuda()
    out = torch.empty((frames, 1080, 1920, 3), dtype=torch.uint8, device='cuda')
    for _ in range(3):
        r.render(p, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        r.render(p, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    gb = frames * 1080 * 1920 * 3 / 1e9
    return {'frames_per_launch': frames, 'triangles': r.n_tris, 'ms_per_launch': ms, 'frames_per_s': frames / ms * 1e3,
            'write_GBps': gb / ms * 1e3, 'bytes_per_frame': 1080 * 1920 * 3}


if __name__ == '__main__':
    res = [measure(64), measure(8), measure(1, reps=200)]
    print(json.dumps(res))
