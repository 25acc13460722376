"""Headless frames for ``capture_image()`` / ``render('rgb_array')``.

The reference asks PyBullet's OpenGL renderer for a 1920x1080 RGB image of the drone over the ground plane
(``envs/base.py:520-544`` capture_image, ``envs/base.py:379-416`` render, used once per control step by
``IWPGAlgorithm.val_once``, ``algs/iwpg/iwpg.py:513-536``).  Graphics is unavailable on this box, so the frames
come from the CUDA rasteriser in ``csrc/cf_raster.cu`` through the C ABI (``cf_raster_*`` in ``include/cf_abi.h``):
same camera numbers, same image shape and dtype, pixels of our own renderer (parity with Bullet's OpenGL output is
UNPINNED and stays so).

Host side here: the camera basis from the reference's yaw / pitch / distance / fov numbers, the triangle soup of the
drone (``proxy_mesh`` = the URDF's collision shapes plus arms, or STL files of the user's asset directory) and the
``Rasteriser`` handle.  Nothing in this module renders on the CPU.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import struct
from typing import Optional, Sequence

import numpy as np

from . import abi

F = np.float32


class Camera(C.Structure):
    """``cf_camera`` of include/cf_abi.h."""
    _fields_ = [('width', C.c_int32), ('height', C.c_int32), ('follow', C.c_int32),
                ('target', C.c_float * 3), ('eye_offset', C.c_float * 3),
                ('right', C.c_float * 3), ('up', C.c_float * 3), ('fwd', C.c_float * 3),
                ('tan_x', C.c_float), ('tan_y', C.c_float), ('near_plane', C.c_float),
                ('light', C.c_float * 3), ('ambient', C.c_float), ('diffuse', C.c_float),
                ('background', C.c_float * 3), ('ground_a', C.c_float * 3), ('ground_b', C.c_float * 3),
                ('inv_checker', C.c_float), ('ground_half', C.c_float)]


def make_camera(width: int = 1920, height: int = 1080, yaw: float = 10.0, pitch: float = -20.0, distance: float = 1.8,
                fov: float = 60.0, near: float = 0.1, follow: bool = True, target: Sequence[float] = (0.0, 0.0, 0.0),
                checker: float = 1.0, ground_half: float = 15.0, roll: float = 0.0) -> Camera:
    """The camera of ``computeViewMatrixFromYawPitchRoll(target, distance, yaw, pitch, roll=0, upAxisIndex=2)`` +
    ``computeProjectionMatrixFOV(fov, width / height, near, far)``: the eye sits at
    ``target + Rz(yaw) Rx(pitch) (0, -distance, 0)`` and looks at the target with +z up; ``fov`` is the vertical
    field of view.  The far plane (100 m) is not modelled: the ground square ends at ``ground_half``."""
    assert width > 0 and width % 4 == 0 and height > 0
    y, p = math.radians(yaw), math.radians(pitch)
    off = np.array([distance * math.cos(p) * math.sin(y), -distance * math.cos(p) * math.cos(y), -distance * math.sin(p)])
    fwd = -off / np.linalg.norm(off)
    up0 = np.array([math.sin(p) * math.sin(y), -math.sin(p) * math.cos(y), math.cos(p)])     # Rz Rx (0, 0, 1)
    right = np.cross(fwd, up0)
    right[2] = 0.0                           # roll 0: the camera's x axis is horizontal (exactly, not to rounding)
    right /= np.linalg.norm(right)
    up = np.cross(right, fwd)
    if roll != 0.0:                          # the reference always passes roll=0
        cr, sr = math.cos(math.radians(roll)), math.sin(math.radians(roll))
        right, up = cr * right + sr * up, cr * up - sr * right
    light = np.array([0.35, -0.35, 0.87])
    light /= np.linalg.norm(light)
    cam = Camera()
    cam.width, cam.height, cam.follow = width, height, int(follow)
    for name, v in (('target', target), ('eye_offset', off), ('right', right), ('up', up), ('fwd', fwd), ('light', light),
                    ('background', (0.78, 0.82, 0.90)), ('ground_a', (0.95, 0.95, 0.95)), ('ground_b', (0.55, 0.68, 0.88))):
        setattr(cam, name, (C.c_float * 3)(*[float(x) for x in v]))
    cam.tan_y = math.tan(math.radians(fov) / 2.0)
    cam.tan_x = math.tan(math.radians(fov) / 2.0) * (float(width) / height)
    cam.near_plane, cam.ambient, cam.diffuse = near, 0.45, 0.55
    cam.inv_checker, cam.ground_half = 1.0 / checker, ground_half
    return cam


def capture_camera(width: int = 1920, height: int = 1080) -> Camera:
    """``capture_image()``: follows the drone, distance 1.8, yaw 10, pitch -20 (envs/base.py:526-536)."""
    return make_camera(width, height, yaw=10.0, pitch=-20.0, distance=1.8, follow=True)


def render_camera(width: int = 1920, height: int = 1080) -> Camera:
    """``render('rgb_array')``: looks at the origin, distance 2.0, yaw 10, pitch -50 (envs/base.py:395-406)."""
    return make_camera(width, height, yaw=10.0, pitch=-50.0, distance=2.0, follow=False)


# ------------------------------------------------------------------------------------------ triangle soups
def box(center, size, yaw: float = 0.0) -> np.ndarray:
    """12 triangles of an axis-aligned box turned by ``yaw`` about z; returns [12, 3, 3] float64."""
    hx, hy, hz = (0.5 * s for s in size)
    v = np.array([[sx * hx, sy * hy, sz * hz] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)])
    c, s = math.cos(yaw), math.sin(yaw)
    v = v @ np.array([[c, s, 0.0], [-s, c, 0.0], [0.0, 0.0, 1.0]]) + np.asarray(center, dtype=np.float64)
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    return np.array([[v[q[0]], v[q[i]], v[q[i + 1]]] for q in quads for i in (1, 2)])


def cylinder(center, radius: float, length: float, segments: int = 24) -> np.ndarray:
    """A z-aligned cylinder (URDF convention): side quads + two caps, [4 * segments, 3, 3] float64."""
    c = np.asarray(center, dtype=np.float64)
    ang = np.linspace(0.0, 2.0 * math.pi, segments + 1)
    ring = np.stack([radius * np.cos(ang), radius * np.sin(ang), np.zeros_like(ang)], axis=1)
    top, bot = ring + c + [0, 0, 0.5 * length], ring + c - [0, 0, 0.5 * length]
    ct, cb = c + [0, 0, 0.5 * length], c - [0, 0, 0.5 * length]
    tris = []
    for i in range(segments):
        tris += [[bot[i], bot[i + 1], top[i + 1]], [bot[i], top[i + 1], top[i]], [ct, top[i], top[i + 1]], [cb, bot[i + 1], bot[i]]]
    return np.array(tris)


def proxy_mesh():
    """Stand-in for the URDF's visual meshes (``assets/cf21x_bullet.urdf``: cf2withoutRotors.dae + four prop STLs),
    built from numbers in that URDF: propeller joints at (+-0.028, +-0.028, 0.0108) carrying the collision cylinders
    (radius 0.015, length 0.005), arms from the centre to the joints, a body block, motor stems.  Returns
    (tris [T, 3, 3] float32, rgb [T, 3] float32, bounding radius)."""
    parts, cols = [], []

    def add(t, colour):
        parts.append(t)
        cols.append(np.tile(np.asarray(colour, dtype=np.float64), (len(t), 1)))

    grey, dark, green = (0.5, 0.5, 0.5), (0.2, 0.2, 0.2), (0.15, 0.45, 0.2)
    add(box((0, 0, 0.0), (0.030, 0.030, 0.006)), green)                  # PCB
    add(box((0, 0, -0.007), (0.032, 0.020, 0.008)), grey)                # battery
    arm = math.hypot(0.028, 0.028)
    for sx, sy in ((1, -1), (-1, -1), (-1, 1), (1, 1)):                  # prop1..prop4 joints of the URDF
        add(box((0.5 * sx * 0.028, 0.5 * sy * 0.028, 0.0), (arm, 0.004, 0.003), yaw=math.atan2(sy, sx)), grey)
        add(cylinder((sx * 0.028, sy * 0.028, 0.0040), 0.0035, 0.0136, segments=12), grey)
        add(cylinder((sx * 0.028, sy * 0.028, 0.0108), 0.015, 0.005, segments=24), dark)
    tris, rgb = np.concatenate(parts), np.concatenate(cols)
    return tris.astype(F), rgb.astype(F), float(np.linalg.norm(tris.reshape(-1, 3), axis=1).max() * 1.001)


def load_stl(path: str, scale=(1.0, 1.0, 1.0), offset=(0.0, 0.0, 0.0)) -> np.ndarray:
    """Triangles of a binary or ASCII STL file (the URDF's ``meshes/cw_prop.stl`` / ``ccw_prop.stl``) as [T, 3, 3]."""
    with open(path, 'rb') as f:
        raw = f.read()
    n = struct.unpack_from('<I', raw, 80)[0] if len(raw) >= 84 else -1
    if n >= 0 and len(raw) == 84 + 50 * n:
        rec = np.frombuffer(raw, dtype=np.dtype([('n', '<f4', 3), ('v', '<f4', (3, 3)), ('a', '<u2')]), count=n, offset=84)
        tris = rec['v'].astype(np.float64)
    else:
        vs = [ln.split()[1:4] for ln in raw.decode('ascii', 'replace').splitlines() if ln.strip().startswith('vertex')]
        tris = np.array(vs, dtype=np.float64).reshape(-1, 3, 3)
    return tris * np.asarray(scale, dtype=np.float64) + np.asarray(offset, dtype=np.float64)


def mesh_with_stl_props(asset_dir: str):
    """The proxy body with the propellers replaced by the STL meshes found under ``asset_dir/meshes`` (the user's
    copy of the reference's assets; nothing of it is shipped here).  The Collada body mesh is not parsed."""
    tris, rgb, _ = proxy_mesh()
    keep = ~np.all(np.isclose(rgb, 0.2), axis=1)
    parts, cols = [tris[keep].astype(np.float64)], [rgb[keep].astype(np.float64)]
    for (sx, sy), name in zip(((1, -1), (-1, -1), (-1, 1), (1, 1)), ('ccw_prop.stl', 'cw_prop.stl', 'ccw_prop.stl', 'cw_prop.stl')):
        t = load_stl(os.path.join(asset_dir, 'meshes', name), offset=(sx * 0.028, sy * 0.028, 0.0108))
        parts.append(t)
        cols.append(np.tile([0.2, 0.2, 0.2], (len(t), 1)))
    tris, rgb = np.concatenate(parts), np.concatenate(cols)
    return tris.astype(F), rgb.astype(F), float(np.linalg.norm(tris.reshape(-1, 3), axis=1).max() * 1.001)


# ------------------------------------------------------------------------------------------ the handle
class Rasteriser:
    """``cf_raster`` handle: a triangle soup resident on one GPU; ``render`` writes uint8 frames into a device tensor."""

    def __init__(self, device: Optional[int] = None, mesh=None, camera: Optional[Camera] = None):
        import torch
        self._L = abi.lib()
        self.device = torch.cuda.current_device() if device is None else int(device)
        tris, rgb, radius = mesh if mesh is not None else proxy_mesh()
        tris = np.ascontiguousarray(tris, dtype=F).reshape(-1, 3, 3)
        rgb = np.ascontiguousarray(rgb, dtype=F).reshape(-1, 3)
        assert len(tris) == len(rgb)
        self.n_tris, self.camera = len(tris), camera if camera is not None else capture_camera()
        self._h = C.c_void_p()
        abi.check(self._L.cf_raster_create(C.byref(self._h), self.device, tris.ctypes.data_as(C.c_void_p),
                                           rgb.ctypes.data_as(C.c_void_p), self.n_tris, float(radius)))

    def render(self, poses, out=None, camera: Optional[Camera] = None):
        """poses: float32 CUDA tensor [F, 7] (position xyz, quaternion xyzw) -> uint8 CUDA tensor [F, H, W, 3];
        asynchronous on torch's current stream."""
        import torch
        cam = camera if camera is not None else self.camera
        assert poses.is_cuda and poses.dtype == torch.float32 and poses.dim() == 2 and poses.shape[1] == 7
        poses = poses.contiguous()
        n = poses.shape[0]
        if out is None:
            out = torch.empty((n, cam.height, cam.width, 3), dtype=torch.uint8, device=poses.device)
        assert out.is_cuda and out.dtype == torch.uint8 and out.is_contiguous() and out.numel() == n * cam.height * cam.width * 3
        st = torch.cuda.current_stream(poses.device).cuda_stream
        abi.check(self._L.cf_raster_render_dev(self._h, C.byref(cam), C.c_void_p(poses.data_ptr()), n,
                                               C.c_void_p(out.data_ptr()), C.c_void_p(st)))
        return out

    def render_host(self, poses: np.ndarray, camera: Optional[Camera] = None) -> np.ndarray:
        """NumPy in, NumPy out (host buffers through ``cf_raster_render_host``)."""
        cam = camera if camera is not None else self.camera
        poses = np.ascontiguousarray(poses, dtype=F).reshape(-1, 7)
        out = np.empty((len(poses), cam.height, cam.width, 3), np.uint8)
        abi.check(self._L.cf_raster_render_host(self._h, C.byref(cam), poses.ctypes.data_as(C.c_void_p), len(poses),
                                                out.ctypes.data_as(C.c_void_p)))
        return out

    def close(self) -> None:
        if getattr(self, '_h', None) is not None and self._h:
            self._L.cf_raster_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
