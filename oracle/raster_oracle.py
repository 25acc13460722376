"""raster_oracle.py - CPU restatement (NumPy, float32) of the frame renderer in csrc/cf_raster.cu.

TEST INFRASTRUCTURE ONLY (tests/, smoke) - never imported by the product package.

PARITY UNPINNED against the reference: its frames come from PyBullet's OpenGL renderer
(`getCameraImage(..., renderer=ER_BULLET_HARDWARE_OPENGL)`, envs/base.py:520-544), and graphics is unavailable on
this box.  What this file pins is the CUDA kernel against an independent statement of the same picture: the camera
of computeViewMatrixFromYawPitchRoll / computeProjectionMatrixFOV with the reference's numbers (target = drone
position, distance 1.8, yaw 10, pitch -20, fov 60, near 0.1, 1920x1080), a z = 0 checker ground, the drone as a
flat-shaded two-sided triangle soup, nearest surface per pixel centre.  Written whole-image / whole-mesh with
NumPy float32 arrays in the kernel's operation order (no fused multiply-add), with none of the kernel's machinery:
no pixel blocks, no bounding-sphere culling, no chunking, no compaction.
"""
from __future__ import annotations

import numpy as np

F = np.float32


def _v(x):
    return np.array(list(x), dtype=F)


def _dot(a, b):
    return (a[..., 0] * b[0] + a[..., 1] * b[1]) + a[..., 2] * b[2]


def _u8(v):
    return np.minimum(F(255), np.maximum(F(0), np.floor(v * F(255) + F(0.5)))).astype(np.uint32)


def _pack(rgb):
    c = _u8(np.asarray(rgb, dtype=F))
    return c[..., 0] | (c[..., 1] << 8) | (c[..., 2] << 16)


def rotation(q):
    x, y, z, w = (F(v) for v in q)
    xx, yy, zz, xy, xz, yz, wx, wy, wz = x * x, y * y, z * z, x * y, x * z, y * z, w * x, w * y, w * z
    one, two = F(1), F(2)
    return np.array([[one - two * (yy + zz), two * (xy - wz), two * (xz + wy)],
                     [two * (xy + wz), one - two * (xx + zz), two * (yz - wx)],
                     [two * (xz - wy), two * (yz + wx), one - two * (xx + yy)]], dtype=F)


def render_frame(cam, tris, rgb, pose):
    """One frame [H, W, 3] uint8.  cam: the ctypes Camera (or anything with its fields); tris [T,3,3], rgb [T,3]."""
    W, H = int(cam.width), int(cam.height)
    right, up, fwd, light = _v(cam.right), _v(cam.up), _v(cam.fwd), _v(cam.light)
    tan_x, tan_y, near = F(cam.tan_x), F(cam.tan_y), F(cam.near_plane)
    amb, dif = F(cam.ambient), F(cam.diffuse)
    pose = np.asarray(pose, dtype=F)
    pos, R = pose[:3], rotation(pose[3:7])
    target = pos if cam.follow else _v(cam.target)
    eye = target + _v(cam.eye_offset)
    half_w, half_h = F(W) / F(2), F(H) / F(2)

    # ---- background / ground ----
    pxc = (np.arange(W, dtype=F) + F(0.5))[None, :]
    pyc = (np.arange(H, dtype=F) + F(0.5))[:, None]
    txv = (pxc * (F(2) / F(W)) - F(1)) * tan_x
    tyv = (F(1) - pyc * (F(2) / F(H))) * tan_y
    d = [(fwd[k] + up[k] * tyv) + right[k] * txv for k in range(3)]
    sg = amb + dif * max(light[2], F(0))
    bg = _pack(_v(cam.background))
    ga, gb = _pack(_v(cam.ground_a) * sg), _pack(_v(cam.ground_b) * sg)
    col = np.full((H, W), bg, dtype=np.uint32)
    best = np.zeros((H, W), dtype=F)
    best_id = np.full((H, W), 0x7fffffff, dtype=np.int64)
    if eye[2] > 0:
        down = d[2] < 0
        with np.errstate(divide='ignore', invalid='ignore', over='ignore'):
            t = eye[2] / -d[2]
            hx, hy = eye[0] + t * d[0], eye[1] + t * d[1]
            hit = down & (np.abs(hx) <= F(cam.ground_half)) & (np.abs(hy) <= F(cam.ground_half))
            ix = np.floor(hx * F(cam.inv_checker)).astype(np.int64)
            iy = np.floor(hy * F(cam.inv_checker)).astype(np.int64)
            col = np.where(hit, np.where(((ix + iy) & 1) == 1, gb, ga), col).astype(np.uint32)
            best = np.where(hit, F(1) / t, best).astype(F)

    # ---- the drone: set-up of all triangles at once ----
    tris = np.asarray(tris, dtype=F).reshape(-1, 3, 3)
    rgb = np.asarray(rgb, dtype=F).reshape(-1, 3)
    T = len(tris)
    if T:
        wv = np.stack([pos[k] + _dot(tris, R[k]) for k in range(3)], axis=-1)           # [T,3 verts,3]
        pc = wv - eye
        xc, yc, zc = _dot(pc, right), _dot(pc, up), _dot(pc, fwd)                        # [T,3]
        with np.errstate(divide='ignore', invalid='ignore', over='ignore'):
            sx = ((xc / zc) / tan_x + F(1)) * half_w
            sy = (F(1) - (yc / zc) / tan_y) * half_h
            iz = F(1) / zc
            area = (sx[:, 2] - sx[:, 0]) * (sy[:, 1] - sy[:, 0]) - (sy[:, 2] - sy[:, 0]) * (sx[:, 1] - sx[:, 0])
            ok = np.all(zc >= near, axis=1) & (area != 0)
            e1, e2, ev = wv[:, 1] - wv[:, 0], wv[:, 2] - wv[:, 0], eye - wv[:, 0]
            n = np.stack([e1[:, 1] * e2[:, 2] - e1[:, 2] * e2[:, 1], e1[:, 2] * e2[:, 0] - e1[:, 0] * e2[:, 2],
                          e1[:, 0] * e2[:, 1] - e1[:, 1] * e2[:, 0]], axis=-1)
            ln = np.sqrt((n[:, 0] * n[:, 0] + n[:, 1] * n[:, 1]) + n[:, 2] * n[:, 2])
            ok &= ln > 0
            ndl = _dot(n, light) / ln
            ndl = np.where(((n[:, 0] * ev[:, 0] + n[:, 1] * ev[:, 1]) + n[:, 2] * ev[:, 2]) < 0, -ndl, ndl)
            sh = amb + dif * np.maximum(ndl, F(0))
            packed = _pack(rgb * sh[:, None])
            ia = F(1) / area
        for i in np.nonzero(ok)[0]:
            x0, x1, x2 = sx[i]
            y0, y1, y2 = sy[i]
            lo_x, hi_x = int(max(0, np.floor(min(x0, x1, x2)) - 2)), int(min(W, np.ceil(max(x0, x1, x2)) + 3))
            lo_y, hi_y = int(max(0, np.floor(min(y0, y1, y2)) - 2)), int(min(H, np.ceil(max(y0, y1, y2)) + 3))
            if lo_x >= hi_x or lo_y >= hi_y:
                continue
            px, py = pxc[:, lo_x:hi_x], pyc[lo_y:hi_y, :]
            with np.errstate(invalid='ignore', over='ignore'):
                b0 = ((px - x1) * (y2 - y1) - (py - y1) * (x2 - x1)) * ia[i]
                b1 = ((px - x2) * (y0 - y2) - (py - y2) * (x0 - x2)) * ia[i]
                b2 = ((px - x0) * (y1 - y0) - (py - y0) * (x1 - x0)) * ia[i]
                inside = (b0 >= 0) & (b1 >= 0) & (b2 >= 0)
                izp = (b0 * iz[i, 0] + b1 * iz[i, 1]) + b2 * iz[i, 2]
            bb, bi = best[lo_y:hi_y, lo_x:hi_x], best_id[lo_y:hi_y, lo_x:hi_x]
            win = inside & ((izp > bb) | ((izp == bb) & (i < bi)))
            best[lo_y:hi_y, lo_x:hi_x] = np.where(win, izp, bb)
            best_id[lo_y:hi_y, lo_x:hi_x] = np.where(win, i, bi)
            col[lo_y:hi_y, lo_x:hi_x] = np.where(win, packed[i], col[lo_y:hi_y, lo_x:hi_x])
    out = np.empty((H, W, 3), np.uint8)
    out[..., 0], out[..., 1], out[..., 2] = col & 0xff, (col >> 8) & 0xff, (col >> 16) & 0xff
    return out


def render(cam, tris, rgb, poses):
    return np.stack([render_frame(cam, tris, rgb, p) for p in np.asarray(poses, dtype=F).reshape(-1, 7)])
