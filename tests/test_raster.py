"""Frames for capture_image() / render('rgb_array').

CPU: camera basis from the reference's numbers, the proxy mesh, the struct layout handed to the C ABI, and the
NumPy oracle's picture (drone at the image centre, ground below the horizon, depth order).
GPU: csrc/cf_raster.cu through the C ABI against the oracle - byte for byte (float32, same operation order).
Parity with PyBullet's OpenGL frames is UNPINNED (graphics is unavailable on this box).
"""
import ctypes as C
import math

import numpy as np
import pytest

import raster_oracle as O
from disturbance_crazyfile_simulation_b200 import raster

F = np.float32


def _quat(roll, pitch, yaw):
    cr, sr, cp, sp, cy, sy = (f(a / 2) for a in (roll, pitch, yaw) for f in (math.cos, math.sin))
    return [sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy, cr * cp * cy + sr * sp * sy]


def _poses():
    return np.array([[0, 0, 1, 0, 0, 0, 1],
                     [0.4, -0.3, 0.6] + _quat(0.5, -0.3, 1.0),
                     [-1.2, 2.0, 0.03] + _quat(0.0, 0.0, 2.5),          # on the ground: props against the plane
                     [0.1, 0.1, 0.005] + _quat(1.2, 0.4, 0.0),          # partly below the plane
                     [14.9, -14.9, 2.0] + _quat(-0.7, 0.9, -2.0)],      # at the corner of the ground square
                    dtype=F)


def test_camera_matches_reference_numbers():
    cam = raster.capture_camera()
    assert C.sizeof(cam) == 148 and (cam.width, cam.height, cam.follow) == (1920, 1080, 1)
    off = np.array(cam.eye_offset[:])
    assert abs(np.linalg.norm(off) - 1.8) < 1e-6                                  # distance 1.8
    assert abs(math.degrees(math.asin(off[2] / 1.8)) - 20.0) < 1e-4              # pitch -20: eye above the target
    assert abs(math.degrees(math.atan2(off[0], -off[1])) - 10.0) < 1e-4          # yaw 10 about +z
    r, u, f = (np.array(v[:]) for v in (cam.right, cam.up, cam.fwd))
    for a, b, want in ((r, r, 1), (u, u, 1), (f, f, 1), (r, u, 0), (r, f, 0), (u, f, 0)):
        assert abs(a @ b - want) < 1e-6
    assert np.allclose(np.cross(r, u), -f, atol=1e-6) and abs(r[2]) < 1e-7 and u[2] > 0    # roll 0, z up
    assert abs(cam.tan_y - math.tan(math.radians(30))) < 1e-7 and abs(cam.tan_x / cam.tan_y - 1920 / 1080) < 1e-6
    rc = raster.render_camera()
    assert rc.follow == 0 and abs(np.linalg.norm(rc.eye_offset[:]) - 2.0) < 1e-6
    assert abs(math.degrees(math.asin(rc.eye_offset[2] / 2.0)) - 50.0) < 1e-4


def test_proxy_mesh_from_urdf_numbers():
    tris, rgb, rad = raster.proxy_mesh()
    assert tris.dtype == F and tris.shape[1:] == (3, 3) and rgb.shape == (len(tris), 3)
    pts = tris.reshape(-1, 3)
    assert rad >= np.linalg.norm(pts, axis=1).max() and rad < 0.07
    dark = np.all(np.isclose(rgb, 0.2), axis=1)                 # the four propeller discs at the URDF's joint origins
    centres = np.round(tris[dark].reshape(-1, 3)[:, :2].reshape(4, -1, 2).mean(axis=1), 3)
    assert sorted(map(tuple, np.abs(centres))) == [(0.028, 0.028)] * 4
    assert abs(tris[dark].reshape(-1, 3)[:, 2].mean() - 0.0108) < 1e-4
    b = raster.box((0, 0, 0), (2, 4, 6))
    assert b.shape == (12, 3, 3) and np.allclose(np.abs(b).max(axis=(0, 1)), (1, 2, 3))


def test_stl_loader_binary_and_ascii(tmp_path):
    tri = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 0, 1], [1, 0, 1], [0, 1, 1]]], dtype='<f4')
    with open(tmp_path / 'b.stl', 'wb') as f:
        f.write(b'\0' * 80 + np.uint32(2).tobytes())
        for t in tri:
            f.write(np.zeros(3, '<f4').tobytes() + t.tobytes() + b'\0\0')
    with open(tmp_path / 'a.stl', 'w') as f:
        f.write('solid x\n')
        for t in tri:
            f.write('facet normal 0 0 1\nouter loop\n' + ''.join('vertex %g %g %g\n' % tuple(v) for v in t) + 'endloop\nendfacet\n')
        f.write('endsolid x\n')
    for name in ('a.stl', 'b.stl'):
        got = raster.load_stl(str(tmp_path / name), scale=(2, 2, 2), offset=(1, 0, 0))
        assert np.allclose(got, tri * 2 + [1, 0, 0])


def test_oracle_picture():
    tris, rgb, _ = raster.proxy_mesh()
    cam = raster.make_camera(256, 144)
    img = O.render_frame(cam, tris, rgb, [0, 0, 1, 0, 0, 0, 1])
    empty = O.render_frame(cam, tris[:0], rgb[:0], [0, 0, 1, 0, 0, 0, 1])
    assert img.shape == (144, 256, 3) and img.dtype == np.uint8
    ys, xs = np.nonzero(np.any(img != empty, axis=2))
    assert len(ys) > 8 and abs(xs.mean() - 127.5) < 2 and abs(ys.mean() - 71.5) < 2      # the camera follows the drone
    assert tuple(empty[0, 0]) == (199, 209, 230)                                           # sky above the horizon
    ground = {tuple(c) for c in empty[-1]}
    assert len(ground) == 2 and (199, 209, 230) not in ground                              # two checker colours below
    # depth order: a drone pushed below the plane disappears, a crate around it hides it
    under = O.render_frame(cam, tris, rgb, [0, 0, -0.5, 0, 0, 0, 1])
    assert np.array_equal(under, O.render_frame(cam, tris[:0], rgb[:0], [0, 0, -0.5, 0, 0, 0, 1]))
    lid = raster.box((0, 0, 0), (0.3, 0.3, 0.3)).astype(F)
    both = O.render_frame(cam, np.concatenate([tris, lid]), np.concatenate([rgb, np.tile(F([1, 0, 0]), (12, 1))]), [0, 0, 1, 0, 0, 0, 1])
    only = O.render_frame(cam, lid, np.tile(F([1, 0, 0]), (12, 1)), [0, 0, 1, 0, 0, 0, 1])
    assert np.array_equal(both, only)


def test_abi_exports_raster_symbols():
    from disturbance_crazyfile_simulation_b200 import abi
    L = abi.lib()
    for name in ('cf_raster_create', 'cf_raster_destroy', 'cf_raster_render_dev', 'cf_raster_render_host', 'cf_get_poses_dev'):
        assert hasattr(L, name), name


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize('w,h', [(256, 144), (132, 50), (1920, 1080)])
def test_gpu_frames_byte_exact(w, h):
    import torch
    tris, rgb, rad = raster.proxy_mesh()
    poses = _poses() if w < 1920 else _poses()[:3]
    cams = [raster.make_camera(w, h), raster.make_camera(w, h, pitch=-50.0, distance=2.0, follow=False)]
    if w < 1920:
        cams.append(raster.make_camera(w, h, roll=25.0, yaw=-140.0))          # rolled camera: per-pixel ground depth
    for cam in cams:
        r = raster.Rasteriser(0, (tris, rgb, rad), cam)
        got = r.render(torch.from_numpy(poses).cuda()).cpu().numpy()
        want = O.render(cam, tris, rgb, poses)
        assert got.shape == want.shape == (len(poses), h, w, 3)
        bad = np.nonzero(np.any(got != want, axis=3))
        assert len(bad[0]) == 0, (len(bad[0]), [b[:5] for b in bad])
        assert np.array_equal(r.render_host(poses[:2]), want[:2])
        r.close()


@pytest.mark.gpu
def test_gpu_many_triangles_and_empty_mesh():
    """More triangles than one 256-chunk per block, coincident surfaces (equal depth -> lower index), no mesh at all."""
    import torch
    rng = np.random.default_rng(5)
    tris, rgb, _ = raster.proxy_mesh()
    extra = (rng.uniform(-0.06, 0.06, size=(1500, 3, 3))).astype(F)
    t2 = np.concatenate([tris, tris, extra])                                  # the proxy twice: every depth ties
    c2 = np.concatenate([rgb, 1 - rgb, rng.uniform(0, 1, size=(1500, 3)).astype(F)])
    cam = raster.make_camera(320, 180, distance=0.5)
    poses = _poses()[:2]
    for T, Cc in ((t2, c2), (tris[:0], rgb[:0])):
        r = raster.Rasteriser(0, (T, Cc, 0.11), cam)
        got = r.render(torch.from_numpy(poses).cuda()).cpu().numpy()
        assert np.array_equal(got, O.render(cam, T, Cc, poses))
        r.close()


@pytest.mark.gpu
def test_env_capture_image_follows_the_drone():
    import torch
    from disturbance_crazyfile_simulation_b200 import make
    env = make('DroneHoverBulletFreeEnvWithAdversary-v0', num_envs=64, seed=3)
    env.reset()
    img = env.capture_image(5)
    assert img.shape == (1080, 1920, 3) and img.dtype == np.uint8
    st = env.get_state()
    pose = np.concatenate([st[5, 0:3], st[5, 4:8]])
    assert np.array_equal(env.poses(torch.tensor([5], dtype=torch.int32)).cpu().numpy()[0], pose)
    tris, rgb, _ = raster.proxy_mesh()
    assert np.array_equal(img, O.render_frame(raster.capture_camera(), tris, rgb, pose))
    frames = env.capture_images(torch.arange(8, dtype=torch.int32))
    assert frames.shape == (8, 1080, 1920, 3) and frames.is_cuda and np.array_equal(frames[5].cpu().numpy(), img)
    rgb_arr = env.render('rgb_array')
    assert rgb_arr.shape == (1080, 1920, 3)
    with pytest.raises(NotImplementedError):
        env.render('human')
    env.close()


@pytest.mark.gpu
def test_trainer_validation_writes_rows_and_gifs(tmp_path):
    """IWPGAlgorithm.validate / val_once (iwpg.py:512-568): deterministic-policy episodes on a twin env, one row each in
    validation_results.csv, one GIF each with a capture_image() frame per control step."""
    import csv
    from PIL import Image
    from disturbance_crazyfile_simulation_b200 import make
    from disturbance_crazyfile_simulation_b200.run_logger import RunLogger
    from disturbance_crazyfile_simulation_b200.trainer import PPOTrainer
    env = make('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', num_envs=256, seed=3, return_terminal_obs=True)
    lg = RunLogger(str(tmp_path))
    cam = raster.make_camera(128, 72)
    t = PPOTrainer(env, rollout_len=8, epochs=4, train_pi_iterations=2, train_v_iterations=1, num_mini_batches=2, seed=0, logger=lg,
                   val_freq=1, val_kwargs=dict(num_evaluations=3, camera=cam, max_steps=40))
    info = t.learn_one_epoch()
    val = info['validation']
    assert val['epoch'] == 0 and len(val['returns']) == 3      # the reference's 0-based epoch index and all(1 <= s <= 40 for s in val['steps'])
    assert all(r < 0 for r in val['returns']) and len(val['gifs']) == 3
    with open(tmp_path / 'validation_results.csv') as f:
        rows = list(csv.reader(f))
    assert rows[0] == ['Training_Epoch', 'disturbance_level', 'Steps', 'Returns', 'Costs'] and len(rows) == 4
    for i, path in enumerate(val['gifs']):
        assert path.endswith(f"0-{i}-{val['disturbance_level'][i]}.gif")
        with Image.open(path) as im:
            assert im.size == (128, 72) and im.n_frames == int(val['steps'][i])
        assert float(rows[1 + i][2]) == val['steps'][i] and abs(float(rows[1 + i][3]) - val['returns'][i]) < 1e-4
    # the twin env does not disturb the training env: a second epoch still runs, and a second validation appends rows
    info = t.learn_one_epoch()
    assert info['validation']['epoch'] == 1
    lg.close()
    with open(tmp_path / 'validation_results.csv') as f:
        assert len(list(csv.reader(f))) == 7
    env.close()
