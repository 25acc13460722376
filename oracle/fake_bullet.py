"""oracle-A, part 1: a float64 stand-in for the PyBullet surface the hover path touches.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is imported by the product
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline
legs may use it, and only as the checker.

PARITY UNPINNED at this boundary: PyBullet 3.21 (environment.yaml:22,178 of the
reference) is where the rigid-body arithmetic of the reference lives, it is not
installable here (no network) and no reference test pins its numbers
(SURVEY.md section 4).  This file restates, from Bullet's published btMultiBody
algorithm, the behaviour of exactly those client calls that the reference makes on
the hover path, so that the reference's *own Python* (envs/base.py, hover_free.py,
hover.py, agents.py, physics.py, sensors.py, distur_gener.py) can run unmodified on
top of it (see ``oracle/ref_harness.py``).

Model assumptions (each is a switch of :class:`BulletModel`; the defaults are what
Bullet's btMultiBody does to the best of our reading, none verifiable here):

B1  semi-implicit Euler: v += a*dt (forces evaluated at the old state), then
    x += v_new*dt and the base quaternion is advanced by the exponential map of
    omega_world*dt (btMultiBody::stepPositionsMultiDof), then re-normalised.
B2  gyroscopic term  omega x (J omega)  in the body frame (m_useGyroTerm = true).
B3  btMultiBody's default damping  m_linearDamping = m_angularDamping = 0.04:
    force  -m v (k + k|v|),  torque  -J w (k + k|w|)  (body frame).
B4  external forces/torques are cleared after every stepSimulation.
B5  LINK_FRAME force at posObj = 0 acts at the link's centre of mass; the four prop
    links sit at (+-0.028, +-0.028, 0.0108) (cf21x_bullet.urdf:59,85,111,136).
B6  torque/force on the zero-mass centre_of_mass link (index 4) passes wholly to the
    base (Featherstone with D = 0 -> invD = 0); its frame is the base frame.
B7  prop links (1e-9 kg, 1e-9 kg m^2) and their velocity motors are ignored
    (setJointMotorControl2 is a no-op): a <= 3e-4 relative effect on Jxx/Jyy that a
    real PyBullet run would show and this model does not.
B8  no contacts (plane, 10x10 room walls): episodes end at z < 0.2 first.
B9  getEulerFromQuaternion with PyBullet's gimbal branch at |sin pitch| >= 0.99999.
"""
from __future__ import annotations

import math
import xml.etree.ElementTree as ET
from dataclasses import dataclass

import numpy as np

# ---- module-level constants mirrored from pybullet (values are irrelevant) ----
GUI = 1
DIRECT = 2
LINK_FRAME = 1
WORLD_FRAME = 2
VELOCITY_CONTROL = 0
URDF_USE_INERTIA_FROM_FILE = 2
COV_ENABLE_GUI = 1
COV_ENABLE_RENDERING = 7
GEOM_SPHERE = 2
ER_BULLET_HARDWARE_OPENGL = 131072
ER_TINY_RENDERER = 65536

ANGULAR_MOTION_THRESHOLD = 0.5 * (math.pi / 2.0)


@dataclass
class BulletModel:
    gyroscopic: bool = True          # B2
    linear_damping: float = 0.04     # B3
    angular_damping: float = 0.04    # B3
    gravity_z: float = -9.81


def getQuaternionFromEuler(rpy):
    """pybullet.getQuaternionFromEuler: [x, y, z, w], normalised."""
    roll, pitch, yaw = (float(v) for v in rpy)
    phi, the, psi = roll / 2.0, pitch / 2.0, yaw / 2.0
    q = [
        math.sin(phi) * math.cos(the) * math.cos(psi) - math.cos(phi) * math.sin(the) * math.sin(psi),
        math.cos(phi) * math.sin(the) * math.cos(psi) + math.sin(phi) * math.cos(the) * math.sin(psi),
        math.cos(phi) * math.cos(the) * math.sin(psi) - math.sin(phi) * math.sin(the) * math.cos(psi),
        math.cos(phi) * math.cos(the) * math.cos(psi) + math.sin(phi) * math.sin(the) * math.sin(psi),
    ]
    n = math.sqrt(sum(c * c for c in q))
    return tuple(c / n for c in q)


def getMatrixFromQuaternion(quat):
    """btMatrix3x3::setRotation, row-major 9-tuple."""
    x, y, z, w = (float(v) for v in quat)
    d = x * x + y * y + z * z + w * w
    s = 2.0 / d
    xs, ys, zs = x * s, y * s, z * s
    wx, wy, wz = w * xs, w * ys, w * zs
    xx, xy, xz = x * xs, x * ys, x * zs
    yy, yz, zz = y * ys, y * zs, z * zs
    return (1.0 - (yy + zz), xy - wz, xz + wy,
            xy + wz, 1.0 - (xx + zz), yz - wx,
            xz - wy, yz + wx, 1.0 - (xx + yy))


def getEulerFromQuaternion(quat):
    """pybullet.getEulerFromQuaternion (B9)."""
    x, y, z, w = (float(v) for v in quat)
    sqx, sqy, sqz, squ = x * x, y * y, z * z, w * w
    sarg = -2.0 * (x * z - w * y)
    if sarg <= -0.99999:
        return (0.0, -0.5 * math.pi, 2.0 * math.atan2(x, -y))
    if sarg >= 0.99999:
        return (0.0, 0.5 * math.pi, 2.0 * math.atan2(-x, y))
    return (math.atan2(2.0 * (y * z + w * x), squ - sqx - sqy + sqz),
            math.asin(sarg),
            math.atan2(2.0 * (x * y + w * z), squ + sqx - sqy - sqz))


def loadURDF(*args, **kwargs):  # module-level pb.loadURDF(room_10x10.urdf): scenery, no dynamics (B8)
    return 99


def _quat_mul(a, b):
    ax, ay, az, aw = a
    bx, by, bz, bw = b
    return np.array([
        aw * bx + ax * bw + ay * bz - az * by,
        aw * by - ax * bz + ay * bw + az * bx,
        aw * bz + ax * by - ay * bx + az * bw,
        aw * bw - ax * bx - ay * by - az * bz,
    ])


class BulletClient:
    """Stand-in for pybullet_utils.bullet_client.BulletClient (one drone body)."""

    # attribute-style constants the reference reads off the client
    COV_ENABLE_RENDERING = COV_ENABLE_RENDERING
    COV_ENABLE_GUI = COV_ENABLE_GUI
    GEOM_SPHERE = GEOM_SPHERE

    model = BulletModel()  # class-level default; tests may replace per instance

    def __init__(self, connection_mode=None):
        self.model = BulletModel(**vars(type(self).model))
        self.dt = 1.0 / 240.0
        self.gravity = np.array([0.0, 0.0, self.model.gravity_z])
        self._next_id = 0
        self.drone_id = None
        self.link_pos = np.zeros((5, 3))
        self.mass = 0.0
        self.inertia = np.ones(3)
        self.pos = np.zeros(3)
        self.quat = np.array([0.0, 0.0, 0.0, 1.0])
        self.vel = np.zeros(3)
        self.omega = np.zeros(3)          # world frame, like Bullet
        self._force_w = np.zeros(3)       # accumulated external force, world
        self._torque_w = np.zeros(3)      # accumulated external torque, world
        self._saved = {}
        self.n_steps = 0

    # ------------------------------------------------------------ no-ops
    def _noop(self, *a, **k):
        return 0

    setAdditionalSearchPath = configureDebugVisualizer = _noop
    resetDebugVisualizerCamera = addUserDebugLine = _noop
    createVisualShape = createMultiBody = _noop
    setJointMotorControl2 = setDefaultContactERP = disconnect = _noop  # B7

    def computeViewMatrixFromYawPitchRoll(self, *a, **k):
        raise NotImplementedError("graphics: unavailable on this box")

    computeProjectionMatrixFOV = getCameraImage = computeViewMatrixFromYawPitchRoll

    # ------------------------------------------------------------ setup
    def setPhysicsEngineParameter(self, fixedTimeStep=None, **kwargs):
        if fixedTimeStep is not None:
            self.dt = float(fixedTimeStep)

    def setGravity(self, x, y, z):
        self.gravity = np.array([x, y, z], dtype=np.float64)

    def loadURDF(self, fileName, basePosition=(0, 0, 0), baseOrientation=(0, 0, 0, 1), **kwargs):
        self._next_id += 1
        if str(fileName).endswith('cf21x_bullet.urdf'):
            root = ET.parse(fileName).getroot()
            base = root.find("link[@name='base_link']/inertial")
            self.mass = float(base.find('mass').attrib['value'])
            ine = base.find('inertia').attrib
            self.inertia = np.array([float(ine['ixx']), float(ine['iyy']), float(ine['izz'])])
            k = 0
            for joint in root.findall('joint'):
                org = joint.find('origin')
                xyz = [0.0, 0.0, 0.0] if org is None else [float(v) for v in org.attrib['xyz'].split()]
                self.link_pos[k] = xyz
                k += 1
            assert k == 5
            self.pos = np.array(basePosition, dtype=np.float64)
            self.quat = np.array(baseOrientation, dtype=np.float64)
            self.drone_id = self._next_id
        return self._next_id

    def changeDynamics(self, bodyUniqueId, linkIndex, mass=None, localInertiaDiagonal=None, **kwargs):
        assert bodyUniqueId == self.drone_id and linkIndex == -1
        if mass is not None:
            self.mass = float(mass)
        if localInertiaDiagonal is not None:
            self.inertia = np.array(localInertiaDiagonal, dtype=np.float64)

    def saveState(self):
        sid = len(self._saved)
        self._saved[sid] = (self.pos.copy(), self.quat.copy(), self.vel.copy(), self.omega.copy())
        return sid

    def restoreState(self, stateId):
        # restores kinematic state only; mass/inertia set by changeDynamics are not part of it
        p, q, v, w = self._saved[stateId]
        self.pos, self.quat, self.vel, self.omega = p.copy(), q.copy(), v.copy(), w.copy()
        self._force_w[:] = 0.0
        self._torque_w[:] = 0.0

    # ------------------------------------------------------------ kinematics
    getQuaternionFromEuler = staticmethod(getQuaternionFromEuler)
    getMatrixFromQuaternion = staticmethod(getMatrixFromQuaternion)
    getEulerFromQuaternion = staticmethod(getEulerFromQuaternion)

    def _R(self):
        return np.array(getMatrixFromQuaternion(self.quat)).reshape(3, 3)

    def getBasePositionAndOrientation(self, bodyUniqueId):
        return tuple(self.pos.tolist()), tuple(self.quat.tolist())

    def getBaseVelocity(self, bodyUniqueId):
        return tuple(self.vel.tolist()), tuple(self.omega.tolist())

    def resetBasePositionAndOrientation(self, bodyUniqueId, posObj, ornObj):
        self.pos = np.array(posObj, dtype=np.float64)
        self.quat = np.array(ornObj, dtype=np.float64)
        self.vel = np.zeros(3)
        self.omega = np.zeros(3)

    def resetBaseVelocity(self, bodyUniqueId, linearVelocity=None, angularVelocity=None):
        if linearVelocity is not None:
            self.vel = np.array(linearVelocity, dtype=np.float64)
        if angularVelocity is not None:
            self.omega = np.array(angularVelocity, dtype=np.float64)

    def getLinkStates(self, bodyUniqueId, linkIndices, computeLinkVelocity=0, computeForwardKinematics=0):
        R = self._R()
        out = []
        for i in linkIndices:
            p = self.pos + R @ self.link_pos[i]
            out.append((tuple(p.tolist()), tuple(self.quat.tolist()), (0, 0, 0), (0, 0, 0, 1),
                        tuple(p.tolist()), tuple(self.quat.tolist()), tuple(self.vel.tolist()),
                        tuple(self.omega.tolist())))
        return out

    # ------------------------------------------------------------ forces (B4-B6)
    def applyExternalForce(self, objectUniqueId, linkIndex, forceObj, posObj, flags):
        assert objectUniqueId == self.drone_id
        f = np.array(forceObj, dtype=np.float64)
        p = np.array(posObj, dtype=np.float64)
        R = self._R()
        r_link = np.zeros(3) if linkIndex < 0 else self.link_pos[linkIndex]
        if flags == LINK_FRAME:
            f_w = R @ f                       # every link frame is parallel to the base frame about z
            arm_w = R @ (r_link + p)
        else:
            f_w = f
            arm_w = p - self.pos
        self._force_w += f_w
        self._torque_w += np.cross(arm_w, f_w)

    def applyExternalTorque(self, objectUniqueId, linkIndex, torqueObj, flags):
        assert objectUniqueId == self.drone_id
        t = np.array(torqueObj, dtype=np.float64)
        self._torque_w += (self._R() @ t) if flags == LINK_FRAME else t

    # ------------------------------------------------------------ integration (B1-B3)
    def stepSimulation(self):
        m, J, dt, mdl = self.mass, self.inertia, self.dt, self.model
        R = self._R()
        w_b = R.T @ self.omega
        v_b = R.T @ self.vel
        # zero-acceleration bias force in the body frame (btMultiBody ABA, base link)
        tau_b = R.T @ self._torque_w
        if mdl.gyroscopic:
            tau_b = tau_b - np.cross(w_b, J * w_b)
        ka = mdl.angular_damping
        tau_b = tau_b - J * w_b * (ka + ka * np.linalg.norm(w_b))
        kl = mdl.linear_damping
        f_b = R.T @ (self._force_w + m * self.gravity) - m * v_b * (kl + kl * np.linalg.norm(v_b))
        self.omega = self.omega + dt * (R @ (tau_b / J))
        self.vel = self.vel + dt * (R @ (f_b / m))
        # positions with the NEW velocities
        self.pos = self.pos + dt * self.vel
        w = self.omega
        ang = float(np.linalg.norm(w))
        if ang * dt > ANGULAR_MOTION_THRESHOLD:
            ang = 0.5 * (math.pi / 2.0) / dt
        if ang < 0.001:
            axis = w * (0.5 * dt - (dt * dt * dt) * 0.020833333333 * ang * ang)
        else:
            axis = w * (math.sin(0.5 * ang * dt) / ang)
        dq = np.array([axis[0], axis[1], axis[2], math.cos(ang * dt * 0.5)])
        q = _quat_mul(dq, self.quat)
        self.quat = q / math.sqrt(float(q @ q))
        self._force_w[:] = 0.0   # B4
        self._torque_w[:] = 0.0
        self.n_steps += 1
