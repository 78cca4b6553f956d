"""Stand-in for the third-party `pybullet` module (TEST INFRASTRUCTURE ONLY).

The reference (`/root/reference/gym_pybullet_drones`) depends on `pybullet ^3.2.5`
(`pyproject.toml:20`), which is not installed in this container and is not vendored.
On the `Physics.DYN` path the reference uses pybullet only as

  * a per-body state store (`resetBasePositionAndOrientation` / `getBasePositionAndOrientation`,
    `resetBaseVelocity` / `getBaseVelocity`; `sb3_envs/BaseAviary.py:713-721,1090-1100,1400`), and
  * three quaternion helpers (`getMatrixFromQuaternion`, `getEulerFromQuaternion`,
    `getQuaternionFromEuler`; `sb3_envs/BaseAviary.py:593,618,714,1064`,
    `control/DSLPIDControl.py:144,187,240-241`).

This module restates Bullet's published formulas for those calls (btMatrix3x3::setRotation,
pybullet.c ZYX Euler extraction with its +-0.99999 gimbal guard, btQuaternion::setEulerZYX) so the
reference's own .py files can be imported UNMODIFIED and run headless.  Everything else the
reference touches at import/construct time (connect, loadURDF, createVisualShape, ...) is an id
allocator or a no-op.  Nothing here is shipped or used by the product path.
"""
import math

GUI = 1
DIRECT = 2
GEOM_SPHERE = 2
LINK_FRAME = 1
WORLD_FRAME = 2
URDF_USE_INERTIA_FROM_FILE = 2
ER_TINY_RENDERER = 0
ER_SEGMENTATION_MASK_OBJECT_AND_LINKINDEX = 1
COV_ENABLE_RGB_BUFFER_PREVIEW = 0
COV_ENABLE_DEPTH_BUFFER_PREVIEW = 1
COV_ENABLE_SEGMENTATION_MARK_PREVIEW = 2
STATE_LOGGING_VIDEO_MP4 = 3

_next_client = [0]
_worlds = {}


class _Body:
    __slots__ = ("pos", "quat", "lin", "ang")

    def __init__(self, pos, quat):
        self.pos = tuple(float(v) for v in pos)
        self.quat = _normalised(quat)
        self.lin = (0.0, 0.0, 0.0)
        self.ang = (0.0, 0.0, 0.0)


def _normalised(q):
    # Bullet stores orientation as a btMatrix3x3 and hands back a unit quaternion.
    x, y, z, w = (float(v) for v in q)
    n = math.sqrt(x * x + y * y + z * z + w * w)
    if n == 0.0:
        return (0.0, 0.0, 0.0, 1.0)
    return (x / n, y / n, z / n, w / n)


def _world(cid):
    return _worlds.setdefault(cid, {"bodies": {}, "next": 0})


def _new_body(cid, pos, quat):
    w = _world(cid)
    bid = w["next"]
    w["next"] += 1
    w["bodies"][bid] = _Body(pos, quat)
    return bid


# ---- connection / world management (ids and no-ops) -------------------------------------------
def connect(mode, options=""):
    cid = _next_client[0]
    _next_client[0] += 1
    _world(cid)
    return cid


def disconnect(physicsClientId=0):
    _worlds.pop(physicsClientId, None)


def resetSimulation(physicsClientId=0):
    _worlds[physicsClientId] = {"bodies": {}, "next": 0}


def setGravity(*a, **k):
    pass


def setRealTimeSimulation(*a, **k):
    pass


def setTimeStep(*a, **k):
    pass


def setAdditionalSearchPath(*a, **k):
    pass


def stepSimulation(*a, **k):
    raise RuntimeError("stand-in pybullet has no solver: only Physics.DYN is supported")


def loadURDF(fileName, basePosition=(0, 0, 0), baseOrientation=(0, 0, 0, 1), physicsClientId=0, **k):
    return _new_body(physicsClientId, basePosition, baseOrientation)


def createVisualShape(*a, **k):
    return 0


def createMultiBody(baseMass=0, baseVisualShapeIndex=-1, basePosition=(0, 0, 0),
                    baseOrientation=(0, 0, 0, 1), physicsClientId=0, **k):
    return _new_body(physicsClientId, basePosition, baseOrientation)


def changeDynamics(*a, **k):
    pass


def configureDebugVisualizer(*a, **k):
    pass


# ---- body state store ---------------------------------------------------------------------------
def resetBasePositionAndOrientation(bodyUniqueId, posObj, ornObj, physicsClientId=0):
    b = _world(physicsClientId)["bodies"][bodyUniqueId]
    b.pos = tuple(float(v) for v in posObj)
    b.quat = _normalised(ornObj)


def getBasePositionAndOrientation(bodyUniqueId, physicsClientId=0):
    b = _world(physicsClientId)["bodies"][bodyUniqueId]
    return b.pos, b.quat


def resetBaseVelocity(objectUniqueId, linearVelocity=None, angularVelocity=None, physicsClientId=0):
    b = _world(physicsClientId)["bodies"][objectUniqueId]
    if linearVelocity is not None:
        b.lin = tuple(float(v) for v in linearVelocity)
    if angularVelocity is not None:
        b.ang = tuple(float(v) for v in angularVelocity)


def getBaseVelocity(bodyUniqueId, physicsClientId=0):
    b = _world(physicsClientId)["bodies"][bodyUniqueId]
    return b.lin, b.ang


# ---- quaternion helpers (x, y, z, w) --------------------------------------------------------------
def getMatrixFromQuaternion(q):
    """btMatrix3x3::setRotation, row-major 9-tuple."""
    x, y, z, w = (float(v) for v in q)
    d = x * x + y * y + z * z + w * w
    s = 2.0 / d
    xs, ys, zs = x * s, y * s, z * s
    wx, wy, wz = w * xs, w * ys, w * zs
    xx, xy, xz = x * xs, x * ys, x * zs
    yy, yz, zz = y * ys, y * zs, z * zs
    return (1.0 - (yy + zz), xy - wz, xz + wy,
            xy + wz, 1.0 - (xx + zz), yz - wx,
            xz - wy, yz + wx, 1.0 - (xx + yy))


def getEulerFromQuaternion(q):
    """pybullet.c pybullet_getEulerFromQuaternion: ZYX with the 0.99999 gimbal guard."""
    x, y, z, w = (float(v) for v in q)
    sqx, sqy, sqz, sqw = x * x, y * y, z * z, w * w
    sarg = -2.0 * (x * z - w * y)
    if sarg <= -0.99999:
        return (0.0, -0.5 * math.pi, 2.0 * math.atan2(x, -y))
    if sarg >= 0.99999:
        return (0.0, 0.5 * math.pi, 2.0 * math.atan2(-x, y))
    roll = math.atan2(2.0 * (y * z + w * x), sqw - sqx - sqy + sqz)
    pitch = math.asin(sarg)
    yaw = math.atan2(2.0 * (x * y + w * z), sqw + sqx - sqy - sqz)
    return (roll, pitch, yaw)


def getQuaternionFromEuler(e):
    """btQuaternion::setEulerZYX(yaw, pitch, roll)."""
    roll, pitch, yaw = (float(v) for v in e)
    cy, sy = math.cos(0.5 * yaw), math.sin(0.5 * yaw)
    cp, sp = math.cos(0.5 * pitch), math.sin(0.5 * pitch)
    cr, sr = math.cos(0.5 * roll), math.sin(0.5 * roll)
    return (sr * cp * cy - cr * sp * sy,
            cr * sp * cy + sr * cp * sy,
            cr * cp * sy - sr * sp * cy,
            cr * cp * cy + sr * sp * sy)


# ---- external forces (PYB_* physics modes; recorded, never integrated) ---------------------------
# The reference's aerodynamic models (`sb3_envs/BaseAviary.py:943-1039`) are plain Python up to a final
# `p.applyExternalForce(...)`.  There is no solver here, so the stand-in RECORDS those calls: the force vectors the
# reference itself computes for a given state become golden values (tests/golden/aero/aero_forces.npz,
# oracle/ref_harness/make_aero_golden.py) for the DYN-path restatement of the same models.
force_log = []

# link frames of assets/cf2x.urdf: links 0-3 are the propellers (inertial origins at :42,54,66,78, fixed joints without
# an origin), link 4 is `center_of_mass_link` at the base origin (:86-95)
_CF2X_LINK_XYZ = ((0.028, -0.028, 0.0), (-0.028, -0.028, 0.0), (-0.028, 0.028, 0.0), (0.028, 0.028, 0.0), (0.0, 0.0, 0.0))


def applyExternalForce(objectUniqueId, linkIndex, forceObj, posObj, flags, physicsClientId=0):
    force_log.append({"kind": "force", "body": int(objectUniqueId), "link": int(linkIndex),
                      "force": tuple(float(v) for v in forceObj), "pos": tuple(float(v) for v in posObj),
                      "flags": int(flags)})


def applyExternalTorque(objectUniqueId, linkIndex, torqueObj, flags, physicsClientId=0):
    force_log.append({"kind": "torque", "body": int(objectUniqueId), "link": int(linkIndex),
                      "torque": tuple(float(v) for v in torqueObj), "flags": int(flags)})


def getLinkStates(bodyUniqueId, linkIndices, computeLinkVelocity=0, computeForwardKinematics=0, physicsClientId=0):
    """linkWorldPosition of each requested link (centre of mass of the link: base position + R(base quat) . inertial
    origin); the remaining entries of a link state are not used by the reference and are returned as zeros."""
    b = _world(physicsClientId)["bodies"][bodyUniqueId]
    R = getMatrixFromQuaternion(b.quat)
    out = []
    for li in linkIndices:
        ox, oy, oz = _CF2X_LINK_XYZ[li]
        wp = (b.pos[0] + R[0] * ox + R[1] * oy + R[2] * oz,
              b.pos[1] + R[3] * ox + R[4] * oy + R[5] * oz,
              b.pos[2] + R[6] * ox + R[7] * oy + R[8] * oz)
        out.append((wp, b.quat, (0.0, 0.0, 0.0), (0.0, 0.0, 0.0, 1.0), wp, b.quat, (0.0, 0.0, 0.0), (0.0, 0.0, 0.0)))
    return out
