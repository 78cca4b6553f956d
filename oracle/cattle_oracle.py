"""CPU oracle for the RL-Cattle-Herding env step (Physics.DYN path)  --  TEST INFRASTRUCTURE ONLY.

This file is a float64 restatement, in plain Python/numpy, of what the reference computes in one
`CattleAviary.step()` / `reset()` on its `Physics.DYN` path.  It exists to check the CUDA kernels
and to serve as the timed CPU baseline in `bench.py`; only `tests/`, `__graft_entry__.smoke()` and
`bench.py`'s cpu_baseline / `--impl reference` legs may import it.  The product path
(`rl-cattle-herding_b200/`) never does.

PARITY PINNING: the reference ships no golden vectors for this path, so the
oracle is pinned against outputs of the reference itself: `oracle/ref_harness/make_golden.py` imports
the unmodified reference from `/root/reference` (stand-in pybullet, see `ref_harness/standins/`) and
writes `tests/golden/*.npz`; `tests/test_oracle_golden.py` replays them through this file
(teacher-forced and free-running) and requires <=1e-9 agreement on floats and exact flags.

Reference files restated (all under /root/reference/gym_pybullet_drones/):
  sb3_envs/BaseAviary.py      :251-277 (drone spawn) :280-331 (reset) :335-465 (step)
                              :547-637 (housekeeping/cow spawn) :704-721 (kinematics refresh)
                              :1043-1117 (_dynamics/_integrateQ) :1352-1400 (_flockingStep)
  sb3_envs/BaseRLAviary.py    :137-238 (_preprocessAction) :272-342 (_computeObs) :348-392 (centroids)
  sb3_envs/CattleAviary.py    :213-332 (reward) :422-492 (terminated) :497-552 (truncated)
                              :572-679 (spacing reward shapes)
  control/DSLPIDControl.py    :37-60 (gains) :149-259 (position + attitude PID)
  utils/flockUtils.py         :11-58 (MathUtils) :116-160,237-374 (flocking terms)
  utils/evaluation.py         :100-138,271-273 (winding-number effectiveness)
  utils/curriculum_learning.py:10-222 (level table, tally)
  assets/cf2x.urdf            :5,11-12 (physical constants)
Third-party semantics restated: pybullet quaternion helpers / state store (Bullet formulas,
SURVEY.md appendix A.1); scipy `Rotation` Euler round trip in DSLPIDControl.py:205,242-244
(an identity on the target rotation, see `_pid`).
"""
import math

import numpy as np

NMAX = 12            # GLOBAL_MAX_NUM_DRONES, BaseAviary.py:112
MAX_NEIGHBORS = 4    # BaseAviary.py:116
ACTIVE_NEIGHBORS = 2  # BaseAviary.py:117
MAX_NEARBY_CATTLE = 16  # BaseAviary.py:118

# ---- cf2x.urdf:5,11-12 ---------------------------------------------------------------------------
M_KG = 0.027
ARM = 0.0397
KF = 3.16e-10
KM = 7.94e-12
T2W = 2.25
MAX_SPEED_KMH = 30.0
IXX, IYY, IZZ = 1.4e-5, 1.4e-5, 2.17e-5
GND_EFF_COEFF = 11.36859
PROP_RADIUS = 2.31348e-2
DRAG_XY, DRAG_Z = 9.1785e-7, 10.311e-7
DW1, DW2, DW3 = 2267.18, 0.16, -0.11
G = 9.8                                   # BaseAviary.py:97
GRAVITY = G * M_KG                        # BaseAviary.py:162
HOVER_RPM = math.sqrt(GRAVITY / (4 * KF))  # BaseAviary.py:163
MAX_RPM = math.sqrt((T2W * GRAVITY) / (4 * KF))  # BaseAviary.py:164
MAX_THRUST = 4 * KF * MAX_RPM ** 2        # BaseAviary.py:165
PROP_XY = ((0.028, -0.028), (-0.028, -0.028), (-0.028, 0.028), (0.028, 0.028))   # cf2x.urdf:42,54,66,78
GND_EFF_H_CLIP = 0.25 * PROP_RADIUS * math.sqrt((15 * MAX_RPM ** 2 * KF * GND_EFF_COEFF) / MAX_THRUST)  # :173
TARGET_ALT = 0.45                         # BaseAviary.py:111
SPEED_LIMIT = 0.3 * MAX_SPEED_KMH * (1000 / 3600)  # BaseRLAviary.py:102

# ---- DSLPIDControl.py:37-53 ----------------------------------------------------------------------
P_FOR = (0.4, 0.4, 1.25)
I_FOR = (0.05, 0.05, 0.05)
D_FOR = (0.2, 0.2, 0.5)
P_TOR = (70000.0, 70000.0, 60000.0)
I_TOR = (0.0, 0.0, 500.0)
D_TOR = (20000.0, 20000.0, 12000.0)
PWM2RPM_SCALE = 0.2685
PWM2RPM_CONST = 4070.3
MIN_PWM, MAX_PWM = 20000.0, 65535.0
MIXER = ((-0.5, -0.5, -1.0), (-0.5, 0.5, 1.0), (0.5, 0.5, -1.0), (0.5, -0.5, 1.0))

# ---- flockUtils.py:13-19,62-74 and BaseAviary.py:51-55,579,1353-1355 ------------------------------
EPS = 0.1
BUMP_H = 0.2
PHI_A, PHI_B = 5.0, 5.0
PHI_C = abs(PHI_A - PHI_B) / math.sqrt(4 * PHI_A * PHI_B)
C2_ALPHA = 2 * math.sqrt(3)
C1_GAMMA = 5.0
C2_GAMMA = 0.2 * math.sqrt(5.0)
C2_BETA = 2 * math.sqrt(20)
ALPHA_RANGE = ALPHA_DIST = 1.2
BETA_RANGE = BETA_DIST = 1.0
SENSING_RANGE = 999.0
DANGER_RANGE = 1.1
CONSENSUS = 1.0
PREDATOR_K = 650000.0
SHEPHERD_R = 1.0
FLOCK_DT = 0.05
MAX_VEL_CATTLE = 0.2

# ---- CattleAviary.py:91-105 ----------------------------------------------------------------------
MISSION_BOUNDARY = 15.0
MAX_FORMATION_DISTANCE = 8.0
COLLISION_THRESHOLD = 0.2
MAX_ALT_ERROR = TARGET_ALT * 0.6
SP_A, SP_B, SP_C, SP_K, SP_D, SP_R0, SP_LAM = 1.2, 2.1, 3.3, 0.2, -1.0, 1.3, 0.8

# ---- curriculum_learning.py:10-194: one row per level ---------------------------------------------
# (desired, tol, hold, approach_min, min_eff, cattle_desired, cattle_tol, min_drones, max_drones,
#  episode_length, w_simple, w_complex, w_survival, w_approach, w_eff, w_c2d, required_tally)
CURRICULUM = {
    0: (0.8, 0.3, 10, None, 0.0, 0.0, 0.0, 3, 3, 40, 1.0, 0.0, 0.0, 0.0, 0.0, 0.0, 100),
    1: (0.8, 0.2, 25, 0.0, 0.0, 0.0, 0.0, 4, 4, 40, 0.0, 1.0, -0.5, 0.0, 0.0, 0.0, 300),
    2: (0.8, 0.2, 15, 0.6, 0.0, 0.0, 0.0, 4, 4, 40, 0.0, 0.8, 0.0, 1.0, 0.0, 0.0, 100),
    3: (0.8, 0.2, 15, 0.3, 0.0, 0.0, 0.0, 4, 4, 40, 0.0, 0.8, -0.5, 1.0, 0.0, 0.0, 400),
    4: (0.8, 0.2, 15, 0.3, 20.0, 0.0, 0.0, 4, 4, 80, 0.0, 0.7, -0.0, 0.8, 1.0, 0.0, 600),
    5: (0.8, 0.2, 15, 0.3, 50.0, 0.8, 0.1, 4, 4, 40, 0.0, 0.7, -0.5, 0.6, 1.0, 0.8, 600),
    6: (0.8, 0.3, 15, 0.2, 50.0, 0.0, 0.0, 4, 12, 80, 0.7, 0.0, -0.0, 0.8, 1.0, 0.0, 600),
    7: (0.8, 0.3, 15, 0.2, 50.0, 0.0, 0.0, 4, 12, 80, 0.0, 0.0, -0.0, 1.0, 1.0, 0.0, 600),
}
(C_DESIRED, C_TOL, C_HOLD, C_APPROACH_MIN, C_MIN_EFF, C_CATTLE_DESIRED, C_CATTLE_TOL, C_MIN_DRONES,
 C_MAX_DRONES, C_EP_LEN, C_W_SIMPLE, C_W_COMPLEX, C_W_SURV, C_W_APPROACH, C_W_EFF, C_W_C2D,
 C_REQ_TALLY) = range(17)


# =================================================================================================
# pybullet helper semantics
# =================================================================================================
def quat_to_matrix(q):
    x, y, z, w = q
    s = 2.0 / (x * x + y * y + z * z + w * w)
    xs, ys, zs = x * s, y * s, z * s
    wx, wy, wz = w * xs, w * ys, w * zs
    xx, xy, xz = x * xs, x * ys, x * zs
    yy, yz, zz = y * ys, y * zs, z * zs
    return ((1.0 - (yy + zz), xy - wz, xz + wy),
            (xy + wz, 1.0 - (xx + zz), yz - wx),
            (xz - wy, yz + wx, 1.0 - (xx + yy)))


def quat_to_euler(q):
    x, y, z, w = q
    sarg = -2.0 * (x * z - w * y)
    if sarg <= -0.99999:
        return (0.0, -0.5 * math.pi, 2.0 * math.atan2(x, -y))
    if sarg >= 0.99999:
        return (0.0, 0.5 * math.pi, 2.0 * math.atan2(-x, y))
    return (math.atan2(2.0 * (y * z + w * x), w * w - x * x - y * y + z * z),
            math.asin(sarg),
            math.atan2(2.0 * (x * y + w * z), w * w + x * x - y * y - z * z))


def quat_normalise(q):
    x, y, z, w = q
    n = math.sqrt(x * x + y * y + z * z + w * w)
    return (x / n, y / n, z / n, w / n)


def _clip(v, lo, hi):
    return lo if v < lo else (hi if v > hi else v)


# =================================================================================================
# State container (mirrors ref_harness.capture_state; NMAX drone slots, M cows)
# =================================================================================================
def aero_terms(flags, pos, R, vel, rpm, drag_rpm, others, k, n):
    """The reference's PYB-only aerodynamic models as force terms for drone `k` (sb3_envs/BaseAviary.py:943-1039),
    pinned to the reference's own outputs by tests/golden/aero/aero_forces.npz (tests/test_oracle_extras.py, 1e-12):
      ground effect (:943-978)  -> per-propeller extra thrust along body z: kf rpm_i^2 c_g (r_p / 4 h_i)^2, h_i = world
                                   height of prop i (getLinkStates; cf2x.urdf:42-78) clipped at GND_EFF_H_CLIP; applied
                                   only if |roll|, |pitch| < pi/2, i.e. R[2][2] = cos(roll) cos(pitch) > 0
      drag (:982-1009)          -> the reference applies R^T (-(c_xy, c_xy, c_z) * sum(2 pi rpm'/60) * v) in the LINK
                                   frame of the centre-of-mass link; in world coordinates that is the vector returned
      downwash (:1013-1039)     -> body-z force: sum over drones i above within 10 m of -a exp(-(dxy/b)^2 / 2),
                                   a = dw1 (r_p / 4 dz)^2, b = dw2 dz + dw3
    Returns (gnd[4], drag_world[3], downwash_z)."""
    gnd = [0.0, 0.0, 0.0, 0.0]
    if (flags & 2) and R[2][2] > 0:
        for i, (ox, oy) in enumerate(PROP_XY):
            h = max(pos[2] + R[2][0] * ox + R[2][1] * oy, GND_EFF_H_CLIP)
            gnd[i] = rpm[i] * rpm[i] * KF * GND_EFF_COEFF * (PROP_RADIUS / (4 * h)) ** 2
    drag = [0.0, 0.0, 0.0]
    if (flags & 1) and drag_rpm is not None:
        om = (2 * math.pi / 60) * (drag_rpm[0] + drag_rpm[1] + drag_rpm[2] + drag_rpm[3])
        drag = [-DRAG_XY * om * vel[0], -DRAG_XY * om * vel[1], -DRAG_Z * om * vel[2]]
    dwz = 0.0
    if (flags & 4) and others is not None:
        for i in range(n):
            dz = others[i][2] - pos[2]
            dxy = math.sqrt((others[i][0] - pos[0]) ** 2 + (others[i][1] - pos[1]) ** 2)
            if i != k and dz > 0 and dxy < 10:
                wgt = math.exp(-0.5 * (dxy / (DW2 * dz + DW3)) ** 2)
                if wgt > 0:
                    dwz = dwz - DW1 * (PROP_RADIUS / (4 * dz)) ** 2 * wgt
    return gnd, drag, dwz


STATE_FIELDS = ("n_drones", "pos", "quat", "rpy", "vel", "ang_v", "rpy_rates", "last_rpm",
                "pid_ipos", "pid_last_rpy", "pid_irpy", "cattle_pos", "cattle_vel", "step_counter",
                "step_counter_A", "prev_cent_dist", "spacing_clock", "level", "tally", "spawn_idx")


class OracleEnv:
    """One environment, CTDE (sb3_envs) packing.  All arithmetic in float64 like the reference."""

    def __init__(self, n_cattle, ctrl_freq=30, pyb_freq=240, start_level=7, spawn_table=None,
                 act="vel", aero_flags=0, cattle_integrate=False):
        if pyb_freq % ctrl_freq != 0:
            raise ValueError("pyb_freq is not divisible by ctrl_freq")  # BaseAviary.py:102-103
        self.M = int(n_cattle)
        self.CTRL_FREQ = ctrl_freq
        self.PYB_FREQ = pyb_freq
        self.K = pyb_freq // ctrl_freq
        self.CTRL_DT = 1.0 / ctrl_freq
        self.PYB_DT = 1.0 / pyb_freq
        self.ABS = int(0.15 * ctrl_freq)                              # BaseRLAviary.py:69
        self.ACTION_SPACE = {"vel": 4, "rpm": 4, "pid": 3, "one_d_rpm": 1, "one_d_pid": 1}[act]  # BaseRLAviary.py:115-123
        self.L = 10 + MAX_NEIGHBORS * 2 + MAX_NEARBY_CATTLE * 2 + self.ABS * self.ACTION_SPACE  # BaseRLAviary.py:262
        self.act = act
        # non-reference extensions, both off on the reference's DYN path
        self.aero_flags = int(aero_flags)            # bit0 drag, bit1 ground effect, bit2 downwash
        self.cattle_integrate = bool(cattle_integrate)
        row = CURRICULUM[start_level]
        # fixed at construction, never refreshed when the level changes (CattleAviary.py:66,72-73)
        self.EPISODE_LEN_SEC = row[C_EP_LEN]
        self.MIN_NUM_DRONES = row[C_MIN_DRONES]
        self.MAX_NUM_DRONES = row[C_MAX_DRONES]
        self.spawn_table = None if spawn_table is None else np.asarray(spawn_table, dtype=np.float64)
        # state
        self.n_drones = 0
        self.pos = np.zeros((NMAX, 3)); self.quat = np.zeros((NMAX, 4)); self.quat[:, 3] = 1.0
        self.rpy = np.zeros((NMAX, 3)); self.vel = np.zeros((NMAX, 3)); self.ang_v = np.zeros((NMAX, 3))
        self.rpy_rates = np.zeros((NMAX, 3)); self.last_rpm = np.zeros((NMAX, 4))
        self.pid_ipos = np.zeros((NMAX, 3)); self.pid_last_rpy = np.zeros((NMAX, 3))
        self.pid_irpy = np.zeros((NMAX, 3))
        self.cattle_pos = np.zeros((self.M, 3)); self.cattle_vel = np.zeros((self.M, 3))
        self.step_counter = 0
        self.step_counter_A = 0
        self.prev_cent_dist = float("nan")      # nan <=> None (CattleAviary.py:89)
        self.spacing_clock = 0.0
        self.level = start_level
        self.tally = 0
        self.spawn_idx = 0                      # BaseAviary.py:121
        # diagnostics of the last step (integer outputs compared bit-exactly by the tests)
        self.last_neighbors = np.full((NMAX, ACTIVE_NEIGHBORS), -1, dtype=np.int32)
        self.last_herded = 0
        self.last_term_in_reward = False

    # ---------------------------------------------------------------------------------------------
    def get_state(self):
        return {k: (np.array(getattr(self, k)) if isinstance(getattr(self, k), np.ndarray) else getattr(self, k))
                for k in STATE_FIELDS}

    def set_state(self, st):
        for k in STATE_FIELDS:
            v = st[k]
            cur = getattr(self, k)
            if isinstance(cur, np.ndarray):
                cur[...] = np.asarray(v, dtype=np.float64).reshape(cur.shape)
            elif isinstance(cur, float):
                setattr(self, k, float(v))
            else:
                setattr(self, k, int(v))

    # ---------------------------------------------------------------------------------------------
    # reset  (BaseAviary.py:280-331, 547-637, 251-277)
    # ---------------------------------------------------------------------------------------------
    @staticmethod
    def drone_spawn(n, spacing=1.75):
        out = np.zeros((n, 3))
        if n <= 4:
            for i in range(n):
                out[i] = (i * spacing, 0.0, TARGET_ALT)
        else:
            n1 = n // 2
            for i in range(n1):
                out[i] = (i * spacing, 0.0, TARGET_ALT)
            for i in range(n - n1):
                out[n1 + i] = (i * spacing, spacing, TARGET_ALT)
        return out

    def reset(self, n_drones, u):
        """`n_drones` is the `random.randint` draw (:307); `u[j] = (u1, u2)` are the two consecutive
        `np.random.rand()` draws per cow (:617,631).  Returns the initial observation."""
        self.step_counter_A = 0
        n = int(n_drones)
        self.n_drones = n
        self.pos[:] = 0; self.quat[:] = 0; self.rpy[:] = 0; self.vel[:] = 0; self.ang_v[:] = 0
        self.rpy_rates[:] = 0; self.last_rpm[:] = 0
        self.step_counter = 0
        self.pos[:n] = self.drone_spawn(n)
        self.quat[:n] = (0.0, 0.0, 0.0, 1.0)
        # spawn row: pre-incremented with wrap (:600-606)
        n_sims = self.spawn_table.shape[0]
        self.spawn_idx += 1
        if self.spawn_idx >= n_sims:
            self.spawn_idx = 0
        rowxy = self.spawn_table[self.spawn_idx]
        u = np.asarray(u, dtype=np.float64).reshape(self.M, 2)
        for j in range(self.M):
            self.cattle_pos[j] = (rowxy[j, 0], rowxy[j, 1], 0.1)
            ang = math.pi * (2.0 * u[j, 1] - 1.0)
            self.cattle_vel[j] = (MAX_VEL_CATTLE * math.cos(ang), MAX_VEL_CATTLE * math.sin(ang), 0.0)
        # NOT reset: PID state, prev_cent_dist, spacing_clock, level/tally
        return self.compute_obs()

    # ---------------------------------------------------------------------------------------------
    # action -> rpm  (BaseRLAviary.py:137-238) and DSL PID (DSLPIDControl.py:149-259)
    # ---------------------------------------------------------------------------------------------
    def _pid(self, k, target_pos, target_yaw, target_vel):
        T = self.CTRL_DT
        p = self.pos[k]; q = self.quat[k]; v = self.vel[k]
        R = quat_to_matrix(q)
        pos_e = [target_pos[i] - p[i] for i in range(3)]
        vel_e = [target_vel[i] - v[i] for i in range(3)]
        ip = self.pid_ipos[k]
        for i in range(3):
            ip[i] = _clip(ip[i] + pos_e[i] * T, -2.0, 2.0)
        ip[2] = _clip(ip[2], -0.15, 0.15)
        F = [P_FOR[i] * pos_e[i] + I_FOR[i] * ip[i] + D_FOR[i] * vel_e[i] for i in range(3)]
        F[2] += GRAVITY
        scalar = max(0.0, F[0] * R[0][2] + F[1] * R[1][2] + F[2] * R[2][2])
        thrust = (math.sqrt(scalar / (4 * KF)) - PWM2RPM_CONST) / PWM2RPM_SCALE
        nF = math.sqrt(F[0] ** 2 + F[1] ** 2 + F[2] ** 2)
        zb = (F[0] / nF, F[1] / nF, F[2] / nF)
        xc = (math.cos(target_yaw), math.sin(target_yaw), 0.0)
        c = (zb[1] * xc[2] - zb[2] * xc[1], zb[2] * xc[0] - zb[0] * xc[2], zb[0] * xc[1] - zb[1] * xc[0])
        nc = math.sqrt(c[0] ** 2 + c[1] ** 2 + c[2] ** 2)
        yb = (c[0] / nc, c[1] / nc, c[2] / nc)
        xb = (yb[1] * zb[2] - yb[2] * zb[1], yb[2] * zb[0] - yb[0] * zb[2], yb[0] * zb[1] - yb[1] * zb[0])
        # target rotation has columns (xb, yb, zb).  The reference pushes it through scipy
        # from_matrix -> as_euler('XYZ') -> from_euler -> as_quat -> from_quat -> as_matrix
        # (DSLPIDControl.py:205,242-244), which reproduces the same rotation (also at gimbal lock,
        # where only the angle split changes) up to ~1e-16; it is restated as the identity.
        Rt = ((xb[0], yb[0], zb[0]), (xb[1], yb[1], zb[1]), (xb[2], yb[2], zb[2]))
        # attitude loop
        rpy = quat_to_euler(q)

        def RtT_R(i, j):   # (Rt^T R)[i][j]
            return Rt[0][i] * R[0][j] + Rt[1][i] * R[1][j] + Rt[2][i] * R[2][j]

        def RT_Rt(i, j):   # (R^T Rt)[i][j]
            return R[0][i] * Rt[0][j] + R[1][i] * Rt[1][j] + R[2][i] * Rt[2][j]

        rot_e = (RtT_R(2, 1) - RT_Rt(2, 1), RtT_R(0, 2) - RT_Rt(0, 2), RtT_R(1, 0) - RT_Rt(1, 0))
        lr = self.pid_last_rpy[k]
        rates_e = [0.0 - (rpy[i] - lr[i]) / T for i in range(3)]
        lr[:] = rpy
        ir = self.pid_irpy[k]
        for i in range(3):
            ir[i] = _clip(ir[i] - rot_e[i] * T, -1500.0, 1500.0)
        ir[0] = _clip(ir[0], -1.0, 1.0)
        ir[1] = _clip(ir[1], -1.0, 1.0)
        tau = [_clip(-P_TOR[i] * rot_e[i] + D_TOR[i] * rates_e[i] + I_TOR[i] * ir[i], -3200.0, 3200.0)
               for i in range(3)]
        rpm = [0.0] * 4
        for m in range(4):
            pwm = thrust + MIXER[m][0] * tau[0] + MIXER[m][1] * tau[1] + MIXER[m][2] * tau[2]
            rpm[m] = PWM2RPM_SCALE * _clip(pwm, MIN_PWM, MAX_PWM) + PWM2RPM_CONST
        return rpm

    def preprocess_action(self, action):
        n = self.n_drones
        rpm = np.zeros((n, 4))
        for k in range(n):
            # Actions arrive as float32 (action_space dtype).  Under numpy>=2 promotion rules the
            # norm, the unit vector and `SPEED_LIMIT*abs(a[3])` stay float32 before being widened
            # (BaseRLAviary.py:188-200); this is also what the fp32 kernel does naturally.
            a32 = np.asarray(action[k], dtype=np.float32)
            a = [float(x) for x in a32]
            f32 = np.float32
            if self.act == "vel":
                h = a32[0:2]
                nrm = f32(np.sqrt(h[0] * h[0] + h[1] * h[1]))
                if nrm != 0:
                    ux, uy = float(h[0] / nrm), float(h[1] / nrm)
                else:
                    ux = uy = 0.0
                scale = float(f32(SPEED_LIMIT) * abs(a32[3]))
                tv = (ux * scale, uy * scale, 0.0 * scale)
                tp = (self.pos[k][0], self.pos[k][1], TARGET_ALT)
                rpm[k] = self._pid(k, tp, self.rpy[k][2], tv)
            elif self.act == "rpm":
                rpm[k] = [HOVER_RPM * float(f32(1) + f32(0.05) * a32[i]) for i in range(4)]
            elif self.act == "one_d_rpm":
                rpm[k] = [HOVER_RPM * float(f32(1) + f32(0.05) * a32[0])] * 4
            elif self.act == "one_d_pid":
                tp = (self.pos[k][0], self.pos[k][1], self.pos[k][2] + 0.1 * a[0])
                rpm[k] = self._pid(k, tp, 0.0, (0.0, 0.0, 0.0))
            elif self.act == "pid":
                # _calculateNextStep with step_size=1 (BaseAviary.py:1318-1348)
                cur = self.pos[k]
                d = [a[i] - cur[i] for i in range(3)]
                dist = math.sqrt(d[0] ** 2 + d[1] ** 2 + d[2] ** 2)
                if dist <= 1:
                    nxt = (a[0], a[1], a[2])
                else:
                    nxt = tuple(cur[i] + d[i] / dist * 1 for i in range(3))
                rpm[k] = self._pid(k, nxt, 0.0, (0.0, 0.0, 0.0))
            else:
                raise ValueError(self.act)
        return rpm

    # ---------------------------------------------------------------------------------------------
    # DYN integrator  (BaseAviary.py:1043-1117) + pybullet store semantics (:704-721)
    # ---------------------------------------------------------------------------------------------
    def _dynamics(self, k, rpm, drag_rpm=None, pos0=None):
        """BaseAviary._dynamics (:1043-1102).  With `aero_flags` the PYB-only models of :943-1039 are added
        as forces of the same explicit step (they have no DYN counterpart in the reference, so this restatement
        defines them):
          ground effect (:943-978)  per-prop extra thrust kf rpm_i^2 c_g (r_p / 4 h_i)^2, h_i = world height of
                                    prop i clipped at GND_EFF_H_CLIP; applied at the props, so it also enters the
                                    x/y torques; gate |roll|,|pitch| < pi/2 <=> R[2][2] > 0
          drag (:982-1009)          world force -(c_xy, c_xy, c_z) * v * sum(2 pi rpm'/60); rpm' is
                                    last_clipped_action (:437,447): the previous control step's RPMs on the first
                                    substep, the current ones afterwards (R^T then LINK_FRAME cancel)
          downwash (:1013-1039)     body-z force -a exp(-(dxy/b)^2/2) from every drone above within 10 m; the
                                    other drones' positions are sampled once per control step (`pos0`)."""
        dt = self.PYB_DT
        pos = self.pos[k]; quat = tuple(self.quat[k]); vel = self.vel[k]; w = self.rpy_rates[k]
        R = quat_to_matrix(quat)
        f = [r * r * KF for r in rpm]
        gnd, drag_w, dwz = aero_terms(self.aero_flags, pos, R, vel, rpm, drag_rpm, pos0, k, self.n_drones)
        for i in range(4):
            f[i] = f[i] + gnd[i]
        thrust = f[0] + f[1] + f[2] + f[3] + dwz
        fw = [R[0][2] * thrust + drag_w[0], R[1][2] * thrust + drag_w[1], R[2][2] * thrust - GRAVITY + drag_w[2]]
        zt = [r * r * KM for r in rpm]
        tz = -zt[0] + zt[1] - zt[2] + zt[3]
        tx = (f[0] + f[1] - f[2] - f[3]) * (ARM / math.sqrt(2))
        ty = (-f[0] + f[1] + f[2] - f[3]) * (ARM / math.sqrt(2))
        Jw = (IXX * w[0], IYY * w[1], IZZ * w[2])
        cr = (w[1] * Jw[2] - w[2] * Jw[1], w[2] * Jw[0] - w[0] * Jw[2], w[0] * Jw[1] - w[1] * Jw[0])
        wd = ((tx - cr[0]) * (1.0 / IXX), (ty - cr[1]) * (1.0 / IYY), (tz - cr[2]) * (1.0 / IZZ))
        for i in range(3):
            vel[i] = vel[i] + dt * (fw[i] / M_KG)
        for i in range(3):
            w[i] = w[i] + dt * wd[i]
        for i in range(3):
            pos[i] = pos[i] + dt * vel[i]
        # _integrateQ
        wn = math.sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2])
        if not (abs(wn) <= 1e-8):                 # np.isclose(wn, 0): atol 1e-8
            p_, q_, r_ = w[0], w[1], w[2]
            th = wn * dt / 2
            c = math.cos(th)
            s = 2 / wn * 0.5 * math.sin(th)
            x, y, z, ww = quat
            quat = (c * x + s * (r_ * y - q_ * z + p_ * ww),
                    c * y + s * (-r_ * x + p_ * z + q_ * ww),
                    c * z + s * (q_ * x - p_ * y + r_ * ww),
                    c * ww + s * (-p_ * x - q_ * y - r_ * z))
        self.quat[k] = quat_normalise(quat)       # Bullet hands back a unit quaternion
        self.ang_v[k] = (R[0][0] * w[0] + R[0][1] * w[1] + R[0][2] * w[2],
                         R[1][0] * w[0] + R[1][1] * w[1] + R[1][2] * w[2],
                         R[2][0] * w[0] + R[2][1] * w[1] + R[2][2] * w[2])

    # ---------------------------------------------------------------------------------------------
    # herd update  (BaseAviary.py:1352-1400, flockUtils.py)
    # ---------------------------------------------------------------------------------------------
    @staticmethod
    def _sigma_norm(z):
        return (np.sqrt(1 + EPS * np.sum(z * z, axis=-1, keepdims=True)) - 1) / EPS

    @staticmethod
    def _sigma_norm_grad(z):
        return z / np.sqrt(1 + EPS * np.sum(z * z, axis=-1, keepdims=True))

    @staticmethod
    def _bump(z):
        ph = np.zeros_like(z)
        m = z <= 1
        ph[m] = (1 + np.cos(np.pi * (z[m] - BUMP_H) / (1 - BUMP_H))) / 2
        ph[z < BUMP_H] = 1
        ph[z < 0] = 0
        return ph

    @staticmethod
    def _sigma_1(z):
        return z / np.sqrt(1 + z * z)

    @classmethod
    def _phi_alpha(cls, z, r, d):
        r_a = (math.sqrt(1 + EPS * r * r) - 1) / EPS
        d_a = (math.sqrt(1 + EPS * d * d) - 1) / EPS
        phi = ((PHI_A + PHI_B) * cls._sigma_1(z - d_a + PHI_C) + (PHI_A - PHI_B)) / 2
        return cls._bump(z / r_a) * phi

    def flocking_step(self):
        n = self.n_drones
        q = self.cattle_pos[:, :2].copy()
        p = self.cattle_vel[:, :2].copy()
        y = self.pos[:n, :2].copy()
        Mc = self.M
        u = np.zeros((Mc, 2))
        r_alpha = (math.sqrt(1 + EPS * ALPHA_RANGE ** 2) - 1) / EPS
        r_beta = (math.sqrt(1 + EPS * BETA_RANGE ** 2) - 1) / EPS
        for i in range(Mc):
            # alpha term over every other cow inside the sensing range (flockUtils.py:237-258,350-354)
            d = q - q[i]
            nb = np.sqrt(np.sum(d * d, axis=1)) <= SENSING_RANGE
            nb[i] = False
            if nb.sum() > 0:
                dj = d[nb]
                sn = self._sigma_norm(dj)
                grad = C2_ALPHA * np.sum(self._phi_alpha(sn, ALPHA_RANGE, ALPHA_DIST) * self._sigma_norm_grad(dj), axis=0)
                cons = C2_ALPHA * np.sum(self._bump(sn / r_alpha) * (p[nb] - p[i]), axis=0)
                u[i] += grad + cons
            # shepherd (delta) term (flockUtils.py:271-323,356-371)
            if n > 0:
                dd = y - q[i]
                inr = np.sqrt(np.sum(dd * dd, axis=1)) <= (SENSING_RANGE + 2)
                if inr.any():
                    qik = []
                    pik = []
                    for kk in np.where(inr)[0]:
                        diff = q[i] - y[kk]
                        dist = math.sqrt(diff[0] ** 2 + diff[1] ** 2) + 1e-6
                        mu = min(dist / SHEPHERD_R, 1.0)
                        ak = diff / dist
                        P = np.eye(2) - np.outer(ak, ak)
                        qik.append(mu * q[i] + (1 - mu) * y[kk])
                        pik.append(mu * (P @ p[i]))
                    qik = np.array(qik); pik = np.array(pik)
                    dj = qik - q[i]
                    sn = self._sigma_norm(dj)
                    grad = C2_BETA * np.sum(self._phi_alpha(sn, BETA_RANGE, BETA_DIST) * self._sigma_norm_grad(dj), axis=0)
                    cons = C2_BETA * np.sum(self._bump(sn / r_beta) * (pik - p[i]), axis=0)
                    u[i] += grad + cons
                # predator repulsion (flockUtils.py:343-348)
                with np.errstate(divide="ignore", invalid="ignore"):
                    for kk in range(n):
                        dv = y[kk] - q[i]
                        dn = math.sqrt(dv[0] ** 2 + dv[1] ** 2)
                        if dn <= DANGER_RANGE:
                            u[i] += -PREDATOR_K * dv / np.float64(dn) ** 3
            # gamma term toward the consensus point (1,1) (flockUtils.py:150-160,340-341)
            u[i] += -C1_GAMMA * self._sigma_1(q[i] - CONSENSUS) - C2_GAMMA * p[i]
        p = p + u * (FLOCK_DT ** 2)
        sp = np.sqrt(np.sum(p * p, axis=1))
        for i in range(Mc):
            if sp[i] > MAX_VEL_CATTLE:
                p[i] *= MAX_VEL_CATTLE / sp[i]
        self.cattle_vel[:, 0:2] = p
        self.cattle_vel[:, 2] = 0.0

    # ---------------------------------------------------------------------------------------------
    # observation  (BaseRLAviary.py:272-342)
    # ---------------------------------------------------------------------------------------------
    def compute_obs(self):
        n = self.n_drones
        L = self.L
        obs = np.zeros((NMAX, L), dtype=np.float32)
        self.last_neighbors[:] = -1
        for i in range(n):
            row = [self.pos[i][2], *self.rpy[i], *self.vel[i], *self.ang_v[i]]
            rel = []
            for j in range(n):
                if j == i:
                    continue
                d = self.pos[j][:2] - self.pos[i][:2]
                rel.append((d, math.sqrt(d[0] ** 2 + d[1] ** 2), j))
            rel.sort(key=lambda t: t[1])          # stable: ties keep index order
            rel = rel[:ACTIVE_NEIGHBORS]
            for s, t in enumerate(rel):
                self.last_neighbors[i, s] = t[2]
            nb = [t[0] for t in rel]
            while len(nb) < NMAX:                 # padded to GLOBAL_MAX_NUM_DRONES, not MAX_NEIGHBORS (:313)
                nb.append(np.zeros(2))
            for v in nb:
                row.extend(v)
            cat = [self.cattle_pos[j][:2] - self.pos[i][:2] for j in range(self.M)]
            while len(cat) < MAX_NEARBY_CATTLE:
                cat.append(np.zeros(2))
            for v in cat[:MAX_NEARBY_CATTLE]:
                row.extend(v)
            # action-buffer region: always the reset-time zeros, n values per slot (:318-319,330-331)
            row.extend([0.0] * (self.ABS * n))
            arr = np.array(row, dtype=np.float32)
            if len(arr) > L:
                arr = arr[:L]
            elif len(arr) < L:
                arr = np.pad(arr, (0, L - len(arr)))
            obs[i] = arr
        return obs

    # ---------------------------------------------------------------------------------------------
    # task layer  (CattleAviary.py, utils/evaluation.py)
    # ---------------------------------------------------------------------------------------------
    def herded_count(self):
        """Winding-number test against the polygon drones[0..n-1] in index order (evaluation.py:100-138)."""
        n = self.n_drones
        poly = self.pos[:n, :2]
        cnt = 0
        for j in range(self.M):
            px, py = self.cattle_pos[j][0], self.cattle_pos[j][1]
            wn = 0
            for i in range(n):
                x1, y1 = poly[i]
                x2, y2 = poly[(i + 1) % n]
                il = (x2 - x1) * (py - y1) - (px - x1) * (y2 - y1)
                if y1 <= py:
                    if y2 > py and il > 0:
                        wn += 1
                else:
                    if y2 <= py and il < 0:
                        wn -= 1
            if wn:
                cnt += 1
        return cnt

    def effectiveness(self):
        c = self.herded_count()
        self.last_herded = c
        return c / self.M * 100 if self.M > 0 else 0

    def _row(self):
        return CURRICULUM[self.level]

    def simple_spacing(self, r):              # CattleAviary.py:594-626
        row = self._row()
        desired = row[C_DESIRED]; tol = desired * row[C_TOL]
        lo, hi = desired - tol, desired + tol
        if lo <= r <= hi:
            return 1.0
        elif r < lo:
            return -1 + (r / lo) * 2
        elif r > hi:
            return 1 - ((r - hi) / (7.0 - hi)) * 2
        return -1.0

    def complex_spacing(self, r):             # CattleAviary.py:630-667
        d_star = self._row()[C_DESIRED]
        with np.errstate(over="ignore", invalid="ignore"):
            gauss = float(np.exp(-0.5 * ((np.float64(r) - d_star) / (0.4 + 1e-9)) ** 2))
        coll = -1.0 * (1.0 - (r / (0.3 + 1e-9))) if r < 0.3 else 0.0
        pull = -0.3 * (r - 1.5) / (5.0 - 1.5) if r > 1.5 else 0.0
        return gauss + coll + pull + 0.1 * (1 - abs(r - d_star))

    @staticmethod
    def cattle_spacing(r):                    # CattleAviary.py:572-592
        def g(x):
            return SP_A * math.exp(-((x - SP_D) ** 2) / (2 * SP_C ** 2)) - SP_B * math.exp(-(x ** 2) / (2 * SP_K ** 2))
        if r <= SP_R0:
            return g(r)
        return g(SP_R0) / math.exp(-SP_LAM * SP_R0) * math.exp(-SP_LAM * r)

    def _centroid_dist(self):
        n = self.n_drones
        cd = np.mean(self.pos[:n, :2], axis=0)
        cc = np.mean(self.cattle_pos[:, :2], axis=0)
        # both centroids carry z = 0.45+0.5 (BaseRLAviary.py:358,382): the 3-D norm equals the xy norm
        return math.sqrt((cd[0] - cc[0]) ** 2 + (cd[1] - cc[1]) ** 2 + 0.0)

    def _two_nearest(self, i):
        n = self.n_drones
        d = np.sqrt(np.sum((self.pos[:n, :2] - self.pos[i, :2]) ** 2, axis=1))
        d[i] = np.inf
        return np.partition(d, 1)[:2]

    def _min_spacing(self):
        n = self.n_drones
        best = np.inf
        for i in range(n):
            for j in range(n):
                if i != j:
                    d = self.pos[i, :2] - self.pos[j, :2]
                    best = min(best, math.sqrt(d[0] ** 2 + d[1] ** 2))
        return best

    def compute_terminated(self):             # CattleAviary.py:422-492 (stateful: spacing clock)
        row = self._row()
        lvl = self.level
        if lvl in (0, 1):
            desired, tol = row[C_DESIRED], row[C_TOL]
            hi = desired + desired * tol
            lo = desired - desired * tol
            ms = self._min_spacing()
            if ms < hi and ms > lo:
                self.spacing_clock += 1 / 240
                if self.spacing_clock >= row[C_HOLD]:
                    return True
            else:
                self.spacing_clock = 0
        elif lvl in (2, 3):
            if self._centroid_dist() < row[C_APPROACH_MIN]:
                return True
        elif lvl in (4, 6):
            if self.effectiveness() > row[C_MIN_EFF]:
                return True
        elif lvl == 5:
            if self.effectiveness() > row[C_MIN_EFF]:
                desired, tol = row[C_CATTLE_DESIRED], row[C_CATTLE_TOL]
                hi = desired + desired * tol
                lo = desired - desired * tol
                ms = self._min_spacing()
                if ms < hi and ms > lo:
                    return True
        return False

    def compute_truncated(self):              # CattleAviary.py:497-552
        n = self.n_drones
        cent = self._centroid_dist()
        for i in range(n):
            if abs(self.pos[i][2] - TARGET_ALT) > MAX_ALT_ERROR:
                return True
        xy = self.pos[:n, :2]
        for i in range(n):
            for j in range(i + 1, n):
                d = xy[i] - xy[j]
                if math.sqrt(d[0] ** 2 + d[1] ** 2) < COLLISION_THRESHOLD:
                    return True
        for i in range(n):
            d = np.sqrt(np.sum((xy - xy[i]) ** 2, axis=1))
            d[i] = np.inf
            if np.all(d > MAX_FORMATION_DISTANCE):
                return True
        if cent > MISSION_BOUNDARY:
            return True
        if self.step_counter / self.CTRL_FREQ > self.EPISODE_LEN_SEC:
            return True
        return False

    def compute_reward(self):                 # CattleAviary.py:213-332
        n = self.n_drones
        row = self._row()
        s_simple = 0.0
        s_complex = 0.0
        per_sp = np.zeros(n)
        for i in range(n):
            for dist in self._two_nearest(i):
                rc = self.complex_spacing(float(dist))
                rs = self.simple_spacing(float(dist))
                s_complex += rc
                s_simple += rs
                per_sp[i] += (rc + rs) / 2.0
        s_complex /= (n * 2.0)
        s_simple /= (n * 2.0)
        approach = 0
        cent = self._centroid_dist()
        if not math.isnan(self.prev_cent_dist):
            change = self.prev_cent_dist - cent
            approach = _clip((change / (SPEED_LIMIT / self.CTRL_FREQ + 1e-6)) * 5, -1.0, 1.0)
        self.prev_cent_dist = cent
        eff = self.effectiveness()
        c2d = 0.0
        per_c = np.zeros(n)
        for i in range(n):
            d = np.sqrt(np.sum((self.cattle_pos[:, :2] - self.pos[i, :2]) ** 2, axis=1))
            rcat = self.cattle_spacing(float(np.min(d)))
            c2d += rcat
            per_c[i] = rcat
        c2d /= n
        r_global = (s_simple * row[C_W_SIMPLE] + s_complex * row[C_W_COMPLEX] + 0.1 * row[C_W_SURV]
                    + approach * row[C_W_APPROACH] + (eff / 100) * row[C_W_EFF] + c2d * row[C_W_C2D])
        with np.errstate(invalid="ignore"):
            diff = (per_sp - np.mean(per_sp)) + (per_c - np.mean(per_c))
            per_drone = r_global + 0.5 * diff
        term = self.compute_terminated()
        self.compute_truncated()
        self.last_term_in_reward = bool(term)
        if term:
            # the +bonus goes to r_global AFTER per_drone was formed and is dropped;
            # only the curriculum side effect survives (curriculum_learning.py:200-222)
            self.tally += 1
            if self.tally >= row[C_REQ_TALLY]:
                self.tally = 0
                self.level += 1
                if self.level >= len(CURRICULUM):
                    self.level = len(CURRICULUM) - 1
        return float(np.mean(per_drone))

    # ---------------------------------------------------------------------------------------------
    # MARL packing  (rllib_envs/BaseMARLAviary.py:253-303, MARLCattleAviary.py:110-383,
    #                rllib_envs/BaseAviary.py:320-438)
    # ---------------------------------------------------------------------------------------------
    def compute_obs_marl(self, i):
        """Per-drone observation (L,): neighbours are padded to MAX_NEIGHBORS = 4 here, so there is no dead gap."""
        n, L = self.n_drones, self.L
        row = [self.pos[i][2], *self.rpy[i], *self.vel[i], *self.ang_v[i]]
        rel = []
        for j in range(n):
            if j == i:
                continue
            d = self.pos[j][:2] - self.pos[i][:2]
            rel.append((d, math.sqrt(d[0] ** 2 + d[1] ** 2), j))
        rel.sort(key=lambda t: t[1])
        rel = rel[:ACTIVE_NEIGHBORS]
        nb = [t[0] for t in rel]
        while len(nb) < MAX_NEIGHBORS:
            nb.append(np.zeros(2))
        for v in nb:
            row.extend(v)
        cat = [self.cattle_pos[j][:2] - self.pos[i][:2] for j in range(self.M)]
        while len(cat) < MAX_NEARBY_CATTLE:
            cat.append(np.zeros(2))
        for v in cat[:MAX_NEARBY_CATTLE]:
            row.extend(v)
        row.extend([0.0] * (self.ABS * self.ACTION_SPACE))     # reset-time zeros, never refreshed (:305,294)
        arr = np.array(row, dtype=np.float32)
        return arr[:L] if len(arr) >= L else np.pad(arr, (0, L - len(arr)))

    def compute_terminated_marl(self):
        """MARLCattleAviary._computeTerminated (:246-321): drone independent, clock += 1/CTRL_FREQ per CALL."""
        row = self._row()
        lvl = self.level
        if lvl in (0, 1):
            desired, tol = row[C_DESIRED], row[C_TOL]
            hi, lo = desired + desired * tol, desired - desired * tol
            ms = self._min_spacing()
            if lo < ms < hi:
                self.spacing_clock += 1 / self.CTRL_FREQ
                if self.spacing_clock >= row[C_HOLD]:
                    return True
            else:
                self.spacing_clock = 0
        elif lvl in (2, 3):
            if self._centroid_dist() < row[C_APPROACH_MIN]:
                return True
        elif lvl in (4, 6):
            if self.effectiveness() > row[C_MIN_EFF]:
                return True
        elif lvl == 5:
            if self.effectiveness() > row[C_MIN_EFF]:
                desired, tol = row[C_CATTLE_DESIRED], row[C_CATTLE_TOL]
                hi, lo = desired + desired * tol, desired - desired * tol
                if lo < self._min_spacing() < hi:
                    return True
        return False

    def compute_truncated_marl(self, i):
        """MARLCattleAviary._computeTruncated (:326-383): own altitude / collision / isolation, shared rest."""
        n = self.n_drones
        cent = self._centroid_dist()
        if abs(self.pos[i][2] - TARGET_ALT) > MAX_ALT_ERROR:
            return True
        xy = self.pos[:n, :2]
        for j in range(n):
            if j != i:
                d = xy[i] - xy[j]
                if math.sqrt(d[0] ** 2 + d[1] ** 2) < COLLISION_THRESHOLD:
                    return True
        d = np.sqrt(np.sum((xy - xy[i]) ** 2, axis=1))
        d[i] = np.inf
        if np.all(d > MAX_FORMATION_DISTANCE):
            return True
        if cent > MISSION_BOUNDARY:
            return True
        if self.step_counter / self.CTRL_FREQ > self.EPISODE_LEN_SEC:
            return True
        return False

    def _end_of_episode_marl(self, i):
        """MARLCattleAviary._endOfEpisodeReward (:180-241)."""
        n = self.n_drones
        row = self._row()
        lvl = self.level
        r = 0.0
        if lvl in (0, 1):
            desired, tol = row[C_DESIRED], row[C_TOL]
            hi, lo = desired + desired * tol, desired - desired * tol
            two = self._two_nearest(i)
            if np.all((two >= lo) & (two <= hi)):
                r += 50.0 / n
        elif lvl in (2, 3):
            if self._centroid_dist() < row[C_APPROACH_MIN]:
                r += 50.0
        elif lvl in (4, 6):
            eff = self.effectiveness()
            cc = np.mean(self.cattle_pos[:, :2], axis=0)
            dd = cc - self.pos[i, :2]
            w = min(max(1.0 - math.sqrt(dd[0] ** 2 + dd[1] ** 2) / 10.0, 0.0), 1.0)
            r += eff * 2 * w
        elif lvl == 5:
            if self.effectiveness() > row[C_MIN_EFF]:
                desired, tol = row[C_CATTLE_DESIRED], row[C_CATTLE_TOL]
                hi, lo = desired + desired * tol, desired - desired * tol
                two = self._two_nearest(i)
                if np.all((two >= lo) & (two <= hi)):
                    r += 50.0 / n
        return r

    def compute_reward_marl(self, i):
        """MARLCattleAviary._computeReward (:110-178).  Stateful: prev_cent_dist (so only the first drone of a
        step sees a non-zero approach term), the spacing clock and the curriculum tally/level."""
        two = self._two_nearest(i)
        s_simple = (self.simple_spacing(float(two[0])) + self.simple_spacing(float(two[1]))) / 2
        s_complex = (self.complex_spacing(float(two[0])) + self.complex_spacing(float(two[1]))) / 2
        r = 0.0
        r += s_simple * self._row()[C_W_SIMPLE]
        r += s_complex * self._row()[C_W_COMPLEX]
        r += 0.1 * self._row()[C_W_SURV]
        cent = self._centroid_dist()
        change = 0.0 if math.isnan(self.prev_cent_dist) else self.prev_cent_dist - cent
        self.prev_cent_dist = cent
        r += _clip((change / (SPEED_LIMIT / self.CTRL_FREQ + 1e-6)) * 5, -1.0, 1.0) * self._row()[C_W_APPROACH]
        r += (self.effectiveness() / 100) * self._row()[C_W_EFF]
        dmin = float(np.min(np.sqrt(np.sum((self.cattle_pos[:, :2] - self.pos[i, :2]) ** 2, axis=1))))
        r += self.cattle_spacing(dmin) * self._row()[C_W_C2D]
        if self.compute_terminated_marl():
            r += self._end_of_episode_marl(i)
            row = self._row()
            self.tally += 1
            if self.tally >= row[C_REQ_TALLY]:
                self.tally = 0
                self.level = min(self.level + 1, len(CURRICULUM) - 1)
        elif self.compute_truncated_marl(i):
            r -= 50
        return float(r)

    def step_marl(self, action):
        """rllib_envs/BaseAviary.step (:320-438): per-drone outputs, `step_counter += 1`.
        Returns (obs (12, L) f32, reward (12,), terminated (12,) bool, truncated (12,) bool)."""
        self.step_counter_A += 1
        n = self.n_drones
        rpm = self.preprocess_action(np.asarray(action))
        pos0 = self.pos.copy() if (self.aero_flags & 4) else None
        for sub in range(self.K):
            for k in range(n):
                self._dynamics(k, rpm[k], self.last_rpm[k] if sub == 0 else rpm[k], pos0)
            self.last_rpm[:n] = rpm
        for k in range(n):
            self.rpy[k] = quat_to_euler(tuple(self.quat[k]))
        if self.step_counter_A % 2 == 0:
            self.flocking_step()
        if self.cattle_integrate:       # non-reference kinematic herd
            self.cattle_pos[:, :2] = self.cattle_pos[:, :2] + self.cattle_vel[:, :2] * self.CTRL_DT
        obs = np.zeros((NMAX, self.L), dtype=np.float32)
        rew = np.zeros(NMAX)
        term = np.zeros(NMAX, dtype=bool)
        trunc = np.zeros(NMAX, dtype=bool)
        for i in range(n):
            obs[i] = self.compute_obs_marl(i)
        for i in range(n):
            rew[i] = self.compute_reward_marl(i)
        for i in range(n):
            term[i] = self.compute_terminated_marl()
        for i in range(n):
            trunc[i] = self.compute_truncated_marl(i)
        self.step_counter += 1
        return obs, rew, term, trunc

    def step_rllib_wrapper(self, action, agents_mask):
        """What the reference's `RLlibMultiAgentWrapper.step` hands to RLlib (rllib_envs/marl_wrapper.py:77-119): it steps
        the env, DISCARDS the outputs and calls the stateful per-agent functions again, interleaved per agent, for the
        agents still in its `agents` list (`agents_mask`, bit i = agent_i); agents whose second-pass `terminated` is set
        leave the list.  Returns (obs, reward, terminated, truncated, new_mask); entries of unlisted agents are zero."""
        obs1, _, _, _ = self.step_marl(action)
        n = self.n_drones
        obs = np.zeros_like(obs1)
        rew = np.zeros(NMAX)
        term = np.zeros(NMAX, dtype=bool)
        trunc = np.zeros(NMAX, dtype=bool)
        mask = int(agents_mask)
        for i in range(n):
            if not (mask >> i) & 1:
                continue
            obs[i] = obs1[i]                              # _computeObs is pure: same row as the step's
            rew[i] = self.compute_reward_marl(i)
            term[i] = self.compute_terminated_marl()
            trunc[i] = self.compute_truncated_marl(i)
            if term[i]:
                mask &= ~(1 << i)
        return obs, rew, term, trunc, mask

    # ---------------------------------------------------------------------------------------------
    # step  (BaseAviary.py:335-465)
    # ---------------------------------------------------------------------------------------------
    def step(self, action):
        self.step_counter_A += 1
        n = self.n_drones
        rpm = self.preprocess_action(np.asarray(action))
        pos0 = self.pos.copy() if (self.aero_flags & 4) else None
        for sub in range(self.K):
            for k in range(n):
                self._dynamics(k, rpm[k], self.last_rpm[k] if sub == 0 else rpm[k], pos0)
            self.last_rpm[:n] = rpm
        for k in range(n):
            self.rpy[k] = quat_to_euler(tuple(self.quat[k]))
        if self.step_counter_A % 2 == 0:
            self.flocking_step()
        if self.cattle_integrate:       # non-reference kinematic herd
            self.cattle_pos[:, :2] = self.cattle_pos[:, :2] + self.cattle_vel[:, :2] * self.CTRL_DT
        obs = self.compute_obs()
        reward = self.compute_reward()
        terminated = self.compute_terminated()
        truncated = self.compute_truncated()
        self.step_counter += self.K
        return obs, reward, bool(terminated), bool(truncated), {"answer": 42}
