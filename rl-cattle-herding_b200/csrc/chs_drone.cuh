// Per-drone work of one control step, entirely in registers:
//   action -> target (BaseRLAviary._preprocessAction, sb3_envs/BaseRLAviary.py:137-238)
//   DSL PID  (control/DSLPIDControl.py:149-259)
//   K explicit substeps of Physics.DYN (sb3_envs/BaseAviary.py:1043-1117) with the pybullet store
//   semantics of :704-721 (quaternion comes back unit-normalised, rpy = euler(quat)).
#pragma once
#include "chs_layout.cuh"

namespace chs {

template <typename T> struct DroneState {
  T p[3], q[4], v[3], w[3];        // pos, quat, vel, rpy_rates (body rates)
  T rpy[3];                        // euler angles of q as stored by the kinematics refresh
  T ip[3], lr[3], ir[3];           // PID: integral_pos_e, last_rpy, integral_rpy_e
  T av[3];                         // ang_v (world), output only
};

template <typename T>
__device__ __forceinline__ void load_drone(const KArgs<T>& a, int d, int e, DroneState<T>& s) {
  const size_t st = (size_t)a.p.D * a.p.E;
  const typename VT<T>::v4* base = a.drone + (size_t)d * a.p.E + e;
  const auto A = base[F_A * st], B = base[F_B * st], C = base[F_C * st], Dd = base[F_D * st];
  const auto Ee = base[F_E * st], Ff = base[F_F * st], Gg = base[F_G * st];
  s.p[0] = A.x; s.p[1] = A.y; s.p[2] = A.z; s.rpy[0] = A.w;
  s.q[0] = B.x; s.q[1] = B.y; s.q[2] = B.z; s.q[3] = B.w;
  s.v[0] = C.x; s.v[1] = C.y; s.v[2] = C.z; s.rpy[1] = C.w;
  s.w[0] = Dd.x; s.w[1] = Dd.y; s.w[2] = Dd.z; s.rpy[2] = Dd.w;
  s.ip[0] = Ee.x; s.ip[1] = Ee.y; s.ip[2] = Ee.z; s.av[0] = Ee.w;
  s.lr[0] = Ff.x; s.lr[1] = Ff.y; s.lr[2] = Ff.z; s.av[1] = Ff.w;
  s.ir[0] = Gg.x; s.ir[1] = Gg.y; s.ir[2] = Gg.z; s.av[2] = Gg.w;
}

template <typename T>
__device__ __forceinline__ void store_drone(const KArgs<T>& a, int d, int e, const DroneState<T>& s) {
  const size_t st = (size_t)a.p.D * a.p.E;
  typename VT<T>::v4* base = a.drone + (size_t)d * a.p.E + e;
  base[F_A * st] = mk4<T>(s.p[0], s.p[1], s.p[2], s.rpy[0]);
  base[F_B * st] = mk4<T>(s.q[0], s.q[1], s.q[2], s.q[3]);
  base[F_C * st] = mk4<T>(s.v[0], s.v[1], s.v[2], s.rpy[1]);
  base[F_D * st] = mk4<T>(s.w[0], s.w[1], s.w[2], s.rpy[2]);
  base[F_E * st] = mk4<T>(s.ip[0], s.ip[1], s.ip[2], s.av[0]);
  base[F_F * st] = mk4<T>(s.lr[0], s.lr[1], s.lr[2], s.av[1]);
  base[F_G * st] = mk4<T>(s.ir[0], s.ir[1], s.ir[2], s.av[2]);
}

// DSLPIDControl._dslPIDPositionControl + _dslPIDAttitudeControl (DSLPIDControl.py:149-259).
// The scipy Euler round trip of the target rotation (:205,242-244) is the identity on the rotation
// (oracle/cattle_oracle.py:_pid), so the target frame is carried as its three columns.
template <typename T>
__device__ __forceinline__ void dsl_pid(DroneState<T>& s, const T tp[3], T target_yaw, bool yaw_is_state, const T tv[3],
                                        T ctrl_dt, T ctrl_freq, T rpm[4]) {
  T R[3][3];
  quat_to_matrix<T>(s.q, R);
  T F[3];
  const T PF[3] = {T(k::P_FOR[0]), T(k::P_FOR[1]), T(k::P_FOR[2])};
  const T IF[3] = {T(k::I_FOR[0]), T(k::I_FOR[1]), T(k::I_FOR[2])};
  const T DF[3] = {T(k::D_FOR[0]), T(k::D_FOR[1]), T(k::D_FOR[2])};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const T pos_e = tp[i] - s.p[i];
    const T vel_e = tv[i] - s.v[i];
    T ip = clipv<T>(s.ip[i] + pos_e * ctrl_dt, T(-2), T(2));
    if (i == 2) ip = clipv<T>(ip, T(-0.15), T(0.15));
    s.ip[i] = ip;
    F[i] = PF[i] * pos_e + IF[i] * ip + DF[i] * vel_e;
  }
  F[2] += T(k::GRAVITY);
  const T scalar = Mth<T>::max(T(0), F[0] * R[0][2] + F[1] * R[1][2] + F[2] * R[2][2]);
  const T thrust = (Mth<T>::sqrt(scalar * T(1.0 / (4 * k::KF))) - T(k::PWM2RPM_CONST)) * T(1.0 / k::PWM2RPM_SCALE);
  const T inF = Mth<T>::rsqrt(F[0] * F[0] + F[1] * F[1] + F[2] * F[2]);
  const T zb[3] = {F[0] * inF, F[1] * inF, F[2] * inF};
  T sy, cy;
  if (!yaw_is_state) {
    sy = T(0); cy = T(1);          // every other PID-driven action type asks for yaw 0
  } else if (sizeof(T) == 4) {
    // VEL: the target yaw is the state's own yaw = atan2(R10, R00) of the stored quaternion (pybullet ZYX), so its
    // sine / cosine are R10, R00 over cos(pitch) -- no trigonometric range reduction on the drone's serial chain.
    // Inside the gimbal guard of getEulerFromQuaternion (cos^2(pitch) <= 2e-5) the stored yaw is used as is.
    const T h2 = R[0][0] * R[0][0] + R[1][0] * R[1][0];
    if (h2 > T(2.1e-5)) { const T ih = Mth<T>::rsqrt(h2); cy = R[0][0] * ih; sy = R[1][0] * ih; }
    else Mth<T>::sincos(target_yaw, &sy, &cy);
  } else {
    Mth<T>::sincos(target_yaw, &sy, &cy);
  }
  // zb x (cy, sy, 0)
  const T c0 = -zb[2] * sy, c1 = zb[2] * cy, c2 = zb[0] * sy - zb[1] * cy;
  const T inc = Mth<T>::rsqrt(c0 * c0 + c1 * c1 + c2 * c2);
  const T yb[3] = {c0 * inc, c1 * inc, c2 * inc};
  const T xb[3] = {yb[1] * zb[2] - yb[2] * zb[1], yb[2] * zb[0] - yb[0] * zb[2], yb[0] * zb[1] - yb[1] * zb[0]};
  // rot_e = vee(Rt^T R - R^T Rt) with Rt = [xb yb zb]:  A_ij = col_i(Rt) . col_j(R)
#define CHS_DOTCOL(v, j) ((v)[0] * R[0][j] + (v)[1] * R[1][j] + (v)[2] * R[2][j])
  const T rot_e[3] = {CHS_DOTCOL(zb, 1) - CHS_DOTCOL(yb, 2),
                      CHS_DOTCOL(xb, 2) - CHS_DOTCOL(zb, 0),
                      CHS_DOTCOL(yb, 0) - CHS_DOTCOL(xb, 1)};
#undef CHS_DOTCOL
  const T PT[3] = {T(k::P_TOR[0]), T(k::P_TOR[1]), T(k::P_TOR[2])};
  const T IT[3] = {T(k::I_TOR[0]), T(k::I_TOR[1]), T(k::I_TOR[2])};
  const T DT[3] = {T(k::D_TOR[0]), T(k::D_TOR[1]), T(k::D_TOR[2])};
  T tau[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const T rates_e = T(0) - (s.rpy[i] - s.lr[i]) * ctrl_freq;   // no angle wrapping (:248)
    s.lr[i] = s.rpy[i];
    T ir = clipv<T>(s.ir[i] - rot_e[i] * ctrl_dt, T(-1500), T(1500));
    if (i < 2) ir = clipv<T>(ir, T(-1), T(1));
    s.ir[i] = ir;
    tau[i] = clipv<T>(-PT[i] * rot_e[i] + DT[i] * rates_e + IT[i] * ir, T(-3200), T(3200));
  }
  // MIXER_MATRIX for CF2X (:48-53), PWM clip and PWM->RPM (:257-259)
  const T mx[4] = {T(-0.5), T(-0.5), T(0.5), T(0.5)};
  const T my[4] = {T(-0.5), T(0.5), T(0.5), T(-0.5)};
  const T mz[4] = {T(-1), T(1), T(-1), T(1)};
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    const T pwm = thrust + mx[m] * tau[0] + my[m] * tau[1] + mz[m] * tau[2];
    rpm[m] = T(k::PWM2RPM_SCALE) * clipv<T>(pwm, T(k::MIN_PWM), T(k::MAX_PWM)) + T(k::PWM2RPM_CONST);
  }
}

// BaseRLAviary._preprocessAction for one drone.  `act` points at this drone's A floats.
// numpy keeps float32 action arithmetic in float32 before widening (oracle preprocess_action), hence the
// explicit single-rounding intrinsics.
template <typename T>
__device__ __forceinline__ void action_to_rpm(const KParams& p, DroneState<T>& s, const float* __restrict__ act,
                                              T ctrl_dt, T rpm[4]) {
  const T cf = T(p.ctrl_freq);
  // every PID-driven action type reduces to (target position, target yaw, target velocity): one dsl_pid call site
  bool pid = true;
  T tp[3] = {s.p[0], s.p[1], s.p[2]}, tv[3] = {T(0), T(0), T(0)}, tyaw = T(0);
  if (p.act == 0) {            // VEL (:185-222)
    const float a0 = act[0], a1 = act[1], a3 = act[3];
    const float nrm = __fsqrt_rn(__fadd_rn(__fmul_rn(a0, a0), __fmul_rn(a1, a1)));
    float ux = 0.f, uy = 0.f;
    if (nrm != 0.f) { ux = __fdiv_rn(a0, nrm); uy = __fdiv_rn(a1, nrm); }
    const float scale = __fmul_rn(2.5f, fabsf(a3));
    tv[0] = T(ux) * T(scale); tv[1] = T(uy) * T(scale);
    tp[2] = T(k::TARGET_ALT);
    tyaw = s.rpy[2];
  } else if (p.act == 1) {     // RPM (:168-169)
    pid = false;
#pragma unroll
    for (int i = 0; i < 4; ++i) rpm[i] = T(k::HOVER_RPM) * T(__fadd_rn(1.0f, __fmul_rn(0.05f, act[i])));
  } else if (p.act == 3) {     // ONE_D_RPM (:223-224)
    pid = false;
    const T r = T(k::HOVER_RPM) * T(__fadd_rn(1.0f, __fmul_rn(0.05f, act[0])));
    rpm[0] = rpm[1] = rpm[2] = rpm[3] = r;
  } else if (p.act == 4) {     // ONE_D_PID (:225-234)
    tp[2] = s.p[2] + T(0.1) * T(act[0]);
  } else {                     // PID waypoint through _calculateNextStep, step_size 1 (BaseAviary.py:1318-1348)
    const T wp[3] = {T(act[0]), T(act[1]), T(act[2])};
    const T d[3] = {wp[0] - s.p[0], wp[1] - s.p[1], wp[2] - s.p[2]};
    const T dist = Mth<T>::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    if (dist <= T(1)) { tp[0] = wp[0]; tp[1] = wp[1]; tp[2] = wp[2]; }
    else {
#pragma unroll
      for (int i = 0; i < 3; ++i) tp[i] = s.p[i] + d[i] / dist;   // rare branch, exact
    }
  }
  if (pid) dsl_pid<T>(s, tp, tyaw, p.act == 0, tv, ctrl_dt, cf, rpm);
}

// K substeps of BaseAviary._dynamics (:1043-1102) + _integrateQ (:1104-1117) with constant rpm.
// The loop is the longest dependent chain of the whole step (K = 8 at 240/30 Hz), so it carries only what the next
// substep needs:
//   * the stored quaternion is unit (Bullet hands back a normalised one, BaseAviary.py:713), so getMatrixFromQuaternion's
//     s = 2/|q|^2 is 2 up to the rounding of that normalisation, and of the rotation matrix only the body z axis (thrust
//     direction) is used inside the loop; the full matrix is built once, after the loop, from the quaternion that entered
//     the last substep (ang_v = R_prev . rpy_rates, :1096-1100);
//   * the body rates do not depend on the attitude (torques are constant over the control step), so their chain runs
//     beside the quaternion chain; _integrateQ's isclose(|w|, 0) early-out (:1106) is a select, not a branch;
//   * the renormalisation 1/sqrt(n2) of a quaternion that one exact rotation step has moved off the unit sphere by
//     rounding only (n2 = 1 + O(1e-7)) is 1.5 - 0.5 n2 to O(1e-14) in the float build.
template <typename T>
__device__ __forceinline__ void dyn_substeps(DroneState<T>& s, const T rpm[4], int K, T dt) {
  // forces and body torques are constant over the K substeps (rpm is)
  T f[4], zt[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) { const T r2 = rpm[i] * rpm[i]; f[i] = r2 * T(k::KF); zt[i] = r2 * T(k::KM); }
  const T acc = (f[0] + f[1] + f[2] + f[3]) * T(1.0 / k::M_KG);      // thrust / m
  const T tz = -zt[0] + zt[1] - zt[2] + zt[3];
  // the sign of tx is the reference's (positive feedback on roll) -- do not "fix"
  const T tx = (f[0] + f[1] - f[2] - f[3]) * T(k::ARM / k::SQRT2);
  const T ty = (-f[0] + f[1] + f[2] - f[3]) * T(k::ARM / k::SQRT2);
  T x = s.q[0], y = s.q[1], z = s.q[2], ww = s.q[3];
  T px = x, py = y, pz = z, pw = ww;          // quaternion entering the last substep
  const T hdt = dt * T(0.5);
#pragma unroll 1
  for (int it = 0; it < K; ++it) {
    px = x; py = y; pz = z; pw = ww;
    // body z axis in world coordinates (third column of getMatrixFromQuaternion)
    const T zx = T(2) * (x * z + ww * y), zy = T(2) * (y * z - ww * x), zz = T(1) - T(2) * (x * x + y * y);
    const T Jw[3] = {T(k::IXX) * s.w[0], T(k::IYY) * s.w[1], T(k::IZZ) * s.w[2]};
    const T cr[3] = {s.w[1] * Jw[2] - s.w[2] * Jw[1], s.w[2] * Jw[0] - s.w[0] * Jw[2], s.w[0] * Jw[1] - s.w[1] * Jw[0]};
    s.v[0] = s.v[0] + dt * (zx * acc);
    s.v[1] = s.v[1] + dt * (zy * acc);
    s.v[2] = s.v[2] + dt * (zz * acc - T(k::GRAVITY / k::M_KG));
    s.w[0] = s.w[0] + dt * ((tx - cr[0]) * T(1.0 / k::IXX));
    s.w[1] = s.w[1] + dt * ((ty - cr[1]) * T(1.0 / k::IYY));
    s.w[2] = s.w[2] + dt * ((tz - cr[2]) * T(1.0 / k::IZZ));
#pragma unroll
    for (int i = 0; i < 3; ++i) s.p[i] = s.p[i] + dt * s.v[i];
    const T pp = s.w[0], qq = s.w[1], rr = s.w[2];
    const T wn2 = pp * pp + qq * qq + rr * rr;
    const bool moving = !(wn2 <= T(1e-16));            // np.isclose(norm, 0): atol 1e-8 (NaN takes the update, like numpy)
    const T iw = Mth<T>::rsqrt(wn2);
    T sn, c;
    Mth<T>::sincos_small(moving ? (wn2 * iw) * hdt : T(0), &sn, &c);     // |w| = 0: 0 * inf stays out of sincos
    const T sc = moving ? sn * iw : T(0);               // (2/|w|) * (1/2) * sin(theta)
    c = moving ? c : T(1);
    const T nx = c * x + sc * (rr * y - qq * z + pp * ww);
    const T ny = c * y + sc * (-rr * x + pp * z + qq * ww);
    const T nz = c * z + sc * (qq * x - pp * y + rr * ww);
    const T nw = c * ww + sc * (-pp * x - qq * y - rr * z);
    const T in = Mth<T>::rsqrt_near_one(nx * nx + ny * ny + nz * nz + nw * nw);   // Bullet returns a unit quaternion
    x = nx * in; y = ny * in; z = nz * in; ww = nw * in;
  }
  s.q[0] = x; s.q[1] = y; s.q[2] = z; s.q[3] = ww;
  // ang_v = R_prev . rpy_rates of the LAST substep (:1096-1100); earlier ones are overwritten
  if (K > 0) {
    const T pq[4] = {px, py, pz, pw};
    T R[3][3];
    quat_to_matrix<T>(pq, R);
#pragma unroll
    for (int i = 0; i < 3; ++i) s.av[i] = R[i][0] * s.w[0] + R[i][1] * s.w[1] + R[i][2] * s.w[2];
  }
  quat_to_euler<T>(s.q, s.rpy);
}

// The same K substeps with the opt-in aerodynamic extras (ChsParams.aero_flags).  The
// reference has these models for its PYB_* modes only (sb3_envs/BaseAviary.py:943-1039); as forces of the explicit
// DYN step they are defined by oracle/cattle_oracle.py:_dynamics, which this function follows term by term:
//   bit1 ground effect (:943-978)  per-prop extra thrust kf rpm_i^2 c_g (r_p/(4 h_i))^2, h_i = world height of
//        prop i (cf2x.urdf:42-78) clipped at GND_EFF_H_CLIP, gate |roll|,|pitch| < pi/2 <=> R[2][2] > 0; applied
//        at the props, so it enters the x/y torques too
//   bit0 drag (:982-1009)          world force -(c_xy,c_xy,c_z) * v * sum(2 pi rpm'/60), rpm' = RPMs of the
//        previous control step on the first substep, the current ones afterwards (:437,447)
//   bit2 downwash (:1013-1039)     body-z force -a exp(-(dxy/b)^2/2) from every drone above within 10 m; the other
//        drones' positions are the snapshot `others[j*G]` taken at the start of the control step
// Kept out of line: the DYN path proper does not pay registers or instruction-cache space for it.
template <typename T>
__device__ __noinline__ void dyn_substeps_aero(DroneState<T>& s, const T* rpm, const T* prev_rpm, int K, T dt, int flags,
                                               const typename VT<T>::v4* others, int G, int n, int self) {
  T f0[4], zt[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) { const T r2 = rpm[i] * rpm[i]; f0[i] = r2 * T(k::KF); zt[i] = r2 * T(k::KM); }
  const T tz = -zt[0] + zt[1] - zt[2] + zt[3];
  const T PX[4] = {T(0.028), T(-0.028), T(-0.028), T(0.028)};
  const T PY[4] = {T(-0.028), T(-0.028), T(0.028), T(0.028)};
  T R[3][3];
  for (int it = 0; it < K; ++it) {
    quat_to_matrix<T>(s.q, R);
    T f[4] = {f0[0], f0[1], f0[2], f0[3]};
    if ((flags & 2) && R[2][2] > T(0)) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const T h = Mth<T>::max(s.p[2] + R[2][0] * PX[i] + R[2][1] * PY[i], T(k::GND_EFF_H_CLIP));
        const T q = T(k::PROP_RADIUS) / (T(4) * h);
        f[i] = f[i] + rpm[i] * rpm[i] * T(k::KF) * T(k::GND_EFF_COEFF) * (q * q);
      }
    }
    T thrust = f[0] + f[1] + f[2] + f[3];
    if (flags & 4) {
      for (int j = 0; j < n; ++j) {
        const auto o = others[j * G];
        const T dz = o.z - s.p[2];
        const T ex = o.x - s.p[0], ey = o.y - s.p[1];
        const T dxy = Mth<T>::sqrt_exact(ex * ex + ey * ey);
        if (j != self && dz > T(0) && dxy < T(10)) {
          const T u = dxy / (T(k::DW2) * dz + T(k::DW3));
          const T wgt = Mth<T>::exp(T(-0.5) * (u * u));
          if (wgt > T(0)) {
            const T q = T(k::PROP_RADIUS) / (T(4) * dz);
            thrust = thrust - T(k::DW1) * (q * q) * wgt;
          }
        }
      }
    }
    T fw[3] = {R[0][2] * thrust, R[1][2] * thrust, R[2][2] * thrust - T(k::GRAVITY)};
    if (flags & 1) {
      const T* dr = (it == 0) ? prev_rpm : rpm;
      const T om = T(2.0 * k::PI / 60.0) * (dr[0] + dr[1] + dr[2] + dr[3]);
      fw[0] = fw[0] - T(k::DRAG_XY) * om * s.v[0];
      fw[1] = fw[1] - T(k::DRAG_XY) * om * s.v[1];
      fw[2] = fw[2] - T(k::DRAG_Z) * om * s.v[2];
    }
    const T tx = (f[0] + f[1] - f[2] - f[3]) * T(k::ARM / k::SQRT2);
    const T ty = (-f[0] + f[1] + f[2] - f[3]) * T(k::ARM / k::SQRT2);
    const T Jw[3] = {T(k::IXX) * s.w[0], T(k::IYY) * s.w[1], T(k::IZZ) * s.w[2]};
    const T cr[3] = {s.w[1] * Jw[2] - s.w[2] * Jw[1], s.w[2] * Jw[0] - s.w[0] * Jw[2], s.w[0] * Jw[1] - s.w[1] * Jw[0]};
    const T wd[3] = {(tx - cr[0]) * T(1.0 / k::IXX), (ty - cr[1]) * T(1.0 / k::IYY), (tz - cr[2]) * T(1.0 / k::IZZ)};
#pragma unroll
    for (int i = 0; i < 3; ++i) s.v[i] = s.v[i] + dt * (fw[i] * T(1.0 / k::M_KG));
#pragma unroll
    for (int i = 0; i < 3; ++i) s.w[i] = s.w[i] + dt * wd[i];
#pragma unroll
    for (int i = 0; i < 3; ++i) s.p[i] = s.p[i] + dt * s.v[i];
    const T wn = Mth<T>::sqrt(s.w[0] * s.w[0] + s.w[1] * s.w[1] + s.w[2] * s.w[2]);
    T x = s.q[0], y = s.q[1], z = s.q[2], ww = s.q[3];
    if (!(wn <= T(1e-8))) {
      const T pp = s.w[0], qq = s.w[1], rr = s.w[2];
      T sn, c;
      Mth<T>::sincos_small(wn * dt * T(0.5), &sn, &c);
      const T sc = sn * Mth<T>::frcp(wn);
      const T nx = c * x + sc * (rr * y - qq * z + pp * ww);
      const T ny = c * y + sc * (-rr * x + pp * z + qq * ww);
      const T nz = c * z + sc * (qq * x - pp * y + rr * ww);
      const T nw = c * ww + sc * (-pp * x - qq * y - rr * z);
      x = nx; y = ny; z = nz; ww = nw;
    }
    const T in = Mth<T>::rsqrt(x * x + y * y + z * z + ww * ww);
    s.q[0] = x * in; s.q[1] = y * in; s.q[2] = z * in; s.q[3] = ww * in;
  }
  if (K > 0) {
#pragma unroll
    for (int i = 0; i < 3; ++i) s.av[i] = R[i][0] * s.w[0] + R[i][1] * s.w[1] + R[i][2] * s.w[2];
  }
  quat_to_euler<T>(s.q, s.rpy);
}

}  // namespace chs
