// Constants of the RL-Cattle-Herding hot path, with the reference line each one comes from.
// (paths relative to /root/reference/gym_pybullet_drones/).  Shared by host and device code.
#pragma once

namespace chs {
namespace k {

// ---- assets/cf2x.urdf:5,11-12 ; sb3_envs/BaseAviary.py:97,162-165 --------------------------------
constexpr double M_KG = 0.027;
constexpr double ARM = 0.0397;
constexpr double KF = 3.16e-10;
constexpr double KM = 7.94e-12;
constexpr double T2W = 2.25;
constexpr double MAX_SPEED_KMH = 30.0;
constexpr double IXX = 1.4e-5, IYY = 1.4e-5, IZZ = 2.17e-5;
constexpr double G_ACC = 9.8;
constexpr double GRAVITY = G_ACC * M_KG;
constexpr double HOVER_RPM = 14468.429183500699;   // sqrt(GRAVITY/(4*KF)), BaseAviary.py:163
constexpr double SQRT2 = 1.4142135623730951;
constexpr double PI = 3.141592653589793;
constexpr double TARGET_ALT = 0.45;                // DRONE_TARGET_ALTITUDE, BaseAviary.py:111
constexpr double SPEED_LIMIT = 0.3 * MAX_SPEED_KMH * (1000.0 / 3600.0);  // BaseRLAviary.py:102
constexpr double DRONE_SPAWN_SPACING = 1.75;       // BaseAviary.py:251
constexpr double COW_SPAWN_Z = 0.1;                // BaseAviary.py:614
// PYB-only aerodynamic extras (cf2x.urdf:11), used only when aero_flags != 0
constexpr double GND_EFF_COEFF = 11.36859;
constexpr double PROP_RADIUS = 2.31348e-2;
constexpr double DRAG_XY = 9.1785e-7, DRAG_Z = 10.311e-7;
constexpr double DW1 = 2267.18, DW2 = 0.16, DW3 = -0.11;
// 0.25 r_p sqrt(15 MAX_RPM^2 kf c_g / MAX_THRUST) = 0.25 r_p sqrt(15 c_g / 4)   (BaseAviary.py:173)
constexpr double GND_EFF_H_CLIP = 0.03776371349209501;

// ---- control/DSLPIDControl.py:37-53 -------------------------------------------------------------
constexpr double P_FOR[3] = {0.4, 0.4, 1.25};
constexpr double I_FOR[3] = {0.05, 0.05, 0.05};
constexpr double D_FOR[3] = {0.2, 0.2, 0.5};
constexpr double P_TOR[3] = {70000.0, 70000.0, 60000.0};
constexpr double I_TOR[3] = {0.0, 0.0, 500.0};
constexpr double D_TOR[3] = {20000.0, 20000.0, 12000.0};
constexpr double PWM2RPM_SCALE = 0.2685;
constexpr double PWM2RPM_CONST = 4070.3;
constexpr double MIN_PWM = 20000.0, MAX_PWM = 65535.0;

// ---- utils/flockUtils.py:13-19,62-74 ; sb3_envs/BaseAviary.py:51-55,579,1353-1355 ------------------
constexpr double EPS = 0.1;
constexpr double BUMP_H = 0.2;
constexpr double PHI_HALF_AB = 5.0;                // ((a+b)/2 with a=b=5; (a-b)/2 = 0; c = 0)
constexpr double C2_ALPHA = 3.4641016151377544;    // 2*sqrt(3)
constexpr double C1_GAMMA = 5.0;
constexpr double C2_GAMMA = 0.447213595499958;    // 0.2*sqrt(5)
constexpr double C2_BETA = 8.94427190999916;       // 2*sqrt(20)
constexpr double ALPHA_RANGE = 1.2;                // == ALPHA_DIST
constexpr double BETA_RANGE = 1.0;                 // == BETA_DIST
constexpr double SIGMA_ALPHA = 0.6957935656967495; // (sqrt(1+0.1*1.2^2)-1)/0.1
constexpr double SIGMA_BETA = 0.4880884817015163; // (sqrt(1+0.1*1.0^2)-1)/0.1
constexpr double DANGER_RANGE = 1.1;
constexpr double CONSENSUS = 1.0;
constexpr double PREDATOR_K = 650000.0;
constexpr double SHEPHERD_R = 1.0;
constexpr double FLOCK_DT2 = 0.05 * 0.05;          // BaseAviary.py:1354-1355
constexpr double MAX_VEL_CATTLE = 0.2;             // BaseAviary.py:579

// ---- sb3_envs/CattleAviary.py:91-105 -------------------------------------------------------------
constexpr double MISSION_BOUNDARY = 15.0;
constexpr double MAX_FORMATION_DISTANCE = 8.0;
constexpr double COLLISION_THRESHOLD = 0.2;
constexpr double MAX_ALT_ERROR = TARGET_ALT * 0.6;
constexpr double SP_A = 1.2, SP_B = 2.1, SP_C = 3.3, SP_K = 0.2, SP_D = -1.0, SP_R0 = 1.3, SP_LAM = 0.8;

// ---- utils/curriculum_learning.py:10-194: one row per level ----------------------------------------
// columns: desired, tol, hold, approach_min, min_eff, cattle_desired, cattle_tol, min_drones, max_drones,
//          episode_length, w_simple, w_complex, w_survival, w_approach, w_eff, w_c2d, required_tally
enum CurrCol { C_DESIRED = 0, C_TOL, C_HOLD, C_APPROACH_MIN, C_MIN_EFF, C_CATTLE_DESIRED, C_CATTLE_TOL,
               C_MIN_DRONES, C_MAX_DRONES, C_EP_LEN, C_W_SIMPLE, C_W_COMPLEX, C_W_SURV, C_W_APPROACH,
               C_W_EFF, C_W_C2D, C_REQ_TALLY, C_NCOLS };
constexpr int NUM_LEVELS = 8;
// level 0 has no `cattle_approach_min_distance` key (typo at curriculum_learning.py:18); it is only read at
// levels 2/3, so any value works there.
#define CHS_CURRICULUM_ROWS                                                                          \
  {0.8, 0.3, 10, 0.0, 0.0, 0.0, 0.0, 3, 3, 40, 1.0, 0.0, 0.0, 0.0, 0.0, 0.0, 100},                   \
  {0.8, 0.2, 25, 0.0, 0.0, 0.0, 0.0, 4, 4, 40, 0.0, 1.0, -0.5, 0.0, 0.0, 0.0, 300},                  \
  {0.8, 0.2, 15, 0.6, 0.0, 0.0, 0.0, 4, 4, 40, 0.0, 0.8, 0.0, 1.0, 0.0, 0.0, 100},                   \
  {0.8, 0.2, 15, 0.3, 0.0, 0.0, 0.0, 4, 4, 40, 0.0, 0.8, -0.5, 1.0, 0.0, 0.0, 400},                  \
  {0.8, 0.2, 15, 0.3, 20.0, 0.0, 0.0, 4, 4, 80, 0.0, 0.7, -0.0, 0.8, 1.0, 0.0, 600},                 \
  {0.8, 0.2, 15, 0.3, 50.0, 0.8, 0.1, 4, 4, 40, 0.0, 0.7, -0.5, 0.6, 1.0, 0.8, 600},                 \
  {0.8, 0.3, 15, 0.2, 50.0, 0.0, 0.0, 4, 12, 80, 0.7, 0.0, -0.0, 0.8, 1.0, 0.0, 600},                \
  {0.8, 0.3, 15, 0.2, 50.0, 0.0, 0.0, 4, 12, 80, 0.0, 0.0, -0.0, 1.0, 1.0, 0.0, 600}

}  // namespace k
}  // namespace chs
