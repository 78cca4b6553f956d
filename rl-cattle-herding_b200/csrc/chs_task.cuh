// Task layer: reward shapes, success / truncation tests and observation pieces of
// sb3_envs/CattleAviary.py:213-332,422-552,572-679 and sb3_envs/BaseRLAviary.py:272-342.
#pragma once
#ifndef CHS_EXP
#define CHS_EXP 0
#endif
#include "chs_layout.cuh"

namespace chs {

// The curriculum rows and 1/n (n = 0..12 drones) live in one float64 table in global memory (KArgs::tab,
// CHS_TAB_DOUBLES entries, filled by the host from CHS_CURRICULUM_ROWS) that every CTA copies to shared memory in
// phase 0, together with its other loads.  A __constant__ table indexed by the per-env level costs a constant-
// cache miss (a DRAM round trip when L2 is cold) on the serial per-env path; the shared copy costs one LDS.
constexpr int CHS_TAB_INV_N = k::NUM_LEVELS * k::C_NCOLS;
constexpr int CHS_TAB_APPROACH = CHS_TAB_INV_N + CHS_MAX_DRONES + 1;   // 5 / (SPEED_LIMIT / ctrl_freq + 1e-6)  (CattleAviary.py:253-256)
constexpr int CHS_TAB_INV_M = CHS_TAB_APPROACH + 1;                    // 1 / num_cattle
constexpr int CHS_TAB_DOUBLES = CHS_TAB_INV_M + 1;

template <typename T> __device__ __forceinline__ T curr(const double* ct, int level, int col) { return T(ct[level * k::C_NCOLS + col]); }

// CattleAviary.SimpleSpacingReward (:594-626)
template <typename T> __device__ __forceinline__ T simple_spacing(T r, T desired, T tolfrac) {
  const T tol = desired * tolfrac;
  const T lo = desired - tol, hi = desired + tol;
  if (lo <= r && r <= hi) return T(1);
  // quotients through Mth::fdiv (MUFU reciprocal in the float build, exact in the double build): the shape is in [-1, 1]
  if (r < lo) return T(-1) + Mth<T>::fdiv(r, lo) * T(2);
  if (r > hi) return T(1) - Mth<T>::fdiv(r - hi, T(7.0) - hi) * T(2);
  return T(-1);
}

// CattleAviary.DroneSpacingRewardFunction (:630-667)
template <typename T> __device__ __forceinline__ T complex_spacing(T r, T d_star) {
  // divisions by constants are multiplications by the reciprocal (one rounding apart from the reference's quotient)
  const T zz = (r - d_star) * T(1.0 / (0.4 + 1e-9));
  const T gauss = Mth<T>::exp(T(-0.5) * zz * zz);
  const T coll = (r < T(0.3)) ? T(-1.0) * (T(1.0) - (r * T(1.0 / (0.3 + 1e-9)))) : T(0);
  const T pull = (r > T(1.5)) ? T(-0.3) * (r - T(1.5)) * T(1.0 / (5.0 - 1.5)) : T(0);
  return gauss + coll + pull + T(0.1) * (T(1) - Mth<T>::abs(r - d_star));
}

// CattleAviary.CattleSpacingRewardFunction (:572-592)
template <typename T> __device__ __forceinline__ T cattle_spacing(T r) {
  if (r <= T(k::SP_R0)) {
    const T a = r - T(k::SP_D);
    return T(k::SP_A) * Mth<T>::exp(-(a * a) * T(1.0 / (2 * k::SP_C * k::SP_C))) -
           T(k::SP_B) * Mth<T>::exp(-(r * r) * T(1.0 / (2 * k::SP_K * k::SP_K)));
  }
  // g(r0)/exp(-lam*r0) * exp(-lam*r); the prefactor is a constant of the shape parameters
  return T(2.662959707043491) * Mth<T>::exp(T(-k::SP_LAM) * r);
}

// Per-drone results handed from the (env, drone) threads to the per-env combine through shared memory.
//   redA = (simple(d1), simple(d2), complex(d1), complex(d2))
//   redB = (cattle_spacing(min cow distance), d1 = distance to the nearest other drone, altitude-violation flag,
//           d2 = distance to the second nearest)
// The obs row (BaseRLAviary.py:286-334) is [z rpy vel ang_v | rel-xy of the 2 nearest drones, stable order |
// zeros up to 12 neighbour slots | rel-xy of cows 0..15 | zeros]; the own-state part is written by the thread
// that integrated the drone, the neighbour part by drone_neighbors, the cattle part by drone_cattle.  The two
// functions touch disjoint shared-memory words, so different threads may run them concurrently.

template <typename T>
__device__ __forceinline__ void drone_neighbors(const KParams& p, const double* ct, int g, int d, int n, int level, bool want_reward,
                                                const typename VT<T>::v4* s_dpos, typename VT<T>::v4* s_redA,
                                                typename VT<T>::v4* s_redB, float* s_stage, int* dbg_nb /* nullable: [2] */,
                                                long long* clk = nullptr /* profiling: three clock64() stamps */) {
  const int G = p.G;
  const auto me = s_dpos[d * G + g];
  // two nearest other drones; strict '<' keeps index order on ties (stable sort, :310-311);
  // a NaN distance never wins '<', and numpy's stable argsort puts NaN last as well
  T d1 = Mth<T>::inf(), d2 = Mth<T>::inf();
  int j1 = -1, j2 = -1;
  // branch free: the drone itself enters with an infinite distance, which never wins a strict '<' (lanes of a warp hold
  // different drones, so a `continue` on j == d splits the warp in every iteration)
#pragma unroll 1
  for (int j = 0; j < n; ++j) {
    const auto o = s_dpos[j * G + g];
    const T dx = o.x - me.x, dy = o.y - me.y;
    const T dist = (j == d) ? Mth<T>::inf() : Mth<T>::sqrt(dx * dx + dy * dy);
    const bool lt1 = dist < d1, lt2 = dist < d2;
    d2 = lt1 ? d1 : (lt2 ? dist : d2);
    j2 = lt1 ? j1 : (lt2 ? j : j2);
    d1 = lt1 ? dist : d1;
    j1 = lt1 ? j : j1;
  }
  if (clk) clk[0] = clock64();
  float* row = s_stage + (size_t)g * p.stage_stride + d * p.L;
  const int nb_slots = (p.L - 10) / 2;       // row truncated to L (BaseRLAviary.py:336-340)
  const bool al8 = (p.L & 1) == 0;
  if (j1 >= 0 && nb_slots > 0) {
    const auto o = s_dpos[j1 * G + g];
    if (al8) *reinterpret_cast<float2*>(row + 10) = make_float2((float)(o.x - me.x), (float)(o.y - me.y));
    else { row[10] = (float)(o.x - me.x); row[11] = (float)(o.y - me.y); }
  }
  if (j2 >= 0 && nb_slots > 1) {
    const auto o = s_dpos[j2 * G + g];
    if (al8) *reinterpret_cast<float2*>(row + 12) = make_float2((float)(o.x - me.x), (float)(o.y - me.y));
    else { row[12] = (float)(o.x - me.x); row[13] = (float)(o.y - me.y); }
  }
  if (dbg_nb) { dbg_nb[0] = j1; dbg_nb[1] = j2; }
  if (clk) clk[1] = clock64();
  if (want_reward) {
    const T desired = curr<T>(ct, level, k::C_DESIRED), tolf = curr<T>(ct, level, k::C_TOL);
    s_redA[d * G + g] = mk4<T>(simple_spacing<T>(d1, desired, tolf), simple_spacing<T>(d2, desired, tolf),
                               complex_spacing<T>(d1, desired), complex_spacing<T>(d2, desired));
    T* rb = reinterpret_cast<T*>(&s_redB[d * G + g]);
    rb[1] = d1;
    rb[2] = (Mth<T>::abs(me.z - T(k::TARGET_ALT)) > T(k::MAX_ALT_ERROR)) ? T(1) : T(0);
    rb[3] = d2;
  }
  if (clk) clk[2] = clock64();
}

// cattle: relative xy of cows 0..15 into the obs row, minimum distance over ALL cows for the reward
template <typename T>
__device__ __forceinline__ void drone_cattle(const KParams& p, int g, int d, bool want_reward,
                                             const typename VT<T>::v4* s_dpos, const typename VT<T>::v2* s_cpos,
                                             typename VT<T>::v4* s_redB, float* s_stage) {
  const int G = p.G;
  const auto me = s_dpos[d * G + g];
  float* row = s_stage + (size_t)g * p.stage_stride + d * p.L;
  T m2 = Mth<T>::inf();
  // CTDE pads the neighbours to 12 slots, not 4 (sb3_envs/BaseRLAviary.py:313-314); the MARL observation pads
  // to MAX_NEIGHBORS = 4 (rllib_envs/BaseMARLAviary.py:276-279)
  const int cbase = (p.packing == CHS_PACK_MARL) ? 10 + 2 * 4 : 10 + 2 * CHS_MAX_DRONES;
  const int nobs = (p.M < 16) ? p.M : 16;
#pragma unroll 4
  for (int j = 0; j < nobs; ++j) {
    const auto c = s_cpos[j * G + g];
    const T dx = c.x - me.x, dy = c.y - me.y;
    const int col = cbase + 2 * j;
    if (col + 1 < p.L) {
      // one 8-byte store per cow where the row is 8-byte aligned (L even): half the shared-memory wavefronts and
      // 2-way instead of 4-way bank conflicts between lanes (envs are a multiple-of-4 stride apart)
      if ((p.L & 1) == 0) *reinterpret_cast<float2*>(row + col) = make_float2((float)dx, (float)dy);
      else { row[col] = (float)dx; row[col + 1] = (float)dy; }
    } else if (col < p.L) { row[col] = (float)dx; }
    m2 = Mth<T>::min(m2, dx * dx + dy * dy);
  }
  if (want_reward) {
#pragma unroll 4
    for (int j = nobs; j < p.M; ++j) {
      const auto c = s_cpos[j * G + g];
      const T dx = c.x - me.x, dy = c.y - me.y;
      m2 = Mth<T>::min(m2, dx * dx + dy * dy);
    }
    reinterpret_cast<T*>(&s_redB[d * G + g])[0] = cattle_spacing<T>(Mth<T>::sqrt(m2));
  }
}

// The same work spread over C = 2^k lanes of one warp per (drone, env) item (lanes sub = 0..C-1 are consecutive): lane
// `sub` takes cows sub, sub + C, ...; the minimum is combined with xor-shuffles, so EVERY lane of the warp has to call
// this function the same number of times (`valid` = false: the lane only takes part in the shuffles).  min() is
// order independent, so the result is bit for bit the serial one; with 100 cows the per-drone loop is the longest chain
// of phase 2 otherwise (BASELINE config 3).
template <typename T>
__device__ __forceinline__ void drone_cattle_coop(const KParams& p, bool valid, int g, int d, int sub, int C, bool want_reward,
                                                  const typename VT<T>::v4* s_dpos, const typename VT<T>::v2* s_cpos,
                                                  typename VT<T>::v4* s_redB, float* s_stage) {
  const int G = p.G;
  T m2 = Mth<T>::inf();
  if (valid) {
    const auto me = s_dpos[d * G + g];
    float* row = s_stage + (size_t)g * p.stage_stride + d * p.L;
    const int cbase = (p.packing == CHS_PACK_MARL) ? 10 + 2 * 4 : 10 + 2 * CHS_MAX_DRONES;
    const int nobs = (p.M < 16) ? p.M : 16;
    const int jend = want_reward ? p.M : nobs;
    const bool al8 = (p.L & 1) == 0;
    if (al8 && cbase + 2 * nobs <= p.L) {
      // the usual shape: every observed cow has both columns inside the row and the row is 8-byte aligned -- one float2 store
      // per cow, no per-cow column tests; cows past the observed 16 only enter the minimum
      float2* r2 = reinterpret_cast<float2*>(row + cbase);
      int j = sub;
      for (; j < nobs; j += C) {
        const auto c = s_cpos[j * G + g];
        const T dx = c.x - me.x, dy = c.y - me.y;
        r2[j] = make_float2((float)dx, (float)dy);
        m2 = Mth<T>::min(m2, dx * dx + dy * dy);
      }
      for (; j < jend; j += C) {
        const auto c = s_cpos[j * G + g];
        const T dx = c.x - me.x, dy = c.y - me.y;
        m2 = Mth<T>::min(m2, dx * dx + dy * dy);
      }
    } else {
      for (int j = sub; j < jend; j += C) {
        const auto c = s_cpos[j * G + g];
        const T dx = c.x - me.x, dy = c.y - me.y;
        if (j < nobs) {
          const int col = cbase + 2 * j;
          if (col + 1 < p.L) {
            if (al8) *reinterpret_cast<float2*>(row + col) = make_float2((float)dx, (float)dy);
            else { row[col] = (float)dx; row[col + 1] = (float)dy; }
          } else if (col < p.L) { row[col] = (float)dx; }
        }
        m2 = Mth<T>::min(m2, dx * dx + dy * dy);
      }
    }
  }
  for (int o = C >> 1; o > 0; o >>= 1) {
    const T other = __shfl_xor_sync(0xffffffffu, m2, o);
    // fmin semantics: a NaN distance loses against a number on either side, like the serial loop
    m2 = Mth<T>::min(m2, other);
  }
  if (valid && want_reward && sub == 0) reinterpret_cast<T*>(&s_redB[d * G + g])[0] = cattle_spacing<T>(Mth<T>::sqrt(m2));
}

// CattleAviary._computeTerminated (:422-492) and MARLCattleAviary._computeTerminated (:246-321); stateful
// through the spacing clock (levels 0/1).  `ms` = minimum pairwise drone xy spacing, `cent` = centroid
// distance, `eff` = effectiveness in percent, `clock_inc` = what one CALL adds to the clock: a hard-coded
// 1/240 in the CTDE class (:441, two calls per step), 1/CTRL_FREQ in the MARL class (:268, 2N calls per step).
template <typename T>
__device__ __forceinline__ bool compute_terminated(const double* ct, int level, T ms, T cent, T eff, T& clock, T clock_inc) {
  if (level <= 1) {
    const T desired = curr<T>(ct, level, k::C_DESIRED), tol = curr<T>(ct, level, k::C_TOL);
    const T hi = desired + desired * tol, lo = desired - desired * tol;
    if (ms < hi && ms > lo) {
      clock += clock_inc;
      if (clock >= curr<T>(ct, level, k::C_HOLD)) return true;
    } else {
      clock = T(0);
    }
  } else if (level <= 3) {
    if (cent < curr<T>(ct, level, k::C_APPROACH_MIN)) return true;
  } else if (level == 4 || level == 6) {
    if (eff > curr<T>(ct, level, k::C_MIN_EFF)) return true;
  } else if (level == 5) {
    if (eff > curr<T>(ct, level, k::C_MIN_EFF)) {
      const T desired = curr<T>(ct, level, k::C_CATTLE_DESIRED), tol = curr<T>(ct, level, k::C_CATTLE_TOL);
      const T hi = desired + desired * tol, lo = desired - desired * tol;
      if (ms < hi && ms > lo) return true;
    }
  }
  return false;   // level 7 has no branch
}

// MARLCattleAviary._endOfEpisodeReward (:180-241) for drone `me`; d1/d2 = its two nearest drone distances.
__device__ __forceinline__ double marl_end_of_episode(const double* ct, int level, int n, double d1, double d2, double cent, double eff,
                                                      double dist_to_herd) {
  if (level <= 1 || level == 5) {
    if (level == 5 && !(eff > ct[level * k::C_NCOLS + (k::C_MIN_EFF)])) return 0.0;
    const double desired = ct[level * k::C_NCOLS + (level == 5 ? k::C_CATTLE_DESIRED : k::C_DESIRED)];
    const double tol = ct[level * k::C_NCOLS + (level == 5 ? k::C_CATTLE_TOL : k::C_TOL)];
    const double hi = desired + desired * tol, lo = desired - desired * tol;
    return (d1 >= lo && d1 <= hi && d2 >= lo && d2 <= hi) ? 50.0 / (double)n : 0.0;
  }
  if (level <= 3) return (cent < ct[level * k::C_NCOLS + (k::C_APPROACH_MIN)]) ? 50.0 : 0.0;
  if (level == 4 || level == 6) return eff * 2 * clipv<double>(1.0 - dist_to_herd / 10.0, 0.0, 1.0);
  return 0.0;
}

}  // namespace chs
