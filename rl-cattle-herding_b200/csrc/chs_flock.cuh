// Herd update: MathUtils + MathematicalFlock terms (utils/flockUtils.py:11-58,116-160,237-374) as driven by
// BaseAviary._flockingStep (sb3_envs/BaseAviary.py:1352-1400).  Scalar device functions; the kernel maps
// one (env, cow) pair per thread with the herd staged in shared memory.
#pragma once
#include "chs_layout.cuh"

namespace chs {

// MathUtils.bump_function (flockUtils.py:31-39), same assignment order: NaN -> 0
template <typename T> __device__ __forceinline__ T bump(T z) {
  if (z < T(0)) return T(0);
  if (z < T(k::BUMP_H)) return T(1);
  if (z <= T(1)) return T(0.5) * (T(1) + Mth<T>::cospi_unit((z - T(k::BUMP_H)) * T(1.0 / (1.0 - k::BUMP_H))));
  return T(0);
}

// One neighbour's contribution to (gradient + consensus) for displacement d = q_j - q_i and relative
// velocity dp = p_j - p_i:  phi_alpha(sigma_norm(d)) * sigma_norm_grad(d) + bump(sigma_norm(d)/r) * dp
// with a = b = 5, c = 0 and r == d (flockUtils.py:41-58,327-337,365-374).
// sigma_norm(d) = (sqrt(1+eps r^2)-1)/eps is evaluated as r^2/(sqrt(1+eps r^2)+1): same value, no cancellation.
// The herd update is scaled by dt^2 = 0.0025 before it reaches the velocity, so the 1-2 ulp MUFU
// approximations of the fp32 build are far inside the parity budget.
template <typename T>
__device__ __forceinline__ void lattice_term(T dx, T dy, T dpx, T dpy, T sigma_r, T inv_sigma_r,
                                             T& gx, T& gy, T& cx, T& cy) {
  const T r2 = dx * dx + dy * dy;
  const T q = T(1) + T(k::EPS) * r2;
  const T rs = Mth<T>::rsqrt(q);                // 1/sqrt(1+eps r^2)
  const T sn = r2 * Mth<T>::frcp(q * rs + T(1));
  const T b = bump<T>(sn * inv_sigma_r);
  const T zz = sn - sigma_r;
  const T w = b * T(k::PHI_HALF_AB) * zz * Mth<T>::rsqrt(T(1) + zz * zz) * rs;
  gx += w * dx; gy += w * dy;
  cx += b * dpx; cy += b * dpy;
}

// alpha term of cow i over every other cow (flockUtils.py:237-258); the herd is read from shared memory
// at stride G (cow-major, env-minor).  Returns C2_alpha*(grad + cons).
// Compact support: bump(sigma_norm(d)/sigma_norm(1.2)) == 0 for |d| >= 1.2 and it reaches 0 continuously, so
// pairs with r^2 >= 1.44 contribute nothing (the sensing range of 999 m never bites).  Lanes of a warp work on
// different envs, so the in-range set differs per lane: a cheap uniform pass collects the in-range cows of a
// 32-cow chunk in a bit mask, then every lane walks only its own set bits (a NaN distance stays in the set and
// propagates like in the reference).
// In-range set of cow i inside the 32-cow chunk starting at j0, as a bit mask (bit jj = cow j0 + jj; cow i itself
// excluded).  `pos(j)` returns the xy of cow j.  One definition for the step kernel and for the per-spawn-row table.
template <typename T, typename Pos>
__device__ __forceinline__ unsigned alpha_chunk_mask(Pos pos, typename VT<T>::v2 qi, int i, int j0, int M) {
  const int jn = (M - j0 < 32) ? (M - j0) : 32;
  unsigned mask = 0;
  for (int jj = 0; jj < jn; ++jj) {
    const auto qj = pos(j0 + jj);
    const T dx = qj.x - qi.x, dy = qj.y - qi.y;
    if (!(dx * dx + dy * dy >= T(k::ALPHA_RANGE * k::ALPHA_RANGE))) mask |= 1u << jj;
  }
  if (i >= j0 && i < j0 + 32) mask &= ~(1u << (i - j0));
  return mask;
}

// `mwords` (nullable): the in-range masks of cow i for this env's spawn row, one word per 32-cow chunk, tabulated by
// chs_set_spawn_table.  Cattle never move on the DYN path, so who is within 1.2 m of whom is a property
// of the spawn row; with the table the O(M^2) distance pass -- most of the herd update's instructions -- becomes M loads.
template <typename T>
__device__ __forceinline__ void alpha_term(const typename VT<T>::v2* s_cpos, const typename VT<T>::v2* s_cvel,
                                           int i, int g, int G, int M, const unsigned* __restrict__ mwords, T& ux, T& uy) {
  const auto qi = s_cpos[i * G + g];
  const auto pi = s_cvel[i * G + g];
  T gx = 0, gy = 0, cx = 0, cy = 0;
  for (int j0 = 0; j0 < M; j0 += 32) {
    unsigned mask;
    if (mwords) mask = __ldg(mwords + (j0 >> 5));
    else mask = alpha_chunk_mask<T>([&](int j) { return s_cpos[j * G + g]; }, qi, i, j0, M);
    while (mask) {
      const int j = j0 + __ffs(mask) - 1;
      mask &= mask - 1;
      const auto qj = s_cpos[j * G + g];
      const auto pj = s_cvel[j * G + g];
      lattice_term<T>(qj.x - qi.x, qj.y - qi.y, pj.x - pi.x, pj.y - pi.y, T(k::SIGMA_ALPHA), T(1.0 / k::SIGMA_ALPHA),
                      gx, gy, cx, cy);
    }
  }
  ux = T(k::C2_ALPHA) * gx + T(k::C2_ALPHA) * cx;
  uy = T(k::C2_ALPHA) * gy + T(k::C2_ALPHA) * cy;
}

// Weights of lattice_term without the accumulation: gradient weight `w` and consensus weight `b` of one displacement.
template <typename T>
__device__ __forceinline__ void lattice_weights(T dx, T dy, T sigma_r, T inv_sigma_r, T& w, T& b) {
  const T r2 = dx * dx + dy * dy;
  const T q = T(1) + T(k::EPS) * r2;
  const T rs = Mth<T>::rsqrt(q);
  const T sn = r2 * Mth<T>::frcp(q * rs + T(1));
  b = bump<T>(sn * inv_sigma_r);
  const T zz = sn - sigma_r;
  w = b * T(k::PHI_HALF_AB) * zz * Mth<T>::rsqrt(T(1) + zz * zz) * rs;
}

// alpha term from the per-spawn-row WEIGHT table (chs_set_spawn_table: alpha_weight_kernel).  Cattle never move on the DYN
// path, so for a herd that sits on its spawn row everything in the alpha term that depends on positions only is a constant
// of (row, cow): the whole gradient sum, and the bump weight of every in-range neighbour.  What is left per neighbour is
// the consensus piece b * (p_j - p_i) with the live velocities: 2 FMAs instead of four MUFU operations and ~40 instructions.
// Neighbours are walked in mask order, the order the table was filled in, so the sums are those of alpha_term.
template <typename T>
__device__ __forceinline__ void alpha_term_tab(const typename VT<T>::v2* s_cvel, int i, int g, int G, int M,
                                               const unsigned* __restrict__ mwords, typename VT<T>::v2 grad,
                                               const T* __restrict__ wts, T& ux, T& uy) {
  const auto pi = s_cvel[i * G + g];
  T cx = 0, cy = 0;
  int kk = 0;
  for (int j0 = 0; j0 < M; j0 += 32) {
    unsigned mask = __ldg(mwords + (j0 >> 5));
    while (mask) {
      const int j = j0 + __ffs(mask) - 1;
      mask &= mask - 1;
      const T b = __ldg(wts + kk);
      ++kk;
      const auto pj = s_cvel[j * G + g];
      cx += b * (pj.x - pi.x); cy += b * (pj.y - pi.y);
    }
  }
  ux = T(k::C2_ALPHA) * grad.x + T(k::C2_ALPHA) * cx;
  uy = T(k::C2_ALPHA) * grad.y + T(k::C2_ALPHA) * cy;
}

// alpha_term for herds of more than 32 cows with the mask table: the first four mask words (up to 128 cows) are fetched
// together, one L2 round trip instead of one per 32-cow chunk (the loads of the plain version are serialised by the walk over
// each chunk's set bits).  Out of line on purpose: the small-herd kernels are sensitive to every instruction added to their
// straight-line path (+2 % at 65 536 envs with this code inlined).
template <typename T>
__device__ __noinline__ void alpha_term_wide(const typename VT<T>::v2* s_cpos, const typename VT<T>::v2* s_cvel,
                                             int i, int g, int G, int M, const unsigned* __restrict__ mwords, T* out) {
  const auto qi = s_cpos[i * G + g];
  const auto pi = s_cvel[i * G + g];
  T gx = 0, gy = 0, cx = 0, cy = 0;
  unsigned mk[4];
  mk[0] = __ldg(mwords); mk[1] = __ldg(mwords + 1);
  mk[2] = (M > 64) ? __ldg(mwords + 2) : 0u;
  mk[3] = (M > 96) ? __ldg(mwords + 3) : 0u;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    unsigned mask = mk[c];
    while (mask) {
      const int j = c * 32 + __ffs(mask) - 1;
      mask &= mask - 1;
      const auto qj = s_cpos[j * G + g];
      const auto pj = s_cvel[j * G + g];
      lattice_term<T>(qj.x - qi.x, qj.y - qi.y, pj.x - pi.x, pj.y - pi.y, T(k::SIGMA_ALPHA), T(1.0 / k::SIGMA_ALPHA),
                      gx, gy, cx, cy);
    }
  }
  for (int j0 = 128; j0 < M; j0 += 32) {
    unsigned mask = __ldg(mwords + (j0 >> 5));
    while (mask) {
      const int j = j0 + __ffs(mask) - 1;
      mask &= mask - 1;
      const auto qj = s_cpos[j * G + g];
      const auto pj = s_cvel[j * G + g];
      lattice_term<T>(qj.x - qi.x, qj.y - qi.y, pj.x - pi.x, pj.y - pi.y, T(k::SIGMA_ALPHA), T(1.0 / k::SIGMA_ALPHA),
                      gx, gy, cx, cy);
    }
  }
  out[0] = T(k::C2_ALPHA) * gx + T(k::C2_ALPHA) * cx;
  out[1] = T(k::C2_ALPHA) * gy + T(k::C2_ALPHA) * cy;
}

// Herd centroid of the G envs of a CTA for herds of 48 cows and more: 8 lanes per env taken from the END of the worker range
// (`wrev` = 0 on the last worker thread, `nw` worker threads, every one of them calls), cows strided over the lanes,
// xor-shuffle tree.  Out of line for the same reason as alpha_term_wide.
template <typename T>
__device__ __noinline__ void centroid_wide(const typename VT<T>::v2* s_cpos, double2* s_ccent, int G, int M, int wrev, int nw) {
  const double inv_m = 1.0 / (double)M;
  for (int base = 0; base < 8 * G; base += nw) {
    const int idx = base + wrev, g = idx >> 3, sub = idx & 7;
    const bool live = idx < 8 * G;
    double cx = 0, cy = 0;
    if (live) {
#pragma unroll 2
      for (int j = sub; j < M; j += 8) { const auto c = s_cpos[j * G + g]; cx += (double)c.x; cy += (double)c.y; }
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
      cx += __shfl_xor_sync(0xffffffffu, cx, o);
      cy += __shfl_xor_sync(0xffffffffu, cy, o);
    }
    if (live && sub == 0) s_ccent[g] = make_double2(cx * inv_m, cy * inv_m);
  }
}

// shepherd (delta) interaction + predator repulsion + gamma term of cow i (flockUtils.py:271-323,343-348,
// 150-160), then the velocity update and speed clip of BaseAviary.py:1384-1395.  `ua` is the alpha term.
template <typename T>
__device__ __forceinline__ typename VT<T>::v2 herd_finish(const typename VT<T>::v4* s_dpos, int n, int g, int G,
                                                          typename VT<T>::v2 qi, typename VT<T>::v2 pi, T uax, T uay) {
  T ux = uax, uy = uay;
  if (n > 0) {
    T gx = 0, gy = 0, cx = 0, cy = 0, px = 0, py = 0;
    for (int kk = 0; kk < n; ++kk) {
      const auto yk = s_dpos[kk * G + g];
      const T ex = qi.x - yk.x, ey = qi.y - yk.y;            // diff = q_i - y_k
      const T r = Mth<T>::sqrt(ex * ex + ey * ey);
      const T dist = r + T(1e-6);
      const T mu = Mth<T>::min(dist * T(1.0 / k::SHEPHERD_R), T(1));
      const T idist = Mth<T>::frcp(dist);
      const T ax = ex * idist, ay = ey * idist;
      const T ap = ax * pi.x + ay * pi.y;
      const T pikx = mu * (pi.x - ax * ap), piky = mu * (pi.y - ay * ap);   // mu * (I - a a^T) p_i
      const T om = T(1) - mu;                                 // q_ik - q_i = (1-mu)(y_k - q_i)
      if (om == T(0)) {
        // drone farther than the shepherd radius (the usual case): the projected point coincides with the cow,
        // sigma_norm = 0, bump = 1 and the gradient weight multiplies a zero displacement -- lattice_term would add
        // exactly (0, 0) to the gradient and 1 * (p_ik - p_i) to the consensus
        cx += pikx - pi.x; cy += piky - pi.y;
      } else {
        lattice_term<T>(-om * ex, -om * ey, pikx - pi.x, piky - pi.y, T(k::SIGMA_BETA), T(1.0 / k::SIGMA_BETA), gx, gy, cx, cy);
      }
      if (r <= T(k::DANGER_RANGE)) {                          // -k (y_k - q_i)/|.|^3 ; r == 0 -> NaN like numpy
        const T inv = T(k::PREDATOR_K) * Mth<T>::frcp(r * r * r);
        px += inv * ex; py += inv * ey;
      }
    }
    ux += T(k::C2_BETA) * gx + T(k::C2_BETA) * cx;
    uy += T(k::C2_BETA) * gy + T(k::C2_BETA) * cy;
    ux += px; uy += py;
  }
  {  // gamma: -C1*sigma_1(q_i - target) - C2*p_i, target = scalar 1 broadcast
    const T zx = qi.x - T(k::CONSENSUS), zy = qi.y - T(k::CONSENSUS);
    ux += -T(k::C1_GAMMA) * (zx * Mth<T>::rsqrt(T(1) + zx * zx)) - T(k::C2_GAMMA) * pi.x;
    uy += -T(k::C1_GAMMA) * (zy * Mth<T>::rsqrt(T(1) + zy * zy)) - T(k::C2_GAMMA) * pi.y;
  }
  T vx = pi.x + ux * T(k::FLOCK_DT2), vy = pi.y + uy * T(k::FLOCK_DT2);
  const T sp = Mth<T>::sqrt(vx * vx + vy * vy);
  if (sp > T(k::MAX_VEL_CATTLE)) { const T f = T(k::MAX_VEL_CATTLE) / sp; vx *= f; vy *= f; }   // exact divide: feeds the state
  return mk2<T>(vx, vy);
}

// evaluate_herding_effectiveness for one cow: winding number w.r.t. the polygon drones[0..n-1] in index
// order (utils/evaluation.py:100-138,271-273).  Returns 1 if the cow is inside.
template <typename T>
__device__ __forceinline__ int cow_inside(const typename VT<T>::v4* s_dpos, int n, int g, int G, T px, T py) {
  int wn = 0;
  if (n <= 0) return 0;
  auto a = s_dpos[g];
  for (int i = 0; i < n; ++i) {
    const int i2 = (i + 1 == n) ? 0 : i + 1;
    const auto b = s_dpos[i2 * G + g];
    const T il = (b.x - a.x) * (py - a.y) - (px - a.x) * (b.y - a.y);
    if (a.y <= py) { if (b.y > py && il > T(0)) ++wn; }
    else           { if (b.y <= py && il < T(0)) --wn; }
    a = b;
  }
  return wn != 0;
}

}  // namespace chs
