// Scalar math layer of the simulator kernels: one code path, two arithmetic types.
// T = float is the product; T = double is the parity instrument (tests compare it with the float64
// oracle at 1e-9).  Quaternions are (x,y,z,w) like pybullet.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "chs_consts.h"

namespace chs {

template <typename T> struct VT;
template <> struct VT<float> { using v2 = float2; using v4 = float4; };
template <> struct VT<double> { using v2 = double2; using v4 = double4; };

template <typename T> __device__ __forceinline__ typename VT<T>::v4 mk4(T x, T y, T z, T w) {
  typename VT<T>::v4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r;
}
template <typename T> __device__ __forceinline__ typename VT<T>::v2 mk2(T x, T y) {
  typename VT<T>::v2 r; r.x = x; r.y = y; return r;
}

template <typename T> struct Mth;
template <> struct Mth<float> {
  // sqrt.approx.f32: MUFU.SQRT, max relative error 2^-23 (IEEE sqrtf costs a ~10 instruction sequence)
  static __device__ __forceinline__ float sqrt(float x) { float r; asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
  static __device__ __forceinline__ float rcp(float x) { return __frcp_rn(x); }
  // 1-2 ulp approximations on the MUFU pipe; the fp32 parity budget is 1e-5 relative
  static __device__ __forceinline__ float rsqrt(float x) { return rsqrtf(x); }
  // 1/sqrt(x) for x = 1 + O(1e-7) (renormalisation of a quaternion that left the unit sphere by rounding only): the first
  // order expansion is exact to O((x-1)^2) and keeps the MUFU latency off the substep chain; anything else takes MUFU.RSQ
  static __device__ __forceinline__ float rsqrt_near_one(float x) {
    return (fabsf(x - 1.0f) < 1e-3f) ? fmaf(-0.5f, x, 1.5f) : rsqrtf(x);
  }
  static __device__ __forceinline__ float fdiv(float a, float b) { return __fdividef(a, b); }
  static __device__ __forceinline__ float frcp(float x) { return __fdividef(1.0f, x); }
  // cos(pi*x) for x in [0,1] (the bump function's argument): MUFU.COS, abs error ~5e-7
  static __device__ __forceinline__ float cospi_unit(float x) { return __cosf(3.14159265358979f * x); }
  // sin/cos of the half rotation angle of one 1/240 s substep (|x| < 0.5 unless the drone spins > 240 rad/s)
  static __device__ __forceinline__ void sincos_small(float x, float* s, float* c) {
    if (fabsf(x) < 0.5f) {
      const float x2 = x * x;
      // Taylor to x^7 / x^8: the first dropped terms are x^9/9! < 5.4e-9 and x^10/10! < 2.7e-10 at |x| = 0.5
      *s = x * (1.0f + x2 * (-1.0f / 6.0f + x2 * (1.0f / 120.0f + x2 * (-1.0f / 5040.0f))));
      *c = 1.0f + x2 * (-0.5f + x2 * (1.0f / 24.0f + x2 * (-1.0f / 720.0f + x2 * (1.0f / 40320.0f))));
    } else {
      sincosf(x, s, c);
    }
  }
  static __device__ __forceinline__ void sincos(float x, float* s, float* c) { sincosf(x, s, c); }
  static __device__ __forceinline__ void sincospi(float x, float* s, float* c) { sincospif(x, s, c); }
  static __device__ __forceinline__ float cospi(float x) { return cospif(x); }
  // ex2.approx based: relative error ~2^-21 for the |x| <~ 10 arguments of the reward shapes
  static __device__ __forceinline__ float exp(float x) { return __expf(x); }
  static __device__ __forceinline__ float atan2(float y, float x) { return atan2f(y, x); }
  static __device__ __forceinline__ float asin(float x) { return asinf(x); }
  static __device__ __forceinline__ float abs(float x) { return fabsf(x); }
  static __device__ __forceinline__ float min(float a, float b) { return fminf(a, b); }
  static __device__ __forceinline__ float max(float a, float b) { return fmaxf(a, b); }
  static __device__ __forceinline__ float sqrt_exact(float x) { return sqrtf(x); }
  static __device__ __forceinline__ float inf() { return __int_as_float(0x7f800000); }
  static __device__ __forceinline__ float nan() { return __int_as_float(0x7fc00000); }
};
template <> struct Mth<double> {
  static __device__ __forceinline__ double sqrt(double x) { return ::sqrt(x); }
  static __device__ __forceinline__ double rcp(double x) { return 1.0 / x; }
  static __device__ __forceinline__ double rsqrt(double x) { return 1.0 / ::sqrt(x); }
  static __device__ __forceinline__ double rsqrt_near_one(double x) { return 1.0 / ::sqrt(x); }
  static __device__ __forceinline__ double fdiv(double a, double b) { return a / b; }
  static __device__ __forceinline__ double frcp(double x) { return 1.0 / x; }
  static __device__ __forceinline__ double cospi_unit(double x) { return ::cospi(x); }
  static __device__ __forceinline__ void sincos_small(double x, double* s, double* c) { ::sincos(x, s, c); }
  static __device__ __forceinline__ void sincos(double x, double* s, double* c) { ::sincos(x, s, c); }
  static __device__ __forceinline__ void sincospi(double x, double* s, double* c) { ::sincospi(x, s, c); }
  static __device__ __forceinline__ double cospi(double x) { return ::cospi(x); }
  static __device__ __forceinline__ double exp(double x) { return ::exp(x); }
  static __device__ __forceinline__ double atan2(double y, double x) { return ::atan2(y, x); }
  static __device__ __forceinline__ double asin(double x) { return ::asin(x); }
  static __device__ __forceinline__ double abs(double x) { return ::fabs(x); }
  static __device__ __forceinline__ double min(double a, double b) { return ::fmin(a, b); }
  static __device__ __forceinline__ double max(double a, double b) { return ::fmax(a, b); }
  static __device__ __forceinline__ double sqrt_exact(double x) { return ::sqrt(x); }
  static __device__ __forceinline__ double inf() { return __longlong_as_double(0x7ff0000000000000LL); }
  static __device__ __forceinline__ double nan() { return __longlong_as_double(0x7ff8000000000000LL); }
};

// python-style clip used by the oracle (`lo if v < lo else (hi if v > hi else v)`): NaN passes through
template <typename T> __device__ __forceinline__ T clipv(T v, T lo, T hi) {
  return v < lo ? lo : (v > hi ? hi : v);
}

// ---- pybullet helper semantics (Bullet btMatrix3x3::setRotation) ----------
// third column of getMatrixFromQuaternion(q): the body z axis in world coordinates
template <typename T> __device__ __forceinline__ void quat_zaxis(const T q[4], T z[3]) {
  const T x = q[0], y = q[1], zz = q[2], w = q[3];
  const T s = T(2) * Mth<T>::frcp(x * x + y * y + zz * zz + w * w);
  const T xs = x * s, ys = y * s, zs = zz * s;
  z[0] = x * zs + w * ys;
  z[1] = y * zs - w * xs;
  z[2] = T(1) - (x * xs + y * ys);
}

template <typename T> __device__ __forceinline__ void quat_to_matrix(const T q[4], T R[3][3]) {
  const T x = q[0], y = q[1], z = q[2], w = q[3];
  const T s = T(2) * Mth<T>::frcp(x * x + y * y + z * z + w * w);
  const T xs = x * s, ys = y * s, zs = z * s;
  const T wx = w * xs, wy = w * ys, wz = w * zs;
  const T xx = x * xs, xy = x * ys, xz = x * zs;
  const T yy = y * ys, yz = y * zs, zz = z * zs;
  R[0][0] = T(1) - (yy + zz); R[0][1] = xy - wz;          R[0][2] = xz + wy;
  R[1][0] = xy + wz;          R[1][1] = T(1) - (xx + zz); R[1][2] = yz - wx;
  R[2][0] = xz - wy;          R[2][1] = yz + wx;          R[2][2] = T(1) - (xx + yy);
}

// getEulerFromQuaternion: ZYX with the +-0.99999 gimbal guard (pybullet.c)
template <typename T> __device__ __forceinline__ void quat_to_euler(const T q[4], T rpy[3]) {
  const T x = q[0], y = q[1], z = q[2], w = q[3];
  const T sarg = T(-2) * (x * z - w * y);
  if (sarg <= T(-0.99999)) {
    rpy[0] = T(0); rpy[1] = T(-0.5 * k::PI); rpy[2] = T(2) * Mth<T>::atan2(x, -y);
  } else if (sarg >= T(0.99999)) {
    rpy[0] = T(0); rpy[1] = T(0.5 * k::PI); rpy[2] = T(2) * Mth<T>::atan2(-x, y);
  } else {
    rpy[0] = Mth<T>::atan2(T(2) * (y * z + w * x), w * w - x * x - y * y + z * z);
    rpy[1] = Mth<T>::asin(sarg);
    rpy[2] = Mth<T>::atan2(T(2) * (x * y + w * z), w * w + x * x - y * y - z * z);
  }
}

// ---- counter-based RNG: Philox4x32-10 (Salmon et al., SC'11) ---------------------------------------
// The reference draws from the process-global Mersenne Twister streams (`random.randint`,
// `np.random.rand`, BaseAviary.py:307,617,631), which cannot be reproduced per environment; resets
// here draw from Philox keyed by (seed) with counter (global env id, episode, index, stream tag).
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)M0 * c[0];
    const uint64_t p1 = (uint64_t)M1 * c[2];
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
    const uint32_t n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
    const uint32_t n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += W0; k1 += W1;
  }
}
// 24-bit uniform in [0,1): exactly representable in fp32 and fp64
__host__ __device__ __forceinline__ double u01_24(uint32_t x) { return (double)(x >> 8) * (1.0 / 16777216.0); }

enum PhiloxTag { TAG_COW = 0, TAG_NDRONES = 1 };

}  // namespace chs
