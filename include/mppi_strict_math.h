/*
 * mppi_strict_math.h — FP32 elementary functions for the `strict_math` flavour (mppi_config_t).
 *
 * north_star: "argmin sample index and collision flags must be bit-exact".  In FP32 that is only
 * well-posed when the GPU and the CPU checker execute the SAME sequence of IEEE-754 operations:
 * glibc's sinf/atanf and CUDA's differ in the last bit, and one flipped bit of a squared distance
 * flips a collision flag.  These functions use nothing but round-to-nearest +, -, *, / and exact
 * float<->int conversions, so any conforming compiler that does not contract a*b+c produces the
 * same bits: on the device the number type maps every operator to __fadd_rn / __fmul_rn /
 * __fdiv_rn (never fused, never approximated); on the host it is plain `float` compiled with
 * -ffp-contract=off.  Accuracy: <= 2 ulp on the argument ranges of the MPPI path (angles of a few
 * radians) — the same class as libm, so strict FP32 costs stay within 1e-4 of the FP64 path.
 *
 * Polynomials after Cephes sinf/cosf/atanf (S. Moshier; public domain).  Header-only, C++11,
 * templated on the number type T; T needs + - * /, unary -, comparison, and the four helpers
 * below (overloaded for float here, for the device wrapper in csrc/mppi_strict.cuh).
 */
#ifndef MPPI_STRICT_MATH_H
#define MPPI_STRICT_MATH_H

#if defined(__CUDACC__)
#define MPPI_SM_HD __host__ __device__ __forceinline__
#else
#define MPPI_SM_HD inline
#endif

namespace mppi_strict {

/* helpers for plain float (host) */
MPPI_SM_HD float sm_abs(float x) { return x < 0.0f ? -x : x; }
MPPI_SM_HD int sm_to_int(float x) { return (int)x; }          /* exact for |x| < 2^24 integral */
MPPI_SM_HD float sm_lit(float, float c) { return c; }          /* literal in the number type */

/* nearest integer (ties to even) of x for |x| < 2^22, by the 1.5 * 2^23 trick: two exact-by-IEEE adds */
template <class T>
MPPI_SM_HD T sm_rint(T x) {
  const T magic = sm_lit(x, 12582912.0f);
  return (x + magic) - magic;
}

/* sin and cos of x (radians).  k = rint(x * 2/pi), r = x - k*pi/2 in three Cody-Waite pieces,
 * kernels on [-pi/4, pi/4], quadrant from k mod 4. */
template <class T>
MPPI_SM_HD void sm_sincos(T x, T* s_out, T* c_out) {
  const T kf = sm_rint(x * sm_lit(x, 0.63661977236758134f));
  T r = x - kf * sm_lit(x, 1.5703125f);
  r = r - kf * sm_lit(x, 4.837512969970703125e-4f);
  r = r - kf * sm_lit(x, 7.54978995489188216e-8f);
  const T z = r * r;
  T ps = sm_lit(x, -1.9515295891e-4f);
  ps = ps * z + sm_lit(x, 8.3321608736e-3f);
  ps = ps * z + sm_lit(x, -1.6666654611e-1f);
  const T sn = ps * z * r + r;
  T pc = sm_lit(x, 2.443315711809948e-5f);
  pc = pc * z + sm_lit(x, -1.388731625493765e-3f);
  pc = pc * z + sm_lit(x, 4.166664568298827e-2f);
  const T cs = pc * z * z - sm_lit(x, 0.5f) * z + sm_lit(x, 1.0f);
  const int q = sm_to_int(kf) & 3;
  if (q == 0) { *s_out = sn; *c_out = cs; }
  else if (q == 1) { *s_out = cs; *c_out = -sn; }
  else if (q == 2) { *s_out = -sn; *c_out = -cs; }
  else { *s_out = -cs; *c_out = sn; }
}
template <class T> MPPI_SM_HD T sm_sin(T x) { T s, c; sm_sincos(x, &s, &c); return s; }
template <class T> MPPI_SM_HD T sm_cos(T x) { T s, c; sm_sincos(x, &s, &c); return c; }
template <class T> MPPI_SM_HD T sm_tan(T x) { T s, c; sm_sincos(x, &s, &c); return s / c; }

/* atan(x): reduce to [0, tan(pi/8)] by x -> -1/x (beyond tan(3pi/8)) or (x-1)/(x+1), odd kernel */
template <class T>
MPPI_SM_HD T sm_atan(T x) {
  const bool neg = x < sm_lit(x, 0.0f);
  T a = sm_abs(x);
  T y0;
  if (a > sm_lit(x, 2.414213562373095f)) { y0 = sm_lit(x, 1.5707963267948966f); a = -(sm_lit(x, 1.0f) / a); }
  else if (a > sm_lit(x, 0.4142135623730950f)) { y0 = sm_lit(x, 0.7853981633974483f); a = (a - sm_lit(x, 1.0f)) / (a + sm_lit(x, 1.0f)); }
  else y0 = sm_lit(x, 0.0f);
  const T z = a * a;
  T p = sm_lit(x, 8.05374449538e-2f);
  p = p * z - sm_lit(x, 1.38776856032e-1f);
  p = p * z + sm_lit(x, 1.99777106478e-1f);
  p = p * z - sm_lit(x, 3.33329491539e-1f);
  const T y = y0 + (p * z * a + a);
  return neg ? -y : y;
}

}  /* namespace mppi_strict */
#endif /* MPPI_STRICT_MATH_H */
