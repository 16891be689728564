// mppi_pack.cuh — two samples per thread on Blackwell's packed FP32 pipe.
//
// sm_100 adds FFMA2 / FADD2 / FMUL2 (PTX fma/add/mul .f32x2): one issue slot performs the FP32
// operation on a 64-bit register pair, and a scalar register can be broadcast to both halves as an
// operand (`R.F32`), so warp-uniform parameters stay single registers.  The rollout kernel is
// issue-bound (84 % issue-active on the dynamic model, profiles/r01_cfg4_*): carrying two samples
// per thread as one f32x2 value halves the issue slots of every add / mul / fma; transcendental
// (MUFU), min/max, compare and integer work stays per lane.  IEEE round-to-nearest, so a lane
// computes bit-for-bit what the one-sample kernel computes for the same sequence of operations.
#pragma once
#include "mppi_device.cuh"

namespace mppi {

struct f32x2 {
  float2 v;
  __device__ __forceinline__ f32x2() {}
  __device__ __forceinline__ f32x2(float s) : v(make_float2(s, s)) {}   // broadcast
  __device__ __forceinline__ f32x2(double s) : v(make_float2(float(s), float(s))) {}
  __device__ __forceinline__ f32x2(int s) : v(make_float2(float(s), float(s))) {}
  __device__ __forceinline__ f32x2(float a, float b) : v(make_float2(a, b)) {}
  __device__ __forceinline__ explicit f32x2(float2 a) : v(a) {}
};
__device__ __forceinline__ f32x2 operator+(f32x2 a, f32x2 b) { return f32x2(__fadd2_rn(a.v, b.v)); }
__device__ __forceinline__ f32x2 operator*(f32x2 a, f32x2 b) { return f32x2(__fmul2_rn(a.v, b.v)); }
__device__ __forceinline__ f32x2 operator-(f32x2 a, f32x2 b) { return f32x2(__ffma2_rn(b.v, make_float2(-1.f, -1.f), a.v)); }   // exact: a + (-b)
__device__ __forceinline__ f32x2 operator-(f32x2 a) { return f32x2(__fmul2_rn(a.v, make_float2(-1.f, -1.f))); }   // exact, one slot
__device__ __forceinline__ f32x2& operator+=(f32x2& a, f32x2 b) { a = a + b; return a; }

struct Mask2 { bool a, b; };
struct Count2 { int a, b; };
__device__ __forceinline__ Mask2 m_and(Mask2 x, Mask2 y) { return Mask2{x.a && y.a, x.b && y.b}; }
__device__ __forceinline__ Mask2 m_or(Mask2 x, Mask2 y) { return Mask2{x.a || y.a, x.b || y.b}; }
__device__ __forceinline__ Count2& operator+=(Count2& x, Count2 y) { x.a += y.a; x.b += y.b; return x; }
__device__ __forceinline__ Count2 operator+(Count2 x, Count2 y) { return Count2{x.a + y.a, x.b + y.b}; }

template <> struct Math<f32x2> {
  using V = f32x2;
  using S = float;
  using F = Math<float>;
  static constexpr int LANES = 2;
  using Mask = Mask2;
  using Count = Count2;
  static __device__ __forceinline__ Mask2 lt(V a, V b) { return Mask2{a.v.x < b.v.x, a.v.y < b.v.y}; }
  static __device__ __forceinline__ Mask2 le(V a, V b) { return Mask2{a.v.x <= b.v.x, a.v.y <= b.v.y}; }
  static __device__ __forceinline__ Mask2 ge(V a, V b) { return Mask2{a.v.x >= b.v.x, a.v.y >= b.v.y}; }
  static __device__ __forceinline__ Mask2 gt(V a, V b) { return Mask2{a.v.x > b.v.x, a.v.y > b.v.y}; }
  static __device__ __forceinline__ Count2 count(Mask2 m) { return Count2{m.a ? 1 : 0, m.b ? 1 : 0}; }
  static __device__ __forceinline__ bool any(Mask2 m) { return m.a || m.b; }
  static __device__ __forceinline__ float lane(V x, int i) { return i ? x.v.y : x.v.x; }
  static __device__ __forceinline__ int lane(Count2 c, int i) { return i ? c.b : c.a; }
  static __device__ __forceinline__ V to_real(Count2 c) { return V(float(c.a), float(c.b)); }
  static __device__ __forceinline__ V make2(float a, float b) { return V(a, b); }

  static __device__ __forceinline__ V fma(V a, V b, V c) { return V(__ffma2_rn(a.v, b.v, c.v)); }
  static __device__ __forceinline__ V fabs(V x) { return V(fabsf(x.v.x), fabsf(x.v.y)); }
  static __device__ __forceinline__ V fmax(V a, V b) { return V(fmaxf(a.v.x, b.v.x), fmaxf(a.v.y, b.v.y)); }
  static __device__ __forceinline__ V fmin(V a, V b) { return V(fminf(a.v.x, b.v.x), fminf(a.v.y, b.v.y)); }
  static __device__ __forceinline__ void sincos(V x, V* s, V* c) {
    float s0, c0, s1, c1;
    __sincosf(x.v.x, &s0, &c0);
    __sincosf(x.v.y, &s1, &c1);
    *s = V(s0, s1); *c = V(c0, c1);
  }
  static __device__ __forceinline__ V sin(V x) { return V(__sinf(x.v.x), __sinf(x.v.y)); }
  static __device__ __forceinline__ V tan(V x) { return V(__tanf(x.v.x), __tanf(x.v.y)); }
  static __device__ __forceinline__ V div(V a, V b) { return V(__fdividef(a.v.x, b.v.x), __fdividef(a.v.y, b.v.y)); }
  static __device__ __forceinline__ V over(V a, float, float inv_b) { return a * V(inv_b); }

  // the odd degree-7 minimax polynomial of Math<float>::atan on both lanes at once
  static __device__ __forceinline__ V atan_poly(V z) {
    const V z2 = z * z;
    V p = V(-0.004054558929055929f);
    p = fma(p, z2, V(0.021862929686903954f));
    p = fma(p, z2, V(-0.05591228976845741f));
    p = fma(p, z2, V(0.09642194956541061f));
    p = fma(p, z2, V(-0.1390862911939621f));
    p = fma(p, z2, V(0.19946566224098206f));
    p = fma(p, z2, V(-0.33329859375953674f));
    p = fma(p, z2, V(0.9999993443489075f));
    return p * z;
  }
  static __device__ __forceinline__ V atan(V x) {
    const float a0 = fabsf(x.v.x), a1 = fabsf(x.v.y);
    const bool b0 = a0 > 1.0f, b1 = a1 > 1.0f;
    const V r = atan_poly(V(b0 ? __fdividef(1.0f, a0) : a0, b1 ? __fdividef(1.0f, a1) : a1));
    const float r0 = b0 ? (1.57079632679f - r.v.x) : r.v.x, r1 = b1 ? (1.57079632679f - r.v.y) : r.v.y;
    return V(copysignf(r0, x.v.x), copysignf(r1, x.v.y));
  }
  static __device__ __forceinline__ V atan_ratio(V n, V d) {
    const float an0 = fabsf(n.v.x), ad0 = fabsf(d.v.x), an1 = fabsf(n.v.y), ad1 = fabsf(d.v.y);
    const bool b0 = an0 > ad0, b1 = an1 > ad1;
    const V r = atan_poly(V(__fdividef(fminf(an0, ad0), fmaxf(an0, ad0)), __fdividef(fminf(an1, ad1), fmaxf(an1, ad1))));
    const float r0 = b0 ? (1.57079632679f - r.v.x) : r.v.x, r1 = b1 ? (1.57079632679f - r.v.y) : r.v.y;
    return V(__uint_as_float(__float_as_uint(r0) | ((__float_as_uint(n.v.x) ^ __float_as_uint(d.v.x)) & 0x80000000u)),
             __uint_as_float(__float_as_uint(r1) | ((__float_as_uint(n.v.y) ^ __float_as_uint(d.v.y)) & 0x80000000u)));
  }
};

template <class V> __device__ __forceinline__ typename Math<V>::Count hits_from_free(int tested, typename Math<V>::Count fr);
template <> __device__ __forceinline__ Count2 hits_from_free<f32x2>(int tested, Count2 fr) { return Count2{tested - fr.a, tested - fr.b}; }

}  // namespace mppi
