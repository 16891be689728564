// mppi_models.cuh — vehicle models, Heun step, obstacle/bound tests, Philox.
#pragma once
#include "mppi_device.cuh"

namespace mppi {

// ---------------------------------------------------------------------------
// Kinematic bicycle — OptimalControl/Parking/src/VehicleModel.jl:1-21.
// State [x, y, psi, ux, sa], control [sr, ax].
// Heun step (MPPIUtils.jl:76-90) with the structure of this model exploited:
//   * ux and sa have constant derivatives over the step, so the Euler predictor
//     ux + ax*dt and sa + sr*dt ARE the corrected values (bit-identical: the
//     corrector is x + dt/2*(k+k)); tan(sa_next) and sincos(psi_next) are
//     carried to the next step's k1, so each step costs 2 sincos + 1 tan.
// ---------------------------------------------------------------------------
template <class R>
struct Kin5 {
  static constexpr int NX = 5;
  static constexpr int IPSI = 2;
  R s[NX];
  R sn, cs;   // sin/cos of s[IPSI], carried between steps
  R tn;       // tan(sa) / L, carried between steps

  __device__ __forceinline__ void init(const StaticParams<R>& p, const double* x0) {
#pragma unroll
    for (int i = 0; i < NX; ++i) s[i] = R(x0[i]);
    Math<R>::sincos(s[IPSI], &sn, &cs);
    tn = Math<R>::over(Math<R>::tan(s[4]), p.Lwb, p.inv_Lwb);
  }
  // stage cost of VehicleDynamics at (state, u): SURVEY 8(a) a6 = dynamic cost with v = r = 0
  template <class P>
  __device__ __forceinline__ R stage_cost(const P& p, R sr, R ax) const {
    return p.w[2] * (ax * ax) + p.w[3] * (sr * sr) + p.w[4] * (s[4] * s[4]) + p.w[5] * (s[1] * s[1]);
  }
  template <class P>
  __device__ __forceinline__ void step(const P& p, R sr, R ax) {
    // k1 at (s, u)
    const R k1x = s[3] * cs, k1y = s[3] * sn, k1p = s[3] * tn;   // ux cosψ, ux sinψ, ux/L tan(sa)
    // predictor x + k1*dt
    const R psit = s[2] + k1p * p.dt;
    const R uxt = s[3] + ax * p.dt;
    const R sat = s[4] + sr * p.dt;
    R st, ct;
    Math<R>::sincos(psit, &st, &ct);
    const R tnt = Math<R>::over(Math<R>::tan(sat), p.Lwb, p.inv_Lwb);
    const R k2x = uxt * ct, k2y = uxt * st, k2p = uxt * tnt;
    // corrector x + dt*(k1+k2)/2
    s[0] = s[0] + p.half_dt * (k1x + k2x);
    s[1] = s[1] + p.half_dt * (k1y + k2y);
    s[2] = s[2] + p.half_dt * (k1p + k2p);
    s[3] = uxt;
    s[4] = sat;
    tn = tnt;
    Math<R>::sincos(s[2], &sn, &cs);
  }
};

// ---------------------------------------------------------------------------
// Dynamic bicycle with Pacejka-style tyres — OptimalControl/MPPI/src/vehicledynamics.jl:1-54.
// State [x, y, v, r, psi, ux, sa], control [sr, ax];
// derivative order [dx, dy, dv, dr, dpsi, du, dsa] (:50) == state order.
// ---------------------------------------------------------------------------
template <class R>
struct Dyn7 {
  static constexpr int NX = 7;
  static constexpr int IPSI = 4;
  R s[NX];
  R sn, cs;

  __device__ __forceinline__ void init(const StaticParams<R>& p, const double* x0) {
#pragma unroll
    for (int i = 0; i < NX; ++i) s[i] = R(x0[i]);
    Math<R>::sincos(s[IPSI], &sn, &cs);
  }
  template <class P>
  __device__ __forceinline__ R stage_cost(const P& p, R sr, R ax) const {  // :52
    return s[2] * s[2] * p.w[0] + p.w[1] * (s[3] * s[3]) + p.w[2] * (ax * ax) + p.w[3] * (sr * sr) +
           p.w[4] * (s[6] * s[6]) + p.w[5] * (s[1] * s[1]);
  }
  // lateral tyre force (:35-38): mu*FZ*D*sin(C*atan(X - E*(X - atan X))), X = B*alpha
  template <class P>
  static __device__ __forceinline__ R tyre(const P& p, R FZ, R alpha) {
    const R X = p.tireB * alpha;
    const R inner = X - p.tireE * (X - Math<R>::atan(X));
    return p.mu * FZ * p.tireD * Math<R>::sin(p.tireC * Math<R>::atan(inner));
  }
  // derivatives of (v, r) and the planar velocity at state q with heading sin/cos (sq, cq)
  template <class P>
  static __device__ __forceinline__ void deriv(const P& p, R v, R r, R ux, R sa, R sq, R cq,
                                               R ax, R* dx, R* dy, R* dv, R* dr) {
    const R t = ax - r * v;
    const R FZF = R(2) * (p.FZF0 - t * p.KFZX);                              // :30
    const R FZR = R(2) * (p.FZR0 + t * p.KFZX);                              // :31
    const R den = ux + R(0.01);
    const R alpha_f = Math<R>::atan_ratio(v + p.la * r, den) - sa;           // :32
    const R alpha_r = Math<R>::atan_ratio(v - p.lb * r, den);                // :33
    const R FY1 = tyre(p, FZF, alpha_f);
    const R FY2 = tyre(p, FZR, alpha_r);
    const R uxc = ux <= R(0) ? R(0) : ux;                                    // :40-42
    *dx = uxc * cq - v * sq;                                                 // :43
    *dy = uxc * sq + v * cq;                                                 // :44
    *dv = Math<R>::div(FY1 + FY2, p.mass) - r * uxc;                         // :46
    *dr = Math<R>::div(FY1 * p.la - FY2 * p.lb, p.Izz);                      // :47
  }
  template <class P>
  __device__ __forceinline__ void step(const P& p, R sr, R ax) {
    R k1x, k1y, k1v, k1r;
    deriv(p, s[2], s[3], s[5], s[6], sn, cs, ax, &k1x, &k1y, &k1v, &k1r);
    const R k1p = s[3];                                                      // dpsi = r (:45)
    const R vt = s[2] + k1v * p.dt, rt = s[3] + k1r * p.dt;
    const R psit = s[4] + k1p * p.dt;
    const R uxt = s[5] + ax * p.dt, sat = s[6] + sr * p.dt;
    R st, ct;
    Math<R>::sincos(psit, &st, &ct);
    R k2x, k2y, k2v, k2r;
    deriv(p, vt, rt, uxt, sat, st, ct, ax, &k2x, &k2y, &k2v, &k2r);
    s[0] = s[0] + p.half_dt * (k1x + k2x);
    s[1] = s[1] + p.half_dt * (k1y + k2y);
    s[2] = s[2] + p.half_dt * (k1v + k2v);
    s[3] = s[3] + p.half_dt * (k1r + k2r);
    s[4] = s[4] + p.half_dt * (k1p + rt);
    s[5] = uxt;   // == ux + dt/2*(ax+ax)
    s[6] = sat;
    Math<R>::sincos(s[4], &sn, &cs);
  }
};

// ---------------------------------------------------------------------------
// Obstacles.  Returns the number of obstacles hit by the pose (each hit costs
// obs_pen and clears the constraint flag: MPPIUtils.jl:125-127).
// ---------------------------------------------------------------------------
// The circle table is padded on the host to a multiple of 4 with never-hit entries
// (r^2 = -1), so the common small counts run as fully unrolled groups whose table
// entries are immediate constant-bank operands (no per-circle load / loop overhead).
template <class R, int Q>
__device__ __forceinline__ int circle_hits_quads(const StaticParams<R>& p, R x, R y) {
  int hits = 0;
#pragma unroll
  for (int o = 0; o < 4 * Q; ++o) {
    const R dx = x - p.circ[o][0], dy = y - p.circ[o][1];
    hits += (dx * dx + dy * dy <= p.circ[o][2]) ? 1 : 0;   // sum((s-o).^2) <= r^2
  }
  return hits;
}
// the first 4*Q circles of a register-resident table (horizon loop of the rollout kernel)
template <class R, int Q, class P>
__device__ __forceinline__ int circle_hits_regs(const P& hp, R px, R py) {
  int h = 0;
#pragma unroll
  for (int o = 0; o < 4 * Q; ++o) {
    const R dx = px - hp.circ[o][0], dy = py - hp.circ[o][1];
    h += (dx * dx + dy * dy <= hp.circ[o][2]) ? 1 : 0;
  }
  return h;
}

template <class R>
__device__ __forceinline__ int circle_hits(const StaticParams<R>& p, R x, R y) {
  switch ((p.n_circles + 3) >> 2) {   // warp-uniform
    case 0: return 0;
    case 1: return circle_hits_quads<R, 1>(p, x, y);
    case 2: return circle_hits_quads<R, 2>(p, x, y);
    default: break;
  }
  int hits = 0;
  for (int o = 0; o < p.n_circles; ++o) {
    const R dx = x - p.circ[o][0], dy = y - p.circ[o][1];
    hits += (dx * dx + dy * dy <= p.circ[o][2]) ? 1 : 0;
  }
  return hits;
}

// Rectangle-vs-rectangle with the reference's ConvexCollision semantics
// (CollisionDetection/src/utils.jl:37-74): "free" needs a separating axis among
// the WALL's edge normals AND one among the VEHICLE's edge normals; touching
// counts as separated (<=).  For rectangles the 4 edge normals reduce to 2 axes
// and the 5-point min/max projections to centre +/- radius, so the literal
// 4 x 10 x 2 projections collapse to 4 dot products and 4 interval tests.
// Vehicle rectangle (hybrid_astar_utils.jl:193): centre = (x, y) + off*(cosψ, sinψ),
// half extents (hl, hw); modπ(ψ) leaves cosψ/sinψ unchanged.
// Broad phase voted across the warp (samples of a warp share x0 and differ only by noise, so
// whole warps skip): one bounding circle around all walls, then per wall when there are many;
// then the exact test.
// `tbl` is the wall table [n][8] (cx, cy, cosΨ, sinΨ, half_len, half_wid, broad-phase d², pad):
// the kernel parameter itself, or — in the rollout kernel — a shared-memory copy read as two
// broadcast 128-bit loads per wall (a parameter table indexed by a loop counter costs an LDC
// with register index per field).  V carries veh_off / veh_hl / veh_hw (StaticParams or HotParams).
template <class R, class V>
__device__ __forceinline__ int block_hits(const V& veh, const R* all4, const R* tbl, int n_blocks, R x, R y, R sn, R cs) {
  using R4 = typename Vec4<R>::type;
  int hits = 0;
  const R vx = Math<R>::fma(veh.veh_off, cs, x), vy = Math<R>::fma(veh.veh_off, sn, y);
  const unsigned mask = __activemask();
  {
    const R ax = vx - all4[0], ay = vy - all4[1];
    if (__ballot_sync(mask, Math<R>::fma(ax, ax, ay * ay) < all4[2]) == 0u) return 0;   // far from every wall
  }
  // exact rectangle-rectangle test against wall b
  auto exact = [&](const R4 w0, const R4 w1, R dx, R dy) {
    const R C = w0.z, S = w0.w, wl = w1.x, ww = w1.y;
    const R ac = Math<R>::fabs(Math<R>::fma(cs, C, sn * S));    // |cos(ψ-Ψ)|
    const R as = Math<R>::fabs(Math<R>::fma(sn, C, -(cs * S))); // |sin(ψ-Ψ)|
    const bool sepW = (Math<R>::fabs(Math<R>::fma(dx, C, dy * S)) >= Math<R>::fma(veh.veh_hl, ac, Math<R>::fma(veh.veh_hw, as, wl))) ||
                      (Math<R>::fabs(Math<R>::fma(dy, C, -(dx * S))) >= Math<R>::fma(veh.veh_hl, as, Math<R>::fma(veh.veh_hw, ac, ww)));
    const bool sepV = (Math<R>::fabs(Math<R>::fma(dx, cs, dy * sn)) >= Math<R>::fma(wl, ac, Math<R>::fma(ww, as, veh.veh_hl))) ||
                      (Math<R>::fabs(Math<R>::fma(dy, cs, -(dx * sn))) >= Math<R>::fma(wl, as, Math<R>::fma(ww, ac, veh.veh_hw)));
    return (sepW && sepV) ? 0 : 1;   // !ConvexCollision => hit (hybrid_astar_utils.jl:198)
  };
  const R4* t4 = reinterpret_cast<const R4*>(tbl);
  if (n_blocks <= 4) {   // few walls: straight loop the compiler can unroll and software-pipeline
#pragma unroll 4
    for (int b = 0; b < n_blocks; ++b) {
      const R4 w0 = t4[2 * b], w1 = t4[2 * b + 1];
      hits += exact(w0, w1, vx - w0.x, vy - w0.y);
    }
  } else {               // many walls: per-wall bounding-circle vote first
    for (int b = 0; b < n_blocks; ++b) {
      const R4 w0 = t4[2 * b], w1 = t4[2 * b + 1];
      const R dx = vx - w0.x, dy = vy - w0.y;
      if (__ballot_sync(mask, Math<R>::fma(dx, dx, dy * dy) < w1.z) == 0u) continue;   // provably free for every lane
      hits += exact(w0, w1, dx, dy);
    }
  }
  return hits;
}
template <class R>
__device__ __forceinline__ int block_hits(const StaticParams<R>& p, R x, R y, R sn, R cs) {
  return block_hits<R>(p, p.blk_all, &p.blk[0][0], p.n_blocks, x, y, sn, cs);
}

// BoundEvaluation (MPPIUtils.jl:135-151): sum of violations |x - XL|, |x - XU|
// (strict inequalities; the slack factor is applied once at the end).
// XL <= XU, so at most one side is violated: max(XL - s, s - XU, 0) equals the sum of
// the two one-sided terms exactly (the other term is +0).
template <class R, int NX, class P>
__device__ __forceinline__ R bound_violation(const P& p, const R* s) {
  R acc = R(0);
#pragma unroll
  for (int i = 0; i < NX; ++i)
    acc += Math<R>::fmax(Math<R>::fmax(p.XL[i] - s[i], s[i] - p.XU[i]), R(0));
  return acc;
}

// ---------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. SC'11), counter = (global k, j/2, iteration,
// instance), key = seed.  One call yields the noise of two horizon steps.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return c;
}

// same function with the 10 round keys precomputed (warp-uniform and identical for every
// call of a launch; 20 UIADD3 issue slots per call otherwise)
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, const uint32_t* key) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ key[2 * r], lo1, hi0 ^ c.w ^ key[2 * r + 1], lo0);
  }
  return c;
}

// Box-Muller: two 32-bit words -> two N(0,1) (same formula as oracle::box_muller)
__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float* z0, float* z1) {
  const float u = __uint2float_rn(a) * 2.3283064365386963e-10f + 1.1641532182693481e-10f;
  const float th = __uint2float_rn(b) * 1.4629180792671596e-09f + (-3.1415926535897931f + 7.3145903963357981e-10f);
  float lg, s;   // sqrt(-2 ln u) = sqrt(-2 ln2 * log2 u); u >= 2^-33 so the .ftz forms are exact-domain
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(u));
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(-1.3862943611198906f * lg));
  float sn, cs;
  __sincosf(th, &sn, &cs);
  *z0 = s * sn;
  *z1 = s * cs;
}

// order-preserving map float/double -> unsigned (NaN, either sign, sorts last,
// like Base.isless used by argmin: types.jl:9)
__device__ __forceinline__ uint32_t ordered_bits(float c) {
  if (c != c) return 0xFFFFFFFFu;
  const uint32_t b = __float_as_uint(c);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ unsigned long long ordered_bits(double c) {
  if (c != c) return 0xFFFFFFFFFFFFFFFFull;
  const unsigned long long b = (unsigned long long)__double_as_longlong(c);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ float unordered_bits(uint32_t k) {
  if (k == 0xFFFFFFFFu) return __uint_as_float(0x7FC00000u);
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}
__device__ __forceinline__ double unordered_bits(unsigned long long k) {
  if (k == 0xFFFFFFFFFFFFFFFFull) return __longlong_as_double(0x7FF8000000000000ll);
  return __longlong_as_double((long long)((k & 0x8000000000000000ull) ? (k & 0x7FFFFFFFFFFFFFFFull) : ~k));
}

}  // namespace mppi
