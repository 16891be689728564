// mppi_models.cuh — vehicle models, Heun step, obstacle/bound tests, Philox.
#pragma once
#include "mppi_device.cuh"
#include "mppi_pack.cuh"

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
// V is the value type a thread carries: float / double (one sample) or f32x2 (two samples on the
// packed FP32 pipe, mppi_pack.cuh); parameters are scalars of type Math<V>::S.  Multiply-adds are
// written as explicit fma so that every instantiation performs the same operation sequence.
template <class V>
struct Kin5 {
  using M = Math<V>;
  using S = typename M::S;
  static constexpr int NX = 5;
  static constexpr int IPSI = 2;
  V s[NX];
  V sn, cs;   // sin/cos of s[IPSI], carried between steps
  V tn;       // tan(sa) / L, carried between steps

  __device__ __forceinline__ void init(const StaticParams<S>& p, const double* x0) {
#pragma unroll
    for (int i = 0; i < NX; ++i) s[i] = V(S(x0[i]));
    M::sincos(s[IPSI], &sn, &cs);
    tn = M::over(M::tan(s[4]), p.Lwb, p.inv_Lwb);
  }
  // stage cost of VehicleDynamics at (state, u): SURVEY 8(a) a6 = dynamic cost with v = r = 0
  template <class P>
  __device__ __forceinline__ V stage_cost(const P& p, V sr, V ax) const {
    return M::fma(V(p.w[5]), s[1] * s[1], M::fma(V(p.w[4]), s[4] * s[4], M::fma(V(p.w[3]), sr * sr, V(p.w[2]) * (ax * ax))));
  }
  template <class P>
  __device__ __forceinline__ void step(const P& p, V sr, V ax) {
    // k1 at (s, u)
    const V k1x = s[3] * cs, k1y = s[3] * sn, k1p = s[3] * tn;   // ux cosψ, ux sinψ, ux/L tan(sa)
    // predictor x + k1*dt
    const V psit = M::fma(k1p, V(p.dt), s[2]);
    const V uxt = M::fma(ax, V(p.dt), s[3]);
    const V sat = M::fma(sr, V(p.dt), s[4]);
    V st, ct;
    M::sincos(psit, &st, &ct);
    const V tnt = M::over(M::tan(sat), p.Lwb, p.inv_Lwb);
    // corrector x + dt*(k1+k2)/2
    s[0] = M::fma(V(p.half_dt), M::fma(uxt, ct, k1x), s[0]);
    s[1] = M::fma(V(p.half_dt), M::fma(uxt, st, k1y), s[1]);
    s[2] = M::fma(V(p.half_dt), M::fma(uxt, tnt, k1p), s[2]);
    s[3] = uxt;
    s[4] = sat;
    tn = tnt;
    M::sincos(s[2], &sn, &cs);
  }
};

// ---------------------------------------------------------------------------
// Dynamic bicycle with Pacejka-style tyres — OptimalControl/MPPI/src/vehicledynamics.jl:1-54.
// State [x, y, v, r, psi, ux, sa], control [sr, ax];
// derivative order [dx, dy, dv, dr, dpsi, du, dsa] (:50) == state order.
// ---------------------------------------------------------------------------
template <class V>
struct Dyn7 {
  using M = Math<V>;
  using S = typename M::S;
  static constexpr int NX = 7;
  static constexpr int IPSI = 4;
  V s[NX];
  V sn, cs;

  __device__ __forceinline__ void init(const StaticParams<S>& p, const double* x0) {
#pragma unroll
    for (int i = 0; i < NX; ++i) s[i] = V(S(x0[i]));
    M::sincos(s[IPSI], &sn, &cs);
  }
  template <class P>
  __device__ __forceinline__ V stage_cost(const P& p, V sr, V ax) const {  // :52
    return M::fma(V(p.w[5]), s[1] * s[1], M::fma(V(p.w[4]), s[6] * s[6], M::fma(V(p.w[3]), sr * sr,
           M::fma(V(p.w[2]), ax * ax, M::fma(V(p.w[1]), s[3] * s[3], V(p.w[0]) * (s[2] * s[2]))))));
  }
  // lateral tyre force (:35-38): mu*FZ*D*sin(C*atan(X - E*(X - atan X))), X = B*alpha
  template <class P>
  static __device__ __forceinline__ V tyre(const P& p, V FZ, V alpha) {
    const V X = V(p.tireB) * alpha;
    const V inner = M::fma(V(-p.tireE), X - M::atan(X), X);
    return (V(p.mu * p.tireD) * FZ) * M::sin(V(p.tireC) * M::atan(inner));
  }
  // derivatives of (v, r) and the planar velocity at state q with heading sin/cos (sq, cq)
  template <class P>
  static __device__ __forceinline__ void deriv(const P& p, V v, V r, V ux, V sa, V sq, V cq,
                                               V ax, V* dx, V* dy, V* dv, V* dr) {
    const V t = M::fma(-r, v, ax);                                           // ax - r*v
    const V FZF = V(S(2)) * M::fma(t, V(-p.KFZX), V(p.FZF0));                // :30
    const V FZR = V(S(2)) * M::fma(t, V(p.KFZX), V(p.FZR0));                 // :31
    const V den = ux + V(S(0.01));
    const V alpha_f = M::atan_ratio(M::fma(V(p.la), r, v), den) - sa;        // :32
    const V alpha_r = M::atan_ratio(M::fma(V(-p.lb), r, v), den);            // :33
    const V FY1 = tyre(p, FZF, alpha_f);
    const V FY2 = tyre(p, FZR, alpha_r);
    const V uxc = M::fmax(ux, V(S(0)));                                      // :40-42
    *dx = M::fma(uxc, cq, -(v * sq));                                        // :43
    *dy = M::fma(uxc, sq, v * cq);                                           // :44
    *dv = M::fma(-r, uxc, M::over(FY1 + FY2, p.mass, p.inv_mass));           // :46
    *dr = M::over(M::fma(FY1, V(p.la), -(FY2 * V(p.lb))), p.Izz, p.inv_Izz); // :47
  }
  template <class P>
  __device__ __forceinline__ void step(const P& p, V sr, V ax) {
    V k1x, k1y, k1v, k1r;
    deriv(p, s[2], s[3], s[5], s[6], sn, cs, ax, &k1x, &k1y, &k1v, &k1r);
    const V k1p = s[3];                                                      // dpsi = r (:45)
    const V vt = M::fma(k1v, V(p.dt), s[2]), rt = M::fma(k1r, V(p.dt), s[3]);
    const V psit = M::fma(k1p, V(p.dt), s[4]);
    const V uxt = M::fma(ax, V(p.dt), s[5]), sat = M::fma(sr, V(p.dt), s[6]);
    V st, ct;
    M::sincos(psit, &st, &ct);
    V k2x, k2y, k2v, k2r;
    deriv(p, vt, rt, uxt, sat, st, ct, ax, &k2x, &k2y, &k2v, &k2r);
    s[0] = M::fma(V(p.half_dt), k1x + k2x, s[0]);
    s[1] = M::fma(V(p.half_dt), k1y + k2y, s[1]);
    s[2] = M::fma(V(p.half_dt), k1v + k2v, s[2]);
    s[3] = M::fma(V(p.half_dt), k1r + k2r, s[3]);
    s[4] = M::fma(V(p.half_dt), k1p + rt, s[4]);
    s[5] = uxt;   // == ux + dt/2*(ax+ax)
    s[6] = sat;
    M::sincos(s[4], &sn, &cs);
  }
};

// ---------------------------------------------------------------------------
// Obstacles.  Returns the number of obstacles hit by the pose (each hit costs
// obs_pen and clears the constraint flag: MPPIUtils.jl:125-127).
// ---------------------------------------------------------------------------
// The circle table is padded on the host to a multiple of 4 with never-hit entries
// (r^2 = -1), so the common small counts run as fully unrolled groups whose table
// entries are immediate constant-bank operands (no per-circle load / loop overhead).
template <class V, int Q>
__device__ __forceinline__ typename Math<V>::Count circle_hits_quads(const StaticParams<typename Math<V>::S>& p, V x, V y) {
  using M = Math<V>;
  typename M::Count hits = M::count(M::lt(V(1), V(0)));   // zero
#pragma unroll
  for (int o = 0; o < 4 * Q; ++o) {
    const V dx = x - V(p.circ[o][0]), dy = y - V(p.circ[o][1]);
    hits += M::count(M::le(M::fma(dx, dx, dy * dy), V(p.circ[o][2])));   // sum((s-o).^2) <= r^2
  }
  return hits;
}
// the first NC circles of a register-resident table (horizon loop of the rollout kernel)
template <class V, int NC, class P>
__device__ __forceinline__ typename Math<V>::Count circle_hits_regs(const P& hp, V px, V py) {
  using M = Math<V>;
  typename M::Count h = M::count(M::lt(V(1), V(0)));
#pragma unroll
  for (int o = 0; o < NC; ++o) {
    const V dx = px - V(hp.circ[o][0]), dy = py - V(hp.circ[o][1]);
    h += M::count(M::le(M::fma(dx, dx, dy * dy), V(hp.circ[o][2])));
  }
  return h;
}

template <class V>
__device__ __forceinline__ typename Math<V>::Count circle_hits(const StaticParams<typename Math<V>::S>& p, V x, V y) {
  using M = Math<V>;
  switch ((p.n_circles + 3) >> 2) {   // warp-uniform
    case 0: return M::count(M::lt(V(1), V(0)));
    case 1: return circle_hits_quads<V, 1>(p, x, y);
    case 2: return circle_hits_quads<V, 2>(p, x, y);
    default: break;
  }
  typename M::Count hits = M::count(M::lt(V(1), V(0)));
  for (int o = 0; o < p.n_circles; ++o) {
    const V dx = x - V(p.circ[o][0]), dy = y - V(p.circ[o][1]);
    hits += M::count(M::le(M::fma(dx, dx, dy * dy), V(p.circ[o][2])));
  }
  return hits;
}

// hits = tested - free, per lane
template <class V> __device__ __forceinline__ typename Math<V>::Count hits_from_free(int tested, typename Math<V>::Count fr);
template <> __device__ __forceinline__ int hits_from_free<float>(int tested, int fr) { return tested - fr; }
template <> __device__ __forceinline__ int hits_from_free<double>(int tested, int fr) { return tested - fr; }

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
// `stat` (diagnostics, profiling mode 2 only, nullptr otherwise): incremented by the number of walls a warp-step tests EXACTLY,
// so that a benchmark can report which share of the rectangle tests the broad phase left standing (mppi_get_phase_clocks()[7]).
// candidate walls of a rectangle centre x (StaticParams::wall_mask): none outside the binned range
template <class S>
__device__ __forceinline__ uint32_t wall_candidates(const StaticParams<S>& p, S vx) {
  const S f = (vx - p.wall_x0) * p.wall_inv_w;
  if (!(f > S(-0.005)) || !(f < S(WALL_BINS) + S(0.005))) return (f != f) ? 0xFFFFFFFFu : 0u;   // NaN pose: test everything
  const int i = min(max(int(f), 0), WALL_BINS - 1);
  return p.wall_mask[i];
}
__device__ __forceinline__ uint32_t wall_candidates(const StaticParams<float>& p, f32x2 vx) {
  return wall_candidates<float>(p, vx.v.x) | wall_candidates<float>(p, vx.v.y);
}

template <class V, class P>
__device__ __forceinline__ typename Math<V>::Count block_hits(const P& veh, const typename Math<V>::S* all4, const typename Math<V>::S* tbl,
                                                              int n_blocks, V x, V y, V sn, V cs, unsigned long long* stat = nullptr,
                                                              const StaticParams<typename Math<V>::S>* bins = nullptr) {
  using M = Math<V>;
  using S = typename M::S;
  using S4 = typename Vec4<S>::type;
  typename M::Count hits = M::count(M::lt(V(1), V(0)));
  const V vx = M::fma(V(veh.veh_off), cs, x), vy = M::fma(V(veh.veh_off), sn, y);
  const unsigned mask = __activemask();
  {
    const V ax = vx - V(all4[0]), ay = vy - V(all4[1]);
    if (__ballot_sync(mask, M::any(M::lt(M::fma(ax, ax, ay * ay), V(all4[2])))) == 0u) return hits;   // far from every wall
  }
  // exact rectangle-rectangle test against one wall
  auto exact = [&](const S4 w0, const S4 w1, V dx, V dy) __attribute__((always_inline)) {
    const V C = V(w0.z), Sn = V(w0.w), wl = V(w1.x), ww = V(w1.y), hl = V(veh.veh_hl), hw = V(veh.veh_hw);
    const V ac = M::fabs(M::fma(cs, C, sn * Sn));     // |cos(ψ-Ψ)|
    const V as = M::fabs(M::fma(sn, C, -(cs * Sn)));  // |sin(ψ-Ψ)|
    const auto sepW = m_or(M::ge(M::fabs(M::fma(dx, C, dy * Sn)), M::fma(hl, ac, M::fma(hw, as, wl))),
                           M::ge(M::fabs(M::fma(dy, C, -(dx * Sn))), M::fma(hl, as, M::fma(hw, ac, ww))));
    const auto sepV = m_or(M::ge(M::fabs(M::fma(dx, cs, dy * sn)), M::fma(wl, ac, M::fma(ww, as, hl))),
                           M::ge(M::fabs(M::fma(dy, cs, -(dx * sn))), M::fma(wl, as, M::fma(ww, ac, hw))));
    return M::count(m_and(sepW, sepV));   // 1 = free; !ConvexCollision => hit (hybrid_astar_utils.jl:198)
  };
  const S4* t4 = reinterpret_cast<const S4*>(tbl);
  typename M::Count freec = hits;
  int tested = 0;
  const bool count_lane = stat != nullptr && int(threadIdx.x & 31) == __ffs(mask) - 1;
  if (count_lane && n_blocks <= 4) atomicAdd(stat, (unsigned long long)n_blocks);
  if (n_blocks <= 4) {   // few walls: straight loop the compiler can unroll and software-pipeline
#pragma unroll 4
    for (int b = 0; b < n_blocks; ++b) {
      const S4 w0 = t4[2 * b], w1 = t4[2 * b + 1];
      freec += exact(w0, w1, vx - V(w0.x), vy - V(w0.y));
    }
    tested = n_blocks;
  } else {               // many walls: x-binned candidates (warp-wide OR), then the per-wall bounding-circle vote
    uint32_t cand = (n_blocks >= 32) ? 0xFFFFFFFFu : ((1u << n_blocks) - 1u);
    if (bins != nullptr) cand &= __reduce_or_sync(mask, wall_candidates(*bins, vx));
    while (cand) {
      const int b = __ffs(cand) - 1;
      cand &= cand - 1u;
      const S4 w0 = t4[2 * b], w1 = t4[2 * b + 1];
      const V dx = vx - V(w0.x), dy = vy - V(w0.y);
      if (__ballot_sync(mask, M::any(M::lt(M::fma(dx, dx, dy * dy), V(w1.z)))) == 0u) continue;   // provably free for every lane
      freec += exact(w0, w1, dx, dy);
      ++tested;
      if (count_lane) atomicAdd(stat, 1ull);
    }
  }
  return hits_from_free<V>(tested, freec);   // hits = walls tested - walls found free
}
template <class V>
__device__ __forceinline__ typename Math<V>::Count block_hits(const StaticParams<typename Math<V>::S>& p, V x, V y, V sn, V cs) {
  return block_hits<V>(p, p.blk_all, &p.blk[0][0], p.n_blocks, x, y, sn, cs, nullptr, &p);
}

// BoundEvaluation (MPPIUtils.jl:135-151): sum of violations |x - XL|, |x - XU|
// (strict inequalities; the slack factor is applied once at the end).
// XL <= XU, so at most one side is violated: max(XL - s, s - XU, 0) equals the sum of
// the two one-sided terms exactly (the other term is +0).
template <class V, int NX, class P>
__device__ __forceinline__ V bound_violation(const P& p, const V* s) {
  using M = Math<V>;
  V acc = V(0);
#pragma unroll
  for (int i = 0; i < NX; ++i)
    acc += M::fmax(M::fmax(V(p.XL[i]) - s[i], s[i] - V(p.XU[i])), V(0));
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
