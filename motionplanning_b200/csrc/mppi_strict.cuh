// mppi_strict.cuh — the K-sample rollout of the `strict_math` FP32 flavour (mppi_config_t::strict_math).
//
// north_star: "argmin sample index and collision flags must be bit-exact".  In FP32 that requires the GPU
// and the CPU checker to execute the same sequence of IEEE-754 operations.  This kernel therefore follows the
// REFERENCE's own evaluation order, statement by statement (files cited per function), in round-to-nearest
// single precision with no fused multiply-add (every operator maps to __fadd_rn / __fmul_rn / __fdiv_rn,
// which nvcc never contracts) and with the elementary functions of include/mppi_strict_math.h.  A float
// restatement of the reference that uses the same functions (the checker's strict mode) then yields the same
// bits: per-sample cost, flags and hence the argmin are compared for EQUALITY in tests/test_gpu_strict.py.
// One sample per thread, nothing clever: a verification flavour (~10x slower than the fast kernels).
// The softmin update and the nominal rollout that follow are the ordinary FP32 kernels.
#pragma once
#include "../../include/mppi_strict_math.h"
#include "mppi_kernels.cuh"

namespace mppi_strict {

// float whose operators can never be contracted or approximated
struct sf {
  float v;
  __device__ __forceinline__ sf() {}
  __device__ __forceinline__ explicit sf(float x) : v(x) {}
};
__device__ __forceinline__ sf operator+(sf a, sf b) { return sf(__fadd_rn(a.v, b.v)); }
__device__ __forceinline__ sf operator-(sf a, sf b) { return sf(__fsub_rn(a.v, b.v)); }
__device__ __forceinline__ sf operator*(sf a, sf b) { return sf(__fmul_rn(a.v, b.v)); }
__device__ __forceinline__ sf operator/(sf a, sf b) { return sf(__fdiv_rn(a.v, b.v)); }
__device__ __forceinline__ sf operator-(sf a) { return sf(-a.v); }
__device__ __forceinline__ bool operator<(sf a, sf b) { return a.v < b.v; }
__device__ __forceinline__ bool operator>(sf a, sf b) { return a.v > b.v; }
__device__ __forceinline__ bool operator<=(sf a, sf b) { return a.v <= b.v; }
__device__ __forceinline__ bool operator>=(sf a, sf b) { return a.v >= b.v; }
__device__ __forceinline__ sf sm_abs(sf x) { return sf(fabsf(x.v)); }
__device__ __forceinline__ int sm_to_int(sf x) { return (int)x.v; }
__device__ __forceinline__ sf sm_lit(sf, float c) { return sf(c); }
__device__ __forceinline__ sf smin(sf a, sf b) { return sf(fminf(a.v, b.v)); }
__device__ __forceinline__ sf smax(sf a, sf b) { return sf(fmaxf(a.v, b.v)); }

}  // namespace mppi_strict

namespace mppi {

using mppi_strict::sf;

// the configuration as the reference's functions see it: every double of mppi_config_t rounded ONCE to float
struct StrictParams {
  int model, n_x, N, K_local, k_offset, n_circles, n_blocks;
  float dt, lambda, Sigma[4], goal[2], XL[MPPI_MAX_NX], XU[MPPI_MAX_NX], CL[2], CU[2];
  float slack_penalty, obstacle_penalty, terminal_scale;
  float la, lb, mass, Izz, g, mu, KFZF, KFZR, KFZX, tire[6], w[6];
  float veh_L, veh_W, veh_off;
  float circles[MPPI_MAX_CIRCLES][3];   // x, y, r
  float blocks[MPPI_MAX_BLOCKS][5];     // cx, cy, psi, half_len, half_wid
};

namespace strict {

constexpr int BS = 128;

// VehicleDynamics: OptimalControl/MPPI/src/vehicledynamics.jl:1-54 (7 states) and
// OptimalControl/Parking/src/VehicleModel.jl:1-21 with the stage cost of SURVEY.md 8(a) a6 (5 states)
template <int MODEL>
__device__ __forceinline__ void vehicle_dynamics(const StrictParams& p, const sf* states, const sf* ctrl, sf* d, sf* cost) {
  const sf w0(p.w[0]), w1(p.w[1]), w2(p.w[2]), w3(p.w[3]), w4(p.w[4]), w5(p.w[5]);
  if (MODEL == MPPI_MODEL_DYNAMIC7) {
    const sf la(p.la), lb(p.lb), M(p.mass), Izz(p.Izz), g(p.g), mu(p.mu), KFZF(p.KFZF), KFZR(p.KFZR), KFZX(p.KFZX);
    const sf B = sf(p.tire[0]) / mu, Cc(p.tire[1]), E(p.tire[3]), D(p.tire[2]);                  // :14-16
    const sf y = states[1], v = states[2], r = states[3], psi = states[4], sa = states[6];    // :20-26
    sf ux = states[5];
    const sf sr = ctrl[0], ax = ctrl[1];                                                      // :27-28
    const sf two(2.0f), c001(0.01f);
    const sf FZF = two * (KFZF * g - (ax - r * v) * KFZX);                                     // :30
    const sf FZR = two * (KFZR * g + (ax - r * v) * KFZX);                                     // :31
    const sf alpha_f = mppi_strict::sm_atan((v + la * r) / (ux + c001)) - sa;                  // :32
    const sf alpha_r = mppi_strict::sm_atan((v - lb * r) / (ux + c001));                       // :33
    const sf X1f = B * alpha_f;                                                                // :35
    const sf FY1 = mu * FZF * D * mppi_strict::sm_sin(Cc * mppi_strict::sm_atan(X1f - E * (X1f - mppi_strict::sm_atan(X1f))));   // :36
    const sf X1r = B * alpha_r;                                                                // :37
    const sf FY2 = mu * FZR * D * mppi_strict::sm_sin(Cc * mppi_strict::sm_atan(X1r - E * (X1r - mppi_strict::sm_atan(X1r))));   // :38
    if (ux <= sf(0.0f)) ux = sf(0.0f);                                                         // :40-42
    sf sn, cs;
    mppi_strict::sm_sincos(psi, &sn, &cs);
    d[0] = ux * cs - v * sn;                                                                   // :43
    d[1] = ux * sn + v * cs;                                                                   // :44
    d[4] = r;                                                                                  // :45
    d[2] = (FY1 + FY2) / M - r * ux;                                                           // :46
    d[3] = (FY1 * la - FY2 * lb) / Izz;                                                        // :47
    d[5] = ax; d[6] = sr;                                                                      // :48-49
    *cost = v * v * w0 + w1 * (r * r) + w2 * (ax * ax) + w3 * (sr * sr) + w4 * (sa * sa) + w5 * (y * y);   // :52
  } else {
    const sf L = sf(p.la) + sf(p.lb);                                                          // VehicleModel.jl:2-4
    const sf y = states[1], psi = states[2], ux = states[3], sa = states[4];                   // :6-10
    const sf sr = ctrl[0], ax = ctrl[1];                                                       // :11-12
    sf sn, cs;
    mppi_strict::sm_sincos(psi, &sn, &cs);
    d[0] = ux * cs;                                                                            // :14
    d[1] = ux * sn;                                                                            // :15
    d[2] = ux / L * mppi_strict::sm_tan(sa);                                                   // :16
    d[3] = ax; d[4] = sr;                                                                      // :17-18
    const sf v(0.0f), r(0.0f);
    *cost = v * v * w0 + w1 * (r * r) + w2 * (ax * ax) + w3 * (sr * sr) + w4 * (sa * sa) + w5 * (y * y);
  }
}

// Rungekutta2, MPPIUtils.jl:76-90 (the model's constraint value is 0, so the slack branch :85-88 is dead)
template <int MODEL, int NX>
__device__ __forceinline__ void rungekutta2(const StrictParams& p, const sf* x, const sf* c, sf dt, sf* x_next, sf* p_cost) {
  sf k1[NX], k2[NX], xt[NX], pc2;
  vehicle_dynamics<MODEL>(p, x, c, k1, p_cost);                             // :77-78
#pragma unroll
  for (int i = 0; i < NX; ++i) xt[i] = x[i] + k1[i] * dt;                   // :79
  vehicle_dynamics<MODEL>(p, xt, c, k2, &pc2);
#pragma unroll
  for (int i = 0; i < NX; ++i) x_next[i] = x[i] + dt * (k1[i] + k2[i]) / sf(2.0f);   // :89
}

// GetRectanglePts, PathPlanning/CollisionDetection/src/utils.jl:14-25
__device__ __forceinline__ void get_rectangle_pts(const sf* block, sf* pts) {
  const sf ox = block[0], oy = block[1], psi = block[2], len = block[3], wid = block[4];
  const sf lx[5] = {-len, -len, len, len, -len};
  const sf ly[5] = {wid, -wid, -wid, wid, wid};
  sf s, c;
  mppi_strict::sm_sincos(psi, &s, &c);
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    pts[2 * i] = (c * lx[i] + (-s) * ly[i]) + ox;
    pts[2 * i + 1] = (s * lx[i] + c * ly[i]) + oy;
  }
}
// SeparatingAxisTheorem, utils.jl:37-62
__device__ __forceinline__ bool separating_axis(const sf* base, const sf* other) {
  const float inf = __int_as_float(0x7f800000);
  for (int b = 0; b < 4; ++b) {                                                      // :39
    const sf bgx = base[2 * b], bgy = base[2 * b + 1];                               // :41
    const sf bvx = base[2 * b + 2] - bgx, bvy = base[2 * b + 3] - bgy;               // :42-44
    const sf nx = -bvy, ny = bvx;                                                    // :46
    sf min_base(inf), max_base(-inf), min_other(inf), max_other(-inf);
#pragma unroll
    for (int i = 0; i < 5; ++i) {                                                    // :48
      const sf dd = (base[2 * i] - bgx) * nx + (base[2 * i + 1] - bgy) * ny;
      min_base = mppi_strict::smin(min_base, dd); max_base = mppi_strict::smax(max_base, dd);
    }
#pragma unroll
    for (int i = 0; i < 5; ++i) {                                                    // :49
      const sf dd = (other[2 * i] - bgx) * nx + (other[2 * i + 1] - bgy) * ny;
      min_other = mppi_strict::smin(min_other, dd); max_other = mppi_strict::smax(max_other, dd);
    }
    if ((max_other <= min_base) || (max_base <= min_other)) return true;             // :57-58
  }
  return false;                                                                      // :61
}
// modπ, PathPlanning/ReedsSheppsCurves/src/ReedsSheppsUtils.jl:32-46 (Julia mod is floored)
__device__ __forceinline__ sf mod_pi(sf angle) {
  const sf pi(float(3.141592653589793));
  if (angle >= -pi && angle <= pi) return angle;
  const sf two_pi = sf(2.0f) * pi;
  sf a(fmodf(angle.v, two_pi.v));
  if (a.v != 0.0f && a < sf(0.0f)) a = a + two_pi;
  if (a < -pi) a = a + two_pi;
  else if (a > pi) a = a - two_pi;
  return a;
}

template <int MODEL>
__global__ void __launch_bounds__(BS) rollout_strict_kernel(const __grid_constant__ StrictParams p, const Buffers<float> buf) {
  constexpr int NX = MODEL == MPPI_MODEL_DYNAMIC7 ? 7 : 5;
  constexpr int IPSI = MODEL == MPPI_MODEL_DYNAMIC7 ? 4 : 2;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  sf* sNom = reinterpret_cast<sf*>(smem_raw);     // [N][2] nominal controls
  sf* sWall = sNom + 2 * p.N;                      // [n_blocks][10] Block2Pts of the walls (utils.jl:28-34)
  __shared__ unsigned long long s_key[BS / 32];
  __shared__ int s_feas[BS / 32];
  const int N = p.N, K = p.K_local, t = threadIdx.x;
  for (int j = t; j < N; j += BS) { sNom[2 * j] = sf(float(buf.U[j])); sNom[2 * j + 1] = sf(float(buf.U[N + j])); }
  for (int b = t; b < p.n_blocks; b += BS) {
    sf blk[5];
    for (int i = 0; i < 5; ++i) blk[i] = sf(p.blocks[b][i]);
    get_rectangle_pts(blk, sWall + 10 * b);
  }
  __syncthreads();
  const int k = blockIdx.x * BS + t;
  unsigned long long key = ~0ull;
  int n_feas = 0;
  if (k < K) {
    const sf dt(p.dt), lambda(p.lambda), slack(p.slack_penalty), pen(p.obstacle_penalty);
    sf x0[NX], states[NX], next[NX];
    for (int i = 0; i < NX; ++i) { x0[i] = sf(float(buf.x0[i])); states[i] = x0[i]; }
    // inv(Sigma), MPPIUtils.jl:45 (2 x 2, closed form; column-major)
    const sf sa_(p.Sigma[0]), sc_(p.Sigma[1]), sb_(p.Sigma[2]), sd_(p.Sigma[3]);
    const sf det = sa_ * sd_ - sb_ * sc_;
    const sf inv0 = sd_ / det, inv1 = -sc_ / det, inv2 = -sb_ / det, inv3 = sa_ / det;
    const float* eps = buf.eps;
    int hit = 0, boundhit = 0;
    sf sum(0.0f);
    auto obstacle_eval = [&](const sf* s, sf* cost) {   // ObstacleEvaluation, MPPIUtils.jl:120-132 + block_collision_check (hybrid_astar_utils.jl:193-201)
      for (int o = 0; o < p.n_circles; ++o) {
        const sf dx = s[0] - sf(p.circles[o][0]), dy = s[1] - sf(p.circles[o][1]);
        const sf d2 = dx * dx + dy * dy, r2 = sf(p.circles[o][2]) * sf(p.circles[o][2]);
        if (d2 <= r2) { hit = 1; *cost = *cost + pen; }                              // :125-127
      }
      if (p.n_blocks > 0) {
        sf vb[5], vp[10], sn, cs;
        mppi_strict::sm_sincos(s[IPSI], &sn, &cs);
        vb[0] = s[0] + sf(p.veh_off) * cs; vb[1] = s[1] + sf(p.veh_off) * sn;        // hybrid_astar_utils.jl:193
        vb[2] = mod_pi(s[IPSI]); vb[3] = sf(p.veh_L) / sf(2.0f); vb[4] = sf(p.veh_W) / sf(2.0f);
        get_rectangle_pts(vb, vp);
        for (int b = 0; b < p.n_blocks; ++b) {
          const sf* wp = sWall + 10 * b;
          const bool free_ = separating_axis(wp, vp) && separating_axis(vp, wp);     // ConvexCollision, utils.jl:64-74
          if (!free_) { hit = 1; *cost = *cost + pen; }                              // :198
        }
      }
    };
    auto bound_eval = [&](const sf* s, sf* cost) {      // BoundEvaluation, MPPIUtils.jl:135-151
#pragma unroll
      for (int i = 0; i < NX; ++i) {
        const sf xl(p.XL[i]), xu(p.XU[i]);
        if (s[i] < xl) { boundhit = 1; *cost = *cost + slack * mppi_strict::sm_abs(s[i] - xl); }
        if (s[i] > xu) { boundhit = 1; *cost = *cost + slack * mppi_strict::sm_abs(s[i] - xu); }
      }
    };
    for (int j = 0; j < N; ++j) {                                                    // TrajectoryRollout, MPPIUtils.jl:39-47
      sf c_cost(0.0f), b_cost(0.0f), d_cost;
      if (j > 0) { obstacle_eval(states, &c_cost); bound_eval(states, &b_cost); }    // :40-41
      const sf un0 = sNom[2 * j], un1 = sNom[2 * j + 1];
      const float2 e = *reinterpret_cast<const float2*>(eps + (size_t(j) * K + k) * 2);
      sf u[2];
      u[0] = mppi_strict::smin(mppi_strict::smax(sf(e.x) + un0, sf(p.CL[0])), sf(p.CU[0]));   // :6-8 sample, PushInBounds :13-20
      u[1] = mppi_strict::smin(mppi_strict::smax(sf(e.y) + un1, sf(p.CL[1])), sf(p.CU[1]));
      rungekutta2<MODEL, NX>(p, states, u, dt, next, &d_cost);                       // :42
#pragma unroll
      for (int i = 0; i < NX; ++i) states[i] = next[i];                              // :43
      const sf r0 = lambda * un0, r1 = lambda * un1;                                 // :45
      const sf t0 = r0 * inv0 + r1 * inv1, t1 = r0 * inv2 + r1 * inv3;
      const sf ctrl_term = t0 * (u[0] - un0) + t1 * (u[1] - un1);
      sum = sum + (d_cost + b_cost + c_cost + ctrl_term + sf(0.0f));                 // cost_his[j], summed left to right (:54)
    }
    {                                                                                // terminal slot, :49-53
      const sf zero_u[2] = {sf(0.0f), sf(0.0f)};
      sf d_cost, c_cost(0.0f), b_cost(0.0f);
      rungekutta2<MODEL, NX>(p, states, zero_u, dt, next, &d_cost);
      obstacle_eval(states, &c_cost);
      bound_eval(states, &b_cost);
      sum = sum + (d_cost + b_cost + c_cost);
    }
    const sf g0(p.goal[0]), g1(p.goal[1]);
    const sf d0 = (x0[0] - g0) * (x0[0] - g0) + (x0[1] - g1) * (x0[1] - g1);
    const sf term = (states[0] - g0) * (states[0] - g0) + (states[1] - g1) * (states[1] - g1);   // TerminalCostEvaluation :113-116
    const sf cost = sum + term / d0 * sf(p.terminal_scale);                          // :54
    const bool feas = !hit && !boundhit;                                             // :55
    buf.costs[k] = cost.v;
    buf.flags[k] = uint8_t((feas ? MPPI_FLAG_FEASIBLE : 0) | (hit ? MPPI_FLAG_COLLIDED : 0) | (boundhit ? MPPI_FLAG_OOB : 0));
    key = make_key(cost.v, (unsigned int)k);
    n_feas = feas ? 1 : 0;
  }
  key = warp_min_u64(key);
  n_feas = warp_sum(n_feas);
  if ((t & 31) == 0) { s_key[t >> 5] = key; s_feas[t >> 5] = n_feas; }
  __syncthreads();
  if (t == 0) {
    int f = 0;
    for (int w = 0; w < BS / 32; ++w) { key = s_key[w] < key ? s_key[w] : key; f += s_feas[w]; }
    atomicMin(&buf.scratch[0].rho_key, key);
    if (f) atomicAdd(&buf.scratch[0].feasible_count, f);
  }
}

}  // namespace strict
}  // namespace mppi
