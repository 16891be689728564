// mppi_oracle_plan.hpp — rollout, softmin weights, MPPIPlan/DWAPlan, Philox noise.
// TEST INFRASTRUCTURE ONLY (see mppi_oracle.hpp header).
#pragma once
#include <algorithm>
#include <chrono>

#include "mppi_oracle.hpp"

namespace oracle {

template <class R>
struct Rollout {
  std::vector<R> states_his;  // (N+1) x n_x, row j at [j*n_x]
  bool feasible = false;
  bool collided = false;  // build addition: any obstacle hit (north star grades it)
  bool oob = false;       // build addition: any bound hit
  R cost_total = 0;
};

template <class R>
void d_cost_extra(const Params<R>& p, const R* states, int* tidx, R* acc) {
  *acc = p.ref_w * track_scan(p, states[0], states[1], tidx);
}

// ---- a3: TrajectoryRollout — MPPIUtils.jl:31-57 ------------------------------
// ctrl_list: N x m, row j at [j*m] (already clamped by the caller, :8,:176).
// nominal:   N x m, same layout (MPPI.s.NominalControl).
// DWA variant (DWAUtils.jl:16-42): ctrl_list is ONE control used at every step
// (:27) and the lambda term is absent (:30).
template <class R>
void trajectory_rollout(const Params<R>& p, const R* x0, const R* ctrl_list, const R* nominal,
                        Rollout<R>* out, Margin<R>* mg, bool keep_traj) {
  const int N = p.N, nx = p.n_x;
  std::vector<R> cost_his(N + 1, R(0));          // :34
  std::vector<int> constraint_his(N + 1, 0);     // :35
  if (keep_traj) out->states_his.assign(size_t(N + 1) * nx, R(0));  // :33
  const R dt = p.dt;                             // :36
  R states[MPPI_MAX_NX], next[MPPI_MAX_NX];
  for (int i = 0; i < nx; ++i) states[i] = x0[i];  // :37
  if (keep_traj) for (int i = 0; i < nx; ++i) out->states_his[i] = states[i];  // :38
  int hit = 0, boundhit = 0;
  const bool track = !p.ref.empty();
  int tidx = track ? track_start(p, states[0], states[1]) : 0;
  // inv(Σ) for the 2x2 case, closed form (MPPIUtils.jl:45 calls inv every step)
  R inv[4] = {0, 0, 0, 0};
  if (p.planner == MPPI_PLANNER_MPPI) {
    const R a = p.Sigma[0], c = p.Sigma[1], b = p.Sigma[2], d = p.Sigma[3];  // col-major
    const R det = a * d - b * c;
    inv[0] = d / det; inv[1] = -c / det; inv[2] = -b / det; inv[3] = a / det;
  }
  for (int j = 0; j < N; ++j) {                  // :39
    int c_cons = 1, b_cons = 1, d_cons = 1;
    R c_cost = 0, b_cost = 0, d_cost = 0;
    if (j > 0) {                                 // :40-41 (skipped at j = 1 in Julia)
      obstacle_evaluation(p, states, &c_cons, &c_cost, &hit, mg);
      bound_evaluation(p, states, &b_cons, &b_cost, mg);
      if (!b_cons) boundhit = 1;
    }
    const R* u = (p.planner == MPPI_PLANNER_DWA) ? ctrl_list : ctrl_list + size_t(j) * p.m;
    R t_cost = 0;   // tracking term rides with the stage cost, at the pre-step state
    if (track) d_cost_extra(p, states, &tidx, &t_cost);
    rungekutta2(p, states, u, dt, next, &d_cons, &d_cost);  // :42
    for (int i = 0; i < nx; ++i) states[i] = next[i];        // :43
    constraint_his[j] = d_cons * c_cons * b_cons;            // :44
    R ctrl_term = 0;
    if (p.planner == MPPI_PLANNER_MPPI) {
      // :45  λ * U_nom[j,:]' * inv(Σ) * (u − U_nom[j,:]), left to right
      const R* un = nominal + size_t(j) * p.m;
      const R r0 = p.lambda * un[0], r1 = p.lambda * un[1];
      const R t0 = r0 * inv[0] + r1 * inv[1];  // row * column 1
      const R t1 = r0 * inv[2] + r1 * inv[3];  // row * column 2
      ctrl_term = t0 * (u[0] - un[0]) + t1 * (u[1] - un[1]);
    }
    cost_his[j] = d_cost + b_cost + c_cost + ctrl_term + t_cost;      // :45 (+ tracking extension)
    if (keep_traj) for (int i = 0; i < nx; ++i) out->states_his[size_t(j + 1) * nx + i] = states[i];  // :46
  }
  {  // terminal slot, :49-53
    const R zero_u[2] = {R(0), R(0)};
    int d_cons = 1, c_cons = 1, b_cons = 1;
    R d_cost = 0, c_cost = 0, b_cost = 0;
    rungekutta2(p, states, zero_u, dt, next, &d_cons, &d_cost);  // :49 (state discarded)
    if (track) { R t = 0; d_cost_extra(p, states, &tidx, &t); d_cost = d_cost + t; }
    obstacle_evaluation(p, states, &c_cons, &c_cost, &hit, mg);  // :50
    bound_evaluation(p, states, &b_cons, &b_cost, mg);           // :51
    if (!b_cons) boundhit = 1;
    constraint_his[N] = d_cons * c_cons * b_cons;                // :52
    cost_his[N] = d_cost + b_cost + c_cost;                      // :53
  }
  R sum = 0;
  for (int j = 0; j <= N; ++j) sum = sum + cost_his[j];          // sum(cost_his)
  const R d0 = (x0[0] - p.goal[0]) * (x0[0] - p.goal[0]) + (x0[1] - p.goal[1]) * (x0[1] - p.goal[1]);
  out->cost_total = sum + terminal_cost_evaluation(p, states) / d0 * p.terminal_scale;  // :54
  bool all_one = true;
  for (int j = 0; j <= N; ++j) all_one = all_one && (constraint_his[j] == 1);
  out->feasible = all_one;                                       // :55
  out->collided = hit != 0;
  out->oob = boundhit != 0;
}

// ---- a11: CalculateMPPIWeights — MPPIUtils.jl:154-167, isless types.jl:9 -----
// argmin over isless(cost): first minimal index, NaN sorts last.
template <class R>
int argmin_isless(const R* cost, int n) {
  int best = 0;
  for (int i = 1; i < n; ++i) {
    const R a = cost[i], b = cost[best];
    const bool less = (a < b) || (!std::isnan(a) && std::isnan(b));  // Base.isless
    if (less) best = i;
  }
  return best;
}

template <class R>
void calculate_mppi_weights(const R* cost, int n, R lambda, std::vector<R>* w, int* argmin,
                            R* rho, R* eta_out) {
  const int idx = argmin_isless(cost, n);                  // :155
  const R r = cost[idx];                                   // :156
  R eta = 0;                                               // :157
  for (int i = 0; i < n; ++i) eta = eta + std::exp(-R(1) / lambda * (cost[i] - r));  // :160
  w->assign(n, R(0));
  for (int i = 0; i < n; ++i) (*w)[i] = R(1) / eta * std::exp(-R(1) / lambda * (cost[i] - r));  // :164
  *argmin = idx; *rho = r; *eta_out = eta;
}

// ---- Philox4x32-10 (Salmon et al., SC'11; Random123) --------------------------
inline void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = uint64_t(0xD2511F53u) * c0, p1 = uint64_t(0xCD9E8D57u) * c2;
    const uint32_t n0 = uint32_t(p1 >> 32) ^ c1 ^ k0, n1 = uint32_t(p1);
    const uint32_t n2 = uint32_t(p0 >> 32) ^ c3 ^ k1, n3 = uint32_t(p0);
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// Box-Muller on two 32-bit words -> two N(0,1) floats.  Same formula as the
// device generator (csrc/philox.cuh); host libm vs device intrinsics differ in
// the last bits, which is why Philox-mode parity replays the DUMPED device eps.
inline void box_muller(uint32_t a, uint32_t b, float* z0, float* z1) {
  const float u = float(a) * 2.3283064365386963e-10f + 1.1641532182693481e-10f;  // (0,1]
  const float th = float(b) * 1.4629180792671596e-09f + (-3.1415926535897931f + 7.3145903963357981e-10f);
  const float s = std::sqrt(-2.0f * std::log(u));
  *z0 = s * std::sin(th);
  *z1 = s * std::cos(th);
}

// a1: eps = chol(Σ)·z (Distributions' MvNormal draw, MPPIUtils.jl:6), layout
// [N][K][m].  Counter = (global k, j/2, iteration, instance), key = seed.
template <class R>
void generate_noise(const double Sigma[4], uint64_t seed, uint32_t iteration, uint32_t instance,
                    int k_offset, int K, int N, R* eps) {
  const double l00 = std::sqrt(Sigma[0]);
  const double l10 = Sigma[1] / l00;
  const double l11 = std::sqrt(Sigma[3] - l10 * l10);
  const R L00 = R(l00), L10 = R(l10), L11 = R(l11);
  const uint32_t key[2] = {uint32_t(seed), uint32_t(seed >> 32)};
  for (int k = 0; k < K; ++k) {
    for (int jp = 0; 2 * jp < N; ++jp) {
      const uint32_t ctr[4] = {uint32_t(k_offset + k), uint32_t(jp), iteration, instance};
      uint32_t r[4];
      philox4x32_10(ctr, key, r);
      for (int h = 0; h < 2; ++h) {
        const int j = 2 * jp + h;
        if (j >= N) break;
        float z0, z1;
        box_muller(r[2 * h], r[2 * h + 1], &z0, &z1);
        R* e = eps + (size_t(j) * K + k) * 2;
        e[0] = L00 * R(z0);
        e[1] = L10 * R(z0) + L11 * R(z1);
      }
    }
  }
}

template <class R>
struct PlanOut {
  std::vector<R> costs;          // per drawn sample (K entries; entries >= R_drawn untouched)
  std::vector<uint8_t> flags;    // MPPI_FLAG_* per sample
  std::vector<R> margins;        // guard-band margin per sample
  std::vector<R> weights;        // R_drawn
  std::vector<R> control;        // U_new, N x m row-major [j*m+c]
  Rollout<R> nominal;            // final nominal rollout (:192)
  R nominal_margin = 0;
  int rollout_count = 0;         // R + 1 (:171,182,197)
  int feasible_count = 0;
  int argmin = 0;                // 0-based
  R rho = 0, eta = 0;
  double seconds = 0;
};

// ---- a12: MPPIPlan — MPPIUtils.jl:169-203, on SUPPLIED noise -----------------
// eps layout [N][K][m].  threads <= 1: the reference's serial loop with its
// data-dependent early exit (:175).  threads > 1: all K rollouts in parallel
// (OpenMP), then the same truncation applied as a prefix count — identical output.
template <class R>
void mppi_plan(const Params<R>& p, const R* x0, const R* nominal, const R* eps, int K,
               int feasibility_count, int threads, bool want_margin, PlanOut<R>* out) {
  const int N = p.N, m = p.m;
  const auto t1 = std::chrono::steady_clock::now();              // :173
  out->costs.assign(K, R(0));
  out->flags.assign(K, 0);
  if (want_margin) out->margins.assign(K, R(0));
  auto one = [&](int k) {
    std::vector<R> ctrl(size_t(N) * m);
    for (int j = 0; j < N; ++j)                                  // :6-7
      for (int c = 0; c < m; ++c) ctrl[j * m + c] = eps[(size_t(j) * K + k) * m + c] + nominal[j * m + c];
    push_in_bounds(p, ctrl.data());                              // :8
    Rollout<R> ro;
    Margin<R> mg;
    trajectory_rollout(p, x0, ctrl.data(), nominal, &ro, want_margin ? &mg : nullptr, false);
    out->costs[k] = ro.cost_total;
    out->flags[k] = uint8_t((ro.feasible ? MPPI_FLAG_FEASIBLE : 0) | (ro.collided ? MPPI_FLAG_COLLIDED : 0) |
                            (ro.oob ? MPPI_FLAG_OOB : 0));
    if (want_margin) out->margins[k] = mg.v;
    return ro.feasible;
  };
  int FeasibilityCount = 0, RolloutCount = 1;                    // :170-171
  if (threads <= 1) {
    while (FeasibilityCount <= feasibility_count && RolloutCount <= K) {  // :175
      if (one(RolloutCount - 1)) FeasibilityCount = FeasibilityCount + 1;  // :176-179
      RolloutCount = RolloutCount + 1;                           // :182
    }
  } else {
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int k = 0; k < K; ++k) one(k);
    while (FeasibilityCount <= feasibility_count && RolloutCount <= K) {
      if (out->flags[RolloutCount - 1] & MPPI_FLAG_FEASIBLE) FeasibilityCount = FeasibilityCount + 1;
      RolloutCount = RolloutCount + 1;
    }
  }
  const int Rn = RolloutCount - 1;                               // :185
  calculate_mppi_weights(out->costs.data(), Rn, p.lambda, &out->weights, &out->argmin, &out->rho,
                         &out->eta);                             // :186
  out->control.assign(size_t(N) * m, R(0));                      // :187
  for (int i = 0; i < Rn; ++i) {                                 // :188-190
    const R w = out->weights[i];
    for (int j = 0; j < N; ++j)
      for (int c = 0; c < m; ++c) {
        R u = eps[(size_t(j) * K + i) * m + c] + nominal[j * m + c];
        u = std::fmin(std::fmax(u, p.CL[c]), p.CU[c]);
        out->control[j * m + c] = out->control[j * m + c] + w * u;
      }
  }
  Margin<R> nm;
  trajectory_rollout(p, x0, out->control.data(), nominal, &out->nominal, &nm, true);  // :192
  out->nominal_margin = nm.v;
  out->rollout_count = RolloutCount;                             // :197
  out->feasible_count = FeasibilityCount;                        // :198
  out->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count();  // :199-200
}

// ---- DWAPlan — DynamicWindow/src/DWAUtils.jl:141-163 --------------------------
// candidates: K x m row-major constant controls (setup.jl:59-85 grid).
template <class R>
void dwa_plan(const Params<R>& p, const R* x0, const R* candidates, int K, int threads,
              PlanOut<R>* out) {
  const auto t1 = std::chrono::steady_clock::now();
  out->costs.assign(K, R(0));
  out->flags.assign(K, 0);
#pragma omp parallel for schedule(static) num_threads(threads > 1 ? threads : 1)
  for (int k = 0; k < K; ++k) {                                  // :146-151
    Rollout<R> ro;
    trajectory_rollout(p, x0, candidates + size_t(k) * p.m, (const R*)nullptr, &ro, (Margin<R>*)nullptr, false);
    out->costs[k] = ro.cost_total;
    out->flags[k] = uint8_t((ro.feasible ? MPPI_FLAG_FEASIBLE : 0) | (ro.collided ? MPPI_FLAG_COLLIDED : 0) |
                            (ro.oob ? MPPI_FLAG_OOB : 0));
  }
  out->argmin = argmin_isless(out->costs.data(), K);             // :152 minimum(collection)
  out->rho = out->costs[out->argmin];
  out->control.assign(size_t(p.N) * p.m, R(0));
  for (int j = 0; j < p.N; ++j)
    for (int c = 0; c < p.m; ++c) out->control[j * p.m + c] = candidates[size_t(out->argmin) * p.m + c];
  trajectory_rollout(p, x0, candidates + size_t(out->argmin) * p.m, (const R*)nullptr, &out->nominal,
                     (Margin<R>*)nullptr, true);                 // :155-158 (BestCand's own rollout)
  out->rollout_count = K + 1;
  out->feasible_count = 0;
  for (int k = 0; k < K; ++k) out->feasible_count += (out->flags[k] & MPPI_FLAG_FEASIBLE) ? 1 : 0;
  out->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count();
}

}  // namespace oracle
