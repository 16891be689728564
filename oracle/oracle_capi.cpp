// oracle_capi.cpp — extern "C" surface of the CPU oracle for ctypes.
// TEST INFRASTRUCTURE ONLY (see mppi_oracle.hpp header): loaded by tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
#include "mppi_oracle_plan.hpp"

using namespace oracle;

// reference path for the tracking extension: process-wide, set by oracle_set_reference_path and
// picked up by every make_params until cleared (n = 0)
static std::vector<double> g_ref;
static double g_ref_w = 0.0;
static int g_ref_window = 0;

template <class R>
static Params<R> make_params(const mppi_config_t* c, const double* circles, int Mc,
                             const double* blocks, int Mp) {
  Params<R> p;
  p.model = c->model; p.n_x = c->n_x; p.m = c->m; p.N = c->N;
  p.dt = R(c->T / c->N);  // setup.jl:53
  p.lambda = R(c->lambda);
  for (int i = 0; i < 4; ++i) p.Sigma[i] = R(c->Sigma[i]);
  for (int i = 0; i < 2; ++i) { p.goal[i] = R(c->goal[i]); p.CL[i] = R(c->CL[i]); p.CU[i] = R(c->CU[i]); }
  for (int i = 0; i < MPPI_MAX_NX; ++i) { p.XL[i] = R(c->XL[i]); p.XU[i] = R(c->XU[i]); }
  p.slack_penalty = R(c->slack_penalty); p.obstacle_penalty = R(c->obstacle_penalty);
  p.terminal_scale = R(c->terminal_scale);
  p.la = R(c->la); p.lb = R(c->lb); p.mass = R(c->mass); p.Izz = R(c->Izz); p.g = R(c->g);
  p.mu = R(c->mu); p.KFZF = R(c->KFZF); p.KFZR = R(c->KFZR); p.KFZX = R(c->KFZX);
  for (int i = 0; i < 6; ++i) { p.tire[i] = R(c->tire[i]); p.w[i] = R(c->stage_w[i]); }
  p.veh_L = R(c->veh_L); p.veh_W = R(c->veh_W); p.veh_off = R(c->veh_center_offset);
  p.planner = c->planner;
  p.circles.resize(size_t(3) * Mc);
  for (int i = 0; i < 3 * Mc; ++i) p.circles[i] = R(circles[i]);
  p.blocks.resize(size_t(5) * Mp);
  for (int i = 0; i < 5 * Mp; ++i) p.blocks[i] = R(blocks[i]);
  p.wall_pts.resize(size_t(10) * Mp);
  for (int i = 0; i < Mp; ++i) get_rectangle_pts(&p.blocks[5 * i], &p.wall_pts[10 * i]);  // Block2Pts
  p.ref.resize(g_ref.size());
  for (size_t i = 0; i < g_ref.size(); ++i) p.ref[i] = R(g_ref[i]);
  p.ref_w = R(g_ref_w); p.ref_window = g_ref_window;
  return p;
}

template <class R>
static void plan_impl(const mppi_config_t* cfg, const double* circles, int Mc, const double* blocks,
                      int Mp, const double* x0, const double* Unom, const void* eps_or_cand,
                      int threads, int want_margin, double* costs, uint8_t* flags, double* margins,
                      double* U_new, double* traj, double* scalars) {
  Params<R> p = make_params<R>(cfg, circles, Mc, blocks, Mp);
  const int N = p.N, m = p.m, nx = p.n_x, K = cfg->K;
  R x0r[MPPI_MAX_NX];
  for (int i = 0; i < nx; ++i) x0r[i] = R(x0[i]);
  std::vector<R> nom(size_t(N) * m, R(0));
  if (Unom)
    for (int j = 0; j < N; ++j)
      for (int c = 0; c < m; ++c) nom[j * m + c] = R(Unom[c * N + j]);  // col-major in
  PlanOut<R> out;
  if (cfg->planner == MPPI_PLANNER_DWA) {
    const double* cand = (const double*)eps_or_cand;  // K x m col-major
    std::vector<R> cr(size_t(K) * m);
    for (int k = 0; k < K; ++k)
      for (int c = 0; c < m; ++c) cr[size_t(k) * m + c] = R(cand[size_t(c) * K + k]);
    dwa_plan(p, x0r, cr.data(), K, threads, &out);
  } else {
    mppi_plan(p, x0r, nom.data(), (const R*)eps_or_cand, K, cfg->feasibility_count, threads,
              want_margin != 0, &out);
  }
  if (costs) for (int k = 0; k < K; ++k) costs[k] = double(out.costs[k]);
  if (flags) for (int k = 0; k < K; ++k) flags[k] = out.flags[k];
  if (margins && want_margin && !out.margins.empty())
    for (int k = 0; k < K; ++k) margins[k] = double(out.margins[k]);
  if (U_new)
    for (int j = 0; j < N; ++j)
      for (int c = 0; c < m; ++c) U_new[c * N + j] = double(out.control[j * m + c]);
  if (traj)
    for (int j = 0; j <= N; ++j)
      for (int i = 0; i < nx; ++i) traj[i * (N + 1) + j] = double(out.nominal.states_his[size_t(j) * nx + i]);
  if (scalars) {
    scalars[0] = double(out.nominal.cost_total);
    scalars[1] = out.nominal.feasible ? 1.0 : 0.0;
    scalars[2] = out.rollout_count;
    scalars[3] = out.feasible_count;
    scalars[4] = out.argmin + 1;  // 1-based like Julia's argmin
    scalars[5] = double(out.rho);
    scalars[6] = double(out.eta);
    scalars[7] = out.seconds;
    scalars[8] = double(out.nominal_margin);
    scalars[9] = (out.nominal.collided ? 2.0 : 0.0) + (out.nominal.oob ? 4.0 : 0.0);
  }
}

template <class R>
static void rollout_impl(const mppi_config_t* cfg, const double* circles, int Mc,
                         const double* blocks, int Mp, const double* x0, const double* Unom,
                         const double* ctrl, double* traj, double* scalars) {
  Params<R> p = make_params<R>(cfg, circles, Mc, blocks, Mp);
  const int N = p.N, m = p.m, nx = p.n_x;
  R x0r[MPPI_MAX_NX];
  for (int i = 0; i < nx; ++i) x0r[i] = R(x0[i]);
  std::vector<R> nom(size_t(N) * m, R(0)), u(size_t(N) * m, R(0));
  for (int j = 0; j < N; ++j)
    for (int c = 0; c < m; ++c) {
      if (Unom) nom[j * m + c] = R(Unom[c * N + j]);
      u[j * m + c] = R(ctrl[c * N + j]);
    }
  Rollout<R> ro;
  Margin<R> mg;
  trajectory_rollout(p, x0r, u.data(), nom.data(), &ro, &mg, true);
  if (traj)
    for (int j = 0; j <= N; ++j)
      for (int i = 0; i < nx; ++i) traj[i * (N + 1) + j] = double(ro.states_his[size_t(j) * nx + i]);
  scalars[0] = double(ro.cost_total);
  scalars[1] = ro.feasible ? 1.0 : 0.0;
  scalars[2] = ro.collided ? 1.0 : 0.0;
  scalars[3] = ro.oob ? 1.0 : 0.0;
  scalars[4] = double(mg.v);
}

extern "C" {

// The reference's own literals, typed here from the reference sources (NOT taken from the product's
// mppi_config_default: tests/test_oracle_config.py checks that the two agree field by field, so a
// wrong constant on either side is caught).  Only the POD layout (include/mppi_b200.h) is shared.
int oracle_config_default(mppi_config_t* c) {
  std::memset(c, 0, sizeof *c);
  c->struct_size = sizeof *c;
  c->model = MPPI_MODEL_DYNAMIC7;
  c->n_x = 7;                                   // OptimalControl/MPPI/main.jl:12  numStates
  c->m = 2;                                     // main.jl:13  numControls
  c->K = 1500;                                  // main.jl:22  SamplingNumber
  c->N = 20;                                    // main.jl:14
  c->T = 3.0;                                   // main.jl:15
  c->lambda = 25.0;                             // main.jl:24
  c->B = 1;
  c->precision = MPPI_FP32; c->noise_mode = MPPI_NOISE_PHILOX; c->planner = MPPI_PLANNER_MPPI;
  c->feasibility_count = 1300;                  // types.jl:24
  c->use_cuda_graph = 1;
  c->seed = 0x5EED0001ull;                      // SURVEY.md 8(d) (the reference is unseeded)
  c->Sigma[0] = 0.05; c->Sigma[3] = 0.1;        // main.jl:27  [0.05 0; 0 0.1]
  const double pi = 3.14159265358979323846;
  const double x0[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 5.0, 0.0};                         // main.jl:7
  const double xl[7] = {-10.0, -20.0, -2.0, -pi / 2, -pi / 2, 1.0, -pi / 9};        // main.jl:8
  const double xu[7] = {130.0, 20.0, 2.0, pi / 2, pi / 2, 10.0, pi / 9};            // main.jl:9
  for (int i = 0; i < 7; ++i) { c->X0[i] = x0[i]; c->XL[i] = xl[i]; c->XU[i] = xu[i]; }
  c->CL[0] = -0.5; c->CL[1] = -2.5;             // main.jl:10
  c->CU[0] = 0.5; c->CU[1] = 2.5;               // main.jl:11
  c->goal[0] = 110.0; c->goal[1] = 0.0;         // main.jl:23
  c->slack_penalty = 100000.0;                  // types.jl:30  SlackPenalty = 1e5
  c->obstacle_penalty = 100.0 * 712.5;          // MPPIUtils.jl:127
  c->terminal_scale = 10000.0;                  // MPPIUtils.jl:54
  c->la = 1.56; c->lb = 1.64;                   // vehicledynamics.jl:2-3
  c->mass = 2020.0; c->Izz = 4095.0;            // :4-5
  c->g = 9.81; c->mu = 0.8;                     // :6-7
  c->KFZF = 1018.28 / 2; c->KFZR = 963.34 / 2;  // :8-9
  c->KFZX = 186.22;                             // :10
  const double tire[6] = {-10.4, 1.3, 1.0, 0.1556, -0.0023, 41.3705};              // :11
  for (int i = 0; i < 6; ++i) c->tire[i] = tire[i];
  // :52  v^2*1 + 1*r^2 + 5*ax^2 + 3*sr^2 + 2*sa^2 + 10*y^2
  c->stage_w[0] = 1; c->stage_w[1] = 1; c->stage_w[2] = 5; c->stage_w[3] = 3; c->stage_w[4] = 2; c->stage_w[5] = 10;
  c->veh_L = 3.0; c->veh_W = 2.0;               // PathPlanning/HybridAstar/main_hybrid_astar.jl:21  vehicle_size = [3, 2]
  c->veh_center_offset = 3.0 / 2;               // hybrid_astar_utils.jl:193  centre L/2 ahead of the rear axle
  return 0;
}

// strict FP32 flavour: the float instantiation takes sin/cos/tan/atan from include/mppi_strict_math.h
// (process-wide switch, set before a plan; see om:: in mppi_oracle.hpp)
void oracle_set_strict_f32(int on) { oracle::strict_f32() = on != 0; }

// MPPIPlan / DWAPlan on supplied noise (MPPI: eps [N][K][m] of float|double by
// `precision`; DWA: candidates K x m col-major double).  Arrays are col-major double.
int oracle_plan(const mppi_config_t* cfg, const double* circles, int Mc, const double* blocks, int Mp,
                const double* x0, const double* Unom, const void* eps_or_cand, int precision,
                int threads, int want_margin, double* costs, uint8_t* flags, double* margins,
                double* U_new, double* traj, double* scalars) {
  if (precision == MPPI_FP32)
    plan_impl<float>(cfg, circles, Mc, blocks, Mp, x0, Unom, eps_or_cand, threads, want_margin, costs,
                     flags, margins, U_new, traj, scalars);
  else
    plan_impl<double>(cfg, circles, Mc, blocks, Mp, x0, Unom, eps_or_cand, threads, want_margin, costs,
                      flags, margins, U_new, traj, scalars);
  return 0;
}

// TrajectoryRollout of one control sequence (N x m col-major, already clamped).
int oracle_rollout(const mppi_config_t* cfg, const double* circles, int Mc, const double* blocks,
                   int Mp, const double* x0, const double* Unom, const double* ctrl, int precision,
                   double* traj, double* scalars) {
  if (precision == MPPI_FP32) rollout_impl<float>(cfg, circles, Mc, blocks, Mp, x0, Unom, ctrl, traj, scalars);
  else rollout_impl<double>(cfg, circles, Mc, blocks, Mp, x0, Unom, ctrl, traj, scalars);
  return 0;
}

// VehicleDynamics(states, ctrl) -> (dstates, constraint, cost), FP64.
int oracle_vehicle_dynamics(const mppi_config_t* cfg, const double* x, const double* u, double* dx,
                            double* cons_cost) {
  Params<double> p = make_params<double>(cfg, nullptr, 0, nullptr, 0);
  vehicle_dynamics(p, x, u, dx, &cons_cost[0], &cons_cost[1]);
  return 0;
}

int oracle_rungekutta2(const mppi_config_t* cfg, const double* x, const double* u, double* x_next,
                       double* cons_cost) {
  Params<double> p = make_params<double>(cfg, nullptr, 0, nullptr, 0);
  int cons; double pc;
  rungekutta2(p, x, u, p.dt, x_next, &cons, &pc);
  cons_cost[0] = cons; cons_cost[1] = pc;
  return 0;
}

// The plant side of the closed loop, OptimalControl/MPPI/main.jl:45-47,65-67,76-83, for ONE MPC
// iteration: controls = Constant{Previous} interpolation of U_new (N x m col-major) over
// time_serial = range(0, T, length = N) sampled at fined_time_serial = range(0, update_time,
// length = steps) (:45-46,65-67); then `steps` explicit-Euler steps of VehicleDynamics at plant_dt
// (:76-78), each appended to the history as [time_idx * dt; states] (:79), stopping inside
// goal_radius of the goal (:80-82).  x is advanced in place; his is (1 + n_x) x steps col-major.
// Returns the number of columns written; *done = 1 if the goal test fired.
int oracle_plant_advance(const mppi_config_t* cfg, double* x, const double* U_new, int steps, double plant_dt,
                         double update_time, double goal_radius, int time_idx0, double* his, int* done) {
  Params<double> p = make_params<double>(cfg, nullptr, 0, nullptr, 0);
  const int N = cfg->N, nx = cfg->n_x;
  const double knot = N > 1 ? cfg->T / double(N - 1) : 0.0;            // step of range(0, T, length = N)
  const double fine = steps > 1 ? update_time / double(steps - 1) : 0.0;
  int ncol = 0;
  *done = 0;
  for (int i = 0; i < steps; ++i) {
    const double tf = double(i) * fine;                                // fined_time_serial[i + 1]
    int j = 0;                                                         // searchsortedlast(time_serial, tf) - 1
    while (j + 1 < N && double(j + 1) * knot <= tf) ++j;
    const double u[2] = {U_new[j], U_new[N + j]};                      // :67 controls[i, :]
    double dx[MPPI_MAX_NX], cons, cost;
    vehicle_dynamics(p, x, u, dx, &cons, &cost);                       // :76
    for (int q = 0; q < nx; ++q) x[q] = x[q] + dx[q] * plant_dt;       // :78
    his[size_t(ncol) * (1 + nx)] = double(time_idx0 + i + 1) * plant_dt;   // :79
    for (int q = 0; q < nx; ++q) his[size_t(ncol) * (1 + nx) + 1 + q] = x[q];
    ++ncol;
    const double gx = x[0] - cfg->goal[0], gy = x[1] - cfg->goal[1];
    if (gx * gx + gy * gy <= goal_radius * goal_radius) { *done = 1; break; }   // :80-82
  }
  return ncol;
}

// xy: 2 x n col-major (x0, y0, x1, y1, ...); n = 0 clears
void oracle_set_reference_path(const double* xy, int n, double weight, int window) {
  g_ref.assign(xy, xy + size_t(2) * n);
  g_ref_w = weight; g_ref_window = window;
}

void oracle_get_rectangle_pts(const double* block, double* pts10) { get_rectangle_pts(block, pts10); }
int oracle_sat(const double* base, int nb, const double* other, int no) {
  return separating_axis_theorem<double>(base, nb, other, no) ? 1 : 0;
}
// ConvexCollision on point sets (2 x n col-major); 1 = collision-free
int oracle_convex_collision(const double* p1, int n1, const double* p2, int n2) {
  return convex_collision<double>(p1, n1, p2, n2, nullptr) ? 1 : 0;
}
double oracle_mod_pi(double a) { return mod_pi(a); }

// block_collision_check per pose: hit[i] = number of obstacles pose i collides with
int oracle_check_poses(const mppi_config_t* cfg, const double* circles, int Mc, const double* blocks,
                       int Mp, const double* poses, int n, int* hit, double* margins) {
  Params<double> p = make_params<double>(cfg, circles, Mc, blocks, Mp);
  for (int i = 0; i < n; ++i) {
    const double s[2] = {poses[3 * i], poses[3 * i + 1]};
    int cons = 1; double cost = 0; Margin<double> mg;
    obstacle_evaluation_circles(p, s, &cons, &cost, &mg);
    obstacle_evaluation_blocks(p, poses[3 * i], poses[3 * i + 1], poses[3 * i + 2], &cons, &cost, &mg);
    hit[i] = int(std::lround(cost / p.obstacle_penalty));
    if (margins) margins[i] = mg.v;
  }
  return 0;
}

void oracle_philox4x32_10(const uint32_t* ctr, const uint32_t* key, uint32_t* out) {
  philox4x32_10(ctr, key, out);
}

int oracle_generate_noise(const double* Sigma, uint64_t seed, uint32_t iteration, uint32_t instance,
                          int k_offset, int K, int N, int precision, void* eps) {
  if (precision == MPPI_FP32) generate_noise<float>(Sigma, seed, iteration, instance, k_offset, K, N, (float*)eps);
  else generate_noise<double>(Sigma, seed, iteration, instance, k_offset, K, N, (double*)eps);
  return 0;
}

int oracle_weights(const double* cost, int n, double lambda, double* w, double* rho_eta, int* argmin) {
  std::vector<double> wv;
  calculate_mppi_weights(cost, n, lambda, &wv, argmin, &rho_eta[0], &rho_eta[1]);
  for (int i = 0; i < n; ++i) w[i] = wv[i];
  return 0;
}

}  // extern "C"
