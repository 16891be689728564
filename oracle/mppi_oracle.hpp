// mppi_oracle.hpp — CPU restatement of the reference MPPI hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// `--impl reference` legs may build, load or call it, and only as the checker
// or the timed CPU baseline.  The product (libmppi_b200.so) never links it.
//
// What it restates (paths relative to the reference root, Julia, all Float64):
//   OptimalControl/MPPI/src/MPPIUtils.jl:5-20,31-57,76-90,113-167,169-203
//   OptimalControl/MPPI/src/vehicledynamics.jl:1-54
//   OptimalControl/MPPI/src/types.jl:9,24,28,30
//   OptimalControl/Parking/src/VehicleModel.jl:1-21
//   PathPlanning/CollisionDetection/src/utils.jl:14-74
//   PathPlanning/HybridAstar/src/hybrid_astar_utils.jl:180-204
//   PathPlanning/ReedsSheppsCurves/src/ReedsSheppsUtils.jl:32-46
//   OptimalControl/DynamicWindow/src/DWAUtils.jl:16-42,105-118,141-163 (DWA mode)
//
// Parity pinning: the reference (Julia) cannot run in the build container and
// ships no tests.  The reference-owned golden data on this path are the two
// closed-loop CSV logs (MPPITrajectory.csv, DWATrajectory.csv).
//   PINNED  (1) rows k,k+1 of both logs pin the derivative output of
//               VehicleDynamics (tests/golden/*_plant_pairs.csv, <= 1e-12);
//           (2) the DWA run is deterministic: all 127 DWAPlan selections and
//               all 12,633 rows of DWATrajectory.csv are reproduced by this
//               restatement (tests/golden/dwa_plan_cases.csv,
//               tests/test_oracle_golden.py).  That pins rollout, Heun step,
//               stage cost, obstacle / bound / terminal terms and arg-min
//               selection, which DWAPlan shares with MPPIPlan.
//   "parity unpinned": the lambda term of the cost, softmin weights, weighted
//           update, cut-off, SAT walls and the kinematic model composition
//           (MPPI samples from an unseeded RNG; no artefact exists): these
//           rest on this file being a line-by-line restatement.
//
// Build: g++ -O2 -ffp-contract=off (no FMA contraction: Julia does not contract
// a*b+c), evaluation order as in the cited lines, sums left to right.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#include "../include/mppi_b200.h"
#include "../include/mppi_strict_math.h"

namespace oracle {

// Elementary functions of the restatement.  Julia's sin/cos/tan/atan are libm-class (<= 1 ulp);
// here: the C library's, in the precision of R.  For the strict FP32 flavour (north_star's
// "bit-exact argmin and flags", checked in tests/test_gpu_strict.py) the float instantiation
// switches to the portable functions of include/mppi_strict_math.h, which the GPU's strict
// rollout executes operation for operation; double is never affected.
inline bool& strict_f32() { static bool on = false; return on; }
namespace om {
inline double sin(double x) { return std::sin(x); }
inline double cos(double x) { return std::cos(x); }
inline double tan(double x) { return std::tan(x); }
inline double atan(double x) { return std::atan(x); }
inline float sin(float x) { return strict_f32() ? mppi_strict::sm_sin(x) : std::sin(x); }
inline float cos(float x) { return strict_f32() ? mppi_strict::sm_cos(x) : std::cos(x); }
inline float tan(float x) { return strict_f32() ? mppi_strict::sm_tan(x) : std::tan(x); }
inline float atan(float x) { return strict_f32() ? mppi_strict::sm_atan(x) : std::atan(x); }
}  // namespace om

// Smallest relative distance of any threshold comparison (obstacle, bound) from
// flipping; used by the tests' guard-band protocol, not part of the reference.
template <class R>
struct Margin {
  R v = std::numeric_limits<R>::infinity();
  void see(R a, R b) {
    R s = std::fmax(std::fmax(std::fabs(a), std::fabs(b)), R(1));
    R d = std::fabs(a - b) / s;
    if (d < v) v = d;
  }
};

template <class R>
struct Params {
  int model, n_x, m, N;
  R dt, lambda;
  R Sigma[4];  // column-major 2x2
  R goal[2], XL[MPPI_MAX_NX], XU[MPPI_MAX_NX], CL[2], CU[2];
  R slack_penalty, obstacle_penalty, terminal_scale;
  R la, lb, mass, Izz, g, mu, KFZF, KFZR, KFZX, tire[6], w[6];
  R veh_L, veh_W, veh_off;
  int planner;
  std::vector<R> circles;  // 3 per circle: x, y, r
  std::vector<R> blocks;   // 5 per block: cx, cy, psi, half_len, half_wid
  std::vector<R> wall_pts; // Block2Pts(blocks): 10 per block, [2][5] col-major
  // reference-path tracking term (an extension, absent from the reference's MPPI)
  std::vector<R> ref;      // 2 per point: x, y
  R ref_w = 0;
  int ref_window = 0;
};

// ---- tracking term: squared distance to the nearest polyline point inside a moving window -------
// The reference's MPPI has only the implicit lane y = 0 (`10*y^2`, vehicledynamics.jl:52); its path
// tracker finds the reference point as the closest of ref_pts[least_idx:maximum_idx], least_idx being
// the previous closest index (PathPlanning/HybridAstar/src/tracker_utils.jl:36-41, main_Tracker.jl:82-83).
// Here: window [idx, min(idx + W, n - 1)], first minimum wins (argmin), idx moves to it.
template <class R>
R track_scan(const Params<R>& p, R x, R y, int* idx) {
  const int n = int(p.ref.size() / 2);
  const int hi = std::min(*idx + p.ref_window, n - 1);
  R best = std::numeric_limits<R>::infinity();
  int bi = *idx;
  for (int i = *idx; i <= hi; ++i) {
    const R dx = x - p.ref[2 * i], dy = y - p.ref[2 * i + 1];
    const R d = dx * dx + dy * dy;
    if (d < best) { best = d; bi = i; }
  }
  *idx = bi;
  return best;
}
// the window starts at the point closest to x0 over the WHOLE path (findclosest with the full range)
template <class R>
int track_start(const Params<R>& p, R x, R y) {
  const int n = int(p.ref.size() / 2);
  R best = std::numeric_limits<R>::infinity();
  int bi = 0;
  for (int i = 0; i < n; ++i) {
    const R dx = x - p.ref[2 * i], dy = y - p.ref[2 * i + 1];
    const R d = dx * dx + dy * dy;
    if (d < best) { best = d; bi = i; }
  }
  return bi;
}

// ---- a5: VehicleDynamics, 7-state dynamic bicycle --------------------------
// OptimalControl/MPPI/src/vehicledynamics.jl:1-54.  State [x,y,v,r,psi,ux,sa],
// control [sr,ax]; derivative order [dx,dy,dv,dr,dpsi,du,dsa] (:50).
template <class R>
void vehicle_dynamics_dyn7(const Params<R>& p, const R* states, const R* ctrl, R* dstates,
                           R* constraint, R* cost) {
  const R la = p.la, lb = p.lb, M = p.mass, Izz = p.Izz, g = p.g, mu = p.mu;  // :2-7
  const R KFZF = p.KFZF, KFZR = p.KFZR, KFZX = p.KFZX;                        // :8-10
  const R B = p.tire[0] / mu;  // :14
  const R C = p.tire[1];       // :15
  const R E = p.tire[3];       // :16   (Sh, Sv :17-18 are read but unused)
  const R y = states[1], v = states[2], r = states[3], psi = states[4];  // :20-26
  R ux = states[5];
  const R sa = states[6];
  const R sr = ctrl[0], ax = ctrl[1];  // :27-28

  const R FZF = R(2) * (KFZF * g - (ax - r * v) * KFZX);                 // :30
  const R FZR = R(2) * (KFZR * g + (ax - r * v) * KFZX);                 // :31
  const R alpha_f = om::atan((v + la * r) / (ux + R(0.01))) - sa;       // :32
  const R alpha_r = om::atan((v - lb * r) / (ux + R(0.01)));            // :33
  const R X1_f = B * alpha_f;                                            // :35
  const R FY1 = mu * FZF * p.tire[2] * om::sin(C * om::atan(X1_f - E * (X1_f - om::atan(X1_f))));  // :36
  const R X1_r = B * alpha_r;                                            // :37
  const R FY2 = mu * FZR * p.tire[2] * om::sin(C * om::atan(X1_r - E * (X1_r - om::atan(X1_r))));  // :38
  if (ux <= R(0)) ux = R(0);  // :40-42 — after the slips, so it only affects dx, dy, dv
  const R dx = ux * om::cos(psi) - v * om::sin(psi);  // :43
  const R dy = ux * om::sin(psi) + v * om::cos(psi);  // :44
  const R dpsi = r;                                      // :45
  const R dv = (FY1 + FY2) / M - r * ux;                 // :46
  const R dr = (FY1 * la - FY2 * lb) / Izz;              // :47
  const R du = ax, dsa = sr;                             // :48-49
  dstates[0] = dx; dstates[1] = dy; dstates[2] = dv; dstates[3] = dr;    // :50
  dstates[4] = dpsi; dstates[5] = du; dstates[6] = dsa;
  *constraint = R(0);                                                    // :51
  // :52  v^2*1 + 1*r^2 + 5*ax^2 + 3*sr^2 + 2*sa^2 + 10*y^2, left to right
  *cost = v * v * p.w[0] + p.w[1] * (r * r) + p.w[2] * (ax * ax) + p.w[3] * (sr * sr) +
          p.w[4] * (sa * sa) + p.w[5] * (y * y);
}

// ---- a6: kinematic bicycle --------------------------------------------------
// OptimalControl/Parking/src/VehicleModel.jl:1-21.  State [x,y,psi,ux,sa],
// control [sr,ax].  The source returns derivatives only; SURVEY.md 8(a) a6 adds
// (identically in oracle and kernel): constraint = 0, stage cost = the dynamic
// model's cost with v = r = 0, no ux clamp.
template <class R>
void vehicle_dynamics_kin5(const Params<R>& p, const R* states, const R* ctrl, R* dstates,
                           R* constraint, R* cost) {
  const R L = p.la + p.lb;  // :2-4
  const R y = states[1], psi = states[2], ux = states[3], sa = states[4];  // :6-10
  const R sr = ctrl[0], ax = ctrl[1];                                       // :11-12
  dstates[0] = ux * om::cos(psi);      // :14
  dstates[1] = ux * om::sin(psi);      // :15
  dstates[2] = ux / L * om::tan(sa);   // :16
  dstates[3] = ax;                      // :17
  dstates[4] = sr;                      // :18
  *constraint = R(0);
  const R v = R(0), r = R(0);
  *cost = v * v * p.w[0] + p.w[1] * (r * r) + p.w[2] * (ax * ax) + p.w[3] * (sr * sr) +
          p.w[4] * (sa * sa) + p.w[5] * (y * y);
}

template <class R>
void vehicle_dynamics(const Params<R>& p, const R* x, const R* u, R* dx, R* cons, R* cost) {
  if (p.model == MPPI_MODEL_DYNAMIC7) vehicle_dynamics_dyn7(p, x, u, dx, cons, cost);
  else vehicle_dynamics_kin5(p, x, u, dx, cons, cost);
}

// ---- a4: Rungekutta2 (Heun) — MPPIUtils.jl:76-90 -----------------------------
template <class R>
void rungekutta2(const Params<R>& p, const R* x, const R* c, R dt, R* x_next, int* cons,
                 R* p_cost) {
  R k1[MPPI_MAX_NX], k2[MPPI_MAX_NX], xt[MPPI_MAX_NX], cv, pc, cv2, pc2;
  vehicle_dynamics(p, x, c, k1, &cv, &pc);                        // :77-78
  for (int i = 0; i < p.n_x; ++i) xt[i] = x[i] + k1[i] * dt;      // :79
  vehicle_dynamics(p, xt, c, k2, &cv2, &pc2);
  if (cv <= R(1)) {                                               // :83-88
    *cons = 1;
  } else {
    *cons = 0;
    pc = (cv - R(1)) * p.slack_penalty * R(10);
  }
  *p_cost = pc;
  for (int i = 0; i < p.n_x; ++i) x_next[i] = x[i] + dt * (k1[i] + k2[i]) / R(2);  // :89
}

// ---- a7: ObstacleEvaluation, circles — MPPIUtils.jl:120-132 ------------------
template <class R>
void obstacle_evaluation_circles(const Params<R>& p, const R* states, int* constraint,
                                 R* cost_value, Margin<R>* mg) {
  for (size_t o = 0; o + 2 < p.circles.size(); o += 3) {
    const R dx = states[0] - p.circles[o], dy = states[1] - p.circles[o + 1];
    const R d2 = dx * dx + dy * dy;                        // sum((s[1:2]-o[1:2]).^2)
    const R r2 = p.circles[o + 2] * p.circles[o + 2];      // o[3]^2
    if (mg) mg->see(d2, r2);
    if (d2 <= r2) {                                        // :125
      *constraint = 0;                                     // :126
      *cost_value = *cost_value + p.obstacle_penalty;      // :127 (100.0*712.5)
    }
  }
}

// ---- a8: polygon collision ---------------------------------------------------
// GetRectanglePts, CollisionDetection/src/utils.jl:14-25: 2x5 closed polygon,
// corners (-l,+w), (-l,-w), (+l,-w), (+l,+w), first again; pts[2*i + {0,1}].
template <class R>
void get_rectangle_pts(const R* block, R* pts) {
  const R ox = block[0], oy = block[1], psi = block[2], len = block[3], wid = block[4];
  const R lx[5] = {-len, -len, len, len, -len};
  const R ly[5] = {wid, -wid, -wid, wid, wid};
  const R c = om::cos(psi), s = om::sin(psi);
  for (int i = 0; i < 5; ++i) {  // R*pts .+ [ox; oy]   (:23-24)
    pts[2 * i] = (c * lx[i] + (-s) * ly[i]) + ox;
    pts[2 * i + 1] = (s * lx[i] + c * ly[i]) + oy;
  }
}

// SeparatingAxisTheorem, utils.jl:37-62.  true = a separating axis was found
// among pts_base's edges.  n_base / n_other = number of columns.
// gap (optional, guard-band bookkeeping only): best normalised separation over the
// edges, >= 0 iff some edge separates; when requested every edge is scanned.
template <class R>
bool separating_axis_theorem(const R* pts_base, int n_base, const R* pts_other, int n_other,
                             R* gap = nullptr) {
  bool found = false;
  R best = -std::numeric_limits<R>::infinity();
  for (int b = 0; b + 1 < n_base; ++b) {                                // :39
    const R bgx = pts_base[2 * b], bgy = pts_base[2 * b + 1];           // :41
    const R bvx = pts_base[2 * b + 2] - bgx, bvy = pts_base[2 * b + 3] - bgy;  // :42-44
    const R nx = -bvy, ny = bvx;                                        // :46
    R min_base = std::numeric_limits<R>::infinity(), max_base = -min_base;
    R min_other = min_base, max_other = -min_base;
    for (int i = 0; i < n_base; ++i) {                                  // :48
      const R d = (pts_base[2 * i] - bgx) * nx + (pts_base[2 * i + 1] - bgy) * ny;
      min_base = std::fmin(min_base, d); max_base = std::fmax(max_base, d);
    }
    for (int i = 0; i < n_other; ++i) {                                 // :49
      const R d = (pts_other[2 * i] - bgx) * nx + (pts_other[2 * i + 1] - bgy) * ny;
      min_other = std::fmin(min_other, d); max_other = std::fmax(max_other, d);
    }
    if (gap) {
      const R sc = std::fmax(std::fmax(std::fabs(min_base), std::fabs(max_base)),
                             std::fmax(std::fmax(std::fabs(min_other), std::fabs(max_other)), R(1)));
      best = std::fmax(best, std::fmax(min_base - max_other, min_other - max_base) / sc);
    }
    if ((max_other <= min_base) || (max_base <= min_other)) {           // :57
      if (!gap) return true;                                            // :58
      found = true;
    }
  }
  if (gap) *gap = best;
  return found;                                                         // :61
}

// ConvexCollision, utils.jl:64-74.  true = collision-FREE; needs a separating
// axis among pts1's edges AND among pts2's edges (stricter than textbook SAT).
// Guard-band margin = how far the OUTCOME is from flipping.
template <class R>
bool convex_collision(const R* pts1, int n1, const R* pts2, int n2, Margin<R>* mg) {
  if (mg) {
    R g1, g2;
    const bool a = separating_axis_theorem(pts1, n1, pts2, n2, &g1);
    const bool b = separating_axis_theorem(pts2, n2, pts1, n1, &g2);
    R m;
    if (a && b) m = std::fmin(g1, g2);                 // free: either direction failing flips it
    else if (a) m = std::fabs(g2);                     // hit because of direction 2 only
    else if (b) m = std::fabs(g1);
    else m = std::fmax(std::fabs(g1), std::fabs(g2));  // both must flip
    if (m < mg->v) mg->v = m;
    return a && b;
  }
  if (separating_axis_theorem(pts1, n1, pts2, n2)) {
    return separating_axis_theorem(pts2, n2, pts1, n1);
  }
  return false;
}

// modπ, ReedsSheppsCurves/src/ReedsSheppsUtils.jl:32-46 (Julia mod = floored)
template <class R>
R mod_pi(R angle) {
  const R pi = R(3.141592653589793);
  if (angle >= -pi && angle <= pi) return angle;
  const R two_pi = R(2) * pi;
  R a = std::fmod(angle, two_pi);
  if (a != R(0) && (a < R(0))) a += two_pi;  // Julia mod(x, y>0) in [0, y)
  if (a < -pi) a = a + two_pi;
  else if (a > pi) a = a - two_pi;
  return a;
}

// Vehicle rectangle from a pose, hybrid_astar_utils.jl:193:
// [x + L/2 cosψ, y + L/2 sinψ, modπ(ψ), L/2, W/2]; the centre offset L/2 is the
// config field veh_center_offset (default veh_L/2).
template <class R>
void vehicle_block(const Params<R>& p, R x, R y, R psi, R* block) {
  block[0] = x + p.veh_off * om::cos(psi);
  block[1] = y + p.veh_off * om::sin(psi);
  block[2] = mod_pi(psi);
  block[3] = p.veh_L / R(2);
  block[4] = p.veh_W / R(2);
}

// Polygon ObstacleEvaluation = block_collision_check (hybrid_astar_utils.jl:196-201)
// for one pose, with a7's penalty per wall hit (SURVEY.md 8(a) a8 definition).
template <class R>
void obstacle_evaluation_blocks(const Params<R>& p, R x, R y, R psi, int* constraint,
                                R* cost_value, Margin<R>* mg) {
  if (p.blocks.empty()) return;
  R vb[5], vp[10];
  vehicle_block(p, x, y, psi, vb);
  get_rectangle_pts(vb, vp);
  const size_t nb = p.blocks.size() / 5;
  for (size_t i = 0; i < nb; ++i) {
    if (!convex_collision(&p.wall_pts[10 * i], 5, vp, 5, mg)) {  // :198
      *constraint = 0;
      *cost_value = *cost_value + p.obstacle_penalty;
    }
  }
}

template <class R>
R heading_of(const Params<R>& p, const R* s) {
  return p.model == MPPI_MODEL_DYNAMIC7 ? s[4] : s[2];
}

template <class R>
void obstacle_evaluation(const Params<R>& p, const R* states, int* constraint, R* cost_value,
                         int* hit, Margin<R>* mg) {
  *constraint = 1;     // :121
  *cost_value = R(0);  // :122
  obstacle_evaluation_circles(p, states, constraint, cost_value, mg);
  obstacle_evaluation_blocks(p, states[0], states[1], heading_of(p, states), constraint,
                             cost_value, mg);
  if (*constraint == 0) *hit = 1;
}

// ---- a9: BoundEvaluation — MPPIUtils.jl:135-151 ------------------------------
template <class R>
void bound_evaluation(const Params<R>& p, const R* states, int* constraint, R* cost_value,
                      Margin<R>* mg) {
  *constraint = 1;
  *cost_value = R(0);
  for (int i = 0; i < p.n_x; ++i) {
    if (mg) { mg->see(states[i], p.XL[i]); mg->see(states[i], p.XU[i]); }
    if (states[i] < p.XL[i]) {  // :139
      *constraint = 0;
      *cost_value = *cost_value + p.slack_penalty * std::fabs(states[i] - p.XL[i]);  // :142
    }
    if (states[i] > p.XU[i]) {  // :144
      *constraint = 0;
      *cost_value = *cost_value + p.slack_penalty * std::fabs(states[i] - p.XU[i]);  // :147
    }
  }
}

// ---- a10: TerminalCostEvaluation — MPPIUtils.jl:113-116 ----------------------
template <class R>
R terminal_cost_evaluation(const Params<R>& p, const R* s) {
  return (s[0] - p.goal[0]) * (s[0] - p.goal[0]) + (s[1] - p.goal[1]) * (s[1] - p.goal[1]);
}

// ---- a2: PushInBounds — MPPIUtils.jl:13-20 -----------------------------------
template <class R>
void push_in_bounds(const Params<R>& p, R* ctrl /* N x m row j at ctrl[j*m] */) {
  for (int i = 0; i < p.N; ++i)
    for (int j = 0; j < p.m; ++j)
      ctrl[i * p.m + j] = std::fmin(std::fmax(ctrl[i * p.m + j], p.CL[j]), p.CU[j]);
}

}  // namespace oracle
