// mppi_plant.cuh — the plant side of the closed loop, on the device.
//
// OptimalControl/MPPI/main.jl:58-83 alternates one MPPIPlan with update_idx = 100 explicit-Euler
// steps of VehicleDynamics at dt = 1 ms under the zero-order-held first controls of the new
// plan.  Run from the host that is a device -> host -> device round trip per iteration for 56
// bytes of state; here the plant is one more kernel node behind finalize, so a whole episode is a
// chain of graph launches and the host only watches a flag.  The plant is sequential (100 dependent
// steps) and runs in FP64 like the reference's: one thread, ~30 us per iteration.
#pragma once
#include "mppi_update.cuh"

namespace mppi {

struct PlantArgs {
  double* x0;                 // [n_x] this instance's X0 in the device input block: read, advanced, written back
  const double* U;            // [2][N] U_new of the iteration that just finished (result block)
  const double* res;          // [RES_SCALARS] its scalars (for the per-iteration log)
  double* his;                // [(1 + n_x)][cap] col-major history, column = [t; states] (main.jl:79)
  int* state;                 // [0] columns written, [1] episode over, [2] iterations run
  volatile int* host_done;    // mapped pinned copy of state[1]: the only thing the host polls
  double* iter_log;           // [4][max_iters]: nominal cost, feasible, RolloutCount, rho
  int N, steps, cap, max_iters;
  double knot_dt;             // spacing of time_serial = range(0, T, length = N) (main.jl:45)
  double fine_dt;             // spacing of fined_time_serial = range(0, update_time, length = steps) (:46)
  double h;                   // plant dt (:19)
  double goal_x, goal_y, r2;  // stop inside this radius of the goal (:80)
};

// dstates = VehicleDynamics(states, controls) in FP64, states in the model's own order
template <class Model> struct PlantDeriv;
template <> struct PlantDeriv<Kin5<double>> {   // Parking/src/VehicleModel.jl:14-18
  static __device__ __forceinline__ void run(const StaticParams<double>& p, const double* s, double sr, double ax, double* d) {
    double sn, cs;
    ::sincos(s[2], &sn, &cs);
    d[0] = s[3] * cs;
    d[1] = s[3] * sn;
    d[2] = s[3] / p.Lwb * ::tan(s[4]);
    d[3] = ax;
    d[4] = sr;
  }
};
template <> struct PlantDeriv<Dyn7<double>> {   // MPPI/src/vehicledynamics.jl:30-50
  static __device__ __forceinline__ void run(const StaticParams<double>& p, const double* s, double sr, double ax, double* d) {
    double sn, cs;
    ::sincos(s[4], &sn, &cs);
    Dyn7<double>::deriv(p, s[2], s[3], s[5], s[6], sn, cs, ax, &d[0], &d[1], &d[2], &d[3]);
    d[4] = s[3];
    d[5] = ax;
    d[6] = sr;
  }
};

template <class Model>
__global__ void plant_kernel(const __grid_constant__ StaticParams<double> p, const PlantArgs a) {
  constexpr int NX = Model::NX;
  if (threadIdx.x != 0 || a.state[1]) return;   // episode over: iterations already queued behind it do nothing here
  int ncol = a.state[0];
  const int it = a.state[2];
  if (a.iter_log && it < a.max_iters) {
    double* lg = a.iter_log + 4 * size_t(it);
    lg[0] = a.res[RES_COST]; lg[1] = a.res[RES_FEASIBLE]; lg[2] = a.res[RES_ROLLOUTS]; lg[3] = a.res[RES_RHO];
  }
  double s[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) s[i] = a.x0[i];
  bool done = false;
  // The held controls change at most a few times per block of plant steps: they are re-read from global memory only when
  // the knot index moves (a dependent L2 round trip per step otherwise, a third of the step), and the knot index comes from
  // a multiplication by the reciprocal spacing, made exact by the two corrections below.
  const double inv_knot = a.knot_dt > 0.0 ? 1.0 / a.knot_dt : 0.0;
  int jc = -1;
  double sr = 0.0, ax = 0.0;
  for (int i = 0; i < a.steps && !done; ++i) {
    // controls[i, :] = Constant{Previous} interpolation of U_new at fined_time_serial[i] (:65-67):
    // the last knot of time_serial that is <= t
    const double tf = double(i) * a.fine_dt;
    int j = int(tf * inv_knot);
    j = max(0, min(j, a.N - 1));
    while (j + 1 < a.N && double(j + 1) * a.knot_dt <= tf) ++j;
    while (j > 0 && double(j) * a.knot_dt > tf) --j;
    if (j != jc) { jc = j; sr = a.U[j]; ax = a.U[a.N + j]; }
    double d[NX];
    PlantDeriv<Model>::run(p, s, sr, ax, d);      // :76
#pragma unroll
    for (int q = 0; q < NX; ++q) s[q] = s[q] + d[q] * a.h;   // :78
    if (ncol < a.cap) {                                        // :79
      double* col = a.his + size_t(ncol) * (1 + NX);
      col[0] = double(it * a.steps + i + 1) * a.h;             // time_idx * dt
#pragma unroll
      for (int q = 0; q < NX; ++q) col[1 + q] = s[q];
      ++ncol;
    }
    const double gx = s[0] - a.goal_x, gy = s[1] - a.goal_y;
    done = gx * gx + gy * gy <= a.r2;                          // :80-82
  }
#pragma unroll
  for (int i = 0; i < NX; ++i) a.x0[i] = s[i];                 // the next iteration's ShiftInitialCondition (:59)
  a.state[0] = ncol;
  a.state[2] = it + 1;
  if (done || it + 1 >= a.max_iters) {
    a.state[1] = 1;
    __threadfence_system();
    *a.host_done = 1;
  }
}

// conventional warm start (an option; the reference re-uses U_new unshifted, main.jl:62):
// U_nom[c][j] = U_new[c][min(j + k, N - 1)]
__global__ void shift_nominal_kernel(double* U_nom, const double* U_new, int N, int k) {
  for (int i = threadIdx.x; i < 2 * N; i += blockDim.x) {
    const int c = i / N, j = i - c * N;
    U_nom[i] = U_new[c * N + min(j + k, N - 1)];
  }
}

}  // namespace mppi
