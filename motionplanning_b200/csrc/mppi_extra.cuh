// mppi_extra.cuh — auxiliary kernels: batched pose collision query, sampled
// trajectory export for plotting, FP32 FMA-chain peak probe.
#pragma once
#include "mppi_update.cuh"

namespace mppi {

// block_collision_check (hybrid_astar_utils.jl:193-201) + circle test for n poses.
// poses: 3 x n col-major (x, y, psi).  FP64 so that hit counts match the oracle.
__global__ void check_poses_kernel(const __grid_constant__ StaticParams<double> p, const double* poses, int n, int* hit) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = poses[3 * i], y = poses[3 * i + 1], psi = poses[3 * i + 2];
  double sn, cs;
  ::sincos(psi, &sn, &cs);
  hit[i] = circle_hits(p, x, y) + block_hits(p, x, y, sn, cs);
}

// xy of every stride-th sampled trajectory (plotRes, MPPIUtils.jl:221-223), re-rolled
// from the stored eps.  out: [n_samples][N+1][2] float.
template <class R, class Model>
__global__ void sample_xy_kernel(const __grid_constant__ StaticParams<R> p, const Buffers<R> buf, int stride, int n_samples, float* out) {
  using R2 = typename Vec2<R>::type;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_samples) return;
  const int k = i * stride, N = p.N;
  Model veh;
  veh.init(p, buf.x0);
  float* o = out + size_t(i) * (N + 1) * 2;
  o[0] = float(veh.s[0]); o[1] = float(veh.s[1]);
  const R2* eps = reinterpret_cast<const R2*>(buf.eps) + k;
  for (int j = 0; j < N; ++j) {
    R u0, u1;
    if (p.planner == MPPI_PLANNER_DWA) {
      u0 = R(buf.cand[k]); u1 = R(buf.cand[size_t(p.K_local) + k]);
    } else {
      const R2 e = eps[size_t(j) * p.K_local];
      u0 = Math<R>::fmin(Math<R>::fmax(R(buf.U[j]) + e.x, p.CL[0]), p.CU[0]);
      u1 = Math<R>::fmin(Math<R>::fmax(R(buf.U[N + j]) + e.y, p.CL[1]), p.CU[1]);
    }
    veh.step(p, u0, u1);
    o[2 * (j + 1)] = float(veh.s[0]); o[2 * (j + 1) + 1] = float(veh.s[1]);
  }
}

// FP32 peak probe: 8 independent FFMA chains per thread, register-resident.
// flops = threads * iters * 8 chains * 2.
__global__ void __launch_bounds__(256) fma_peak_kernel(float* out, int iters, float a, float b) {
  float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
#pragma unroll 1
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
      x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
    }
  }
  const float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
  if (s == 123.456f) out[0] = s;   // never true; keeps the chains alive
}

}  // namespace mppi
