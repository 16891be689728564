// mppi_track.cuh — reference-path tracking term of the stage cost.
//
// The reference's MPPI tracks the implicit lane y = 0 (`10*y^2`, vehicledynamics.jl:52).  Its path
// tracker finds the reference point as the closest of ref_pts[least_idx:maximum_idx], least_idx
// being the previous closest index (HybridAstar/src/tracker_utils.jl:36-41, main_Tracker.jl:82-83).
// Generalisation used here (restated on the CPU for the tests as track_scan / track_start):
//   cost_j += ref_w * min_{i in [idx, min(idx + W, n-1)]} |(x_j, y_j) - ref_i|^2,  idx <- argmin (first)
// evaluated wherever the stage cost is (pre-step states x_0..x_{N-1} and the terminal slot x_N),
// idx starting at the point closest to x_0 over the whole path.  The polyline (<= MPPI_MAX_REF
// points) is staged in shared memory once per CTA; the window scan is W+1 fma pairs per step.
// The rollout kernel carries it as a separate instantiation (TRACK), so the plain kernels keep
// their register budget.
#pragma once
#include "mppi_device.cuh"

namespace mppi {

template <class S>
__device__ __forceinline__ S track_scan(const S* ref, int n, int W, S x, S y, int& idx) {
  const int hi = min(idx + W, n - 1);
  S best = S(INFINITY);
  int bi = idx;
  for (int i = idx; i <= hi; ++i) {
    const S dx = x - ref[2 * i], dy = y - ref[2 * i + 1];
    const S d = dx * dx + dy * dy;
    if (d < best) { best = d; bi = i; }
  }
  idx = bi;
  return best;
}

// index of the point closest to (x, y) over the whole path, first minimum; all threads of the block
// call it, all get the answer.  s_d / s_i: shared scratch for one value per warp.
template <class S>
__device__ __forceinline__ int track_start(const S* ref, int n, S x, S y, S* s_d, int* s_i) {
  const int t = threadIdx.x, nt = blockDim.x, lane = t & 31, wid = t >> 5, nw = (nt + 31) >> 5;
  S best = S(INFINITY);
  int bi = 0x7FFFFFFF;
  for (int i = t; i < n; i += nt) {
    const S dx = x - ref[2 * i], dy = y - ref[2 * i + 1];
    const S d = dx * dx + dy * dy;
    if (d < best) { best = d; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const S od = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (od < best || (od == best && oi < bi)) { best = od; bi = oi; }
  }
  if (lane == 0) { s_d[wid] = best; s_i[wid] = bi; }
  __syncthreads();
  best = s_d[0]; bi = s_i[0];
  for (int w = 1; w < nw; ++w)
    if (s_d[w] < best || (s_d[w] == best && s_i[w] < bi)) { best = s_d[w]; bi = s_i[w]; }
  __syncthreads();
  return bi == 0x7FFFFFFF ? 0 : bi;
}

// what a rollout thread carries: nothing when the term is off
template <class V, bool ON> struct Tracker;
template <class V> struct Tracker<V, false> {
  __device__ __forceinline__ void add(V&, V, V) {}
};
template <class V> struct Tracker<V, true> {
  using M = Math<V>;
  using S = typename M::S;
  const S* ref;
  int n, W;
  S w;
  int idx[M::LANES];
  __device__ __forceinline__ void add(V& stage, V x, V y) {
    S c[M::LANES];
#pragma unroll
    for (int l = 0; l < M::LANES; ++l) c[l] = track_scan(ref, n, W, M::lane(x, l), M::lane(y, l), idx[l]);
    stage += V(w) * M::make2(c[0], c[M::LANES - 1]);
  }
};

}  // namespace mppi
