// mppi_kernels.cuh — the kernels of one MPPI iteration (MPPIPlan, MPPIUtils.jl:169-203):
//   rollout_kernel   K-sample rollout + fused stage/obstacle/bound/terminal cost,
//                    noise (supplied | Philox | DWA grid), per-block min -> atomicMin
//   cutoff_kernel    optional emulation of the reference's sequential feasibility cut-off
//   update_kernel    softmin weights on the fly + weighted control sum (HBM-bound stream)
//   finalize_kernel  partial -> (multi-rank) combine -> U_new -> nominal rollout
#pragma once
#include "mppi_models.cuh"

namespace mppi {

enum { SRC_SUPPLIED = 0, SRC_PHILOX = 1, SRC_DWA = 2 };
constexpr int ROLLOUT_BS = 64;   // 65,536 samples = 1024 CTAs = 6.9 per SM: one balanced wave on 148 SMs
constexpr int UPDATE_BS = 256;
constexpr int UPDATE_JG = 4;
constexpr int FINAL_BS = 256;
constexpr int CUTOFF_BS = 1024;


__device__ __forceinline__ unsigned long long shfl_xor_u64(unsigned long long v, int m) {
  return __shfl_xor_sync(0xffffffffu, v, m);
}
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long t = shfl_xor_u64(v, o);
    v = t < v ? t : v;
  }
  return v;
}
template <class T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// pack (cost, local sample id) so that an unsigned min gives (min cost, first index)
__device__ __forceinline__ unsigned long long make_key(float c, unsigned int k) {
  return ((unsigned long long)ordered_bits(c) << 32) | k;
}
__device__ __forceinline__ unsigned long long make_key(double c, unsigned int) { return ordered_bits(c); }
__device__ __forceinline__ float key_cost(unsigned long long key, float) { return unordered_bits((uint32_t)(key >> 32)); }
__device__ __forceinline__ double key_cost(unsigned long long key, double) { return unordered_bits(key); }

// ---------------------------------------------------------------------------
// rollout + cost.  One sample per thread; state, running cost and flags live in
// registers; U_nom[j] and g_j = lambda*inv(Sigma)*U_nom[j] are broadcast from
// shared memory; bounds, obstacle tables and vehicle constants are constant-bank
// operands (StaticParams is a __grid_constant__ argument).
// Schedule (MPPIUtils.jl:39-53): states x_1..x_N are each tested once against
// obstacles and bounds (x_0 never is); stage cost is taken at the pre-step state.
// ---------------------------------------------------------------------------
// Step prologue for B == 1: pull the caller's x0 / U_nom (mapped pinned host memory, read over
// PCIe by one small CTA) into the device input block.  Replaces an H2D memcpy node in the graph.
__global__ void __launch_bounds__(128) begin_kernel(const double* __restrict__ host_in, double* __restrict__ dev_in, int first, int count) {
  for (int i = threadIdx.x; i < count; i += 128) dev_in[first + i] = host_in[first + i];
}

// Per-sample accumulators of the rollout (registers).
template <class R>
struct RolloutAcc {
  R stage = R(0);   // stage + control cost (MPPIUtils.jl:45, vehicledynamics.jl:52)
  R viol = R(0);    // sum of bound violations (slack applied once at the end)
  int hits = 0;     // obstacle hits
};

// The horizon loop for one sample, specialised on the number of circle quads held in
// registers (NQ >= 0) or reading the circle table (NQ < 0).
template <class R, class Model, int SRC, int NQ>
__device__ __forceinline__ void rollout_horizon(const StaticParams<R>& p, const Buffers<R>& buf, Model& veh,
                                                RolloutAcc<R>& acc, const HotParams<R, Model::NX>* sHot,
                                                const R* sU, const R* sBlk, int b, int k) {
  using R2 = typename Vec2<R>::type;
  using R4 = typename Vec4<R>::type;
  const int N = p.N;
  const HotParams<R, Model::NX> hp = *sHot;   // smem -> registers (unused fields are dead)
  const bool walls = p.n_blocks > 0;
  const size_t estride = size_t(p.K_local);
  R2* eps = reinterpret_cast<R2*>(buf.eps) + size_t(b) * N * estride + k;
  const R4* nu_p = reinterpret_cast<const R4*>(sU);
  R cu0 = R(0), cu1 = R(0);
  if (SRC == SRC_DWA) {
    cu0 = R(buf.cand[k]);
    cu1 = R(buf.cand[estride + k]);
  }
  // one horizon step: PushInBounds, control cost, stage cost, Heun step, tests of x_{j+1}
  auto advance = [&](R4 nu, R2 e) {
    R u0, u1;
    if (SRC == SRC_DWA) {
      u0 = cu0; u1 = cu1;
    } else {
      u0 = Math<R>::fmin(Math<R>::fmax(nu.x + e.x, hp.CL[0]), hp.CU[0]);   // PushInBounds (:13-20)
      u1 = Math<R>::fmin(Math<R>::fmax(nu.y + e.y, hp.CL[1]), hp.CU[1]);
      acc.stage += nu.z * (u0 - nu.x) + nu.w * (u1 - nu.y);                // lambda*U_nom'*inv(Sigma)*(u-U_nom) (:45)
    }
    acc.stage += veh.stage_cost(hp, u0, u1);       // VehicleDynamics cost at the pre-step state
    veh.step(hp, u0, u1);                          // Rungekutta2 (:76-90)
    if (NQ >= 0) acc.hits += circle_hits_regs<R, (NQ > 0 ? NQ : 0)>(hp, veh.s[0], veh.s[1]);   // x_{j+1}: ObstacleEvaluation
    else acc.hits += circle_hits(p, veh.s[0], veh.s[1]);
    if (walls) acc.hits += block_hits<R>(hp, hp.blk_all, sBlk, p.n_blocks, veh.s[0], veh.s[1], veh.sn, veh.cs);
    acc.viol += bound_violation<R, Model::NX>(hp, veh.s);                 // BoundEvaluation
  };
  if (SRC == SRC_PHILOX) {
    // one Philox4x32-10 call feeds two horizon steps; eps = chol(Sigma) * z (MPPIUtils.jl:6)
    // is kept for the update kernel / oracle replay
    const R L00 = hp.L00, L10 = hp.L10, L11 = hp.L11;
    const uint32_t gk = uint32_t(p.k_offset + k), iter = *buf.iter;
    // software pipeline: the noise of pair j+2 is generated while pair j integrates, so the
    // integer Philox rounds and the MUFU work of Box-Muller fill the FP32 recurrence's issue gaps
    auto draw = [&](int j, R2& ea, R2& eb) {
      const uint4 rnd = philox4x32_10(make_uint4(gk, uint32_t(j >> 1), iter, uint32_t(b)), hp.key);
      float z0, z1;
      box_muller(rnd.x, rnd.y, &z0, &z1);
      ea.x = L00 * R(z0);
      ea.y = L10 * R(z0) + L11 * R(z1);
      box_muller(rnd.z, rnd.w, &z0, &z1);
      eb.x = L00 * R(z0);
      eb.y = L10 * R(z0) + L11 * R(z1);
    };
    R2 na, nb;
    draw(0, na, nb);
    for (int j = 0; j < N; j += 2) {
      const R2 ea = na, eb = nb;
      if (j + 2 < N) draw(j + 2, na, nb);
      *eps = ea;
      advance(nu_p[j], ea);
      if (j + 1 < N) {
        eps[estride] = eb;
        advance(nu_p[j + 1], eb);
      }
      eps += 2 * estride;
    }
  } else if (SRC == SRC_SUPPLIED) {
    R2 e_next = *eps;
    for (int j = 0; j < N; ++j) {
      const R2 e = e_next;
      eps += estride;
      if (j + 1 < N) e_next = *eps;   // prefetch next step's noise
      advance(nu_p[j], e);
    }
  } else {
    for (int j = 0; j < N; ++j) advance(R4{R(0), R(0), R(0), R(0)}, R2{R(0), R(0)});
  }
}

template <class R, class Model, int SRC>
__global__ void __launch_bounds__(ROLLOUT_BS)
rollout_kernel(const __grid_constant__ StaticParams<R> p, const Buffers<R> buf) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  R* sU = reinterpret_cast<R*>(smem_raw);  // [N][4]: u0, u1, g0, g1
  R* sBlk = sU + 4 * p.N;                  // [n_blocks][8]: wall table
  __shared__ unsigned long long s_key[ROLLOUT_BS / 32];
  __shared__ int s_feas[ROLLOUT_BS / 32];
  __shared__ __align__(16) HotParams<R, Model::NX> sHot;

  const int b = blockIdx.y;
  const int N = p.N;
  const int k = blockIdx.x * ROLLOUT_BS + threadIdx.x;
  trace_begin(buf.dbg, buf.trace, 0);
  if (threadIdx.x == ROLLOUT_BS - 1) sHot.fill(p);
  for (int i = threadIdx.x; i < 8 * p.n_blocks; i += ROLLOUT_BS) sBlk[i] = (&p.blk[0][0])[i];
  const bool active = k < p.K_local;
  const double* Ub = buf.U + size_t(b) * 2 * N;
  if (SRC != SRC_DWA) {
    for (int j = threadIdx.x; j < N; j += ROLLOUT_BS) {
      const R u0 = R(Ub[j]), u1 = R(Ub[N + j]);
      sU[4 * j + 0] = u0;
      sU[4 * j + 1] = u1;
      sU[4 * j + 2] = p.G00 * u0 + p.G01 * u1;
      sU[4 * j + 3] = p.G10 * u0 + p.G11 * u1;
    }
  }
  __syncthreads();

  R cost = R(0);
  RolloutAcc<R> acc;
  if (active) {
    const double* x0 = buf.x0 + size_t(b) * MPPI_MAX_NX;
    Model veh;
    veh.init(p, x0);
    const R gx0 = veh.s[0] - p.goal[0], gy0 = veh.s[1] - p.goal[1];
    const R d0 = gx0 * gx0 + gy0 * gy0;
    const int nq = (p.n_circles + 3) >> 2;   // warp-uniform: the whole horizon loop is specialised
    if (sizeof(R) == 8) rollout_horizon<R, Model, SRC, -1>(p, buf, veh, acc, &sHot, sU, sBlk, b, k);
    else if (nq == 0) rollout_horizon<R, Model, SRC, 0>(p, buf, veh, acc, &sHot, sU, sBlk, b, k);
    else if (nq == 1) rollout_horizon<R, Model, SRC, 1>(p, buf, veh, acc, &sHot, sU, sBlk, b, k);
    else if (nq == 2) rollout_horizon<R, Model, SRC, 2>(p, buf, veh, acc, &sHot, sU, sBlk, b, k);
    else rollout_horizon<R, Model, SRC, -1>(p, buf, veh, acc, &sHot, sU, sBlk, b, k);
    acc.stage += veh.stage_cost(p, R(0), R(0));      // terminal slot: VehicleDynamics(x_N, [0,0]) (:49,:53)
    const R gx = veh.s[0] - p.goal[0], gy = veh.s[1] - p.goal[1];
    const R term = Math<R>::div(gx * gx + gy * gy, d0) * p.term_scale;   // (:54)
    cost = acc.stage + p.slack * acc.viol + p.obs_pen * R(acc.hits) + term;
    buf.costs[size_t(b) * p.K_local + k] = cost;
    buf.flags[size_t(b) * p.K_local + k] =
        uint8_t(((acc.hits == 0 && !(acc.viol > R(0))) ? MPPI_FLAG_FEASIBLE : 0) | (acc.hits ? MPPI_FLAG_COLLIDED : 0) |
                ((acc.viol > R(0)) ? MPPI_FLAG_OOB : 0));
  }
  // block epilogue: (min cost, first index) and feasible count -> one atomic each
  unsigned long long key = active ? make_key(cost, (unsigned int)k) : 0xFFFFFFFFFFFFFFFFull;
  key = warp_min_u64(key);
  const unsigned feas_mask = __ballot_sync(0xffffffffu, active && acc.hits == 0 && !(acc.viol > R(0)));
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { s_key[wid] = key; s_feas[wid] = __popc(feas_mask); }
  __syncthreads();
  if (threadIdx.x == 0) {
    int f = 0;
#pragma unroll
    for (int w = 0; w < ROLLOUT_BS / 32; ++w) { key = s_key[w] < key ? s_key[w] : key; f += s_feas[w]; }
    atomicMin(&buf.scratch[b].rho_key, key);
    if (f) atomicAdd(&buf.scratch[b].feasible_count, f);
  }
  trace_end(buf.dbg, buf.trace, 0);
}

// ---------------------------------------------------------------------------
// Feasibility cut-off (MPPIUtils.jl:175,185): the reference stops sampling right
// after the (feasibility_count+1)-th feasible sample; emulate by a prefix count
// in sample order, then recompute (min, argmin) over the surviving prefix.
// One block per instance.  Only launched when feasibility_count < K.
// ---------------------------------------------------------------------------
template <class R>
__global__ void __launch_bounds__(CUTOFF_BS) cutoff_kernel(const __grid_constant__ StaticParams<R> p, const Buffers<R> buf) {
  __shared__ int s_cnt[CUTOFF_BS];
  __shared__ int s_keff, s_feas;
  __shared__ unsigned long long s_key[CUTOFF_BS / 32];
  __shared__ unsigned int s_arg[CUTOFF_BS / 32];
  const int b = blockIdx.x, t = threadIdx.x;
  const int K = p.K_local;
  const uint8_t* fl = buf.flags + size_t(b) * K;
  const R* cs = buf.costs + size_t(b) * K;
  const int chunk = (K + CUTOFF_BS - 1) / CUTOFF_BS;
  const int k0 = min(t * chunk, K), k1 = min(k0 + chunk, K);
  int cnt = 0;
  for (int k = k0; k < k1; ++k) cnt += fl[k] & MPPI_FLAG_FEASIBLE;
  s_cnt[t] = cnt;
  if (t == 0) { s_keff = K; s_feas = -1; }
  __syncthreads();
  // inclusive scan (Hillis-Steele; one block, not on the fast path)
  for (int o = 1; o < CUTOFF_BS; o <<= 1) {
    const int v = (t >= o) ? s_cnt[t - o] : 0;
    __syncthreads();
    s_cnt[t] += v;
    __syncthreads();
  }
  const int incl = s_cnt[t], excl = incl - cnt;
  const int target = p.feasibility_count + 1;   // loop exits once FeasibilityCount > feasibility_count
  if (excl < target && target <= incl) {
    int seen = excl;
    for (int k = k0; k < k1; ++k) {
      seen += fl[k] & MPPI_FLAG_FEASIBLE;
      if (seen == target) { s_keff = k + 1; s_feas = target; break; }
    }
  }
  __syncthreads();
  const int keff = s_keff;
  unsigned long long key = 0xFFFFFFFFFFFFFFFFull;
  unsigned int arg = 0xFFFFFFFFu;
  for (int k = k0; k < min(k1, keff); ++k) {
    const unsigned long long kk = make_key(cs[k], (unsigned int)k);
    if (kk < key) { key = kk; arg = k; }
  }
  // block min of (key, arg) — lexicographic so ties resolve to the first index
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long ok = shfl_xor_u64(key, o);
    const unsigned int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (ok < key || (ok == key && oa < arg)) { key = ok; arg = oa; }
  }
  if ((t & 31) == 0) { s_key[t >> 5] = key; s_arg[t >> 5] = arg; }
  __syncthreads();
  if (t == 0) {
    for (int w = 0; w < CUTOFF_BS / 32; ++w)
      if (s_key[w] < key || (s_key[w] == key && s_arg[w] < arg)) { key = s_key[w]; arg = s_arg[w]; }
    Scratch& sc = buf.scratch[b];
    sc.rho_key = key;
    sc.argmin = arg;
    sc.k_eff = keff;
    sc.feasible_count = (s_feas >= 0) ? s_feas : s_cnt[CUTOFF_BS - 1];
  }
}

}  // namespace mppi
