// mppi_kernels.cuh — the kernels of one MPPI iteration (MPPIPlan, MPPIUtils.jl:169-203):
//   rollout_kernel   K-sample rollout + fused stage/obstacle/bound/terminal cost,
//                    noise (supplied | Philox | DWA grid), per-block min -> atomicMin
//   cutoff_kernel    optional emulation of the reference's sequential feasibility cut-off
//   update_kernel    softmin weights on the fly + weighted control sum (HBM-bound stream)
//   finalize_kernel  partial -> (multi-rank) combine -> U_new -> nominal rollout
#pragma once
#include "mppi_models.cuh"
#include "mppi_track.cuh"

namespace mppi {

enum { SRC_SUPPLIED = 0, SRC_PHILOX = 1, SRC_DWA = 2 };
constexpr int EPS_GROUP = 8;         // horizon rows per mbarrier when the eps slab is staged in shared memory
constexpr int EPS_MAX_GROUPS = 16;
constexpr int EPS_SLAB_MAX = 52 * 1024;   // bytes per block, packed threads (4 blocks per SM); one-sample threads stage up to 32 KB (7 per SM)
constexpr int ROLLOUT_BS = 64;   // 65,536 samples = 1024 CTAs = 6.9 per SM: one balanced wave on 148 SMs
constexpr int UPDATE_BS = 256;
constexpr int UPDATE_JG = 4;
constexpr int FINAL_BS = 1024;  // 32 warps: the partial tree covers 128 columns per pass with 4 per warp
constexpr int FINAL_BS_BATCH = 128;   // batched instances (B >= FINAL_BATCH_MIN_B): occupancy over latency
constexpr int FINAL_BATCH_MIN_B = 64;
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
  pdl_trigger();   // the rollout may start its input-independent prologue (parameters, eps slab) while the PCIe read is in flight
  for (int i = threadIdx.x; i < count; i += 128) dev_in[first + i] = host_in[first + i];
}

// Per-sample accumulators of the rollout (registers); V carries one or two samples.
template <class V>
struct RolloutAcc {
  V stage = V(0);   // stage + control cost (MPPIUtils.jl:45, vehicledynamics.jl:52)
  V viol = V(0);    // sum of bound violations (slack applied once at the end)
  typename Math<V>::Count hits = Math<V>::count(Math<V>::lt(V(1), V(0)));   // obstacle hits
};

template <class V> struct Ctl { V x, y; };   // the two controls (sr, ax) / their noise

// the same model for another value type
template <class Model, class V> struct Rebind;
template <class R, class V> struct Rebind<Kin5<R>, V> { using type = Kin5<V>; };
template <class R, class V> struct Rebind<Dyn7<R>, V> { using type = Dyn7<V>; };

// The horizon loop for the LANES samples of one thread, specialised on the number of circle
// circles held in registers (NQ = their number, padded entries never hit) or reading the circle table (NQ < 0).
template <class V, class Model, int SRC, int NQ, bool TRACK>
__device__ __forceinline__ void rollout_horizon(const StaticParams<typename Math<V>::S>& p, const Buffers<typename Math<V>::S>& buf,
                                                Model& veh, RolloutAcc<V>& acc, const HotParams<typename Math<V>::S, Model::NX>* sHot,
                                                const typename Math<V>::S* sU, const typename Math<V>::S* sBlk, int b, int k, Tracker<V, TRACK>& trk,
                                                const void* sEps = nullptr, uint64_t* sBar = nullptr) {
  using M = Math<V>;
  using S = typename M::S;
  using S2 = typename Vec2<S>::type;
  using S4 = typename Vec4<S>::type;
  constexpr int LANES = M::LANES;
  const int N = p.N;
  const HotParams<S, Model::NX> hp = *sHot;   // smem -> registers (unused fields are dead)
  const bool walls = p.n_blocks > 0;
  unsigned long long* const wall_stat = buf.trace ? buf.dbg + 7 : nullptr;   // profiling mode 2: exact wall tests are counted
  const size_t estride = size_t(p.K_local);
  S2* eps = reinterpret_cast<S2*>(buf.eps) + size_t(b) * N * estride + k;
  const S4* nu_p = reinterpret_cast<const S4*>(sU);
  auto store_eps = [&](S2* dst, Ctl<V> e) __attribute__((always_inline)) {
    if (LANES == 1) *dst = S2{M::lane(e.x, 0), M::lane(e.y, 0)};
    else *reinterpret_cast<S4*>(dst) = S4{M::lane(e.x, 0), M::lane(e.y, 0), M::lane(e.x, 1), M::lane(e.y, 1)};   // two adjacent samples
  };
  auto load_eps = [&](const S2* src) __attribute__((always_inline)) {
    Ctl<V> e;
    if (LANES == 1) { const S2 t = *src; e.x = V(t.x); e.y = V(t.y); }
    else { const S4 t = *reinterpret_cast<const S4*>(src); e.x = M::make2(t.x, t.z); e.y = M::make2(t.y, t.w); }
    return e;
  };
  V cu0 = V(0), cu1 = V(0);
  if (SRC == SRC_DWA) {
    cu0 = M::make2(S(buf.cand[k]), S(buf.cand[k + LANES - 1]));
    cu1 = M::make2(S(buf.cand[estride + k]), S(buf.cand[estride + k + LANES - 1]));
  }
  // one horizon step: PushInBounds, control cost, stage cost, Heun step, tests of x_{j+1}
  auto advance = [&](S4 nu, Ctl<V> e) __attribute__((always_inline)) {
    V u0, u1;
    if (SRC == SRC_DWA) {
      u0 = cu0; u1 = cu1;
    } else {
      u0 = M::fmin(M::fmax(V(nu.x) + e.x, V(hp.CL[0])), V(hp.CU[0]));   // PushInBounds (:13-20)
      u1 = M::fmin(M::fmax(V(nu.y) + e.y, V(hp.CL[1])), V(hp.CU[1]));
      acc.stage += M::fma(V(nu.z), u0 - V(nu.x), V(nu.w) * (u1 - V(nu.y)));   // lambda*U_nom'*inv(Sigma)*(u-U_nom) (:45)
    }
    acc.stage += veh.stage_cost(hp, u0, u1);       // VehicleDynamics cost at the pre-step state
    trk.add(acc.stage, veh.s[0], veh.s[1]);        // + tracking term (nothing unless TRACK)
    veh.step(hp, u0, u1);                          // Rungekutta2 (:76-90)
    if (NQ >= 0) acc.hits += circle_hits_regs<V, (NQ > 0 ? NQ : 0)>(hp, veh.s[0], veh.s[1]);   // x_{j+1}: ObstacleEvaluation
    else acc.hits += circle_hits<V>(p, veh.s[0], veh.s[1]);
    if (walls) acc.hits += block_hits<V>(hp, hp.blk_all, sBlk, p.n_blocks, veh.s[0], veh.s[1], veh.sn, veh.cs, wall_stat, &p);
    acc.viol += bound_violation<V, Model::NX>(hp, veh.s);                 // BoundEvaluation
  };
  if (SRC == SRC_PHILOX) {
    // one Philox4x32-10 call per sample feeds two horizon steps; eps = chol(Sigma) * z
    // (MPPIUtils.jl:6) is kept for the update kernel / oracle replay
    const S L00 = hp.L00, L10 = hp.L10, L11 = hp.L11;
    const uint32_t gk = uint32_t(p.k_offset + k), iter = *buf.iter;
    // software pipeline: the noise of pair j+2 is generated while pair j integrates, so the
    // integer Philox rounds and the MUFU work of Box-Muller fill the FP32 recurrence's issue gaps
    auto draw = [&](int j, Ctl<V>& ea, Ctl<V>& eb) __attribute__((always_inline)) {
      float za0[2], za1[2], zb0[2], zb1[2];
#pragma unroll
      for (int l = 0; l < LANES; ++l) {
        const uint4 rnd = philox4x32_10(make_uint4(gk + l, uint32_t(j >> 1), iter, uint32_t(b)), hp.key);
        box_muller(rnd.x, rnd.y, &za0[l], &za1[l]);
        box_muller(rnd.z, rnd.w, &zb0[l], &zb1[l]);
      }
      const V a0 = M::make2(S(za0[0]), S(za0[LANES - 1])), a1 = M::make2(S(za1[0]), S(za1[LANES - 1]));
      const V b0 = M::make2(S(zb0[0]), S(zb0[LANES - 1])), b1 = M::make2(S(zb1[0]), S(zb1[LANES - 1]));
      ea.x = V(L00) * a0; ea.y = M::fma(V(L11), a1, V(L10) * a0);
      eb.x = V(L00) * b0; eb.y = M::fma(V(L11), b1, V(L10) * b0);
    };
    Ctl<V> na, nb;
    draw(0, na, nb);
    for (int j = 0; j < N; j += 2) {
      const Ctl<V> ea = na, eb = nb;
      if (j + 2 < N) draw(j + 2, na, nb);
      store_eps(eps, ea);
      advance(nu_p[j], ea);
      if (j + 1 < N) {
        store_eps(eps + estride, eb);
        advance(nu_p[j + 1], eb);
      }
      eps += 2 * estride;
    }
  } else if (SRC == SRC_SUPPLIED && sEps != nullptr) {
    // eps slab of this block staged in shared memory by bulk copies (rollout_kernel): row j = the block's samples at
    // horizon step j, EPS_GROUP rows per mbarrier; a thread waits once per group and reads its own samples
    const S2* row = reinterpret_cast<const S2*>(sEps) + threadIdx.x * LANES;
    for (int j = 0; j < N; ++j) {
      if ((j % EPS_GROUP) == 0) mbar_wait(sBar + j / EPS_GROUP, 0);
      advance(nu_p[j], load_eps(row));
      row += ROLLOUT_BS * LANES;
    }
  } else if (SRC == SRC_SUPPLIED) {
    // eps comes from memory (host-supplied, or pre-generated by noise_kernel): a 3-deep register queue
    // keeps the loads of steps j+1..j+3 in flight, enough to cover an HBM miss (~1.5 steps at cfg 2)
    Ctl<V> q0 = load_eps(eps), q1 = q0, q2 = q0;
    if (1 < N) q1 = load_eps(eps + estride);
    if (2 < N) q2 = load_eps(eps + 2 * estride);
    eps += 3 * estride;
    for (int j = 0; j < N; ++j) {
      const Ctl<V> e = q0;
      q0 = q1; q1 = q2;
      if (j + 3 < N) q2 = load_eps(eps);
      eps += estride;
      advance(nu_p[j], e);
    }
  } else {
    for (int j = 0; j < N; ++j) advance(S4{S(0), S(0), S(0), S(0)}, Ctl<V>{V(0), V(0)});
  }
}

// ---------------------------------------------------------------------------
// Pre-generated noise (step variant V_PREGEN, mppi_pipeline.cuh).  Philox is counter based, so the eps
// tensor of iteration i+1 depends on nothing iteration i computes: it is drawn by this kernel on a
// second graph branch while the one-block finalize of iteration i leaves 147 SMs idle, and the rollout
// of iteration i+1 reads it back (SRC_SUPPLIED instantiation) instead of spending ~60 of its ~150 issue
// slots per sample-step on Philox rounds and Box-Muller.  Same counters, same arithmetic, same bits as
// the in-rollout generator (`draw` above): one thread = one Philox call = two horizon steps of one sample.
// *iter_ptr is the iteration to draw for: the handle's counter itself when the kernel runs ahead of the
// rollout in the same stream, or the slot the rollout publishes (iter[1] = iter[0] + 1) when it runs
// beside finalize, which increments iter[0] at an unknown moment.
// ---------------------------------------------------------------------------
constexpr int NOISE_BS = 256;
constexpr int NOISE_PAIRS = 4;   // Philox calls per thread: the index / parameter prologue is a fifth of a one-call thread
template <class R>
__global__ void __launch_bounds__(NOISE_BS) noise_kernel(const __grid_constant__ StaticParams<R> p, const Buffers<R> buf,
                                                         const unsigned int* __restrict__ iter_ptr, int delta) {
  using R2 = typename Vec2<R>::type;
  const int k = blockIdx.x * NOISE_BS + threadIdx.x, b = blockIdx.z;
  const int K = p.K_local, N = p.N;
  trace_begin(buf.dbg, buf.trace, 3);
  if (k >= K) { pdl_wait(); return; }
  const uint32_t iter = __ldcg(iter_ptr) + uint32_t(delta);   // L2: the producer may still be running (programmatic dependent launch)
  const R L00 = p.L00, L10 = p.L10, L11 = p.L11;
#pragma unroll
  for (int q = 0; q < NOISE_PAIRS; ++q) {
    const int j = 2 * (blockIdx.y * NOISE_PAIRS + q);
    if (j < N) {
      const uint4 rnd = philox4x32_10(make_uint4(uint32_t(p.k_offset + k), uint32_t(j >> 1), iter, uint32_t(b)), p.seed_lo, p.seed_hi);
      float a0, a1, b0, b1;
      box_muller(rnd.x, rnd.y, &a0, &a1);
      box_muller(rnd.z, rnd.w, &b0, &b1);
      R2* eps = reinterpret_cast<R2*>(buf.eps) + (size_t(b) * N + j) * size_t(K) + k;
      eps[0] = R2{L00 * R(a0), Math<R>::fma(L11, R(a1), L10 * R(a0))};
      if (j + 1 < N) eps[K] = R2{L00 * R(b0), Math<R>::fma(L11, R(b1), L10 * R(b0))};
    }
  }
  trace_end(buf.dbg, buf.trace, 3);
  // Launched as a programmatic dependent of scan_rollout_kernel this grid runs beside that kernel's tail; it must not
  // COMPLETE before it (the next step's kernels are ordered behind this grid only).  A no-op in a normal launch.
  pdl_wait();
}

// Bulk copies (TMA) of the block's eps slab into shared memory: one copy per horizon row (the block's samples at step j,
// contiguous in the [N][K][2] tensor), EPS_GROUP rows per mbarrier, groups [g0, g1).  Called by ONE thread.  All HBM misses
// of the block are in flight together; the horizon loop starts as soon as its first group has landed.
template <class R, int LANES>
__device__ __forceinline__ void issue_eps_rows(const Buffers<R>& buf, const StaticParams<R>& p, int b, const unsigned char* sEps,
                                               uint64_t* s_bar, int g0, int g1) {
  using R2 = typename Vec2<R>::type;
  const int N = p.N;
  const int k0 = blockIdx.x * ROLLOUT_BS * LANES;
  const int n_here = min(ROLLOUT_BS * LANES, p.K_local - k0);
  const uint32_t row_bytes = uint32_t(n_here) * sizeof(R2);
  const R2* src = reinterpret_cast<const R2*>(buf.eps) + size_t(b) * N * p.K_local + k0;
  for (int g = g0; g < g1 && g * EPS_GROUP < N; ++g) {
    const int j0 = g * EPS_GROUP, j1 = min(j0 + EPS_GROUP, N);
    mbar_arrive_expect_tx(&s_bar[g], row_bytes * uint32_t(j1 - j0));
    for (int j = j0; j < j1; ++j)
      bulk_load(const_cast<unsigned char*>(sEps) + size_t(j) * ROLLOUT_BS * LANES * sizeof(R2), src + size_t(j) * p.K_local, row_bytes, &s_bar[g]);
  }
}

// V = float / double: one sample per thread.  V = f32x2: two adjacent samples per thread on the
// packed FP32 pipe (K_local must be even; chosen by the host for large sample counts).
// One-sample FP32 threads are capped at 146 registers (7 CTAs of 64 per SM = 448 threads) so that
// 65,536 samples fit ONE wave on 148 SMs; a second, nearly empty wave would cost a full CTA time.
#ifndef MPPI_PACKED_KIN_MINB
#define MPPI_PACKED_KIN_MINB 7
#endif
template <class V, class Model, int SRC, bool TRACK = false>
__global__ void __launch_bounds__(ROLLOUT_BS, sizeof(typename Math<V>::S) == 8 ? 1 : (Math<V>::LANES == 1 ? 7 : (Model::NX == 5 ? MPPI_PACKED_KIN_MINB : 5)))
rollout_kernel(const __grid_constant__ StaticParams<typename Math<V>::S> p, const Buffers<typename Math<V>::S> buf) {
  using M = Math<V>;
  using R = typename M::S;
  constexpr int LANES = M::LANES;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  R* sU = reinterpret_cast<R*>(smem_raw);  // [N][4]: u0, u1, g0, g1
  R* sBlk = sU + 4 * p.N;                  // [n_blocks][8]: wall table
  R* sRef = sBlk + 8 * p.n_blocks;         // [n_ref][2]: reference polyline (TRACK)
  __shared__ unsigned long long s_key[ROLLOUT_BS / 32];
  __shared__ int s_feas[ROLLOUT_BS / 32];
  __shared__ __align__(16) HotParams<R, Model::NX> sHot;

  const int b = blockIdx.y;
  const int N = p.N;
  const int k = (blockIdx.x * ROLLOUT_BS + threadIdx.x) * LANES;   // first sample of this thread
  if (buf.halt && *buf.halt) return;   // closed loop already at its goal (uniform across the grid)
  trace_begin(buf.dbg, buf.trace, 0);
  __shared__ __align__(8) uint64_t s_bar[EPS_MAX_GROUPS];
  const unsigned char* sEps = nullptr;
  if (SRC == SRC_SUPPLIED) {
    if ((blockIdx.x | blockIdx.y | threadIdx.x) == 0) buf.iter[1] = buf.iter[0] + 1u;   // for noise_kernel beside finalize
    if (buf.eps_stage) {
      sEps = smem_raw + ((size_t(4 * N + 8 * p.n_blocks + 2 * p.n_ref) * sizeof(R) + 127) & ~size_t(127));
      if (threadIdx.x == 0) {
        for (int g = 0; g * EPS_GROUP < N; ++g) mbar_init(&s_bar[g], 1);
        mbar_init_fence();
        // The same thread asks for the first group of rows right away (the other threads first wait on a barrier
        // after the __syncthreads below, so they see it initialised); the rest of the slab follows after that barrier.
        issue_eps_rows<R, LANES>(buf, p, b, sEps, s_bar, 0, 1);
      }
    }
  }
  if (threadIdx.x == ROLLOUT_BS - 1) sHot.fill(p);
  for (int i = threadIdx.x; i < 8 * p.n_blocks; i += ROLLOUT_BS) sBlk[i] = (&p.blk[0][0])[i];
  if (TRACK) for (int i = threadIdx.x; i < 2 * p.n_ref; i += ROLLOUT_BS) sRef[i] = buf.ref[i];
  const bool active = k < p.K_local;
  const double* Ub = buf.U + size_t(b) * 2 * N;
  pdl_wait();   // launched as a programmatic dependent of begin_kernel: x0 / U_nom are read from here on (a no-op otherwise)
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
  if (SRC == SRC_SUPPLIED && sEps != nullptr && threadIdx.x == 0) issue_eps_rows<R, LANES>(buf, p, b, sEps, s_bar, 1, EPS_MAX_GROUPS);
  Tracker<V, TRACK> trk;
  if constexpr (TRACK) {   // every sample of the instance starts its window at the point closest to x0
    __shared__ R s_td[ROLLOUT_BS / 32];
    __shared__ int s_ti[ROLLOUT_BS / 32];
    const double* x0t = buf.x0 + size_t(b) * MPPI_MAX_NX;
    const int i0 = track_start(sRef, p.n_ref, R(x0t[0]), R(x0t[1]), s_td, s_ti);
    trk.ref = sRef; trk.n = p.n_ref; trk.W = p.ref_window; trk.w = p.ref_w;
#pragma unroll
    for (int l = 0; l < LANES; ++l) trk.idx[l] = i0;
  }

  unsigned long long key = 0xFFFFFFFFFFFFFFFFull;
  int n_feas = 0;
  if (active) {
    const double* x0 = buf.x0 + size_t(b) * MPPI_MAX_NX;
    RolloutAcc<V> acc;
    Model veh;
    veh.init(p, x0);
    const V gx0 = veh.s[0] - V(p.goal[0]), gy0 = veh.s[1] - V(p.goal[1]);
    const V d0 = M::fma(gx0, gx0, gy0 * gy0);
    if constexpr (TRACK || sizeof(R) == 8) {
      rollout_horizon<V, Model, SRC, -1, TRACK>(p, buf, veh, acc, &sHot, sU, sBlk, b, k, trk, sEps, s_bar);
    } else {
      // specialised on the circle count: none, the reference scene's three (main.jl:30), up to four, up to eight
      const int nc = p.n_circles;
      if (nc == 0) rollout_horizon<V, Model, SRC, 0, false>(p, buf, veh, acc, &sHot, sU, sBlk, b, k, trk, sEps, s_bar);
      else if (nc <= 3) rollout_horizon<V, Model, SRC, 3, false>(p, buf, veh, acc, &sHot, sU, sBlk, b, k, trk, sEps, s_bar);
      else if (nc == 4) rollout_horizon<V, Model, SRC, 4, false>(p, buf, veh, acc, &sHot, sU, sBlk, b, k, trk, sEps, s_bar);
      else if (nc <= 8) rollout_horizon<V, Model, SRC, 8, false>(p, buf, veh, acc, &sHot, sU, sBlk, b, k, trk, sEps, s_bar);
      else rollout_horizon<V, Model, SRC, -1, false>(p, buf, veh, acc, &sHot, sU, sBlk, b, k, trk, sEps, s_bar);
    }
    acc.stage += veh.stage_cost(p, V(0), V(0));      // terminal slot: VehicleDynamics(x_N, [0,0]) (:49,:53)
    trk.add(acc.stage, veh.s[0], veh.s[1]);
    const V gx = veh.s[0] - V(p.goal[0]), gy = veh.s[1] - V(p.goal[1]);
    const V term = M::div(M::fma(gx, gx, gy * gy), d0) * V(p.term_scale);   // (:54)
    const V cost = M::fma(V(p.slack), acc.viol, M::fma(V(p.obs_pen), M::to_real(acc.hits), acc.stage)) + term;
#pragma unroll
    for (int l = 0; l < LANES; ++l) {
      if (k + l < p.K_local) {
        const R c = M::lane(cost, l), vl = M::lane(acc.viol, l);
        const int hl = M::lane(acc.hits, l);
        const bool feas = hl == 0 && !(vl > R(0));
        buf.costs[size_t(b) * p.K_local + k + l] = c;
        buf.flags[size_t(b) * p.K_local + k + l] =
            uint8_t((feas ? MPPI_FLAG_FEASIBLE : 0) | (hl ? MPPI_FLAG_COLLIDED : 0) | ((vl > R(0)) ? MPPI_FLAG_OOB : 0));
        const unsigned long long kk = make_key(c, (unsigned int)(k + l));
        key = kk < key ? kk : key;
        n_feas += feas ? 1 : 0;
      }
    }
  }
  // block epilogue: (min cost, first index) and feasible count -> one atomic each
  key = warp_min_u64(key);
  n_feas = warp_sum(n_feas);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { s_key[wid] = key; s_feas[wid] = n_feas; }
  __syncthreads();
  if (threadIdx.x == 0) {
    int f = 0;
#pragma unroll
    for (int w = 0; w < ROLLOUT_BS / 32; ++w) { key = s_key[w] < key ? s_key[w] : key; f += s_feas[w]; }
    atomicMin(&buf.scratch[b].rho_key, key);
    if (f) atomicAdd(&buf.scratch[b].feasible_count, f);
  }
  trace_end(buf.dbg, buf.trace, 0);
  pdl_trigger();
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
  if (buf.halt && *buf.halt) return;
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
  // Sample-sharded step (B == 1, peer exchange): the reference's loop walks the samples in GLOBAL order (MPPIUtils.jl:175,185), so
  // this rank's cut index depends on how many feasible samples the LOWER ranks hold.  One all-gather of the per-rank feasible
  // counts over the peer windows (LL lines, flag = step number, as in finalize's exchange) gives every rank its local target:
  // ranks wholly before the crossing keep all their samples, the crossing rank cuts, ranks behind it contribute nothing.
  int before = 0;
  if (buf.n_ranks > 1 && buf.peers != nullptr) {
    __shared__ int s_rank_cnt[PEER_MAX_RANKS];
    const int G = buf.n_ranks, me = buf.rank;
    const unsigned long long step = *buf.seq - buf.seq_base + 1ull;
    const int par = int(step & 1ull);
    const uint32_t flag = uint32_t(step);
    if (t < G) {
      ll_store(&buf.peer_w[t]->cnt[par][me], double(s_cnt[CUTOFF_BS - 1]), flag);
      const unsigned long long t0 = globaltimer_ns();
      double v = 0.0;
      while (!ll_try_load(&buf.peer_w[me]->cnt[par][t], flag, &v)) {
        if (globaltimer_ns() - t0 > 2000000000ull) { buf.peer_w[me]->timeout = step; break; }   // finalize's exchange reports the failure
      }
      s_rank_cnt[t] = int(v);
    }
    __syncthreads();
    for (int g = 0; g < me; ++g) before += s_rank_cnt[g];
  }
  const int target = p.feasibility_count + 1 - before;   // loop exits once FeasibilityCount > feasibility_count
  if (target <= 0) {
    if (t == 0) { s_keff = 0; s_feas = 0; }   // the cut fell on a lower rank
  } else if (excl < target && target <= incl) {
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
