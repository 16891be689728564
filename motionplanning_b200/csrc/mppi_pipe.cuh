// mppi_pipe.cuh — the K-sample rollout of the kinematic bicycle as a two-stage warp pipeline.
//
// Why: at the headline size (K = 65,536, N = 50) the one-thread-per-sample rollout leaves 1.7 warps on every scheduler
// and each of them walks one dependent instruction stream: issue-active 47 %, top stall `wait` (profiles/r02_ncu_cfg2_*).
// The samples alone do not hold enough parallelism for 148 SMs; the horizon does, but a scan-form rollout pays for it with
// 2.4x the instructions (mppi_scan.cuh: 210 issue slots per sample-step against 87).  The model
// (OptimalControl/Parking/src/VehicleModel.jl:14-18 under the Heun step of MPPIUtils.jl:76-90) is triangular:
//   stage A   (sa, ux) are running sums of the clamped controls; tan(sa)/L, the control cost (MPPIUtils.jl:45), the effort
//             and sa^2 stage terms (vehicledynamics.jl:52), the bounds of sa and ux (:135-151) need nothing else;
//   stage B   psi, x, y integrate quantities stage A produced; sincos, the y^2 term, the circle tests (:120-132), the
//             bounds of x, y, psi and the terminal term (:54) need stage A's (ux, tan(sa)/L) of the step and nothing else.
// So each sample pair is walked by TWO threads in two warps: a producer running stage A over the whole horizon and a
// consumer running stage B a few steps behind, the values in between handed over through shared memory.  Twice the warps,
// each with half the dependent chain, no recomputation and no padding: the instruction count stays that of the sequential
// kernel (+ one 16-byte store / load per step).
//
// Hand-over without extra shared memory: the block's eps slab (TMA-staged, one row of 256 samples per horizon step) is
// consumed by stage A exactly once, 16 bytes per pair and step, and stage A's output for stage B is 16 bytes per pair and
// step, so the producer overwrites the eps it has just read IN PLACE.  A producer warp feeds exactly one consumer warp (same
// lanes, same addresses), so the only synchronisation is "producer warp w has finished hand-over unit g of PIPE_HAND steps": one
// mbarrier per warp pair and unit (32 arrivals, release) that the consumer waits on (acquire).
// Measured at cfg 2 on B200 (L2 flushed): rollout span 17.9 -> 15.3 us; per-CTA timeline: 2.5 us until the first eps rows have
// landed, then 50 steps in 8-12 us with the schedulers at ~60 % issue (two co-resident CTAs halve each other's speed: the SM
// is then throughput-, no longer latency-bound).  Hand-over every 4 steps and a loop body of 2 steps were the best of the
// grid {1, 2, 4, 8} x {1, 2, 4}.
//
// CTA = 8 warps = 4 producer / consumer pairs over 256 samples.  Warp i and warp i + 4 form a pair and sit on the same SM
// sub-partition (warp id mod 4), so every scheduler holds both stages and the two instruction streams interleave.
// FP32 packed (two samples per thread, mppi_pack.cuh), circles only, eps read from memory (host-supplied or pre-generated).
// Differences from rollout_kernel in the result: the running cost is the sum of two partial sums (stage A's and stage B's
// terms) instead of one, i.e. a different summation order — costs agree to FP32 rounding, flags are identical.
#pragma once
#include "mppi_kernels.cuh"

namespace mppi {

constexpr int PIPE_BS = 256;
constexpr int PIPE_PAIRS_W = 4;                       // producer / consumer warp pairs per CTA
constexpr int PIPE_SAMPLES = PIPE_PAIRS_W * 64;       // samples per CTA (two per thread)
constexpr int PIPE_SLAB_MAX = 112 * 1024;             // N * PIPE_SAMPLES * 8 bytes: two CTAs per SM up to N = 56
#ifndef MPPI_PIPE_HAND
#define MPPI_PIPE_HAND 4                              // horizon steps per producer -> consumer hand-over (divides EPS_GROUP)
#endif
#ifndef MPPI_PIPE_SPIN
#define MPPI_PIPE_SPIN 0   // bit 0: spin (test_wait) on the TMA barriers, bit 1: on the hand-over barriers; measured equal to try_wait (15.6 vs 15.3 us), so the parking wait stays
#endif
#ifndef MPPI_PIPE_UNROLL
#define MPPI_PIPE_UNROLL 2                            // steps per loop body: lets the scheduler overlap the MUFU chains of neighbouring steps
#endif
constexpr int PIPE_HAND = MPPI_PIPE_HAND;
constexpr int PIPE_UNROLL = MPPI_PIPE_UNROLL;
constexpr int PIPE_MAX_HANDS = EPS_GROUP * EPS_MAX_GROUPS / PIPE_HAND;

struct PipeHot {
  float CL0, CU0, CL1, CU1, w2, w3, w4, w5;
  float dt, half_dt, inv_Lwb, pad0;
  float XL[5], XU[5];
  float goal0, goal1, slack, obs_pen, term_scale, pad1;
  float circ[HOT_CIRCLES][3];
};

// float index of PipeHot field i (i < 28) inside StaticParams<float>; -1 = padding
__device__ __forceinline__ int pipe_hot_offset(int i) {
  using P = StaticParams<float>;
  constexpr int CL = offsetof(P, CL) >> 2, CU = offsetof(P, CU) >> 2, W = offsetof(P, w) >> 2, XL = offsetof(P, XL) >> 2, XU = offsetof(P, XU) >> 2;
  constexpr int tab[28] = {CL, CU, CL + 1, CU + 1, W + 2, W + 3, W + 4, W + 5,
                           int(offsetof(P, dt) >> 2), int(offsetof(P, half_dt) >> 2), int(offsetof(P, inv_Lwb) >> 2), -1,
                           XL, XL + 1, XL + 2, XL + 3, XL + 4, XU, XU + 1, XU + 2, XU + 3, XU + 4,
                           int(offsetof(P, goal) >> 2), int(offsetof(P, goal) >> 2) + 1, int(offsetof(P, slack) >> 2), int(offsetof(P, obs_pen) >> 2),
                           int(offsetof(P, term_scale) >> 2), -1};
  int r = -1;
#pragma unroll
  for (int k = 0; k < 28; ++k) r = (i == k) ? tab[k] : r;   // select chain over immediates: no memory, no branches
  return r;
}
static_assert(sizeof(PipeHot) == (28 + 3 * HOT_CIRCLES) * 4, "PipeHot is filled field by field through a float view");
// Spin on the barrier's phase without suspending: mbarrier.try_wait parks the thread for an implementation-defined time, which is
// the right thing for long waits and the wrong one for a hand-over that is expected within a few hundred cycles.
__device__ __forceinline__ void mbar_spin(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// FOLD tail: this block's share of the weighted control update, all eight warps.  Called at the END of the producer branch and at
// the end of the consumer branch (two call sites, each followed by the thread's exit, instead of one behind the re-joined
// branches: with a common continuation the compiler wraps both stage loops in convergence barriers and the 50-step walk gets
// 2 us slower).  The block barrier is the same barrier 0 from both sites.
__device__ __forceinline__ void pipe_fold_tail(const StaticParams<float>& p, const Buffers<float>& buf, const PipeHot& sh, const float4* sU, float4* slab,
                                               uint64_t* s_bar, const float2* s_w, const float* s_eta, const unsigned long long* s_key,
                                               const float2* src, uint32_t row_bytes) {
  const int N = p.N, K = p.K_local, t = threadIdx.x, lane = t & 31, wid = t >> 5;
  // Every consumer warp is past its last read of the slab (barrier 1, all 256 threads, at the call sites): fetch the eps again, rows
  // in the same groups on the same barriers (second phase).  ALL eight warps issue, a group of 8 rows each: a bulk copy issued by a
  // lane costs ~60 cycles of its warp, and 50 of them from one warp put 1.7 us in front of the sums.
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  for (int g = wid; g * EPS_GROUP < N; g += PIPE_BS / 32) {
    const int j0 = g * EPS_GROUP, j1 = min(j0 + EPS_GROUP, N);
    if (lane == 0) mbar_arrive_expect_tx(&s_bar[g], row_bytes * uint32_t(j1 - j0));
    __syncwarp();
    const int j = j0 + lane;
    if (j < j1) bulk_load(slab + size_t(j) * (PIPE_SAMPLES / 2), src + size_t(j) * K, row_bytes, &s_bar[g]);
  }
  __syncthreads();                                        // weights and eta are in shared memory
  trace_begin(buf.dbg, buf.trace, 1);
  {
    const int n_cta = gridDim.x, cta = blockIdx.x;
    float* part = buf.cta_part + cta;
    const float CL0 = sh.CL0, CU0 = sh.CU0, CL1 = sh.CL1, CU1 = sh.CU1;
    float2 wv[PIPE_PAIRS_W];
#pragma unroll
    for (int i = 0; i < PIPE_PAIRS_W; ++i) wv[i] = s_w[lane + 32 * i];
    constexpr int NWARP = PIPE_BS / 32;
    static_assert(NWARP == EPS_GROUP, "warp w sums rows w, w + 8, ...: one row of every copy group");
    for (int j = wid; j < N; j += NWARP) {
      mbar_wait(&s_bar[j / EPS_GROUP], 1);
      const float4 nu = sU[j];
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int i = 0; i < PIPE_PAIRS_W; ++i) {
        const float4 e = slab[size_t(j) * (PIPE_SAMPLES / 2) + lane + 32 * i];
        // the CLAMPED controls are averaged (:187-190 after PushInBounds :13-20); a lane past the last sample carries weight 0 and
        // whatever the hand-over left in its slot, which max / min turn into a finite number
        const float ua0 = fminf(fmaxf(nu.x + e.x, CL0), CU0), ua1 = fminf(fmaxf(nu.y + e.y, CL1), CU1);
        const float ub0 = fminf(fmaxf(nu.x + e.z, CL0), CU0), ub1 = fminf(fmaxf(nu.y + e.w, CL1), CU1);
        a0 = fmaf(wv[i].y, ub0, fmaf(wv[i].x, ua0, a0));
        a1 = fmaf(wv[i].y, ub1, fmaf(wv[i].x, ua1, a1));
      }
      a0 = warp_sum(a0);
      a1 = warp_sum(a1);
      if (lane == 0) { part[size_t(2 * j) * n_cta] = a0; part[size_t(2 * j + 1) * n_cta] = a1; }
    }
    if (t == 0) {
      float eta = 0.f;
      unsigned long long kc = s_key[0];
#pragma unroll
      for (int i = 0; i < PIPE_PAIRS_W; ++i) { eta += s_eta[i]; kc = s_key[i] < kc ? s_key[i] : kc; }
      part[size_t(2 * N) * n_cta] = eta;
      part[size_t(2 * N + 1) * n_cta] = key_cost(kc, 0.f);
    }
  }
  trace_end(buf.dbg, buf.trace, 1);
}

// FOLD: the block also does its share of the softmin update (MPPIUtils.jl:154-167,186-190) before it leaves, so the step needs no
// update_kernel: weights against the BLOCK's own minimum, the eps slab fetched a second time into the same shared memory (the
// hand-over has overwritten it; the lines are L2 hits, the block read them a few microseconds ago), every warp sums a few horizon
// rows over the block's 256 samples, and the block leaves one column of Buffers::cta_part ([2N + 2][n_cta]: weighted control sums,
// eta, rho of the block) that finalize folds against the global minimum (PH_FOLD) — the softmin combine of the multi-GPU step,
// applied once more inside the GPU.  No cut-off (the cut needs all costs before any weight).
template <int NQ, bool FOLD>
__global__ void __launch_bounds__(PIPE_BS, 2)
rollout_pipe_kernel(const __grid_constant__ StaticParams<float> p, const Buffers<float> buf) {
  using V = f32x2;
  using M = Math<V>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float4* sU = reinterpret_cast<float4*>(smem_raw);                                        // [N]: u0, u1, g0, g1
  float4* slab = reinterpret_cast<float4*>(smem_raw + ((size_t(p.N) * 16 + 127) & ~size_t(127)));   // [N][PIPE_SAMPLES / 2]: one float4 per pair and step
  __shared__ __align__(8) uint64_t s_bar[EPS_MAX_GROUPS];                     // TMA: rows of group g have landed
  __shared__ __align__(8) uint64_t s_prog[PIPE_PAIRS_W][PIPE_MAX_HANDS];      // producer warp w has finished hand-over unit g (PIPE_HAND steps)
  __shared__ __align__(16) PipeHot sh;
  __shared__ float4 s_handA[PIPE_BS / 2];                                     // producer's partial sums: stage (2 samples), violation (2 samples)
  __shared__ unsigned long long s_key[PIPE_PAIRS_W];
  __shared__ int s_feas[PIPE_PAIRS_W];
  __shared__ float s_x0[8];
  __shared__ float2 s_w[FOLD ? PIPE_BS / 2 : 1];                              // FOLD: softmin weights of the block's samples, per pair
  __shared__ float s_eta[PIPE_PAIRS_W];

  const int N = p.N, K = p.K_local, b = blockIdx.y;
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
  if (buf.halt && *buf.halt) return;
  trace_begin(buf.dbg, buf.trace, 0);
  if ((blockIdx.x | blockIdx.y | t) == 0) buf.iter[1] = buf.iter[0] + 1u;   // for noise_kernel beside finalize (V_PREGEN)
  const int k0 = blockIdx.x * PIPE_SAMPLES;
  const int n_here = min(PIPE_SAMPLES, K - k0);
  const int n_groups = (N + EPS_GROUP - 1) / EPS_GROUP;
  const uint32_t row_bytes = uint32_t(n_here) * 8u;
  const float2* src = reinterpret_cast<const float2*>(buf.eps) + size_t(b) * N * K + k0;
  const int n_hands = (N + PIPE_HAND - 1) / PIPE_HAND;
  // prologue spread over the block: barrier initialisation one per thread, parameters one field per thread; the bulk copies are
  // issued by a CONSUMER lane (thread PIPE_BS / 2), whose warp has nothing to do until its producer has finished the first steps
  // eps does not depend on begin_kernel: the whole slab is asked for before anything else, by the first CONSUMER warp (idle
  // until its producer has finished the first steps).  Group by group, so that group 0 lands first; within a group lane r issues
  // row r, i.e. one instruction per group instead of one per row (50 copies issued one after another by one thread took 3 us).
  if (wid == PIPE_PAIRS_W) {
    if (lane < n_groups) mbar_init(&s_bar[lane], 1);
    mbar_init_fence();
    __syncwarp();
    for (int g = 0; g < n_groups; ++g) {
      const int j0 = g * EPS_GROUP, j1 = min(j0 + EPS_GROUP, N);
      if (lane == 0) mbar_arrive_expect_tx(&s_bar[g], row_bytes * uint32_t(j1 - j0));
      const int j = j0 + lane;
      if (j < j1) bulk_load(slab + size_t(j) * (PIPE_SAMPLES / 2), src + size_t(j) * K, row_bytes, &s_bar[g]);
    }
  }
  for (int i = t; i < PIPE_PAIRS_W * n_hands; i += PIPE_BS) mbar_init(&s_prog[i / n_hands][i % n_hands], 32);
  // hot parameters: thread 64 + i copies field i of PipeHot from the kernel parameter block through an offset table (ONE indexed
  // constant load per thread; a chain of selects cost 1.8 us of cold constant-cache misses in front of the block barrier)
  {
    constexpr int NHOT = sizeof(PipeHot) / 4;
    if (t >= 64 && t < 64 + NHOT) {
      const int i = t - 64;
      const float* pf = reinterpret_cast<const float*>(&p);
      float v;
      if (i < 28) {
        const int off = pipe_hot_offset(i);
        v = off >= 0 ? pf[off] : 0.f;
      } else {
        const int o = (i - 28) / 3, c = (i - 28) % 3;
        v = o < p.n_circles ? pf[(offsetof(StaticParams<float>, circ) >> 2) + 3 * o + c] : (c == 2 ? -1.f : 0.f);   // d^2 <= -1 never holds
      }
      reinterpret_cast<float*>(&sh)[i] = v;
    }
  }
  pdl_wait();   // programmatic dependent of begin_kernel: x0 / U_nom are read from here on
  // x0 and U_nom are cold in HBM after an L2 flush (~1 us a miss): x0 is fetched by another warp than U_nom so that the two misses
  // overlap instead of following each other in one warp (they used to put the block barrier at 3.3 us)
  const double* Ub = buf.U + size_t(b) * 2 * N;
  if (t >= PIPE_BS - 32 && t < PIPE_BS - 32 + 5) s_x0[t - (PIPE_BS - 32)] = float(buf.x0[size_t(b) * MPPI_MAX_NX + t - (PIPE_BS - 32)]);
  for (int j = t; j < N; j += PIPE_BS) {
    const double d0 = Ub[j], d1 = Ub[N + j];
    const float u0 = float(d0), u1 = float(d1);
    sU[j] = make_float4(u0, u1, p.G00 * u0 + p.G01 * u1, p.G10 * u0 + p.G11 * u1);   // g_j = lambda inv(Sigma) U_nom[j] (MPPIUtils.jl:45, hoisted)
  }
  __syncthreads();

  const int w = wid & (PIPE_PAIRS_W - 1);                 // the pair this warp belongs to
  const int pair = w * 32 + lane;                         // sample pair within the CTA
  float4* row = slab + pair;
  constexpr int RS = PIPE_SAMPLES / 2;                    // float4 per horizon row

  if (wid < PIPE_PAIRS_W) {
    // ---------------- stage A: controls, (sa, ux), tan(sa)/L, their cost terms and bounds ----------------
    const float CL0 = sh.CL0, CU0 = sh.CU0, CL1 = sh.CL1, CU1 = sh.CU1, w2 = sh.w2, w3 = sh.w3, w4 = sh.w4;
    const float dt = sh.dt, inv_Lwb = sh.inv_Lwb, XL3 = sh.XL[3], XU3 = sh.XU[3], XL4 = sh.XL[4], XU4 = sh.XU[4];
    V sa = V(s_x0[4]), ux = V(s_x0[3]);
    V stage = V(0.f), viol = V(0.f);
#pragma unroll PIPE_UNROLL
    for (int j = 0; j < N; ++j) {
      if ((j % EPS_GROUP) == 0) { if (MPPI_PIPE_SPIN & 1) mbar_spin(&s_bar[j / EPS_GROUP], 0); else mbar_wait(&s_bar[j / EPS_GROUP], 0); }
      const float4 e = row[size_t(j) * RS];               // eps of the pair: (e0, e1) of sample a, (e0, e1) of sample b
      const float4 nu = sU[j];
      const V u0 = M::fmin(M::fmax(V(nu.x) + V(e.x, e.z), V(CL0)), V(CU0));   // PushInBounds (MPPIUtils.jl:13-20)
      const V u1 = M::fmin(M::fmax(V(nu.y) + V(e.y, e.w), V(CL1)), V(CU1));
      stage += M::fma(V(nu.z), u0 - V(nu.x), V(nu.w) * (u1 - V(nu.y)));       // lambda U_nom' inv(Sigma) (u - U_nom) (:45)
      stage += M::fma(V(w4), sa * sa, M::fma(V(w3), u0 * u0, V(w2) * (u1 * u1)));   // pre-step state's sa^2, effort (vehicledynamics.jl:52)
      sa = M::fma(u0, V(dt), sa);                         // Heun with a constant derivative: x + dt (k + k) / 2
      ux = M::fma(u1, V(dt), ux);
      const V tn = M::tan(sa) * V(inv_Lwb);
      viol += M::fmax(M::fmax(V(XL3) - ux, ux - V(XU3)), V(0.f));             // BoundEvaluation of the new state (:135-151)
      viol += M::fmax(M::fmax(V(XL4) - sa, sa - V(XU4)), V(0.f));
      row[size_t(j) * RS] = make_float4(ux.v.x, ux.v.y, tn.v.x, tn.v.y);      // in place of the eps just consumed
      if ((j % PIPE_HAND) == PIPE_HAND - 1 && j != N - 1) mbar_arrive(&s_prog[w][j / PIPE_HAND]);
    }
    stage += V(w4) * (sa * sa);                           // terminal slot: VehicleDynamics(x_N, [0, 0]) (:49,:53), sa part
    s_handA[pair] = make_float4(stage.v.x, stage.v.y, viol.v.x, viol.v.y);
    mbar_arrive(&s_prog[w][(N - 1) / PIPE_HAND]);         // last unit + the partial sums
    pdl_trigger();
    if (FOLD) {
      asm volatile("bar.sync 1, %0;" ::"n"(PIPE_BS) : "memory");   // the consumers have finished with the slab
      pipe_fold_tail(p, buf, sh, sU, slab, s_bar, s_w, s_eta, s_key, src, row_bytes);
    }
    return;
  }

  // ---------------- stage B: psi, x, y, their cost terms, circle tests, bounds, terminal term ----------------
  const float dt = sh.dt, half_dt = sh.half_dt, w5 = sh.w5, inv_Lwb = sh.inv_Lwb;
  const float XL0 = sh.XL[0], XU0 = sh.XU[0], XL1 = sh.XL[1], XU1 = sh.XU[1], XL2 = sh.XL[2], XU2 = sh.XU[2];
  float cx[NQ > 0 ? NQ : 1], cy[NQ > 0 ? NQ : 1], cr[NQ > 0 ? NQ : 1];
#pragma unroll
  for (int o = 0; o < NQ; ++o) { cx[o] = sh.circ[o][0]; cy[o] = sh.circ[o][1]; cr[o] = sh.circ[o][2]; }
  V x = V(s_x0[0]), y = V(s_x0[1]), psi = V(s_x0[2]);
  V ux = V(s_x0[3]), tn = V(__tanf(s_x0[4]) * inv_Lwb);
  V sn, cs;
  M::sincos(psi, &sn, &cs);
  const V gx0 = x - V(sh.goal0), gy0 = y - V(sh.goal1);
  const V d0 = M::fma(gx0, gx0, gy0 * gy0);
  V stage = V(0.f), viol = V(0.f);
  Count2 hits{0, 0};
#pragma unroll PIPE_UNROLL
  for (int j = 0; j < N; ++j) {
    if ((j % PIPE_HAND) == 0) { if (MPPI_PIPE_SPIN & 2) mbar_spin(&s_prog[w][j / PIPE_HAND], 0); else mbar_wait(&s_prog[w][j / PIPE_HAND], 0); }
    const float4 q = row[size_t(j) * RS];
    const V uxn = V(q.x, q.y), tnn = V(q.z, q.w);
    stage += V(w5) * (y * y);                             // pre-step state's y^2
    const V k1p = ux * tn;
    const V psit = M::fma(k1p, V(dt), psi);               // predictor heading
    V st, ct;
    M::sincos(psit, &st, &ct);
    x = M::fma(V(half_dt), M::fma(uxn, ct, ux * cs), x);  // corrector x + dt (k1 + k2) / 2 (MPPIUtils.jl:76-90)
    y = M::fma(V(half_dt), M::fma(uxn, st, ux * sn), y);
    psi = M::fma(V(half_dt), M::fma(uxn, tnn, k1p), psi);
    ux = uxn; tn = tnn;
    M::sincos(psi, &sn, &cs);
#pragma unroll
    for (int o = 0; o < NQ; ++o) {                        // ObstacleEvaluation of x_{j+1} (:120-132)
      const V dx = x - V(cx[o]), dy = y - V(cy[o]);
      hits += M::count(M::le(M::fma(dx, dx, dy * dy), V(cr[o])));
    }
    viol += M::fmax(M::fmax(V(XL0) - x, x - V(XU0)), V(0.f));
    viol += M::fmax(M::fmax(V(XL1) - y, y - V(XU1)), V(0.f));
    viol += M::fmax(M::fmax(V(XL2) - psi, psi - V(XU2)), V(0.f));
  }
  stage += V(w5) * (y * y);                               // terminal slot, y part
  const float4 a = s_handA[pair];                         // ordered behind the producer's last arrival (waited on in the last group)
  stage += V(a.x, a.y);
  viol += V(a.z, a.w);
  const V gx = x - V(sh.goal0), gy = y - V(sh.goal1);
  const V term = M::div(M::fma(gx, gx, gy * gy), d0) * V(sh.term_scale);   // (:54)
  const V cost = M::fma(V(sh.slack), viol, M::fma(V(sh.obs_pen), M::to_real(hits), stage)) + term;
  unsigned long long key = ~0ull;
  int n_feas = 0;
  const int k = k0 + 2 * pair;
#pragma unroll
  for (int l = 0; l < 2; ++l) {
    if (k + l < K) {
      const float c = M::lane(cost, l), vl = M::lane(viol, l);
      const int hl = M::lane(hits, l);
      const bool feas = hl == 0 && !(vl > 0.f);
      buf.costs[size_t(b) * K + k + l] = c;
      buf.flags[size_t(b) * K + k + l] = uint8_t((feas ? MPPI_FLAG_FEASIBLE : 0) | (hl ? MPPI_FLAG_COLLIDED : 0) | ((vl > 0.f) ? MPPI_FLAG_OOB : 0));
      const unsigned long long kk = make_key(c, (unsigned int)(k + l));
      key = kk < key ? kk : key;
      n_feas += feas ? 1 : 0;
    }
  }
  key = warp_min_u64(key);
  n_feas = warp_sum(n_feas);
  if (lane == 0) { s_key[w] = key; s_feas[w] = n_feas; }
  if (FOLD) asm volatile("bar.sync 1, %0;" ::"n"(PIPE_BS) : "memory");   // with the producers, which wait here to help with the tail
  else asm volatile("bar.sync 1, 128;" ::: "memory");     // the four consumer warps only: the producers have left
  if (FOLD) {
    unsigned long long kc = s_key[0];
#pragma unroll
    for (int i = 1; i < PIPE_PAIRS_W; ++i) kc = s_key[i] < kc ? s_key[i] : kc;
    const float rho_c = key_cost(kc, 0.f);
    const float w0 = (k < K) ? Math<float>::exp((rho_c - M::lane(cost, 0)) * p.inv_lambda) : 0.f;       // CalculateMPPIWeights (:154-167)
    const float w1 = (k + 1 < K) ? Math<float>::exp((rho_c - M::lane(cost, 1)) * p.inv_lambda) : 0.f;
    s_w[pair] = make_float2(w0, w1);
    const float es = warp_sum(w0 + w1);
    if (lane == 0) s_eta[w] = es;
  }
  if (t == PIPE_BS / 2) {
    int f = 0;
#pragma unroll
    for (int i = 0; i < PIPE_PAIRS_W; ++i) { key = s_key[i] < key ? s_key[i] : key; f += s_feas[i]; }
    atomicMin(&buf.scratch[b].rho_key, key);
    if (f) atomicAdd(&buf.scratch[b].feasible_count, f);
    if (buf.trace) atomicMax(&buf.dbg[9], globaltimer_ns());
  }
  pdl_trigger();
  if (FOLD) pipe_fold_tail(p, buf, sh, sU, slab, s_bar, s_w, s_eta, s_key, src, row_bytes);
}

}  // namespace mppi
