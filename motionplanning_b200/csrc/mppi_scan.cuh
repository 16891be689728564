// mppi_scan.cuh — the MPPI step of the kinematic bicycle as ONE horizon-parallel kernel.
//
// Why: at the headline size (K = 65,536, N = 50) a one-sample-per-thread rollout puts 1.7 warps on every
// scheduler and walks 50 dependent Heun steps in each: latency-bound at 0.18 of the FP32 pipe (profiles/r01_*).
// The kinematic model (OptimalControl/Parking/src/VehicleModel.jl:14-18) is triangular:
//   (sa, ux)   running sums of the clamped controls                      x + dt*(k+k)/2 with k constant
//   psi        a running sum of terms in (sa, ux) only                   d psi = ux/L tan(sa)
//   (x, y)     running sums of terms in (psi, ux) only                   d x = ux cos(psi), d y = ux sin(psi)
// so the Heun recurrence (MPPIUtils.jl:76-90) splits into three prefix sums with all transcendental work,
// cost terms (MPPIUtils.jl:45, vehicledynamics.jl:52) and obstacle / bound tests (:120-151) done for every
// horizon step independently.  K*N = 3.3 M independent work items instead of 65,536 chains.
//
// Layout: SCAN_L = 8 lanes own one sample, lane l the C consecutive steps [l*C, l*C + C); a warp carries 4
// samples.  Per phase a lane sums its chunk sequentially (C - 1 adds) and an 8-wide segmented shuffle scan
// supplies the offset of the chunk.  States differ from the sequential form by summation order only.
//
// Fused behind the rollout (same registers, no second pass over eps):
//   * softmin weighting + weighted control sum (MPPIUtils.jl:154-167, :187-190) as an ONLINE reduction: every
//     warp keeps (rho_w, eta_w, S_w[N][2]) against its running minimum and rescales by exp((rho_new - rho_old)/lambda)
//     when a tile lowers it (the flash-attention recurrence, deterministic: tile -> warp assignment is fixed);
//   * per-CTA fold of the warps' partials, one [2N + 2] column per CTA in global memory;
//   * the LAST CTA to arrive (ticket counter) runs the tail of the step — fold over CTAs, multi-GPU exchange,
//     combine, U_new, nominal rollout, result block — through finalize_block (mppi_update.cuh).
// The eps tensor is therefore read exactly once per step (update_kernel's second pass is gone) and one MPPI
// step is one kernel.  FP32, MPPI planner, circles, no cut-off, no tracking term; everything else keeps the
// sequential kernels.
#pragma once
#include "mppi_update.cuh"

namespace mppi {

constexpr int SCAN_L = 8;                 // lanes per sample
constexpr int SCAN_SPW = 32 / SCAN_L;     // samples per warp
constexpr int SCAN_BS = 256;
constexpr int SCAN_NW = SCAN_BS / 32;

// parameters the tile loop reads from shared memory (one LDS per use and phase, amortised over the chunk)
struct ScanHot {
  float CL0, CU0, CL1, CU1;
  float w2, w3, w4, w5;
  float dt, half_dt, inv_Lwb, inv_lambda;
  float XL[5], XU[5];
  float goal0, goal1, term_over_d0, slack, obs_pen;
  float x0[5];
  float kp0, c1x0, c1y0;      // state 0: ux0 tan(sa0)/L, ux0 cos(psi0), ux0 sin(psi0)
  float cost0;                // stage cost of state 0's own terms: w5 y0^2 + w4 sa0^2
  float circ[8][3];
};

__device__ __forceinline__ float seg_excl_scan(float total, int l) {   // exclusive prefix over the 8 lanes of a sample
  float inc = total;
#pragma unroll
  for (int o = 1; o < SCAN_L; o <<= 1) {
    const float n = __shfl_up_sync(0xffffffffu, inc, o, SCAN_L);
    if (l >= o) inc += n;
  }
  return inc - total;
}
__device__ __forceinline__ float seg_sum(float v) {                    // sum over the 8 lanes of a sample, in every lane
#pragma unroll
  for (int o = 1; o < SCAN_L; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o, SCAN_L);
  return v;
}
// value of lane l - 1 (lane 0 of a sample gets `first`)
__device__ __forceinline__ float seg_prev(float v, float first, int l) {
  const float n = __shfl_up_sync(0xffffffffu, v, 1, SCAN_L);
  return l == 0 ? first : n;
}

template <int C, int NC, int PHASES>
__global__ void __launch_bounds__(SCAN_BS, 3)
scan_rollout_kernel(const __grid_constant__ StaticParams<float> p, const Buffers<float> buf) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float4* sU = reinterpret_cast<float4*>(smem_raw);                       // [SCAN_L * C]: n0, n1, g0, g1 (zero beyond N)
  float* sW = reinterpret_cast<float*>(sU + SCAN_L * C);                  // [SCAN_NW][2N + 2] per-warp partials
  __shared__ ScanHot sh;
  __shared__ unsigned long long s_key[SCAN_NW];
  __shared__ int s_feas[SCAN_NW];
  __shared__ unsigned int s_ticket;

  const int N = p.N, K = p.K_local;
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
  const int l = lane & (SCAN_L - 1), sgrp = lane >> 3;
  if (buf.halt && *buf.halt) return;
  trace_begin(buf.dbg, buf.trace, 0);
  if ((blockIdx.x | t) == 0) { buf.iter[1] = buf.iter[0] + 1u; __threadfence(); }   // for noise_kernel running behind this kernel (V_PREGEN)

  // ---- block prologue: static parameters (before the dependency wait), then x0 / U_nom ----
  if (t == 0) {
    sh.CL0 = p.CL[0]; sh.CU0 = p.CU[0]; sh.CL1 = p.CL[1]; sh.CU1 = p.CU[1];
    sh.w2 = p.w[2]; sh.w3 = p.w[3]; sh.w4 = p.w[4]; sh.w5 = p.w[5];
    sh.dt = p.dt; sh.half_dt = p.half_dt; sh.inv_Lwb = p.inv_Lwb; sh.inv_lambda = p.inv_lambda;
    for (int i = 0; i < 5; ++i) { sh.XL[i] = p.XL[i]; sh.XU[i] = p.XU[i]; }
    sh.goal0 = p.goal[0]; sh.goal1 = p.goal[1]; sh.slack = p.slack; sh.obs_pen = p.obs_pen;
    for (int o = 0; o < 8; ++o) {
      const bool on = o < p.n_circles;
      sh.circ[o][0] = on ? p.circ[o][0] : 0.f; sh.circ[o][1] = on ? p.circ[o][1] : 0.f; sh.circ[o][2] = on ? p.circ[o][2] : -1.f;   // d^2 <= -1 never holds
    }
  }
  pdl_wait();   // programmatic dependent of begin_kernel: x0 / U_nom are read from here on
  for (int j = t; j < SCAN_L * C; j += SCAN_BS) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (j < N) {
      const float u0 = float(buf.U[j]), u1 = float(buf.U[N + j]);
      v = make_float4(u0, u1, p.G00 * u0 + p.G01 * u1, p.G10 * u0 + p.G11 * u1);   // g_j = lambda inv(Sigma) U_nom[j] (MPPIUtils.jl:45, hoisted)
    }
    sU[j] = v;
  }
  if (t == 32) {
    float x0[5];
    for (int i = 0; i < 5; ++i) { x0[i] = float(buf.x0[i]); sh.x0[i] = x0[i]; }
    float sn, cs;
    __sincosf(x0[2], &sn, &cs);
    sh.kp0 = x0[3] * (__tanf(x0[4]) * p.inv_Lwb);
    sh.c1x0 = x0[3] * cs; sh.c1y0 = x0[3] * sn;
    sh.cost0 = fmaf(p.w[5], x0[1] * x0[1], p.w[4] * (x0[4] * x0[4]));
    const float gx = x0[0] - p.goal[0], gy = x0[1] - p.goal[1];
    sh.term_over_d0 = __fdividef(p.term_scale, fmaf(gx, gx, gy * gy));                 // 1e4 / |x0 - goal|^2 (:54)
  }
  __syncthreads();

  // ---- per-warp online softmin state ----
  float rho_w = __int_as_float(0x7f800000);   // +inf
  float eta_l = 0.f;                          // sum of weights of this lane's sample slot (same in its 8 lanes)
  float S0[C], S1[C];
#pragma unroll
  for (int i = 0; i < C; ++i) { S0[i] = 0.f; S1[i] = 0.f; }
  unsigned long long key = ~0ull;
  int n_feas = 0;

  const int j0 = l * C;
  const int n_tiles = (K + SCAN_SPW - 1) / SCAN_SPW;
  const int warps_total = gridDim.x * SCAN_NW;
  const float2* eps = reinterpret_cast<const float2*>(buf.eps);

  for (int tile = blockIdx.x * SCAN_NW + wid; tile < n_tiles; tile += warps_total) {
    const int k = tile * SCAN_SPW + sgrp;
    const bool live = k < K;
    // ---- phase 1: PushInBounds (:13-20), control cost (:45), effort; (sa, ux) = running sums of the controls ----
    float u0[C], u1[C], A[C], B[C];
    float2 e[C];
#pragma unroll
    for (int i = 0; i < C; ++i) {
      const int j = j0 + i;
      e[i] = (live && j < N) ? __ldg(eps + size_t(j) * K + k) : make_float2(0.f, 0.f);
    }
    float stage = 0.f, viol = 0.f;
    float r0 = 0.f, r1 = 0.f;
#pragma unroll
    for (int i = 0; i < C; ++i) {
      const int j = j0 + i;
      const float4 nu = sU[j];
      float a = fminf(fmaxf(nu.x + e[i].x, sh.CL0), sh.CU0);
      float c = fminf(fmaxf(nu.y + e[i].y, sh.CL1), sh.CU1);
      if (j >= N) { a = 0.f; c = 0.f; }
      u0[i] = a; u1[i] = c;
      stage += fmaf(nu.z, a - nu.x, nu.w * (c - nu.y));
      stage += fmaf(sh.w3, a * a, sh.w2 * (c * c));
      r0 += a; r1 += c;
      A[i] = r0; B[i] = r1;           // inclusive local prefix
    }
    {
      const float o0 = seg_excl_scan(r0, l), o1 = seg_excl_scan(r1, l);
      const float sa0 = sh.x0[4], ux0 = sh.x0[3], dt = sh.dt;
#pragma unroll
      for (int i = 0; i < C; ++i) {
        A[i] = fmaf(dt, o0 + A[i], sa0);   // sa of state j + 1
        B[i] = fmaf(dt, o1 + B[i], ux0);   // ux of state j + 1
      }
    }
    // bounds of sa, ux; the sa^2 stage term of states 1..N (pre-step states 1..N-1 and the terminal slot, :49-53)
#pragma unroll
    for (int i = 0; i < C; ++i) {
      if (j0 + i < N) {
        stage = fmaf(sh.w4, A[i] * A[i], stage);
        viol += fmaxf(fmaxf(sh.XL[4] - A[i], A[i] - sh.XU[4]), 0.f);
        viol += fmaxf(fmaxf(sh.XL[3] - B[i], B[i] - sh.XU[3]), 0.f);
      }
    }
    // ---- phase 2: psi = running sum of dt/2 (kp_j + kp_{j+1}), kp = ux tan(sa) / L ----
    float P[C];
#pragma unroll
    for (int i = 0; i < C; ++i) A[i] = B[i] * (__tanf(A[i]) * sh.inv_Lwb);   // A := kp of state j + 1
    const float kp_prev = seg_prev(A[C - 1], sh.kp0, l);
    {
      float r = 0.f;
#pragma unroll
      for (int i = 0; i < C; ++i) {
        r += (i == 0 ? kp_prev : A[i - 1]) + A[i];
        P[i] = r;
      }
      const float o = seg_excl_scan(r, l);
      const float psi0 = sh.x0[2], hdt = sh.half_dt;
#pragma unroll
      for (int i = 0; i < C; ++i) P[i] = fmaf(hdt, o + P[i], psi0);   // psi of state j + 1
    }
#pragma unroll
    for (int i = 0; i < C; ++i)
      if (j0 + i < N) viol += fmaxf(fmaxf(sh.XL[2] - P[i], P[i] - sh.XU[2]), 0.f);
    // ---- phase 3: (x, y) = running sums of dt/2 (ux_j cos psi_j + ux_{j+1} cos psit_j), psit_j = psi_j + dt kp_j ----
    const float psi_prev = seg_prev(P[C - 1], sh.x0[2], l);
    float X[C], Y[C];
    float c1x_last, c1y_last;
    {
      // first pass: sin/cos of every produced state -> c1 = ux (cos, sin) psi, kept in X / Y for the moment
#pragma unroll
      for (int i = 0; i < C; ++i) {
        float sn, cs;
        __sincosf(P[i], &sn, &cs);
        X[i] = B[i] * cs; Y[i] = B[i] * sn;
      }
      c1x_last = X[C - 1]; c1y_last = Y[C - 1];
    }
    const float c1x_prev = seg_prev(c1x_last, sh.c1x0, l), c1y_prev = seg_prev(c1y_last, sh.c1y0, l);
    {
      float rx = 0.f, ry = 0.f;
      float kx = c1x_prev, ky = c1y_prev, ps = psi_prev, kq = kp_prev;
      const float dt = sh.dt;
#pragma unroll
      for (int i = 0; i < C; ++i) {
        float st, ct;
        __sincosf(fmaf(dt, kq, ps), &st, &ct);      // predictor heading psi_j + dt * k1p
        const float nkx = X[i], nky = Y[i];
        rx += fmaf(B[i], ct, kx);                    // k1x + k2x of step j
        ry += fmaf(B[i], st, ky);
        X[i] = rx; Y[i] = ry;
        kx = nkx; ky = nky; ps = P[i]; kq = A[i];
      }
      const float ox = seg_excl_scan(rx, l), oy = seg_excl_scan(ry, l);
      const float x00 = sh.x0[0], y00 = sh.x0[1], hdt = sh.half_dt;
#pragma unroll
      for (int i = 0; i < C; ++i) {
        X[i] = fmaf(hdt, ox + X[i], x00);
        Y[i] = fmaf(hdt, oy + Y[i], y00);
      }
    }
    // ---- per-state tests of x_1..x_N (x_0 never is, :40-41): circles (:120-132), bounds (:135-151), y^2 stage term, terminal (:54) ----
    int hits = 0;
#pragma unroll
    for (int i = 0; i < C; ++i) {
      if (j0 + i < N) {
        const float x = X[i], y = Y[i];
        stage = fmaf(sh.w5, y * y, stage);
        viol += fmaxf(fmaxf(sh.XL[0] - x, x - sh.XU[0]), 0.f);
        viol += fmaxf(fmaxf(sh.XL[1] - y, y - sh.XU[1]), 0.f);
#pragma unroll
        for (int o = 0; o < NC; ++o) {
          const float dx = x - sh.circ[o][0], dy = y - sh.circ[o][1];
          hits += (fmaf(dx, dx, dy * dy) <= sh.circ[o][2]) ? 1 : 0;
        }
        if (j0 + i == N - 1) {
          const float gx = x - sh.goal0, gy = y - sh.goal1;
          stage = fmaf(fmaf(gx, gx, gy * gy), sh.term_over_d0, stage);
        }
      }
    }
    // ---- per-sample reduction over the 8 lanes ----
    const unsigned hb = __ballot_sync(0xffffffffu, hits > 0), vb = __ballot_sync(0xffffffffu, viol > 0.f);
    const bool any_hit = ((hb >> (sgrp * SCAN_L)) & 0xFFu) != 0u, any_viol = ((vb >> (sgrp * SCAN_L)) & 0xFFu) != 0u;
    float cost = fmaf(sh.slack, viol, fmaf(sh.obs_pen, float(hits), stage));
    cost = seg_sum(cost) + sh.cost0;
    if (live) {
      if (l == 0) {
        const bool feas = !any_hit && !any_viol;
        buf.costs[k] = cost;
        buf.flags[k] = uint8_t((feas ? MPPI_FLAG_FEASIBLE : 0) | (any_hit ? MPPI_FLAG_COLLIDED : 0) | (any_viol ? MPPI_FLAG_OOB : 0));
        n_feas += feas ? 1 : 0;
      }
      const unsigned long long kk = make_key(cost, (unsigned int)k);
      key = kk < key ? kk : key;
    }
    // ---- online softmin: running minimum of the warp, rescale on improvement, accumulate ----
    float m = live ? cost : __int_as_float(0x7f800000);
    m = fminf(m, __shfl_xor_sync(0xffffffffu, m, 8));
    m = fminf(m, __shfl_xor_sync(0xffffffffu, m, 16));
    if (m < rho_w) {   // warp-uniform
      const float sc = __expf((m - rho_w) * sh.inv_lambda);   // rho_w = +inf the first time: exp(-inf) = 0 on zero accumulators
      eta_l *= sc;
#pragma unroll
      for (int i = 0; i < C; ++i) { S0[i] *= sc; S1[i] *= sc; }
      rho_w = m;
    }
    const float w = live ? __expf((rho_w - cost) * sh.inv_lambda) : 0.f;
    eta_l += w;
#pragma unroll
    for (int i = 0; i < C; ++i) { S0[i] = fmaf(w, u0[i], S0[i]); S1[i] = fmaf(w, u1[i], S1[i]); }
  }

  // ---- warp epilogue: fold the 4 sample slots, leave (S_w, eta_w, rho_w) in shared memory ----
#pragma unroll
  for (int i = 0; i < C; ++i) {
    S0[i] += __shfl_xor_sync(0xffffffffu, S0[i], 8);  S0[i] += __shfl_xor_sync(0xffffffffu, S0[i], 16);
    S1[i] += __shfl_xor_sync(0xffffffffu, S1[i], 8);  S1[i] += __shfl_xor_sync(0xffffffffu, S1[i], 16);
  }
  eta_l += __shfl_xor_sync(0xffffffffu, eta_l, 8);
  eta_l += __shfl_xor_sync(0xffffffffu, eta_l, 16);
  key = warp_min_u64(key);
  n_feas = warp_sum(n_feas);
  const int W = 2 * N + 2;
  float* mine = sW + wid * W;
  if (sgrp == 0) {
#pragma unroll
    for (int i = 0; i < C; ++i)
      if (j0 + i < N) { mine[2 * (j0 + i)] = S0[i]; mine[2 * (j0 + i) + 1] = S1[i]; }
  }
  if (lane == 0) { mine[2 * N] = eta_l; mine[2 * N + 1] = rho_w; s_key[wid] = key; s_feas[wid] = n_feas; }
  __syncthreads();
  // ---- CTA fold against the CTA's minimum, one column of the per-CTA partial table ----
  {
    float rho_c = sW[2 * N + 1];
#pragma unroll
    for (int w = 1; w < SCAN_NW; ++w) rho_c = fminf(rho_c, sW[w * W + 2 * N + 1]);
    const int n_cta = gridDim.x;
    for (int col = t; col <= 2 * N; col += SCAN_BS) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < SCAN_NW; ++w) {
        const float rw = sW[w * W + 2 * N + 1];
        const float sc = (rw < __int_as_float(0x7f800000)) ? __expf((rho_c - rw) * sh.inv_lambda) : 0.f;   // a warp without tiles holds +inf
        s = fmaf(sW[w * W + col], sc, s);
      }
      buf.cta_part[size_t(col) * n_cta + blockIdx.x] = s;
    }
    if (t == 0) {
      buf.cta_part[size_t(2 * N + 1) * n_cta + blockIdx.x] = rho_c;
      int f = 0;
      unsigned long long kmin = ~0ull;
#pragma unroll
      for (int w = 0; w < SCAN_NW; ++w) { kmin = s_key[w] < kmin ? s_key[w] : kmin; f += s_feas[w]; }
      atomicMin(&buf.scratch[0].rho_key, kmin);
      if (f) atomicAdd(&buf.scratch[0].feasible_count, f);
    }
  }
  trace_end(buf.dbg, buf.trace, 0);
  pdl_trigger();   // noise_kernel for the next iteration may take the SMs this block leaves
  if (!buf.tail_in_kernel) return;
  // ---- ticket: the last CTA to arrive runs the tail of the step ----
  __threadfence();
  __syncthreads();
  if (t == 0) s_ticket = atomicAdd(buf.ticket, 1u);
  __syncthreads();
  if (s_ticket != gridDim.x - 1) return;
  if (t == 0) *buf.ticket = 0u;   // re-armed for the next step
  __threadfence();
  finalize_block<float, Kin5<float>, SRC_SUPPLIED, PHASES, SCAN_BS>(p, buf, smem_raw, 0, false);
}

}  // namespace mppi
