// mppi_update.cuh — softmin weighting + weighted control update, and the
// finalize (partial -> combine -> nominal rollout) kernel.
#pragma once
#include "mppi_kernels.cuh"

namespace mppi {

// 2 consecutive samples x 2 controls of one horizon step: [k][0..1], [k+1][0..1]
template <class R> struct Pair;
template <> struct Pair<float> {
  float4 v;
  __device__ __forceinline__ void load(const float* p) { v = __ldcs(reinterpret_cast<const float4*>(p)); }
  __device__ __forceinline__ void load1(const float* p) { const float2 t = __ldcs(reinterpret_cast<const float2*>(p)); v = make_float4(t.x, t.y, 0.f, 0.f); }
};
template <> struct Pair<double> {
  double4 v;
  __device__ __forceinline__ void load(const double* p) {
    const double2 a = __ldcs(reinterpret_cast<const double2*>(p)), c = __ldcs(reinterpret_cast<const double2*>(p) + 1);
    v = make_double4(a.x, a.y, c.x, c.y);
  }
  __device__ __forceinline__ void load1(const double* p) { const double2 a = __ldcs(reinterpret_cast<const double2*>(p)); v = make_double4(a.x, a.y, 0., 0.); }
};

// ---------------------------------------------------------------------------
// update_kernel: S[j][c] = sum_k exp(-(c_k - rho)/lambda) * clamp(U_nom[j][c] + eps[j][k][c])
// (MPPIUtils.jl:154-167 weights, :187-190 weighted sum of the CLAMPED controls;
// the 1/eta normalisation is applied once in finalize).  A K x (N*m) streaming
// reduction: each eps element is read exactly once, 16 B per lane per load,
// UPDATE_JG independent loads in flight per lane; weights are recomputed from
// the (L1/L2-resident) costs, never materialised.  HBM/L2-bound; no tensor cores —
// this is a GEMV-shaped contraction with arithmetic intensity ~0.5 flop/B.
// grid = (ksplit, jsplit, B): a block owns one sample chunk and walks the groups of UPDATE_JG
// horizon rows jg = blockIdx.y, blockIdx.y + jsplit, ...; per group each warp leaves its tree sum
// in shared memory and streams on (no barrier), one barrier and one fold at the end.  With one
// instance jsplit = the number of groups (one group per block, most blocks in flight); with
// thousands of small instances jsplit = 1, so a block streams its instance's whole eps slab instead
// of paying launch + prologue + reduction per 32 KB.  Deterministic: fixed partition, fixed
// tree, partials summed in fixed order by finalize (no float atomics).
// ---------------------------------------------------------------------------
template <class R>
__global__ void __launch_bounds__(UPDATE_BS) update_kernel(const __grid_constant__ StaticParams<R> p, const Buffers<R> buf) {
  extern __shared__ __align__(16) unsigned char upd_smem[];
  constexpr int NW = UPDATE_BS / 32, COLS = 2 * UPDATE_JG;
  R* s_red = reinterpret_cast<R*>(upd_smem);   // [local groups][NW][COLS]
  __shared__ R s_eta[NW];
  const int b = blockIdx.z, N = p.N, K = p.K_local;
  if (buf.halt && *buf.halt) return;
  pdl_wait();   // everything below reads what the rollout wrote
  trace_begin(buf.dbg, buf.trace, 1);
  const Scratch sc = buf.scratch[b];
  const int keff = sc.k_eff;
  const R rho = key_cost(sc.rho_key, R(0));
  int chunk = (K + gridDim.x - 1) / gridDim.x;
  chunk = (chunk + 1) & ~1;
  const int k0 = blockIdx.x * chunk, k1 = min(k0 + chunk, keff);
  const bool vec_ok = (K & 1) == 0;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int ngroups = (N + UPDATE_JG - 1) / UPDATE_JG;
  const double* Ub = buf.U + size_t(b) * 2 * N;
  const R* cs = buf.costs + size_t(b) * K;
  const R* eb = buf.eps + size_t(b) * N * K * 2;
  R eta = R(0);

  int gl = 0;
  for (int jg = blockIdx.y; jg < ngroups; jg += gridDim.y, ++gl) {
    const int j0 = jg * UPDATE_JG;
    const bool lead = (jg == 0);   // the block that owns group 0 also sums the weights
    R n0[UPDATE_JG], n1[UPDATE_JG], acc0[UPDATE_JG], acc1[UPDATE_JG];
#pragma unroll
    for (int jj = 0; jj < UPDATE_JG; ++jj) {
      const int j = min(j0 + jj, N - 1);
      n0[jj] = R(Ub[j]); n1[jj] = R(Ub[N + j]);
      acc0[jj] = R(0); acc1[jj] = R(0);
    }
    for (int k = k0 + 2 * threadIdx.x; k < k1; k += 2 * UPDATE_BS) {
      const bool two = (k + 1 < k1);
      const R c0 = cs[k], c1 = two ? cs[k + 1] : R(0);
      const R w0 = Math<R>::exp((rho - c0) * p.inv_lambda);
      const R w1 = two ? Math<R>::exp((rho - c1) * p.inv_lambda) : R(0);
      if (lead) {
        eta += w0 + w1;
        if (sizeof(R) == 8) {   // FP64 key carries no index: first index with cost == rho
          if (c0 == rho) atomicMin(&buf.scratch[b].argmin, (unsigned int)k);
          if (two && c1 == rho) atomicMin(&buf.scratch[b].argmin, (unsigned int)(k + 1));
        }
      }
      Pair<R> e[UPDATE_JG];
#pragma unroll
      for (int jj = 0; jj < UPDATE_JG; ++jj) {
        const int j = min(j0 + jj, N - 1);
        const R* src = eb + (size_t(j) * K + k) * 2;
        if (two && vec_ok) e[jj].load(src);
        else { e[jj].load1(src); if (two) { Pair<R> t; t.load1(src + 2); e[jj].v.z = t.v.x; e[jj].v.w = t.v.y; } }
      }
#pragma unroll
      for (int jj = 0; jj < UPDATE_JG; ++jj) {
        const R ua0 = Math<R>::fmin(Math<R>::fmax(n0[jj] + e[jj].v.x, p.CL[0]), p.CU[0]);
        const R ua1 = Math<R>::fmin(Math<R>::fmax(n1[jj] + e[jj].v.y, p.CL[1]), p.CU[1]);
        const R ub0 = Math<R>::fmin(Math<R>::fmax(n0[jj] + e[jj].v.z, p.CL[0]), p.CU[0]);
        const R ub1 = Math<R>::fmin(Math<R>::fmax(n1[jj] + e[jj].v.w, p.CL[1]), p.CU[1]);
        acc0[jj] += w0 * ua0 + w1 * ub0;
        acc1[jj] += w0 * ua1 + w1 * ub1;
      }
    }
#pragma unroll
    for (int jj = 0; jj < UPDATE_JG; ++jj) {
      acc0[jj] = warp_sum(acc0[jj]);
      acc1[jj] = warp_sum(acc1[jj]);
    }
    if (lane == 0) {
      R* dst = s_red + (size_t(gl) * NW + wid) * COLS;
#pragma unroll
      for (int jj = 0; jj < UPDATE_JG; ++jj) { dst[2 * jj] = acc0[jj]; dst[2 * jj + 1] = acc1[jj]; }
    }
  }
  const bool has_lead = (blockIdx.y == 0);
  if (has_lead) {
    eta = warp_sum(eta);
    if (lane == 0) s_eta[wid] = eta;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < gl * COLS; i += UPDATE_BS) {   // fold the warps' sums, one output per thread
    const int g = i / COLS, col = i - g * COLS;
    R s = R(0);
#pragma unroll
    for (int w = 0; w < NW; ++w) s += s_red[(size_t(g) * NW + w) * COLS + col];
    const int j = (blockIdx.y + g * gridDim.y) * UPDATE_JG + (col >> 1);
    if (j < N) buf.Spart[((size_t(b) * gridDim.x + blockIdx.x) * N + j) * 2 + (col & 1)] = s;
  }
  if (has_lead && threadIdx.x == 0) {
    R s = R(0);
#pragma unroll
    for (int w = 0; w < NW; ++w) s += s_eta[w];
    buf.eta_part[size_t(b) * gridDim.x + blockIdx.x] = s;
  }
  trace_end(buf.dbg, buf.trace, 1);
  pdl_trigger();
}

// ---------------------------------------------------------------------------
// States x_0..x_N of the nominal rollout into shared memory rows [state | sin ψ | cos ψ].
// Generic form: one thread integrates the Heun recurrence (the only sequential part).
// Kinematic bicycle: the model is triangular — (sa, ux) are running sums of the controls, ψ a
// running sum of terms that depend on (sa, ux) only, (x, y) running sums of terms that depend on
// (ψ, ux) only — so the recurrence splits into three short scans (one fma per step) with the
// transcendental work done for all steps in parallel in between.  The scans are warp-shuffle prefix
// sums, so states differ from the sequential form by summation order only (a few ulp).
// `sT` is scratch [N+1][3].  Must be called by all threads of the block.
// ---------------------------------------------------------------------------
template <class Model>
struct NominalStates {
  using R = typename Model::S;
  static constexpr int NX = Model::NX, ROW = Model::NX + 2;
  static __device__ __forceinline__ void run(const StaticParams<R>& p, const double* x0, const R* sNew, R* sX, R* /*sT*/, int N, int t) {
    if (t == 0) {
      Model veh;
      veh.init(p, x0);
#pragma unroll
      for (int i = 0; i < NX; ++i) sX[i] = veh.s[i];
      sX[NX] = veh.sn; sX[NX + 1] = veh.cs;
      for (int j = 0; j < N; ++j) {
        veh.step(p, sNew[2 * j], sNew[2 * j + 1]);
        R* row = sX + (j + 1) * ROW;
#pragma unroll
        for (int i = 0; i < NX; ++i) row[i] = veh.s[i];
        row[NX] = veh.sn; row[NX + 1] = veh.cs;
      }
    }
    __syncthreads();
  }
};

template <class R>
struct NominalStates<Kin5<R>> {
  using M = Math<R>;
  static constexpr int NX = 5, ROW = 7;   // x, y, psi, ux, sa, sin psi, cos psi
  // out[(j+1) * ROW] = init + sum_{i<=j} coef * in[i * stride]: one warp, 32 steps per pass as a
  // shuffle prefix sum (5 shuffle+add levels instead of 32 dependent shared-memory round trips)
  static __device__ __forceinline__ void warp_scan(const R* in, int stride, R coef, R init, R* out, int N, int lane) {
    R carry = init;
    for (int j0 = 0; j0 < N; j0 += 32) {
      const int j = j0 + lane;
      R v = (j < N) ? coef * in[j * stride] : R(0);
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const R n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
      }
      v += carry;
      if (j < N) out[(j + 1) * ROW] = v;
      carry = __shfl_sync(0xffffffffu, v, 31);
    }
  }
  static __device__ __forceinline__ void sub_barrier() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
  // ---- N <= 64: the whole recurrence in ONE warp's registers (lane l owns steps l and l + 32) ----
  // Same arithmetic as the shared-memory form below, operation for operation (the scans are the same shuffle prefix sums with
  // the same carry between the two 32-step chunks), so the states are bit-identical; what disappears are five block-level
  // barriers and every shared-memory round trip between the phases: 3,980 -> ~1,500 SM cycles at N = 50.
  static __device__ __forceinline__ void scan2(R v[2], R init, int lane) {
    R carry = init;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      R x = v[c];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const R n = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += n;
      }
      x += carry;
      v[c] = x;
      carry = __shfl_sync(0xffffffffu, x, 31);
    }
  }
  // value of the previous horizon step: lane - 1 of the same chunk, lane 31 of chunk 0 for lane 0 of chunk 1, `first` for step 0
  static __device__ __forceinline__ void prev2(const R v[2], R first, int lane, R out[2]) {
    const R last0 = __shfl_sync(0xffffffffu, v[0], 31);
    const R p0 = __shfl_up_sync(0xffffffffu, v[0], 1), p1 = __shfl_up_sync(0xffffffffu, v[1], 1);
    out[0] = lane == 0 ? first : p0;
    out[1] = lane == 0 ? last0 : p1;
  }
  static __device__ __forceinline__ void run_warp(const StaticParams<R>& p, const double* x0, const R* sNew, R* sX, int N, int lane) {
    const R X0 = R(x0[0]), Y0 = R(x0[1]), PSI0 = R(x0[2]), UX0 = R(x0[3]), SA0 = R(x0[4]);
    R sa[2], ux[2], tn[2], ps[2], sn[2], cs[2], xs[2], ys[2];
    bool on[2];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const int j = lane + 32 * c;
      on[c] = j < N;
      sa[c] = on[c] ? p.dt * sNew[2 * j] : R(0);
      ux[c] = on[c] ? p.dt * sNew[2 * j + 1] : R(0);
    }
    scan2(sa, SA0, lane);   // sa_{j+1}
    scan2(ux, UX0, lane);   // ux_{j+1}
    const R tn0 = M::over(M::tan(SA0), p.Lwb, p.inv_Lwb);
#pragma unroll
    for (int c = 0; c < 2; ++c) tn[c] = M::over(M::tan(sa[c]), p.Lwb, p.inv_Lwb);   // tan(sa_{j+1}) / L
    R uxp[2], tnp[2];
    prev2(ux, UX0, lane, uxp);   // ux_j
    prev2(tn, tn0, lane, tnp);   // tan(sa_j) / L
#pragma unroll
    for (int c = 0; c < 2; ++c) ps[c] = on[c] ? p.half_dt * M::fma(ux[c], tn[c], uxp[c] * tnp[c]) : R(0);   // dt/2 (k1p + k2p)
    scan2(ps, PSI0, lane);  // psi_{j+1}
    R sn0, cs0;
    M::sincos(PSI0, &sn0, &cs0);
#pragma unroll
    for (int c = 0; c < 2; ++c) M::sincos(ps[c], &sn[c], &cs[c]);
    R psp[2], snp[2], csp[2];
    prev2(ps, PSI0, lane, psp);
    prev2(sn, sn0, lane, snp);
    prev2(cs, cs0, lane, csp);
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const R psit = M::fma(uxp[c] * tnp[c], p.dt, psp[c]);   // predictor heading
      R st, ct;
      M::sincos(psit, &st, &ct);
      xs[c] = on[c] ? p.half_dt * M::fma(ux[c], ct, uxp[c] * csp[c]) : R(0);   // dt/2 (k1x + k2x)
      ys[c] = on[c] ? p.half_dt * M::fma(ux[c], st, uxp[c] * snp[c]) : R(0);
    }
    scan2(xs, X0, lane);
    scan2(ys, Y0, lane);
    if (lane == 0) { sX[0] = X0; sX[1] = Y0; sX[2] = PSI0; sX[3] = UX0; sX[4] = SA0; sX[5] = sn0; sX[6] = cs0; }
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      if (on[c]) {
        R* row = sX + (lane + 32 * c + 1) * ROW;
        row[0] = xs[c]; row[1] = ys[c]; row[2] = ps[c]; row[3] = ux[c]; row[4] = sa[c]; row[5] = sn[c]; row[6] = cs[c];
      }
    }
  }
  static __device__ __forceinline__ void run(const StaticParams<R>& p, const double* x0, const R* sNew, R* sX, R* sT, int N, int t) {
    if (N <= 64) {
      if (t < 32) run_warp(p, x0, sNew, sX, N, t);
      __syncthreads();
      return;
    }
    // At most 257 rows: four warps do all of it and meet on a NAMED barrier (bar.sync 1, 128); a block-wide barrier of
    // 32 warps costs ~250 cycles and this routine needs six of them.  The other warps go straight to the closing barrier.
    constexpr int nt = 128;
    if (t < nt) {
    // scan 1: sa (warp 0) and ux (warp 1) are running sums of the controls: s + u*dt
    const int lane = t & 31;
    // (row 0 is written by a third warp: nvcc threads a lane-dependent `if` ahead of a scan into the
    // scan's own warp test, the scanning warp then runs split in two, and a split warp takes the
    // slow divergent form of every shuffle: ~3,000 cycles per scan instead of ~700)
    if (t < 32) warp_scan(sNew, 2, p.dt, R(x0[4]), sX + 4, N, lane);
    else if (t < 64) warp_scan(sNew + 1, 2, p.dt, R(x0[3]), sX + 3, N, lane);
    else if (t < 64 + NX) sX[t - 64] = R(x0[t - 64]);
    sub_barrier();
    for (int j = t; j <= N; j += nt) sT[3 * j] = M::over(M::tan(sX[j * ROW + 4]), p.Lwb, p.inv_Lwb);   // tan(sa_j)/L
    sub_barrier();
    for (int j = t; j < N; j += nt)   // k1p + k2p of step j
      sT[3 * j + 1] = M::fma(sX[(j + 1) * ROW + 3], sT[3 * (j + 1)], sX[j * ROW + 3] * sT[3 * j]);
    sub_barrier();
    // scan 2: psi_{j+1} = psi_j + dt/2 (k1p + k2p)
    if (t < 32) warp_scan(sT + 1, 3, p.half_dt, sX[2], sX + 2, N, lane);
    sub_barrier();
    for (int j = t; j <= N; j += nt) {
      R* row = sX + j * ROW;
      M::sincos(row[2], &row[5], &row[6]);                         // carried sin/cos of psi_j
      if (j < N) {
        const R psit = M::fma(row[3] * sT[3 * j], p.dt, row[2]);   // predictor heading
        R st, ct;
        M::sincos(psit, &st, &ct);
        const R uxt = sX[(j + 1) * ROW + 3];
        sT[3 * j + 1] = M::fma(uxt, ct, row[3] * row[6]);          // k1x + k2x
        sT[3 * j + 2] = M::fma(uxt, st, row[3] * row[5]);          // k1y + k2y
      }
    }
    sub_barrier();
    // scan 3: x (warp 0) and y (warp 1)
    if (t < 32) warp_scan(sT + 1, 3, p.half_dt, sX[0], sX + 0, N, lane);
    else if (t < 64) warp_scan(sT + 2, 3, p.half_dt, sX[1], sX + 1, N, lane);
    }
    __syncthreads();
  }
};

// ---------------------------------------------------------------------------
// finalize_kernel.  One block per instance: BS = FINAL_BS threads for a single instance (latency:
// 32 warps cover the partial tree in one pass), FINAL_BS_BATCH when there are more instances than
// SMs (throughput: many small blocks per SM instead of one 1024-thread block at a time).
//   PH_PARTIAL: sum the update partials in fixed order -> this rank's
//               (rho, eta, argmin, counts, S) against its LOCAL rho.
//   PH_COMBINE: fold all ranks' partials in rank order with the rescale
//               e^{-(rho_g - rho)/lambda} (one-shot softmin combine), U_new = S/eta,
//               then the final nominal rollout (MPPIUtils.jl:192-198) on thread 0.
// DWA (DWAUtils.jl:152-158): U_new = the arg-min candidate, replicated over the horizon.
// ---------------------------------------------------------------------------
enum { PH_PARTIAL = 1, PH_COMBINE = 2, PH_XCHG = 4,
       PH_FOLD = 8 };   // PH_FOLD: this rank's partial comes from the per-CTA partials of scan_rollout_kernel (mppi_scan.cuh), not from update_kernel's
// phase clocks, written out with the results; stamped by a lane of the LAST warp: an `if (t == 0)` ahead
// of a warp-collective section gets threaded into that section's warp test and splits warp 0 (see NominalStates)
#define MPPI_STAMP(i) do { if (t == BS - 32) s_clk[i] = clock64(); } while (0)
__device__ __forceinline__ Scratch fresh_scratch(int K_local) {
  Scratch fresh;
  fresh.rho_key = ~0ull; fresh.argmin = ~0u; fresh.feasible_count = 0; fresh.k_eff = K_local; fresh.pad = 0;
  return fresh;
}
// The block-level body: called by finalize_kernel (one block per instance) and by the LAST block of
// scan_rollout_kernel to finish (mppi_scan.cuh), which saves the launch and the inter-kernel gap.
// `wait_prev`: issue griddepcontrol.wait before the first read of the previous kernel's output.
template <class R, class Model, int SRC, int PHASES, int BS>
__device__ __forceinline__ void finalize_block(const StaticParams<R>& p, const Buffers<R>& buf, unsigned char* smem_raw, const int b, const bool wait_prev) {
  // Everything this block produces is staged in shared memory and leaves in one burst at the end:
  // a global store followed by a barrier, or a read-back of a value the block just stored, costs
  // an L2 round trip (~1,000 cycles, more across the die boundary), and the old form chained six.
  const int N = p.N, t = threadIdx.x;
  constexpr int NX = Model::NX, ROW = Model::NX + 2;   // state, sin(heading), cos(heading)
  const int plen = PART_HDR + 2 * N;
  const int n_res = RES_SCALARS + 2 * N + (N + 1) * NX;
  double* sRS = reinterpret_cast<double*>(smem_raw);   // result block, laid out like the host block:
  double* sRC = sRS + RES_SCALARS;                     //   scalars | U_new [2][N] | trajectory (N+1) x n_x col-major
  double* sTR = sRC + 2 * N;
  double* sOwn = sRS + n_res;                          // [plen] this rank's partial
  double* sGat = sOwn + plen;                          // [G][plen] all ranks' partials (PH_XCHG only)
  const bool xchg = (PHASES & PH_XCHG) && buf.n_ranks > 1 && buf.peers != nullptr;   // compiled in, decided per launch
  R* sU = reinterpret_cast<R*>(sGat + ((PHASES & PH_XCHG) ? buf.n_ranks * plen : 0));   // [N][4]: U_nom0, U_nom1, g0, g1
  R* sNew = sU + 4 * N;                                // [N][2]: U_new
  __shared__ double s_scale[8];
  __shared__ double s_hdr[4];
  __shared__ double s_x0[MPPI_MAX_NX];
  __shared__ unsigned long long s_clk[8];
  double* part = buf.partial + size_t(b) * plen;
  if (buf.halt && *buf.halt) {   // closed loop already at its goal: keep the step count in line with the host, nothing else
    if (t == 0 && b == 0 && (PHASES & PH_COMBINE)) { const unsigned long long done = *buf.seq + 1ull; *buf.seq = done; *buf.host_seq = done; }
    return;
  }
  MPPI_STAMP(0);
  trace_begin(buf.dbg, buf.trace, 2);
  // loads that do not depend on this step's kernels go first, so their latency hides behind the partial tree
  unsigned long long seq0 = 0;
  unsigned int iter0 = 0;
  if (t == 0) { seq0 = *buf.seq; iter0 = *buf.iter; }
  const unsigned long long seq_all = *buf.seq;   // every thread stamps its result lines with the step number
  if (PHASES & PH_COMBINE) {
    if (t < MPPI_MAX_NX) s_x0[t] = buf.x0[size_t(b) * MPPI_MAX_NX + t];
    // old nominal + g for the nominal rollout's control-cost term (evaluated against the OLD
    // U_nom: MPPIUtils.jl:192 runs before main.jl:62 replaces it)
    const double* Ub = buf.U + size_t(b) * 2 * N;
    for (int j = t; j < N; j += BS) {
      const R u0 = (SRC == SRC_DWA) ? R(0) : R(Ub[j]), u1 = (SRC == SRC_DWA) ? R(0) : R(Ub[N + j]);
      sU[4 * j + 0] = u0; sU[4 * j + 1] = u1;
      sU[4 * j + 2] = p.G00 * u0 + p.G01 * u1;
      sU[4 * j + 3] = p.G10 * u0 + p.G11 * u1;
    }
  }
  if (wait_prev) pdl_wait();   // the loads above do not depend on the update kernel; everything below does
  // PH_FOLD: the per-block partial table and its rho row are asked for BEFORE the scratch is read, so that the three fetches share
  // one L2 round trip.  Staged form (the table fits beside this block's own shared memory: buf.fold_off != 0): ONE bulk copy brings
  // the whole table into shared memory.  (Lanes across CTAs with a load loop per column took 8.7 us at 256 CTAs: eight dependent
  // round trips per column, four columns per warp.)
  __shared__ __align__(8) uint64_t s_fbar;
  R fold_rc0 = R(0);
  if (PHASES & PH_FOLD) {
    const uint32_t bytes = (uint32_t(2 * N + 2) * uint32_t(buf.n_cta) * uint32_t(sizeof(R)) + 15u) & ~15u;   // the buffer is FOLD_MAX_CTAS wide: the round-up stays inside
    if (buf.fold_off != 0 && t == 0) {
      mbar_init(&s_fbar, 1);
      mbar_init_fence();
      mbar_arrive_expect_tx(&s_fbar, bytes);
      bulk_load(smem_raw + buf.fold_off, buf.cta_part, bytes, &s_fbar);
    }
    if (t < buf.n_cta) fold_rc0 = buf.cta_part[size_t(2 * N + 1) * buf.n_cta + t];
  }
  const Scratch sc = buf.scratch[b];
  constexpr bool PART_TO_GLOBAL = !(PHASES & PH_COMBINE);   // only a partial-only launch has a consumer outside this block
  double* pdst = PART_TO_GLOBAL ? part : sOwn;

  if (PHASES & PH_FOLD) {
    // Per-CTA partials of scan_rollout_kernel, laid out [2N + 2][n_cta]: 2N weighted control sums, eta, rho_cta, each against
    // the CTA's own minimum.  Fold them against the global minimum (already in the scratch key): a warp per column,
    // lanes across CTAs, fixed order -> deterministic.
    __shared__ double s_cscale[FOLD_MAX_CTAS];
    const int n_cta = buf.n_cta;
    const R* Cp = buf.cta_part;
    const double rho = double(key_cost(sc.rho_key, R(0)));
    const int lane = t & 31, wid = t >> 5;
    constexpr int NW = BS / 32;
    const bool staged = buf.fold_off != 0;
    const R* tab = staged ? reinterpret_cast<const R*>(smem_raw + buf.fold_off) : Cp;
    for (int c = t; c < n_cta; c += BS) {
      const double rc = double(c == t ? fold_rc0 : Cp[size_t(2 * N + 1) * n_cta + c]);
      s_cscale[c] = (rc < 1e300 && rc == rc) ? exp(-(rc - rho) / double(p.lambda)) : 0.0;   // a CTA that saw no sample carries rho = +inf
    }
    __syncthreads();
    if (staged) mbar_wait(&s_fbar, 0);
    for (int col = wid; col <= 2 * N; col += NW) {
      const R* src = tab + size_t(col) * n_cta;
      double s0 = 0.0;
      for (int c = lane; c < n_cta; c += 32) s0 += double(src[c]) * s_cscale[c];
      s0 = warp_sum(s0);
      if (lane == 0) {
        if (col < 2 * N) pdst[PART_HDR + ((col & 1) * N + (col >> 1))] = s0;   // CTA partials are [j][c] interleaved; the partial is col-major N x m
        else pdst[PART_ETA] = s0;
      }
    }
    if (t == 0) {
      pdst[PART_RHO] = rho;
      pdst[PART_ARGMIN] = double((unsigned int)(sc.rho_key & 0xFFFFFFFFull)) + double(p.k_offset);
      pdst[PART_ROLLOUTS] = double(sc.k_eff);
      pdst[PART_FEASIBLE] = double(sc.feasible_count);
    }
    __syncthreads();
  }
  if (PHASES & PH_PARTIAL) {
    if (SRC == SRC_DWA) {
      // arg-min candidate (DWAUtils.jl:152).  FP32 keys carry the index; FP64 keys do not:
      // first index whose cost equals rho.
      __shared__ unsigned int s_arg;
      const R rho = key_cost(sc.rho_key, R(0));
      if (t == 0) s_arg = (sizeof(R) == 4) ? (unsigned int)(sc.rho_key & 0xFFFFFFFFull) : 0xFFFFFFFFu;
      __syncthreads();
      if (sizeof(R) == 8) {
        const R* cs = buf.costs + size_t(b) * p.K_local;
        for (int k = t; k < p.K_local; k += BS)
          if (cs[k] == rho) atomicMin(&s_arg, (unsigned int)k);
        __syncthreads();
      }
      if (t == 0) {
        pdst[PART_RHO] = double(rho);
        pdst[PART_ETA] = 1.0;
        pdst[PART_ARGMIN] = double(s_arg) + double(p.k_offset);
        pdst[PART_ROLLOUTS] = double(sc.k_eff);
        pdst[PART_FEASIBLE] = double(sc.feasible_count);
      }
    } else {
      // Fixed-order tree over the ksplit partials: one warp per output column (2N control
      // sums + eta), lanes across ksplit, all loads of a warp's columns independent.
      {
        const int lane = t & 31, wid = t >> 5, ks_n = buf.ksplit;
        const R* Sp = buf.Spart + size_t(b) * ks_n * N * 2;
        const R* Ep = buf.eta_part + size_t(b) * ks_n;
        constexpr int NW = BS / 32, CB = 4;    // CB columns per warp in flight (independent loads + trees)
        for (int col0 = wid; col0 <= 2 * N; col0 += NW * CB) {   // col = c*N + j; col == 2N: eta
          double s[CB];
#pragma unroll
          for (int u = 0; u < CB; ++u) {
            const int col = col0 + u * NW;
            // ksplit <= 64: two predicated loads per lane, no loop, so the CB columns' loads overlap
            const bool eta_col = (col == 2 * N);
            const int c = col >= N ? 1 : 0, j = col - c * N;
            R v0 = R(0), v1 = R(0);
            if (col <= 2 * N) {
              if (lane < ks_n) v0 = eta_col ? Ep[lane] : Sp[(size_t(lane) * N + j) * 2 + c];
              if (lane + 32 < ks_n) v1 = eta_col ? Ep[lane + 32] : Sp[(size_t(lane + 32) * N + j) * 2 + c];
            }
            s[u] = double(v0) + double(v1);
          }
#pragma unroll
          for (int u = 0; u < CB; ++u) s[u] = warp_sum(s[u]);
          if (lane == 0) {
#pragma unroll
            for (int u = 0; u < CB; ++u) {
              const int col = col0 + u * NW;
              if (col < 2 * N) pdst[PART_HDR + col] = s[u];
              else if (col == 2 * N) pdst[PART_ETA] = s[u];
            }
          }
        }
      }
      if (t == 0) {
        pdst[PART_RHO] = double(key_cost(sc.rho_key, R(0)));
        // FP32 key carries the index unless the cut-off kernel wrote argmin explicitly
        const unsigned int arg = (sizeof(R) == 4 && sc.argmin == 0xFFFFFFFFu) ? (unsigned int)(sc.rho_key & 0xFFFFFFFFull) : sc.argmin;
        pdst[PART_ARGMIN] = double(arg) + double(p.k_offset);
        pdst[PART_ROLLOUTS] = double(sc.k_eff);
        pdst[PART_FEASIBLE] = double(sc.feasible_count);
      }
    }
    __syncthreads();   // block-scope visibility of the partial for the combine phase
  }
  MPPI_STAMP(1);
  if (!(PHASES & PH_COMBINE)) {
    if (t == 0) buf.scratch[b] = fresh_scratch(p.K_local);   // last consumer of the scratch: re-arm it for the next step
    return;
  }

  const int G = buf.n_ranks;
  // gathered partials [G][B][plen]: an earlier kernel's output (COMBINE-only launch), or shared memory
  const double* gat = (PHASES & (PH_PARTIAL | PH_FOLD)) ? sOwn - size_t(b) * plen : buf.gathered;
  size_t gstride = size_t(p.B) * plen;
  bool xchg_timeout = false;
  if (xchg) {
    // In-kernel all-gather over NVLink peer memory (B == 1), LL protocol (see PeerWindow).
    const int me = buf.rank;
    // exchange steps are counted from mppi_peer_init (seq_base), so ranks whose handles ran different numbers of
    // steps before the exchange was set up still agree on the flag; from then on they step in lockstep
    const unsigned long long step = seq_all - buf.seq_base + 1ull;
    const int par = int(step & 1ull);
    const uint32_t flag = uint32_t(step);
    for (int i = t; i < plen; i += BS) {
      const double v = sOwn[i];
      for (int g = 0; g < G; ++g) ll_store(&buf.peer_w[g]->line[par][me][i], v, flag);
    }
    __shared__ int s_timeout;
    if (t == 0) s_timeout = 0;
    __syncthreads();
    const unsigned long long t0 = globaltimer_ns();
    for (int i = t; i < G * plen; i += BS) {
      const int g = i / plen, e = i - g * plen;
      const uint4* src = &buf.peer_w[me]->line[par][g][e];
      double v = 0.0;
      while (!ll_try_load(src, flag, &v)) {
        if (globaltimer_ns() - t0 > 2000000000ull) { s_timeout = 1; buf.peer_w[me]->timeout = step; break; }
      }
      sGat[i] = v;
    }
    __syncthreads();
    xchg_timeout = s_timeout != 0;
    gat = sGat;
    gstride = plen;
  }
  if (t == 0) {
    // global rho = isless-min over ranks (NaN last); ties -> lowest rank == lowest global index
    int best = 0;
    double rho = gat[b * plen + PART_RHO];
    for (int g = 1; g < G; ++g) {
      const double r = gat[g * gstride + b * plen + PART_RHO];
      if (r < rho || (r == r && rho != rho)) { rho = r; best = g; }
    }
    double eta = 0.0, rollouts = 0.0, feas = 0.0;
    for (int g = 0; g < G; ++g) {
      const double* pg = gat + g * gstride + b * plen;
      // a rank whose samples all lie behind the feasibility cut-off holds no sample: rho = NaN, eta = 0, S = 0
      const double sc_g = (G == 1) ? 1.0 : ((pg[PART_RHO] == pg[PART_RHO]) ? exp(-(pg[PART_RHO] - rho) / double(p.lambda)) : 0.0);
      s_scale[g] = sc_g;
      eta += pg[PART_ETA] * sc_g;
      rollouts += pg[PART_ROLLOUTS];
      feas += pg[PART_FEASIBLE];
    }
    s_hdr[0] = eta;
    sRS[RES_ROLLOUTS] = xchg_timeout ? -1.0 : rollouts + 1.0;   // RolloutCount starts at 1 (MPPIUtils.jl:171,197); -1: exchange timed out
    sRS[RES_FEASCOUNT] = feas;
    sRS[RES_ARGMIN] = gat[best * gstride + b * plen + PART_ARGMIN] + 1.0;   // 1-based like Julia
    sRS[RES_RHO] = rho;
    sRS[RES_ETA] = eta;
  }
  __syncthreads();
  MPPI_STAMP(2);
  if (SRC == SRC_DWA) {
    const size_t arg = size_t(sRS[RES_ARGMIN] - 1.0) - p.k_offset;
    for (int i = t; i < 2 * N; i += BS) {
      const int c = i / N, j = i - c * N;
      const double u = buf.cand[size_t(c) * p.K_local + arg];
      sRC[i] = u;
      sNew[2 * j + c] = R(u);
    }
  } else {
    const double eta = s_hdr[0];
    for (int i = t; i < 2 * N; i += BS) {
      const int c = i / N, j = i - c * N;
      double s = 0.0;
      for (int g = 0; g < G; ++g) s += gat[g * gstride + b * plen + PART_HDR + i] * s_scale[g];
      const double u = s / eta;
      sRC[i] = u;
      sNew[2 * j + c] = R(u);
    }
  }
  __syncthreads();
  MPPI_STAMP(3);
  // Final nominal rollout, TrajectoryRollout(MPPI, MPPICtrl) (MPPIUtils.jl:192).  Only the
  // state recurrence is sequential (NominalStates); then the N+1 states are costed / tested /
  // exported in parallel and summed in horizon order.
  R* sX = sNew + 2 * N;                 // [N+1][ROW]
  R* sC = sX + (N + 1) * ROW;           // [N+1][3]: stage+control cost, bound violation, hits
  R* sTrk = sC + 3 * (N + 1);           // [N+1]: tracking term (only with a reference path)
  R* sRef = sTrk + (N + 1);             // [n_ref][2]
  const bool track = p.n_ref > 0;
  if (track) for (int i = t; i < 2 * p.n_ref; i += BS) sRef[i] = buf.ref[i];   // barrier: inside run()
  NominalStates<Model>::run(p, s_x0, sNew, sX, sC, N, t);
  if (track) {   // the window index is sequential along the horizon: one thread walks it (N+1 scans of W+1 points)
    __shared__ R s_td[32];
    __shared__ int s_ti[32];
    const int i0 = track_start(sRef, p.n_ref, R(s_x0[0]), R(s_x0[1]), s_td, s_ti);
    if (t == 0) {
      int idx = i0;
      for (int j = 0; j <= N; ++j) sTrk[j] = p.ref_w * track_scan(sRef, p.n_ref, p.ref_window, sX[j * ROW], sX[j * ROW + 1], idx);
    }
    __syncthreads();
  }
  MPPI_STAMP(4);
  for (int j = t; j <= N; j += BS) {
    const R* row = sX + j * ROW;
    Model veh;
#pragma unroll
    for (int i = 0; i < NX; ++i) { veh.s[i] = row[i]; sTR[size_t(i) * (N + 1) + j] = double(row[i]); }
    R c = R(0), v = R(0);
    int hits = 0;
    if (j < N) {
      const R u0 = sNew[2 * j], u1 = sNew[2 * j + 1];
      if (SRC != SRC_DWA) c = sU[4 * j + 2] * (u0 - sU[4 * j]) + sU[4 * j + 3] * (u1 - sU[4 * j + 1]);
      c += veh.stage_cost(p, u0, u1);
    } else {
      c = veh.stage_cost(p, R(0), R(0));   // terminal slot (:49,:53)
    }
    if (track) c += sTrk[j];
    if (j > 0) {   // x_0 is never tested (:40-41)
      hits = circle_hits(p, row[0], row[1]) + block_hits(p, row[0], row[1], row[NX], row[NX + 1]);
      v = bound_violation<R, NX>(p, row);
    }
    sC[3 * j] = c; sC[3 * j + 1] = v; sC[3 * j + 2] = R(hits);
  }
  __syncthreads();
  MPPI_STAMP(5);
  if (t < 32) {   // fixed-shape tree over the horizon: lanes stride over j, then a shuffle tree
    R stage = R(0), viol = R(0), fh = R(0);
    for (int j = t; j <= N; j += 32) { stage += sC[3 * j]; viol += sC[3 * j + 1]; fh += sC[3 * j + 2]; }
    stage = warp_sum(stage); viol = warp_sum(viol); fh = warp_sum(fh);
    if (t == 0) {
      const R gx0 = sX[0] - p.goal[0], gy0 = sX[1] - p.goal[1];
      const R d0 = gx0 * gx0 + gy0 * gy0;
      const int hits = int(fh);
      const R gx = sX[N * ROW] - p.goal[0], gy = sX[N * ROW + 1] - p.goal[1];
      const R term = Math<R>::div(gx * gx + gy * gy, d0) * p.term_scale;
      const R cost = stage + p.slack * viol + p.obs_pen * R(hits) + term;
      const bool feas = (hits == 0 && !(viol > R(0)));
      sRS[RES_COST] = double(cost);
      sRS[RES_FEASIBLE] = feas ? 1.0 : 0.0;
      sRS[RES_FLAGS] = double((feas ? MPPI_FLAG_FEASIBLE : 0) | (hits ? MPPI_FLAG_COLLIDED : 0) | ((viol > R(0)) ? MPPI_FLAG_OOB : 0));
    }
  }
  __syncthreads();
  MPPI_STAMP(6);
  // one burst out: device result arrays, and for B == 1 the caller-visible mapped host block
  // (posted PCIe writes, no D2H memcpy node; the host spins on host_seq instead of a stream synchronise)
  double* rs = buf.res_scalars + size_t(b) * RES_SCALARS;
  double* rc = buf.res_control + size_t(b) * 2 * N;
  double* tr = buf.res_traj + size_t(b) * (N + 1) * NX;   // col-major (N+1) x n_x
  const bool to_host = p.B == 1 && buf.host_res != nullptr;
  const uint32_t flag = uint32_t(seq_all + 1ull);   // host side: wait_step (mppi_pipeline.cuh)
  for (int i = t; i < n_res; i += BS) {
    const double v = sRS[i];
    if (to_host) ll_store(buf.host_res + i, v, flag);
    if (i < RES_SCALARS) rs[i] = v;
    else if (i < RES_SCALARS + 2 * N) rc[i - RES_SCALARS] = v;
    else tr[i - RES_SCALARS - 2 * N] = v;
  }
  if (t == 0) {
    if (PHASES & (PH_PARTIAL | PH_FOLD)) buf.scratch[b] = fresh_scratch(p.K_local);   // last consumer of the scratch: re-arm it for the next step
    if (b == 0) {
      *buf.iter = iter0 + 1u;   // next step draws fresh Philox noise
      *buf.seq = seq0 + 1ull;
      *buf.host_seq = seq0 + 1ull;
    }
  }
  if (t == BS - 32 && b == 0) {
#pragma unroll
    for (int i = 0; i < 7; ++i) buf.dbg[i] = s_clk[i];
  }
  trace_end(buf.dbg, buf.trace, 2);
}

template <class R, class Model, int SRC, int PHASES, int BS>
__global__ void __launch_bounds__(BS) finalize_kernel(const __grid_constant__ StaticParams<R> p, const Buffers<R> buf) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  finalize_block<R, Model, SRC, PHASES, BS>(p, buf, smem_raw, blockIdx.x, true);
}

}  // namespace mppi
