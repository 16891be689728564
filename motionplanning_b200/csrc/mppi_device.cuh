// mppi_device.cuh — device-side building blocks of the B200 MPPI engine:
// math traits (FP32 fast path / FP64 verification path), the two vehicle models,
// circle and rectangle (separating-axis) obstacle tests, bounds, Philox noise.
//
// Reference semantics (paths relative to the reference root):
//   dynamic bicycle    OptimalControl/MPPI/src/vehicledynamics.jl:1-54
//   kinematic bicycle  OptimalControl/Parking/src/VehicleModel.jl:1-21
//   Heun step          OptimalControl/MPPI/src/MPPIUtils.jl:76-90
//   circles            OptimalControl/MPPI/src/MPPIUtils.jl:120-132
//   rectangles / SAT   PathPlanning/CollisionDetection/src/utils.jl:14-74,
//                      PathPlanning/HybridAstar/src/hybrid_astar_utils.jl:193-201
//   bounds             OptimalControl/MPPI/src/MPPIUtils.jl:135-151
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/mppi_b200.h"

namespace mppi {

template <class R> struct Vec2;
template <> struct Vec2<float> { using type = float2; };
template <> struct Vec2<double> { using type = double2; };
template <class R> struct Vec4;
template <> struct Vec4<float> { using type = float4; };
template <> struct Vec4<double> { using type = double4; };

// ---------------------------------------------------------------------------
// Static parameters: passed BY VALUE as a __grid_constant__ kernel argument, so
// every field is a constant-bank operand (no load instruction, warp-uniform).
// Rebuilt (and the step graph re-captured) only when the configuration changes.
// ---------------------------------------------------------------------------
template <class R>
struct StaticParams {
  int K, K_local, k_offset, N, B;
  int n_circles, n_blocks, planner;
  int n_ref, ref_window;   // reference-path tracking term (mppi_track.cuh); n_ref == 0: off
  R ref_w;
  int feasibility_count, n_x;
  uint32_t seed_lo, seed_hi;
  R dt, half_dt, lambda, inv_lambda;
  R L00, L10, L11;            // chol(Sigma), lower
  R G00, G01, G10, G11;       // lambda * inv(Sigma)  (g_j = G * U_nom[j])
  R goal[2];
  R XL[MPPI_MAX_NX], XU[MPPI_MAX_NX];
  R CL[2], CU[2];
  R slack, obs_pen, term_scale;
  // vehicle (vehicledynamics.jl:2-18), pre-folded where the folding is exact
  R la, lb, mass, inv_mass, Izz, inv_Izz, mu, KFZX;
  R FZF0, FZR0;               // KFZF*g, KFZR*g
  R tireB, tireC, tireD, tireE;  // B = tire[0]/mu, C = tire[1], D = tire[2], E = tire[3]
  R Lwb, inv_Lwb;             // la + lb
  R w[6];                     // stage cost weights v2, r2, ax2, sr2, sa2, y2
  R veh_off, veh_hl, veh_hw;  // rectangle: centre offset, half length, half width
  R circ[MPPI_MAX_CIRCLES][3];  // x, y, r^2
  R blk_all[4];                 // bounding circle of ALL walls: cx, cy, broad-phase d^2, pad
  R blk[MPPI_MAX_BLOCKS][8];    // cx, cy, cos(psi), sin(psi), half_len, half_wid, broad-phase d^2, pad
  // many walls (n_blocks > 4): x-binned candidate masks.  Bit b of wall_mask[i] is set iff wall b's broad-phase circle can be
  // reached from SOME pose whose rectangle centre has x in bin i (any y); a wall whose bit is clear is provably free for
  // every such pose by the same criterion the per-wall vote applies, so only the set bits are voted on (WALL_BINS bins over
  // [wall_x0, wall_x0 + WALL_BINS / wall_inv_w]; outside that range no wall is reachable).
  R wall_x0, wall_inv_w;
  uint32_t wall_mask[32];
};
constexpr int WALL_BINS = 32;

// per-instance scratch; initialised at create and re-armed by finalize once it has been consumed
struct Scratch {
  unsigned long long rho_key;  // atomicMin: FP32 (ordered cost bits << 32 | local k); FP64 ordered cost bits
  unsigned int argmin;         // FP64 path / cut-off path: min local k with cost == rho
  int feasible_count;          // atomicAdd
  int k_eff;                   // samples that participate (cut-off); K_local when disabled
  int pad;
};

// layout of the per-rank partial exchanged between GPUs (doubles)
enum { PART_RHO = 0, PART_ETA = 1, PART_ARGMIN = 2, PART_ROLLOUTS = 3, PART_FEASIBLE = 4, PART_HDR = 8 };
// layout of the per-instance result scalars (doubles)
enum { RES_COST = 0, RES_FEASIBLE = 1, RES_ROLLOUTS = 2, RES_FEASCOUNT = 3, RES_ARGMIN = 4, RES_RHO = 5,
       RES_ETA = 6, RES_FLAGS = 7, RES_SCALARS = 8 };

// Exchange window of one rank for the in-kernel softmin all-gather over NVLink peer memory
// (B == 1), low-latency "LL" protocol: every double travels as one 16-byte line
// {lo32, flag, hi32, flag} written with a single vector store, flag = step number, so data and
// arrival notice are the same transaction — no fence, no separate flag write: one NVLink store
// latency end to end.  Rank r stores its partial into slot [parity][r] of EVERY rank's window;
// a rank polls its own window until every line carries this step's flag.  Two parities: a rank
// can be at most one step ahead of the slowest.
constexpr int FOLD_MAX_CTAS = 640;    // scan_rollout_kernel grid limit (per-CTA partials folded by one block)
constexpr int PEER_MAX_RANKS = 8;
constexpr int PEER_PLEN = 8 + 2 * MPPI_MAX_N;   // PART_HDR + 2 N at the largest horizon
struct PeerWindow {
  uint4 line[2][PEER_MAX_RANKS][PEER_PLEN];
  uint4 cnt[2][PEER_MAX_RANKS];   // feasible-sample counts of all ranks (cross-GPU feasibility cut-off, cutoff_kernel), same protocol
  unsigned long long timeout;   // set to the step number when a wait gave up
};
__device__ __forceinline__ void ll_store(uint4* dst, double v, uint32_t flag) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "r"(uint32_t(b)), "r"(flag), "r"(uint32_t(b >> 32)), "r"(flag) : "memory");
}
__device__ __forceinline__ bool ll_try_load(const uint4* src, uint32_t flag, double* v) {
  uint4 l;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(l.x), "=r"(l.y), "=r"(l.z), "=r"(l.w) : "l"(src) : "memory");
  if (l.y != flag || l.w != flag) return false;
  *v = __longlong_as_double((long long)(((unsigned long long)l.z << 32) | l.x));
  return true;
}

template <class R>
struct Buffers {
  const double* x0;     // [B][MPPI_MAX_NX]
  const double* U;      // [B][2*N]  col-major N x m (host layout)
  R* eps;               // [B][N][K_local][2]
  R* costs;             // [B][K_local]
  uint8_t* flags;       // [B][K_local]
  Scratch* scratch;     // [B]
  unsigned int* iter;   // Philox iteration counter (device-incremented)
  R* Spart;             // [B][ksplit][N][2]  unnormalised weighted control sums
  R* eta_part;          // [B][ksplit]
  double* partial;      // [B][PART_HDR + 2N]   this rank's partial
  const double* gathered;  // [G][B][PART_HDR + 2N]  all ranks' partials (G == 1: == partial)
  double* res_scalars;  // [B][RES_SCALARS]
  double* res_control;  // [B][2*N]  col-major N x m
  double* res_traj;     // [B][(N+1)*n_x] col-major (N+1) x n_x
  const double* cand;   // DWA: [K][2] col-major K x m candidates
  const R* ref;         // [n_ref][2] reference polyline (tracking term)
  uint4* host_res;          // mapped pinned host mirror of the result block (B == 1) as LL lines {lo, flag, hi, flag}, or nullptr
  unsigned long long* seq;  // device count of completed steps
  volatile unsigned long long* host_seq;  // mapped pinned host copy of *seq: the completion flag the host spins on
  unsigned long long* dbg;  // [16] SM-clock stamps at the phase boundaries of finalize (instance 0)
  int ksplit;
  int n_ranks;
  int rank;
  PeerWindow* const* peers;   // non-null: the in-kernel peer exchange is set up (device copy of the table below; only tested against nullptr)
  PeerWindow* peer_w[PEER_MAX_RANKS];   // exchange windows of all ranks (own + peer-mapped), BY VALUE in the parameter block: a kernel reaches a
                                        // window through one indexed constant load instead of a dependent global load that is cold after an L2 flush
  unsigned long long seq_base;   // *seq when the peer exchange was set up: exchange flags count steps from there
  const int* halt;      // closed loop: nonzero once the episode is over; the step's kernels then do nothing (nullptr otherwise)
  int trace;            // 1: kernels stamp %globaltimer at first-CTA start / last-CTA end into dbg[8..]
  int eps_stage;        // rollout reading eps (SRC_SUPPLIED): 1 = the block's eps slab is bulk-copied (TMA) into shared memory
  // scan-form fused step (mppi_scan.cuh)
  R* cta_part;          // [2N + 2][n_cta] per-CTA softmin partials: 2N weighted control sums ([j][c] interleaved), eta, rho_cta
  int n_cta;            // CTAs of scan_rollout_kernel (columns of cta_part)
  int fold_off;         // PH_FOLD: byte offset in finalize's dynamic shared memory where cta_part is staged by one bulk copy; 0 = read it from global memory
  unsigned int* ticket; // arrival counter: the last CTA to finish runs the step's tail (fold, exchange, combine, nominal rollout)
  int tail_in_kernel;   // 1: that tail runs inside scan_rollout_kernel's last CTA; 0: finalize_kernel<PH_FOLD|...> follows as its own launch
};

// ---------------------------------------------------------------------------
// Hot parameters held in per-thread registers for the horizon loop.
// sm_100 ALU instructions take no constant-bank operands: every use of a kernel
// parameter inside the loop costs an LDCU (uniform load) issue slot, and ptxas
// re-issues them every iteration (~25 of 243 slots per step in profiles/r01a).
// The block therefore stages the hot parameters ONCE in shared memory and every
// thread copies them to registers before its horizon loop; a value that arrived
// through LDS cannot be rematerialised from the constant bank.
// Same field names as StaticParams so the model code is generic over both.
// ---------------------------------------------------------------------------
constexpr int HOT_CIRCLES = 8;   // circles (2 quads) that can ride in registers

template <class R, int NX>
struct HotParams {
  R dt, half_dt, Lwb, inv_Lwb;
  R XL[NX], XU[NX], CL[2], CU[2], w[6];
  R la, lb, mass, inv_mass, Izz, inv_Izz, mu, KFZX, FZF0, FZR0, tireB, tireC, tireD, tireE;
  R L00, L10, L11;
  R veh_off, veh_hl, veh_hw;
  R blk_all[4];
  R circ[HOT_CIRCLES][3];
  uint32_t key[20];   // Philox round keys
  __device__ __forceinline__ void fill(const StaticParams<R>& p) {
    dt = p.dt; half_dt = p.half_dt; Lwb = p.Lwb; inv_Lwb = p.inv_Lwb;
    for (int i = 0; i < NX; ++i) { XL[i] = p.XL[i]; XU[i] = p.XU[i]; }
    for (int i = 0; i < 2; ++i) { CL[i] = p.CL[i]; CU[i] = p.CU[i]; }
    for (int i = 0; i < 6; ++i) w[i] = p.w[i];
    la = p.la; lb = p.lb; mass = p.mass; inv_mass = p.inv_mass; Izz = p.Izz; inv_Izz = p.inv_Izz; mu = p.mu; KFZX = p.KFZX;
    FZF0 = p.FZF0; FZR0 = p.FZR0;
    tireB = p.tireB; tireC = p.tireC; tireD = p.tireD; tireE = p.tireE;
    L00 = p.L00; L10 = p.L10; L11 = p.L11;
    veh_off = p.veh_off; veh_hl = p.veh_hl; veh_hw = p.veh_hw;
    for (int i = 0; i < 4; ++i) blk_all[i] = p.blk_all[i];
    for (int o = 0; o < HOT_CIRCLES; ++o)
      for (int c = 0; c < 3; ++c) circ[o][c] = p.circ[o][c];
    uint32_t k0 = p.seed_lo, k1 = p.seed_hi;
    for (int r = 0; r < 10; ++r) { key[2 * r] = k0; key[2 * r + 1] = k1; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
  }
};

// ---------------------------------------------------------------------------
// math traits
// ---------------------------------------------------------------------------
// kernel span tracing (profiling mode 2): dbg[8 + 2*slot] = min start, dbg[9 + 2*slot] = max end, in ns
// Programmatic dependent launch (sm_90+): a kernel launched with the programmatic-stream-serialization
// attribute may start while its predecessor still runs; it orders itself with pdl_wait() (returns when
// the predecessor grid has completed and its writes are visible; a no-op without the attribute), and
// the predecessor lets it go early with pdl_trigger().  Hides the ~0.8 us launch gap between kernels.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;"); }

// ---- bulk asynchronous copies (TMA, 1-D) completing on a shared-memory mbarrier --------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t arrivals) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {   // make the initialised barriers visible to the async proxy (bulk copies complete on them)
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned; completion is counted on `bar`
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (!done);
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void trace_begin(unsigned long long* dbg, int trace, int slot) {
  if (trace && threadIdx.x == 0) atomicMin(&dbg[8 + 2 * slot], globaltimer_ns());
}
__device__ __forceinline__ void trace_end(unsigned long long* dbg, int trace, int slot) {
  if (trace && threadIdx.x == 0) atomicMax(&dbg[9 + 2 * slot], globaltimer_ns());
}

template <class R> struct Math;

template <> struct Math<float> {
  using S = float;                 // scalar type of parameters / buffers
  static constexpr int LANES = 1;  // samples carried per thread
  using Mask = bool;
  using Count = int;
  static __device__ __forceinline__ bool lt(float a, float b) { return a < b; }
  static __device__ __forceinline__ bool le(float a, float b) { return a <= b; }
  static __device__ __forceinline__ bool ge(float a, float b) { return a >= b; }
  static __device__ __forceinline__ bool gt(float a, float b) { return a > b; }
  static __device__ __forceinline__ int count(bool m) { return m ? 1 : 0; }
  static __device__ __forceinline__ bool any(bool m) { return m; }
  static __device__ __forceinline__ float lane(float v, int) { return v; }
  static __device__ __forceinline__ int lane(int c, int) { return c; }
  static __device__ __forceinline__ float to_real(int c) { return float(c); }
  static __device__ __forceinline__ float make2(float a, float) { return a; }
  // Fast path: MUFU-based intrinsics.  |abs err| of sin/cos.approx <= 2^-21.4 on
  // [-pi, pi]; well inside the 1e-4 relative cost tolerance of FP32 mode.
  static __device__ __forceinline__ void sincos(float x, float* s, float* c) { __sincosf(x, s, c); }
  static __device__ __forceinline__ float tan(float x) { return __tanf(x); }
  static __device__ __forceinline__ float div(float a, float b) { return __fdividef(a, b); }
  // division by a parameter whose reciprocal was taken on the host
  static __device__ __forceinline__ float over(float a, float, float inv_b) { return a * inv_b; }
  static __device__ __forceinline__ float exp(float x) { return __expf(x); }
  static __device__ __forceinline__ float fma(float a, float b, float c) { return fmaf(a, b, c); }
  static __device__ __forceinline__ float sin(float x) { return __sinf(x); }
  static __device__ __forceinline__ float fabs(float x) { return fabsf(x); }
  static __device__ __forceinline__ float fmax(float a, float b) { return fmaxf(a, b); }
  static __device__ __forceinline__ float fmin(float a, float b) { return fminf(a, b); }
  // atan: odd minimax polynomial on [0,1] + reciprocal reduction (no MUFU atan exists)
  static __device__ __forceinline__ float atan(float x) {
    const float ax = fabsf(x);
    const bool big = ax > 1.0f;
    const float z = big ? __fdividef(1.0f, ax) : ax;
    const float z2 = z * z;
    // minimax fit of atan(z)/z in z^2, degree 7 (Lawson-reweighted LSQ);
    // max abs error 1.5e-7 on [0,1] evaluated in FP32.
    float p = -0.004054558929055929f;
    p = fmaf(p, z2, 0.021862929686903954f);
    p = fmaf(p, z2, -0.05591228976845741f);
    p = fmaf(p, z2, 0.09642194956541061f);
    p = fmaf(p, z2, -0.1390862911939621f);
    p = fmaf(p, z2, 0.19946566224098206f);
    p = fmaf(p, z2, -0.33329859375953674f);
    p = fmaf(p, z2, 0.9999993443489075f);
    float r = p * z;
    r = big ? (1.57079632679f - r) : r;
    return copysignf(r, x);
  }
  // atan(n/d) with ONE reciprocal: reduce on max(|n|,|d|) directly instead of dividing first and
  // taking a second reciprocal when |n/d| > 1.  Sign = sign(n) xor sign(d), as for atan(n/d).
  static __device__ __forceinline__ float atan_ratio(float n, float d) {
    const float an = fabsf(n), ad = fabsf(d);
    const bool big = an > ad;
    const float z = __fdividef(fminf(an, ad), fmaxf(an, ad));
    const float z2 = z * z;
    float p = -0.004054558929055929f;
    p = fmaf(p, z2, 0.021862929686903954f);
    p = fmaf(p, z2, -0.05591228976845741f);
    p = fmaf(p, z2, 0.09642194956541061f);
    p = fmaf(p, z2, -0.1390862911939621f);
    p = fmaf(p, z2, 0.19946566224098206f);
    p = fmaf(p, z2, -0.33329859375953674f);
    p = fmaf(p, z2, 0.9999993443489075f);
    float r = p * z;
    r = big ? (1.57079632679f - r) : r;
    return __uint_as_float(__float_as_uint(r) | ((__float_as_uint(n) ^ __float_as_uint(d)) & 0x80000000u));
  }
};

template <> struct Math<double> {
  using S = double;
  static constexpr int LANES = 1;
  using Mask = bool;
  using Count = int;
  static __device__ __forceinline__ bool lt(double a, double b) { return a < b; }
  static __device__ __forceinline__ bool le(double a, double b) { return a <= b; }
  static __device__ __forceinline__ bool ge(double a, double b) { return a >= b; }
  static __device__ __forceinline__ bool gt(double a, double b) { return a > b; }
  static __device__ __forceinline__ int count(bool m) { return m ? 1 : 0; }
  static __device__ __forceinline__ bool any(bool m) { return m; }
  static __device__ __forceinline__ double lane(double v, int) { return v; }
  static __device__ __forceinline__ int lane(int c, int) { return c; }
  static __device__ __forceinline__ double to_real(int c) { return double(c); }
  static __device__ __forceinline__ double make2(double a, double) { return a; }
  static __device__ __forceinline__ double atan_ratio(double n, double d) { return ::atan(n / d); }
  static __device__ __forceinline__ void sincos(double x, double* s, double* c) { ::sincos(x, s, c); }
  static __device__ __forceinline__ double tan(double x) { return ::tan(x); }
  static __device__ __forceinline__ double div(double a, double b) { return a / b; }
  static __device__ __forceinline__ double over(double a, double b, double) { return a / b; }
  static __device__ __forceinline__ double exp(double x) { return ::exp(x); }
  static __device__ __forceinline__ double fma(double a, double b, double c) { return ::fma(a, b, c); }
  static __device__ __forceinline__ double sin(double x) { return ::sin(x); }
  static __device__ __forceinline__ double atan(double x) { return ::atan(x); }
  static __device__ __forceinline__ double fabs(double x) { return ::fabs(x); }
  static __device__ __forceinline__ double fmax(double a, double b) { return ::fmax(a, b); }
  static __device__ __forceinline__ double fmin(double a, double b) { return ::fmin(a, b); }
};

// mask combinators usable with bool and with the packed two-lane mask
__device__ __forceinline__ bool m_and(bool a, bool b) { return a && b; }
__device__ __forceinline__ bool m_or(bool a, bool b) { return a || b; }

}  // namespace mppi
