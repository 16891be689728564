// mppi_api.cu — extern "C" entry points declared in include/mppi_b200.h.
// Each function cites the reference interface it replaces in the header.
#include <cstdlib>

#include "mppi_extra.cuh"
#include "mppi_pipeline.cuh"

using namespace mppi;

#ifndef MPPI_PACK_MIN_SAMPLES
#define MPPI_PACK_MIN_SAMPLES 262144
#endif
#ifndef MPPI_PACK_MIN_SAMPLES_PREGEN
#define MPPI_PACK_MIN_SAMPLES_PREGEN 65536
#endif

#define H_CHECK(h)            \
  if (!(h)) return MPPI_EINVAL; \
  cudaSetDevice((h)->cfg.device)

// single-process multi-GPU: repeat a configuration call on the other ranks' handles, then make rank 0's device current again
#define FORWARD_SUBS(h, expr)                                   \
  do {                                                          \
    for (mppi_handle* s_ : (h)->subs) {                         \
      const int rc_ = (expr);                                   \
      if (rc_) { (h)->err = s_->err; cudaSetDevice((h)->cfg.device); return rc_; } \
    }                                                           \
    if (!(h)->subs.empty()) cudaSetDevice((h)->cfg.device);     \
  } while (0)
#define NOT_MULTI(h, what) \
  if (!(h)->subs.empty() || (h)->spmg_member) return fail(h, MPPI_EINVAL, what " is not available on a multi-GPU handle (n_gpus > 1)")

static int fail(mppi_handle* h, int code, const std::string& msg) {
  if (h) h->err = msg; else create_error() = msg;
  return code;
}

extern "C" {

int32_t mppi_version(void) { return MPPI_B200_VERSION; }

int32_t mppi_config_default(mppi_config_t* c) {
  if (!c) return MPPI_EINVAL;
  memset(c, 0, sizeof(*c));
  c->struct_size = sizeof(mppi_config_t);
  c->model = MPPI_MODEL_DYNAMIC7; c->n_x = 7; c->m = 2;        // main.jl:12-13
  c->K = 1500; c->N = 20; c->B = 1;                            // main.jl:22,14
  c->precision = MPPI_FP32; c->noise_mode = MPPI_NOISE_PHILOX; c->planner = MPPI_PLANNER_MPPI;
  c->feasibility_count = 1300;                                 // types.jl:24
  c->use_cuda_graph = 1; c->device = 0; c->k_offset = 0; c->K_local = 0;
  c->seed = 0x5EED0001ull;
  c->T = 3.0; c->lambda = 25.0;                                // main.jl:15,24
  c->Sigma[0] = 0.05; c->Sigma[1] = 0.0; c->Sigma[2] = 0.0; c->Sigma[3] = 0.1;   // main.jl:27
  const double pi = 3.141592653589793;
  const double X0[7] = {0, 0, 0, 0, 0, 5, 0};                                   // main.jl:7
  const double XL[7] = {-10, -20, -2, -pi / 2, -pi / 2, 1, -pi / 9};           // main.jl:8
  const double XU[7] = {130, 20, 2, pi / 2, pi / 2, 10, pi / 9};               // main.jl:9
  for (int i = 0; i < 7; ++i) { c->X0[i] = X0[i]; c->XL[i] = XL[i]; c->XU[i] = XU[i]; }
  c->goal[0] = 110; c->goal[1] = 0;                                             // main.jl:23
  c->CL[0] = -0.5; c->CL[1] = -2.5; c->CU[0] = 0.5; c->CU[1] = 2.5;             // main.jl:10-11
  c->slack_penalty = 1e5;                                                       // types.jl:30
  c->obstacle_penalty = 100.0 * 712.5;                                          // MPPIUtils.jl:127
  c->terminal_scale = 10000.0;                                                  // MPPIUtils.jl:54
  c->la = 1.56; c->lb = 1.64; c->mass = 2020; c->Izz = 4095; c->g = 9.81; c->mu = 0.8;   // vehicledynamics.jl:2-7
  c->KFZF = 1018.28 / 2; c->KFZR = 963.34 / 2; c->KFZX = 186.22;                          // :8-10
  const double tire[6] = {-10.4, 1.3, 1.0, 0.1556, -0.0023, 41.3705};                     // :11
  for (int i = 0; i < 6; ++i) c->tire[i] = tire[i];
  const double w[6] = {1, 1, 5, 3, 2, 10};                                                // :52
  for (int i = 0; i < 6; ++i) c->stage_w[i] = w[i];
  c->veh_L = 3; c->veh_W = 2; c->veh_center_offset = 1.5;   // main_hybrid_astar.jl:21, hybrid_astar_utils.jl:193
  return MPPI_OK;
}

const char* mppi_last_error(const mppi_handle_t* h) { return h ? h->err.c_str() : create_error().c_str(); }

int64_t mppi_launch_count(const mppi_handle_t* h) {
  if (!h) return 0;
  int64_t n = h->launches;
  for (const mppi_handle* s : h->subs) n += s->launches;
  return n;
}

// MvNormal(zeros(m), Sigma) (setup.jl:57,77) throws unless Sigma is symmetric positive definite; the Cholesky factor
// build_params takes must exist: s00 > 0, s11 - s10^2 / s00 > 0, s01 == s10.
static bool sigma_ok(const double* S) {
  const double s00 = S[0], s10 = S[1], s01 = S[2], s11 = S[3];
  if (!(s00 > 0) || !(s11 > 0)) return false;
  if (!(std::fabs(s10 - s01) <= 1e-12 * std::max(1.0, std::max(std::fabs(s10), std::fabs(s01))))) return false;
  return s11 - s10 * s10 / s00 > 0;
}

static int validate(const mppi_config_t* c, std::string* why) {
  char buf[256];
  // same checks, same wording as setup.jl:19-39 where they apply to a fixed-size POD
  if (c->struct_size != sizeof(mppi_config_t)) { snprintf(buf, sizeof buf, "struct_size (%u) must be %zu", c->struct_size, sizeof(mppi_config_t)); *why = buf; return MPPI_EINVAL; }
  if (c->m <= 0) { snprintf(buf, sizeof buf, "Controls (%d) must be > 0", c->m); *why = buf; return MPPI_EINVAL; }
  if (c->n_x <= 0) { snprintf(buf, sizeof buf, "States (%d) must be > 0", c->n_x); *why = buf; return MPPI_EINVAL; }
  if (c->m != 2) { snprintf(buf, sizeof buf, "Controls (%d) must be 2 (sr, ax) for the bicycle models", c->m); *why = buf; return MPPI_EINVAL; }
  const int want = c->model == MPPI_MODEL_DYNAMIC7 ? 7 : 5;
  if (c->model != MPPI_MODEL_DYNAMIC7 && c->model != MPPI_MODEL_KINEMATIC5) { *why = "unknown model"; return MPPI_EINVAL; }
  if (c->n_x != want) { snprintf(buf, sizeof buf, "States (%d) must match the model (%d)", c->n_x, want); *why = buf; return MPPI_EINVAL; }
  if (c->K <= 0) { snprintf(buf, sizeof buf, "SamplingNumber (%d) must be > 0", c->K); *why = buf; return MPPI_EINVAL; }
  if (c->N <= 0 || c->N > MPPI_MAX_N) { snprintf(buf, sizeof buf, "N (%d) must be in 1..%d", c->N, MPPI_MAX_N); *why = buf; return MPPI_EINVAL; }
  if (c->B <= 0) { snprintf(buf, sizeof buf, "B (%d) must be > 0", c->B); *why = buf; return MPPI_EINVAL; }
  if (!(c->T > 0)) { *why = "T must be > 0"; return MPPI_EINVAL; }
  if (!(c->lambda > 0)) { *why = "lambda must be > 0"; return MPPI_EINVAL; }
  if (c->precision != MPPI_FP32 && c->precision != MPPI_FP64) { *why = "unknown precision"; return MPPI_EINVAL; }
  if (c->noise_mode != MPPI_NOISE_SUPPLIED && c->noise_mode != MPPI_NOISE_PHILOX) { *why = "unknown noise_mode"; return MPPI_EINVAL; }
  if (c->planner != MPPI_PLANNER_MPPI && c->planner != MPPI_PLANNER_DWA) { *why = "unknown planner"; return MPPI_EINVAL; }
  if (c->rollout_variant < 0 || c->rollout_variant > 7) { *why = "rollout_variant must be 0..7"; return MPPI_EINVAL; }
  if (c->noise_policy < 0 || c->noise_policy > 2) { *why = "noise_policy must be 0 (auto), 1 (latency) or 2 (throughput)"; return MPPI_EINVAL; }
  if (c->strict_math != 0 && c->strict_math != 1) { *why = "strict_math must be 0 or 1"; return MPPI_EINVAL; }
  if (c->strict_math && (c->precision != MPPI_FP32 || c->planner != MPPI_PLANNER_MPPI || c->B != 1)) { *why = "strict_math is the FP32 verification flavour of a single-instance MPPI step"; return MPPI_EINVAL; }
  if (c->n_gpus < 0 || c->n_gpus > MPPI_MAX_GPUS) { *why = "n_gpus must be in 0..8"; return MPPI_EINVAL; }
  if (c->n_gpus >= 2) {
    if (c->B != 1 || c->planner != MPPI_PLANNER_MPPI) { *why = "n_gpus > 1 shards the samples of ONE MPPI instance (B = 1, MPPI planner)"; return MPPI_EINVAL; }
    if (c->k_offset != 0 || (c->K_local != 0 && c->K_local != c->K)) { *why = "n_gpus > 1: the handle owns all K samples (k_offset = 0, K_local = 0)"; return MPPI_EINVAL; }
    if (c->K < c->n_gpus) { *why = "n_gpus > 1 needs at least one sample per GPU"; return MPPI_EINVAL; }
    for (int g = 0; g < c->n_gpus; ++g)
      if (c->device_ids[g] < 0) { *why = "device_ids must be CUDA device ordinals"; return MPPI_EINVAL; }
  }
  const int Kl = c->K_local > 0 ? c->K_local : c->K;
  if (c->k_offset < 0 || c->k_offset + Kl > c->K) { *why = "sample shard [k_offset, k_offset+K_local) must lie inside [0, K)"; return MPPI_EINVAL; }
  if (!sigma_ok(c->Sigma)) { *why = "Sigma must be symmetric positive definite"; return MPPI_EINVAL; }
  for (int i = 0; i < c->n_x; ++i)
    if (c->XL[i] > c->XU[i]) { snprintf(buf, sizeof buf, "XL[%d] (%g) must be <= XU[%d] (%g)", i, c->XL[i], i, c->XU[i]); *why = buf; return MPPI_EINVAL; }
  if (c->planner == MPPI_PLANNER_DWA && (Kl != c->K || c->B != 1)) { *why = "DWA planner runs unsharded, B = 1"; return MPPI_EINVAL; }
  return MPPI_OK;
}

int32_t mppi_destroy(mppi_handle_t* h) {
  if (!h) return MPPI_EINVAL;
  for (mppi_handle* s : h->subs) { s->subs.clear(); mppi_destroy(s); }
  h->subs.clear();
  cudaSetDevice(h->cfg.device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  drop_graphs(h);
  if (h->comm && nccl_api().CommDestroy) nccl_api().CommDestroy(h->comm);
  for (int g = 0; g < 8; ++g)
    if (h->peers_ipc && h->peer_ptr[g] && h->peer_ptr[g] != h->d_window) cudaIpcCloseMemHandle(h->peer_ptr[g]);
  if (h->d_window) cudaFree(h->d_window);
  if (h->d_peers) cudaFree(h->d_peers);
  void* dev[] = {h->d_in, h->d_eps, h->d_costs, h->d_flags, h->d_scratch, h->d_scratch_tmpl, h->d_iter, h->d_Spart,
                 h->d_eta_part, h->d_partial, h->d_gathered, h->d_res, h->d_cand, h->d_dbg, h->d_seq, h->d_his, h->d_iterlog, h->d_pstate, h->d_ref,
                 h->d_cta_part, h->d_ticket};
  for (void* p : dev) if (p) cudaFree(p);
  void* host[] = {h->h_in, h->h_res, h->h_partial, h->h_gathered, h->h_seq, h->h_ll, h->h_pdone};
  for (void* p : host) if (p) cudaFreeHost(p);
  for (auto& e : h->ev) if (e) cudaEventDestroy(e);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  if (h->stream2) cudaStreamDestroy(h->stream2);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return MPPI_OK;
}

static int create_impl(mppi_handle* h) {
  const mppi_config_t& c = h->cfg;
  int ndev = 0;
  MPPI_CUDA(h, cudaGetDeviceCount(&ndev));
  if (c.device < 0 || c.device >= ndev) { h->err = "device ordinal out of range"; return MPPI_EINVAL; }
  MPPI_CUDA(h, cudaSetDevice(c.device));
  MPPI_CUDA(h, cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  for (auto& e : h->ev) MPPI_CUDA(h, cudaEventCreate(&e));
  MPPI_CUDA(h, cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking));
  MPPI_CUDA(h, cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, c.device));
  MPPI_CUDA(h, cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
  MPPI_CUDA(h, cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
  const size_t B = c.B, N = c.N, K = h->K_local, rs = real_size(h);
  h->plen = PART_HDR + 2 * c.N;
  int ks = (h->K_local + 2047) / 2048;
  h->ksplit = ks < 1 ? 1 : (ks > 64 ? 64 : ks);
  {   // enough update blocks for ~2 waves of 4 per SM, and no more: extra blocks only repeat the prologue
    const int ngroups = (c.N + UPDATE_JG - 1) / UPDATE_JG;
    const long long per_j = (long long)c.B * h->ksplit;
    long long js = (1184 + per_j - 1) / per_j;
    h->jsplit = int(js < 1 ? 1 : (js > ngroups ? ngroups : js));
    if (const char* e = getenv("MPPI_UPDATE_JSPLIT")) { const int v = atoi(e); if (v >= 1 && v <= ngroups) h->jsplit = v; }
  }
  h->cutoff = c.feasibility_count < c.K;
  {
    int variant = c.rollout_variant;
    if (const char* e = getenv("MPPI_ROLLOUT_VARIANT")) variant = atoi(e);
    if (const char* e = getenv("MPPI_PDL")) h->pdl = atoi(e) != 0;
    if (const char* e = getenv("MPPI_STEP_EVENTS")) h->step_events = atoi(e) != 0;
    const bool can_pack = c.precision == MPPI_FP32 && (h->K_local % 2) == 0;
    // pre-generated noise (variants 0-2): single-instance Philox MPPI steps; 3 / 4 keep Philox inside the rollout
    // noise_policy: latency = pre-generate (one more HBM pass over eps, the result arrives earlier), throughput = Philox inside
    // the rollout.  auto: latency up to 131,072 samples per GPU (the eps tensor stays L2-sized and the step is latency-bound).
    const long long samples_ = (long long)h->K_local * c.B;
    const bool want_pregen = c.noise_policy == 1 || (c.noise_policy == 0 && (samples_ <= 131072 || c.model == MPPI_MODEL_DYNAMIC7));
    h->pregen_ok = c.planner == MPPI_PLANNER_MPPI && c.noise_mode == MPPI_NOISE_PHILOX && c.B == 1 && variant != 3 && variant != 4 &&
                   (want_pregen || variant == 5) && !c.strict_math;
    if (const char* e = getenv("MPPI_PREGEN")) h->pregen_ok = h->pregen_ok && atoi(e) != 0;
    // scan-form fused step (mppi_scan.cuh): FP32 kinematic MPPI, one instance, no cut-off; obstacle kind / reference path are
    // checked per step (scan_eligible).
    h->scan_ok = c.precision == MPPI_FP32 && c.model == MPPI_MODEL_KINEMATIC5 && c.planner == MPPI_PLANNER_MPPI && c.B == 1 &&
                 !h->cutoff && !c.strict_math && scan_chunk(c.N) != 0;
    // opt-in (rollout_variant = 5, or MPPI_SCAN=1 with the auto variant): on B200 the sequential packed rollout + update + finalize
    // is still the faster step at every BASELINE size (profiles/r02_scan_vs_sequential.md), so auto keeps it
    bool want_scan = variant == 5;
    if (const char* e = getenv("MPPI_SCAN")) want_scan = want_scan || (variant == 0 && atoi(e) != 0);
    h->scan_ok = h->scan_ok && want_scan;
    if (variant == 5 && !h->scan_ok) { h->err = "rollout_variant 5 (scan-form fused step) needs FP32, the kinematic model, the MPPI planner, B = 1, no cut-off, N <= 104"; return MPPI_EINVAL; }
    // auto: packing halves the warps, so it needs enough samples to keep every SM sub-partition busy; with the
    // per-lane Philox / Box-Muller work out of the loop it pays from 65,536 samples (tools/sweep_pregen.py)
    const long long samples = (long long)h->K_local * c.B;
    const long long pack_min = h->pregen_ok ? MPPI_PACK_MIN_SAMPLES_PREGEN : MPPI_PACK_MIN_SAMPLES;
    h->packed = !c.strict_math && can_pack && (variant == 2 || variant == 4 || variant == 6 || variant == 7 || (variant == 0 && samples >= pack_min));
    // two-stage warp pipeline (mppi_pipe.cuh): packed FP32 kinematic rollout that READS eps; obstacle kind / reference path are
    // checked per step (pipe_eligible).  auto uses it at EVERY sample count (its 50-step chain is the shortest of the three
    // rollouts from 64 samples up: 9.3 us against 10.8 one sample per thread and 12.4 packed, tools/dbg/spans.py); where a step
    // does not qualify (walls, reference path) a handle below the packing threshold falls back to one sample per thread.
    const bool pipe_cfg = !c.strict_math && can_pack && c.model == MPPI_MODEL_KINEMATIC5 && c.planner == MPPI_PLANNER_MPPI && c.B == 1 &&
                          (c.noise_mode == MPPI_NOISE_SUPPLIED || h->pregen_ok) && size_t(c.N) * PIPE_SAMPLES * 8 <= size_t(PIPE_SLAB_MAX) &&
                          c.N <= EPS_GROUP * EPS_MAX_GROUPS;
    h->pipe = pipe_cfg && (variant == 0 || variant == 6 || variant == 7);
    h->packed_seq = h->packed;            // the sequential rollout packs (what a step that cannot take the pipeline runs)
    h->packed = h->packed || h->pipe;
    if (const char* e = getenv("MPPI_PIPE")) h->pipe = h->pipe && atoi(e) != 0;
    // ... with the softmin update done by the rollout's own blocks (no update_kernel; rollout_variant 7, or MPPI_PIPE_FOLD=1 with the
    // auto variant).  Opt-in: on B200 the device-timed step gets 1.2 us shorter (the next iteration's noise starts 2.8 us earlier)
    // and the result reaches the host 1.3 us LATER (folding 256 per-block partials in finalize costs more than the 64-way tree
    // behind update_kernel), DESIGN.md 4.5
    bool want_fold = variant == 7;
    if (const char* e = getenv("MPPI_PIPE_FOLD")) want_fold = want_fold || (variant == 0 && atoi(e) != 0);
    h->pipe_fold = h->pipe && want_fold && !h->cutoff && (h->K_local + PIPE_SAMPLES - 1) / PIPE_SAMPLES <= FOLD_MAX_CTAS;
    if (variant == 7 && !h->pipe_fold) { h->err = "rollout_variant 7 (warp pipeline with the update in its blocks) needs what variant 6 needs, no feasibility cut-off (FeasibilityCount >= SamplingNumber) and at most 163,840 samples per GPU"; return MPPI_EINVAL; }
    if (variant == 6 && !h->pipe) { h->err = "rollout_variant 6 (two-stage warp pipeline) needs FP32, even K_local, the kinematic model, the MPPI planner, eps read from memory (SUPPLIED or pre-generated noise) and N <= 56"; return MPPI_EINVAL; }
  }
  MPPI_CUDA(h, cudaMalloc(&h->d_in, B * (MPPI_MAX_NX + 2 * N) * sizeof(double)));
  if (c.planner == MPPI_PLANNER_MPPI) MPPI_CUDA(h, cudaMalloc(&h->d_eps, B * N * K * 2 * rs));
  MPPI_CUDA(h, cudaMalloc(&h->d_costs, B * K * rs));
  MPPI_CUDA(h, cudaMalloc(&h->d_flags, B * K));
  MPPI_CUDA(h, cudaMalloc(&h->d_scratch, B * sizeof(Scratch)));
  MPPI_CUDA(h, cudaMalloc(&h->d_scratch_tmpl, B * sizeof(Scratch)));
  MPPI_CUDA(h, cudaMalloc(&h->d_iter, 2 * sizeof(unsigned int)));   // [0] iteration counter, [1] = [0] + 1 published by the rollout
  MPPI_CUDA(h, cudaMalloc(&h->d_Spart, B * h->ksplit * N * 2 * rs));
  MPPI_CUDA(h, cudaMalloc(&h->d_eta_part, B * h->ksplit * rs));
  MPPI_CUDA(h, cudaMalloc(&h->d_partial, (B + 8) * h->plen * sizeof(double)));   // + room for the peers' copies (in-kernel exchange)
  MPPI_CUDA(h, cudaMalloc(&h->d_gathered, 8 * B * h->plen * sizeof(double)));
  // results: [B][RES_SCALARS] scalars | [B][2N] U_new | [B][(N+1)*n_x] nominal trajectory, one D2H copy
  MPPI_CUDA(h, cudaMalloc(&h->d_res, B * (RES_SCALARS + 2 * N + (N + 1) * c.n_x) * sizeof(double)));
  h->d_traj = h->d_res + B * (RES_SCALARS + 2 * N);
  if (c.planner == MPPI_PLANNER_DWA) MPPI_CUDA(h, cudaMalloc(&h->d_cand, K * 2 * sizeof(double)));
  if (h->pipe && !h->scan_ok) MPPI_CUDA(h, cudaMalloc(&h->d_cta_part, size_t(2 * N + 2) * FOLD_MAX_CTAS * sizeof(float)));
  if (h->scan_ok) {
    const int tiles = (h->K_local + SCAN_SPW - 1) / SCAN_SPW;
    const int want = h->sm_count * 3;   // __launch_bounds__(SCAN_BS, 3): one resident wave
    h->scan_ctas = std::max(1, std::min(std::min(want, FOLD_MAX_CTAS), (tiles + SCAN_NW - 1) / SCAN_NW));
    if (const char* e = getenv("MPPI_SCAN_CTAS")) { const int v = atoi(e); if (v >= 1 && v <= FOLD_MAX_CTAS) h->scan_ctas = v; }
    MPPI_CUDA(h, cudaMalloc(&h->d_cta_part, size_t(2 * N + 2) * FOLD_MAX_CTAS * sizeof(float)));
    MPPI_CUDA(h, cudaMalloc(&h->d_ticket, sizeof(unsigned int)));
    MPPI_CUDA(h, cudaMemset(h->d_ticket, 0, sizeof(unsigned int)));
  }
  MPPI_CUDA(h, cudaMalloc(&h->d_dbg, 16 * sizeof(unsigned long long)));
  MPPI_CUDA(h, cudaMemset(h->d_dbg, 0, 16 * sizeof(unsigned long long)));
  // inputs and results are MAPPED pinned memory: for B == 1 the step's kernels read / write them
  // directly over PCIe (no memcpy nodes in the step graph); every 16-byte result line carries the
  // step's sequence number, so the host needs no separate completion flag (and the GPU no system fence)
  MPPI_CUDA(h, cudaHostAlloc(&h->h_in, B * (MPPI_MAX_NX + 2 * N) * sizeof(double), cudaHostAllocMapped));
  MPPI_CUDA(h, cudaHostAlloc(&h->h_seq, sizeof(unsigned long long), cudaHostAllocMapped));
  *h->h_seq = 0;
  MPPI_CUDA(h, cudaMalloc(&h->d_seq, sizeof(unsigned long long)));
  MPPI_CUDA(h, cudaMemset(h->d_seq, 0, sizeof(unsigned long long)));
  MPPI_CUDA(h, cudaHostAlloc(&h->h_res, (B * (RES_SCALARS + 2 * N) + (N + 1) * c.n_x) * sizeof(double), cudaHostAllocMapped));
  h->h_traj = h->h_res + B * (RES_SCALARS + 2 * N);
  h->n_ll = RES_SCALARS + 2 * N + (N + 1) * c.n_x;
  MPPI_CUDA(h, cudaHostAlloc(&h->h_ll, size_t(h->n_ll) * 16, cudaHostAllocMapped));
  memset(h->h_ll, 0, size_t(h->n_ll) * 16);
  MPPI_CUDA(h, cudaMallocHost(&h->h_partial, B * h->plen * sizeof(double)));
  MPPI_CUDA(h, cudaMallocHost(&h->h_gathered, 8 * B * h->plen * sizeof(double)));
  // inputs: X0 from the config for every instance, zero nominal (setup.jl:55)
  memset(h->h_in, 0, B * (MPPI_MAX_NX + 2 * N) * sizeof(double));
  for (size_t b = 0; b < B; ++b)
    for (int i = 0; i < c.n_x; ++i) h->h_in[b * MPPI_MAX_NX + i] = c.X0[i];
  MPPI_CUDA(h, cudaMemcpy(h->d_in, h->h_in, B * (MPPI_MAX_NX + 2 * N) * sizeof(double), cudaMemcpyHostToDevice));
  std::vector<Scratch> tmpl(B);
  for (auto& s : tmpl) { s.rho_key = ~0ull; s.argmin = ~0u; s.feasible_count = 0; s.k_eff = h->K_local; s.pad = 0; }
  MPPI_CUDA(h, cudaMemcpy(h->d_scratch_tmpl, tmpl.data(), B * sizeof(Scratch), cudaMemcpyHostToDevice));
  MPPI_CUDA(h, cudaMemcpy(h->d_scratch, tmpl.data(), B * sizeof(Scratch), cudaMemcpyHostToDevice));
  MPPI_CUDA(h, cudaMemset(h->d_iter, 0, 2 * sizeof(unsigned int)));
  MPPI_CUDA(h, cudaMemset(h->d_res, 0, B * (RES_SCALARS + 2 * N) * sizeof(double)));
  rebuild_params(h);
  return MPPI_OK;
}

// ONE handle, ONE host thread, G GPUs (cfg.n_gpus, cfg.device_ids): the reference's caller is a single Julia task
// (OptimalControl/MPPI/main.jl:55-83), so a drop-in that needs one process per GPU would leave it on one GPU.  Rank g owns the
// contiguous samples shard_range(K, g, G) on device_ids[g]; every member is an ordinary sharded handle whose exchange window is
// a plain device allocation reachable from the other GPUs through peer access (no CUDA IPC inside one process); the step
// launches all members' graphs before it waits for any (the in-kernel all-gather spins for its peers).  The same ordinal
// may appear more than once in device_ids: the shards then share that GPU on separate streams (logical shards; tests).
static int create_multi(const mppi_config_t* cfg, mppi_handle_t** out) {
  const int G = cfg->n_gpus;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess) return fail(nullptr, MPPI_ECUDA, "cudaGetDeviceCount failed");
  for (int g = 0; g < G; ++g)
    if (cfg->device_ids[g] >= ndev) return fail(nullptr, MPPI_EINVAL, "device_ids: device ordinal out of range");
  std::vector<mppi_handle*> hs;
  auto bail = [&](int rc, const std::string& why) {
    create_error() = why;
    for (mppi_handle* x : hs) mppi_destroy(x);
    return rc;
  };
  const int base = cfg->K / G, rem = cfg->K % G;   // = motionplanning_b200/dist.py shard_range
  for (int g = 0; g < G; ++g) {
    mppi_handle* h = new (std::nothrow) mppi_handle();
    if (!h) return bail(MPPI_ENOMEM, "out of host memory");
    hs.push_back(h);
    h->cfg = *cfg;
    h->cfg.device = cfg->device_ids[g];
    h->cfg.n_gpus = g == 0 ? G : 0;
    h->cfg.k_offset = g * base + std::min(g, rem);
    h->K_local = base + (g < rem ? 1 : 0);
    h->cfg.K_local = h->K_local;
    h->spmg_member = true;
    const int rc = create_impl(h);
    if (rc != MPPI_OK) return bail(rc, h->err);
  }
  for (int a = 0; a < G; ++a) {
    for (int b = 0; b < G; ++b) {
      const int da = hs[a]->cfg.device, db = hs[b]->cfg.device;
      if (da == db) continue;
      int can = 0;
      cudaSetDevice(da);
      if (cudaDeviceCanAccessPeer(&can, da, db) != cudaSuccess || !can) return bail(MPPI_EINVAL, "n_gpus > 1 needs peer access between the devices (NVLink / NVSwitch)");
      const cudaError_t e = cudaDeviceEnablePeerAccess(db, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return bail(MPPI_ECUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
      cudaGetLastError();
    }
  }
  for (int g = 0; g < G; ++g) {
    mppi_handle* h = hs[g];
    cudaSetDevice(h->cfg.device);
    if (cudaMalloc(&h->d_window, sizeof(PeerWindow)) != cudaSuccess || cudaMemset(h->d_window, 0, sizeof(PeerWindow)) != cudaSuccess)
      return bail(MPPI_ENOMEM, "exchange window allocation failed");
  }
  for (int g = 0; g < G; ++g) {
    mppi_handle* h = hs[g];
    cudaSetDevice(h->cfg.device);
    for (int q = 0; q < G; ++q) h->peer_ptr[q] = hs[q]->d_window;
    if (cudaMalloc(&h->d_peers, PEER_MAX_RANKS * sizeof(PeerWindow*)) != cudaSuccess ||
        cudaMemcpy(h->d_peers, h->peer_ptr, PEER_MAX_RANKS * sizeof(PeerWindow*), cudaMemcpyHostToDevice) != cudaSuccess)
      return bail(MPPI_ECUDA, "peer table upload failed");
    h->seq_base = 0;
    h->rank = g; h->n_ranks = G; h->p2p = true;
  }
  mppi_handle* root = hs[0];
  root->subs.assign(hs.begin() + 1, hs.end());
  cudaSetDevice(root->cfg.device);
  *out = root;
  return MPPI_OK;
}

int32_t mppi_create(const mppi_config_t* cfg, mppi_handle_t** out) {
  if (!cfg || !out) return fail(nullptr, MPPI_EINVAL, "null argument");
  *out = nullptr;
  std::string why;
  int rc = validate(cfg, &why);
  if (rc != MPPI_OK) return fail(nullptr, rc, why);
  if (cfg->n_gpus >= 2) return create_multi(cfg, out);
  mppi_handle* h = new (std::nothrow) mppi_handle();
  if (!h) return fail(nullptr, MPPI_ENOMEM, "out of host memory");
  h->cfg = *cfg;
  h->K_local = cfg->K_local > 0 ? cfg->K_local : cfg->K;
  h->cfg.K_local = h->K_local;
  rc = create_impl(h);
  if (rc != MPPI_OK) { create_error() = h->err; mppi_destroy(h); return rc; }
  *out = h;
  return MPPI_OK;
}

static int quiesce(mppi_handle* h) {   // configuration is about to change: the stream must be idle
  if (h->in_flight) { MPPI_CUDA(h, cudaStreamSynchronize(h->stream)); h->in_flight = h->inputs_busy = false; }
  return MPPI_OK;
}
static int quiesce_inputs(mppi_handle* h) {   // pinned inputs are about to change: no step may still read them
  if (h->inputs_busy) { MPPI_CUDA(h, cudaStreamSynchronize(h->stream)); h->in_flight = h->inputs_busy = false; }
  return MPPI_OK;
}

int32_t mppi_set_x0(mppi_handle_t* h, const double* x0, int32_t n) {
  H_CHECK(h);
  char buf[160];
  if (!x0) return fail(h, MPPI_EINVAL, "null x0");
  if (n != h->cfg.n_x) {   // setup.jl:25-27
    snprintf(buf, sizeof buf, "Length of X0 (%d) must match number of states (%d)", n, h->cfg.n_x);
    return fail(h, MPPI_EINVAL, buf);
  }
  FORWARD_SUBS(h, mppi_set_x0(s_, x0, n));
  int rc = quiesce_inputs(h); if (rc) return rc;
  for (int b = 0; b < h->cfg.B; ++b)
    for (int i = 0; i < n; ++i) h->h_in[size_t(b) * MPPI_MAX_NX + i] = x0[size_t(b) * n + i];
  h->x0_dirty = true;   // uploaded by the next step's graph (one H2D node for x0 + U_nom)
  return MPPI_OK;
}

int32_t mppi_set_nominal(mppi_handle_t* h, const double* U, int32_t N, int32_t m) {
  H_CHECK(h);
  if (!U) return fail(h, MPPI_EINVAL, "null U");
  if (N != h->cfg.N || m != h->cfg.m) return fail(h, MPPI_EINVAL, "NominalControl must be N x numControls");
  FORWARD_SUBS(h, mppi_set_nominal(s_, U, N, m));
  int rc = quiesce_inputs(h); if (rc) return rc;
  const size_t B = h->cfg.B, off = B * MPPI_MAX_NX, n = B * 2 * N;
  memcpy(h->h_in + off, U, n * sizeof(double));
  h->u_dirty = true;
  return MPPI_OK;
}

int32_t mppi_set_sigma_lambda(mppi_handle_t* h, const double* Sigma, double lambda) {
  H_CHECK(h);
  if (!Sigma) return fail(h, MPPI_EINVAL, "null Sigma");
  if (!sigma_ok(Sigma) || !(lambda > 0)) return fail(h, MPPI_EINVAL, "Sigma must be symmetric positive definite and lambda > 0");
  FORWARD_SUBS(h, mppi_set_sigma_lambda(s_, Sigma, lambda));
  int rc = quiesce(h); if (rc) return rc;
  for (int i = 0; i < 4; ++i) h->cfg.Sigma[i] = Sigma[i];
  h->cfg.lambda = lambda;
  rebuild_params(h);
  drop_graphs(h);
  h->noise_ready = false;   // pre-generated eps was drawn with the old chol(Sigma)
  return MPPI_OK;
}

int32_t mppi_set_circles(mppi_handle_t* h, const double* xyr, int32_t M_c) {
  H_CHECK(h);
  if (M_c < 0 || M_c > MPPI_MAX_CIRCLES || (M_c > 0 && !xyr)) return fail(h, MPPI_EINVAL, "circle count out of range");
  FORWARD_SUBS(h, mppi_set_circles(s_, xyr, M_c));
  int rc = quiesce(h); if (rc) return rc;
  h->circles.assign(xyr, xyr + size_t(3) * M_c);
  rebuild_params(h);
  drop_graphs(h);
  return MPPI_OK;
}

int32_t mppi_set_blocks(mppi_handle_t* h, const double* blocks, int32_t M_p) {
  H_CHECK(h);
  if (M_p < 0 || M_p > MPPI_MAX_BLOCKS || (M_p > 0 && !blocks)) return fail(h, MPPI_EINVAL, "block count out of range");
  FORWARD_SUBS(h, mppi_set_blocks(s_, blocks, M_p));
  int rc = quiesce(h); if (rc) return rc;
  h->blocks.assign(blocks, blocks + size_t(5) * M_p);
  rebuild_params(h);
  drop_graphs(h);
  return MPPI_OK;
}

int32_t mppi_set_goal(mppi_handle_t* h, const double* goal2) {
  H_CHECK(h);
  if (!goal2) return fail(h, MPPI_EINVAL, "null goal");
  FORWARD_SUBS(h, mppi_set_goal(s_, goal2));
  int rc = quiesce(h); if (rc) return rc;
  h->cfg.goal[0] = goal2[0]; h->cfg.goal[1] = goal2[1];
  rebuild_params(h);
  drop_graphs(h);
  return MPPI_OK;
}

int32_t mppi_set_reference_path(mppi_handle_t* h, const double* xy, int32_t n, double weight, int32_t window) {
  H_CHECK(h);
  if (n < 0 || n > MPPI_MAX_REF) { char b[96]; snprintf(b, sizeof b, "reference path: n (%d) must be in 0..%d", n, MPPI_MAX_REF); return fail(h, MPPI_EINVAL, b); }
  if (n > 0 && (!xy || !(weight >= 0) || window < 1)) return fail(h, MPPI_EINVAL, "reference path: null points, negative weight or window < 1");
  if (n > 0 && h->cfg.strict_math) return fail(h, MPPI_EINVAL, "the reference-path term is not part of the strict_math flavour");
  if (n > 0 && h->cfg.planner == MPPI_PLANNER_MPPI && h->cfg.noise_mode != MPPI_NOISE_PHILOX)
    return fail(h, MPPI_EINVAL, "a reference path needs PHILOX noise (the tracking rollout is not built for supplied noise)");
  FORWARD_SUBS(h, mppi_set_reference_path(s_, xy, n, weight, window));
  int rc = quiesce(h); if (rc) return rc;
  h->ref_path.assign(xy, xy + size_t(2) * n);
  h->ref_w = n > 0 ? weight : 0.0;
  h->ref_window = n > 0 ? window : 0;
  if (n > 0) {
    if (!h->d_ref) MPPI_CUDA(h, cudaMalloc(&h->d_ref, size_t(2) * MPPI_MAX_REF * sizeof(double)));
    if (h->cfg.precision == MPPI_FP64) {
      MPPI_CUDA(h, cudaMemcpy(h->d_ref, xy, size_t(2) * n * sizeof(double), cudaMemcpyHostToDevice));
    } else {
      std::vector<float> f(size_t(2) * n);
      for (size_t i = 0; i < f.size(); ++i) f[i] = float(xy[i]);
      MPPI_CUDA(h, cudaMemcpy(h->d_ref, f.data(), f.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
  }
  rebuild_params(h);
  drop_graphs(h);
  return MPPI_OK;
}

int32_t mppi_set_candidates(mppi_handle_t* h, const double* ctrl, int32_t K, int32_t m) {
  H_CHECK(h);
  if (h->cfg.planner != MPPI_PLANNER_DWA) return fail(h, MPPI_ESTATE, "candidates are only used by the DWA planner");
  if (!ctrl || K != h->K_local || m != 2) return fail(h, MPPI_EINVAL, "candidates must be K x 2");
  int rc = quiesce(h); if (rc) return rc;
  MPPI_CUDA(h, cudaMemcpy(h->d_cand, ctrl, size_t(K) * 2 * sizeof(double), cudaMemcpyHostToDevice));
  h->have_candidates = true;
  return MPPI_OK;
}

}  // extern "C"

#include "mppi_api_step.inc"

// ---------------------------------------------------------------------------------------------------
// mppi_check_poses_static: mppi_check_poses without any allocation.  No counterpart in the reference;
// added for callers that cannot call cudaMalloc (stream capture,
// tight serving loops, harnesses that forbid it).  mppi_check_poses stages the caller's poses and the
// hit counts in two cudaMalloc'ed scratch buffers per call; here that scratch is ONE file-scope device
// array of fixed size, held by every GPU that loads the library until the process exits
// (MPPI_STATIC_POSES_MAX poses: 3 doubles in + 1 int out each = 448 KiB).
// In host code `g_pose_scratch` is only a stand-in for the symbol, not a device address: the address
// is resolved once per device, with that device current, by cudaGetSymbolAddress and that pointer is
// what the kernel receives.  All calls share the array and it keeps the last call's contents: the hit
// region is cleared on the stream before the kernel runs; calls that overlap on the GPU (two handles
// on different streams, or two host threads) would race on it.
// ---------------------------------------------------------------------------------------------------
struct PoseScratch {
  double poses[3 * MPPI_STATIC_POSES_MAX];
  int hit[MPPI_STATIC_POSES_MAX];
};
static __device__ __align__(256) PoseScratch g_pose_scratch;

extern "C" int32_t mppi_check_poses_static(mppi_handle_t* h, const double* poses, int32_t n, int32_t* hit) {
  H_CHECK(h);
  if (n < 0 || (n > 0 && (!poses || !hit))) return fail(h, MPPI_EINVAL, "bad pose buffer");
  if (n > MPPI_STATIC_POSES_MAX) {
    char b[160];
    snprintf(b, sizeof b, "mppi_check_poses_static: n (%d) exceeds the static scratch (MPPI_STATIC_POSES_MAX = %d); split the batch or use mppi_check_poses",
             n, MPPI_STATIC_POSES_MAX);
    return fail(h, MPPI_EINVAL, b);
  }
  if (n == 0) return MPPI_OK;
  static PoseScratch* dev_ptr[64] = {nullptr};   // per device: resolved at first use with that device current (H_CHECK set it)
  const int dev = h->cfg.device;
  if (dev < 0 || dev >= 64) return fail(h, MPPI_EINVAL, "device ordinal out of range for the static scratch table");
  if (!dev_ptr[dev]) {
    PoseScratch* p = nullptr;
    MPPI_CUDA(h, cudaGetSymbolAddress((void**)&p, g_pose_scratch));
    dev_ptr[dev] = p;
  }
  PoseScratch* sc = dev_ptr[dev];
  MPPI_CUDA(h, cudaMemcpyAsync(sc->poses, poses, size_t(3) * n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  MPPI_CUDA(h, cudaMemsetAsync(sc->hit, 0, size_t(n) * sizeof(int), h->stream));   // the array keeps the previous call's hits
  check_poses_kernel<<<(n + 127) / 128, 128, 0, h->stream>>>(h->pd, sc->poses, n, sc->hit);
  h->launches++;
  MPPI_CUDA(h, cudaGetLastError());
  MPPI_CUDA(h, cudaMemcpyAsync(hit, sc->hit, size_t(n) * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  MPPI_CUDA(h, cudaStreamSynchronize(h->stream));
  return MPPI_OK;
}
