// mppi_handle.cuh — host-side handle: configuration, device/pinned buffers,
// static parameter blocks, NCCL (dlopen'ed lazily; no link-time dependency).
#pragma once
#include <dlfcn.h>
#include <nccl.h>   // declarations only; symbols are resolved with dlsym at mppi_comm_init

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "mppi_scan.cuh"
#include "mppi_pipe.cuh"
#include "mppi_strict.cuh"
#include "mppi_plant.cuh"

namespace mppi {

struct NcclApi {
  void* lib = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclAllGather) AllGather = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  std::string err;
  bool load() {
    if (lib) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (lib) break;
    }
    if (!lib) { err = std::string("cannot dlopen libnccl: ") + dlerror(); return false; }
#define MPPI_NCCL_SYM(f) f = (decltype(f))dlsym(lib, "nccl" #f); if (!f) { err = "missing symbol nccl" #f; lib = nullptr; return false; }
    MPPI_NCCL_SYM(GetUniqueId) MPPI_NCCL_SYM(CommInitRank) MPPI_NCCL_SYM(AllGather)
    MPPI_NCCL_SYM(CommDestroy) MPPI_NCCL_SYM(GetErrorString)
#undef MPPI_NCCL_SYM
    return true;
  }
};
inline NcclApi& nccl_api() { static NcclApi a; return a; }

}  // namespace mppi

using mppi::Scratch;

struct mppi_handle {
  mppi_config_t cfg;
  std::string err;
  cudaStream_t stream = nullptr;
  int K_local = 0, ksplit = 1, jsplit = 1, plen = 0;   // update_kernel grid = (ksplit, jsplit, B)
  bool cutoff = false;
  bool packed = false;   // rollout carries two samples per thread as f32x2 (FP32, even K_local)
  bool packed_seq = false;   // the SEQUENTIAL rollout packs (a handle may use the pipeline below the packing threshold and fall back to one-sample threads)
  bool pipe_fold = false;   // ... whose blocks also do their share of the softmin update (no update_kernel in the step; pipe_folds)
  bool pipe = false;     // ... and, where the step qualifies (pipe_eligible), as the two-stage warp pipeline of mppi_pipe.cuh
  std::vector<double> circles, blocks;   // host copies (3 x Mc, 5 x Mp col-major)
  // device
  double* d_in = nullptr;        // [B][8] x0 then [B][2N] U  (d_x0 = d_in, d_U = d_in + 8B)
  void* d_eps = nullptr;
  void* d_costs = nullptr;
  uint8_t* d_flags = nullptr;
  Scratch* d_scratch = nullptr;
  Scratch* d_scratch_tmpl = nullptr;
  unsigned int* d_iter = nullptr;
  void* d_Spart = nullptr;
  void* d_eta_part = nullptr;
  double* d_partial = nullptr;   // [B][plen]
  double* d_gathered = nullptr;  // [G][B][plen]
  double* d_res = nullptr;       // [B][RES_SCALARS] then [B][2N]
  double* d_traj = nullptr;      // [B][(N+1)*n_x], inside the d_res allocation
  double* d_cand = nullptr;      // DWA candidates K x 2 col-major
  unsigned long long* d_dbg = nullptr;   // [16] finalize phase clocks
  unsigned long long* d_seq = nullptr;   // completed-step counter
  unsigned long long* h_seq = nullptr;   // mapped pinned: completion flag written by finalize
  // closed loop (mppi_run_closed_loop): plant kernel arguments baked into the V_PLANT graphs
  std::vector<double> ref_path;   // 2 per point; tracking term
  double ref_w = 0.0; int ref_window = 0;
  void* d_ref = nullptr;          // the polyline in the handle's precision
  bool pdl = true;                // programmatic dependent launches inside the step (env MPPI_PDL=0 disables)
  // pre-generated noise (V_PREGEN): second capture branch for noise_kernel, and whether d_eps already holds
  // the noise of the iteration the device counter points at
  bool pregen_ok = false, noise_ready = false;
  // scan-form fused step (mppi_scan.cuh): eligible configuration, grid size, per-CTA partial table, arrival counter
  bool scan_ok = false;
  int scan_ctas = 0;
  float* d_cta_part = nullptr;
  unsigned int* d_ticket = nullptr;
  int sm_count = 148;
  cudaStream_t stream2 = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  mppi::PlantArgs plant = {};
  int shift_k = 0;
  double* d_his = nullptr; double* d_iterlog = nullptr; int* d_pstate = nullptr; int* h_pdone = nullptr;
  int his_cap = 0, log_cap = 0;
  unsigned long long seq_launched = 0;   // steps (finalize<COMBINE>) enqueued so far
  // pinned host mirrors
  double* h_in = nullptr;
  double* h_res = nullptr;
  unsigned long long* h_ll = nullptr;    // mapped pinned: the B == 1 result block as LL lines, two 8-byte words {value lo32 | flag, value hi32 | flag} per double
  int n_ll = 0;
  double* h_traj = nullptr;      // inside the h_res allocation
  double* h_partial = nullptr;
  double* h_gathered = nullptr;
  // static params
  mppi::StaticParams<float> pf;
  mppi::StaticParams<double> pd;
  mppi::StrictParams sp;   // strict_math flavour: the configuration rounded once to float, as the reference's functions see it
  // graphs, keyed by variant
  std::map<int, cudaGraphExec_t> graphs;
  // multi-GPU
  ncclComm_t comm = nullptr;
  int rank = 0, n_ranks = 1;
  mppi::PeerWindow* d_window = nullptr;          // this rank's exchange window (cudaMalloc: IPC-exportable)
  mppi::PeerWindow* peer_ptr[8] = {nullptr};     // all ranks' windows as seen from this device
  mppi::PeerWindow** d_peers = nullptr;          // device copy of peer_ptr
  bool p2p = false;                              // exchange in-kernel over peer memory instead of NCCL
  bool peers_ipc = false;                        // peer_ptr[] hold CUDA-IPC mappings (mppi_peer_init) rather than plain pointers of this process
  // single-process multi-GPU (cfg.n_gpus >= 2; the reference's caller is ONE Julia task, main.jl:55-83): this handle is rank 0
  // and drives the other ranks' handles from the caller's thread.  Members take the FULL eps tensor [N][K][m] in SUPPLIED mode.
  std::vector<mppi_handle*> subs;                // ranks 1..G-1 (empty on ordinary handles and on the members themselves)
  bool spmg_member = false;                      // rank 0..G-1 of a single-process group
  unsigned long long seq_base = 0;               // device step counter at mppi_peer_init
  // profiling
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // step begin | rollout end | update end | finalize end | step end | inputs staged
  int profiling = 0;   // 0 off | 1 eager launches bracketed by events | 2 graph replay, kernels stamp %globaltimer
  float us[4] = {0, 0, 0, 0};
  int64_t launches = 0;
  bool stepped = false;
  bool timings_pending = false;   // events of the last step not yet read back
  bool step_events = false;       // graph replays are bracketed by ev[0] / ev[4] (armed by the first mppi_get_timings / mppi_set_profiling call)
  bool step_timed = false;        // ... and the last replay was
  bool x0_dirty = false, u_dirty = false;   // pinned x0 / U_nom changed since the last upload
  bool in_flight = false;   // work enqueued on the stream and not yet synchronised
  bool inputs_busy = false; // a step that reads the pinned input block may not have consumed it yet
  bool have_candidates = false;
};

namespace mppi {

inline std::string& create_error() { static std::string e; return e; }

#define MPPI_CUDA(h, call)                                                                   \
  do {                                                                                       \
    cudaError_t e_ = (call);                                                                 \
    if (e_ != cudaSuccess) {                                                                 \
      (h)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                         \
      return (e_ == cudaErrorMemoryAllocation) ? MPPI_ENOMEM : MPPI_ECUDA;                   \
    }                                                                                        \
  } while (0)

template <class R>
void build_params(const mppi_handle* h, StaticParams<R>* p) {
  const mppi_config_t& c = h->cfg;
  memset(p, 0, sizeof(*p));
  p->K = c.K; p->K_local = h->K_local; p->k_offset = c.k_offset; p->N = c.N; p->B = c.B;
  p->n_circles = int(h->circles.size() / 3); p->n_blocks = int(h->blocks.size() / 5);
  p->n_ref = int(h->ref_path.size() / 2); p->ref_window = h->ref_window; p->ref_w = R(h->ref_w);
  p->planner = c.planner; p->feasibility_count = c.feasibility_count; p->n_x = c.n_x;
  p->seed_lo = uint32_t(c.seed); p->seed_hi = uint32_t(c.seed >> 32);
  const double dt = c.T / c.N;   // setup.jl:53
  p->dt = R(dt); p->half_dt = R(dt) / R(2); p->lambda = R(c.lambda); p->inv_lambda = R(1.0 / c.lambda);
  // Sigma col-major [s00, s10, s01, s11]; chol lower; lambda*inv(Sigma)
  const double s00 = c.Sigma[0], s10 = c.Sigma[1], s01 = c.Sigma[2], s11 = c.Sigma[3];
  const double l00 = std::sqrt(s00), l10 = s10 / l00, l11 = std::sqrt(s11 - l10 * l10);
  p->L00 = R(l00); p->L10 = R(l10); p->L11 = R(l11);
  const double det = s00 * s11 - s01 * s10;
  const double i00 = s11 / det, i10 = -s10 / det, i01 = -s01 / det, i11 = s00 / det;   // inv(r,c)
  p->G00 = R(c.lambda * i00); p->G01 = R(c.lambda * i10); p->G10 = R(c.lambda * i01); p->G11 = R(c.lambda * i11);
  for (int i = 0; i < 2; ++i) { p->goal[i] = R(c.goal[i]); p->CL[i] = R(c.CL[i]); p->CU[i] = R(c.CU[i]); }
  for (int i = 0; i < MPPI_MAX_NX; ++i) { p->XL[i] = R(c.XL[i]); p->XU[i] = R(c.XU[i]); }
  p->slack = R(c.slack_penalty); p->obs_pen = R(c.obstacle_penalty); p->term_scale = R(c.terminal_scale);
  p->la = R(c.la); p->lb = R(c.lb); p->mass = R(c.mass); p->Izz = R(c.Izz); p->inv_mass = R(1) / R(c.mass); p->inv_Izz = R(1) / R(c.Izz); p->mu = R(c.mu); p->KFZX = R(c.KFZX);
  p->FZF0 = R(c.KFZF) * R(c.g); p->FZR0 = R(c.KFZR) * R(c.g);
  p->tireB = R(c.tire[0]) / R(c.mu); p->tireC = R(c.tire[1]); p->tireD = R(c.tire[2]); p->tireE = R(c.tire[3]);
  p->Lwb = R(c.la) + R(c.lb); p->inv_Lwb = R(1) / p->Lwb;
  for (int i = 0; i < 6; ++i) p->w[i] = R(c.stage_w[i]);
  p->veh_off = R(c.veh_center_offset); p->veh_hl = R(c.veh_L) / R(2); p->veh_hw = R(c.veh_W) / R(2);
  for (int o = 0; o < p->n_circles; ++o) {
    const R r = R(h->circles[3 * o + 2]);
    p->circ[o][0] = R(h->circles[3 * o]); p->circ[o][1] = R(h->circles[3 * o + 1]); p->circ[o][2] = r * r;
  }
  for (int o = p->n_circles; o < ((p->n_circles + 3) & ~3) && o < MPPI_MAX_CIRCLES; ++o) {
    p->circ[o][0] = R(0); p->circ[o][1] = R(0); p->circ[o][2] = R(-1);   // padding: d^2 <= -1 never holds
  }
  {  // one bounding circle around every wall: a single test per pose culls the whole table
    double cx = 0, cy = 0, Rc = 0;
    for (int b = 0; b < p->n_blocks; ++b) { cx += h->blocks[5 * b] / p->n_blocks; cy += h->blocks[5 * b + 1] / p->n_blocks; }
    for (int b = 0; b < p->n_blocks; ++b) {
      const double* bl = &h->blocks[5 * b];
      Rc = std::max(Rc, std::hypot(bl[0] - cx, bl[1] - cy) + std::hypot(bl[3], bl[4]));
    }
    const double Rv = std::hypot(c.veh_L / 2, c.veh_W / 2);
    p->blk_all[0] = R(cx); p->blk_all[1] = R(cy); p->blk_all[2] = R(2.0 * (Rv + Rc) * (Rv + Rc) * (1.0 + 1e-5)); p->blk_all[3] = R(0);
  }
  for (int b = 0; b < p->n_blocks; ++b) {
    const double* bl = &h->blocks[5 * b];
    p->blk[b][0] = R(bl[0]); p->blk[b][1] = R(bl[1]);
    p->blk[b][2] = R(std::cos(bl[2])); p->blk[b][3] = R(std::sin(bl[2]));
    p->blk[b][4] = R(bl[3]); p->blk[b][5] = R(bl[4]);
    // broad phase: with centre distance d, max(|d.n1|, |d.n2|) >= d/sqrt(2) for either rectangle's
    // orthonormal axes and every projection radius is <= the bounding-circle radius, so
    // d^2 >= 2 (R_veh + R_wall)^2 guarantees a separating axis among BOTH axis pairs (= free
    // under the reference's ConvexCollision semantics).  1e-5 covers the FP32 rounding of d^2.
    const double Rv = std::hypot(c.veh_L / 2, c.veh_W / 2), Rw = std::hypot(bl[3], bl[4]);
    p->blk[b][6] = R(2.0 * (Rv + Rw) * (Rv + Rw) * (1.0 + 1e-5)); p->blk[b][7] = R(0);
  }
  if (p->n_blocks > 4) {   // x-binned candidate masks (StaticParams::wall_mask); reach = radius of the broad-phase circle, with slack for FP32 binning
    const double Rv = std::hypot(c.veh_L / 2, c.veh_W / 2);
    double lo = 1e300, hi = -1e300;
    std::vector<double> reach(p->n_blocks);
    for (int b = 0; b < p->n_blocks; ++b) {
      const double* bl = &h->blocks[5 * b];
      const double Rw = std::hypot(bl[3], bl[4]);
      reach[b] = std::sqrt(2.0 * (Rv + Rw) * (Rv + Rw) * (1.0 + 1e-5)) * (1.0 + 1e-4) + 1e-3;
      lo = std::min(lo, bl[0] - reach[b]); hi = std::max(hi, bl[0] + reach[b]);
    }
    const double w = std::max((hi - lo) / WALL_BINS, 1e-6);
    p->wall_x0 = R(lo); p->wall_inv_w = R(1.0 / w);
    for (int i = 0; i < WALL_BINS; ++i) {
      const double b_lo = lo + (i - 0.01) * w, b_hi = lo + (i + 1.01) * w;   // 1 % overlap: the device bins in FP32
      uint32_t m = 0;
      for (int b = 0; b < p->n_blocks; ++b)
        if (h->blocks[5 * b] + reach[b] >= b_lo && h->blocks[5 * b] - reach[b] <= b_hi) m |= 1u << b;
      p->wall_mask[i] = m;
    }
  }
}

inline void build_strict_params(const mppi_handle* h, StrictParams* p) {
  const mppi_config_t& c = h->cfg;
  memset(p, 0, sizeof(*p));
  p->model = c.model; p->n_x = c.n_x; p->N = c.N; p->K_local = h->K_local; p->k_offset = c.k_offset;
  p->n_circles = int(h->circles.size() / 3); p->n_blocks = int(h->blocks.size() / 5);
  p->dt = float(c.T / c.N);   // setup.jl:53, rounded once
  p->lambda = float(c.lambda);
  for (int i = 0; i < 4; ++i) p->Sigma[i] = float(c.Sigma[i]);
  for (int i = 0; i < 2; ++i) { p->goal[i] = float(c.goal[i]); p->CL[i] = float(c.CL[i]); p->CU[i] = float(c.CU[i]); }
  for (int i = 0; i < MPPI_MAX_NX; ++i) { p->XL[i] = float(c.XL[i]); p->XU[i] = float(c.XU[i]); }
  p->slack_penalty = float(c.slack_penalty); p->obstacle_penalty = float(c.obstacle_penalty); p->terminal_scale = float(c.terminal_scale);
  p->la = float(c.la); p->lb = float(c.lb); p->mass = float(c.mass); p->Izz = float(c.Izz); p->g = float(c.g); p->mu = float(c.mu);
  p->KFZF = float(c.KFZF); p->KFZR = float(c.KFZR); p->KFZX = float(c.KFZX);
  for (int i = 0; i < 6; ++i) { p->tire[i] = float(c.tire[i]); p->w[i] = float(c.stage_w[i]); }
  p->veh_L = float(c.veh_L); p->veh_W = float(c.veh_W); p->veh_off = float(c.veh_center_offset);
  for (int o = 0; o < p->n_circles; ++o) for (int q = 0; q < 3; ++q) p->circles[o][q] = float(h->circles[3 * o + q]);
  for (int b = 0; b < p->n_blocks; ++b) for (int q = 0; q < 5; ++q) p->blocks[b][q] = float(h->blocks[5 * b + q]);
}

inline void rebuild_params(mppi_handle* h) {
  build_params(h, &h->pf);
  build_params(h, &h->pd);
  if (h->cfg.strict_math) build_strict_params(h, &h->sp);
}

inline void drop_graphs(mppi_handle* h) {
  for (auto& kv : h->graphs) cudaGraphExecDestroy(kv.second);
  h->graphs.clear();
}

template <class R>
Buffers<R> make_buffers(const mppi_handle* h) {
  Buffers<R> b;
  const int B = h->cfg.B;
  b.x0 = h->d_in;
  b.U = h->d_in + size_t(B) * MPPI_MAX_NX;
  b.eps = (R*)h->d_eps; b.costs = (R*)h->d_costs; b.flags = h->d_flags;
  b.scratch = h->d_scratch; b.iter = h->d_iter;
  b.Spart = (R*)h->d_Spart; b.eta_part = (R*)h->d_eta_part;
  b.partial = h->d_partial;
  b.gathered = h->d_gathered;
  b.res_scalars = h->d_res;
  b.res_control = h->d_res + size_t(B) * RES_SCALARS;
  b.res_traj = h->d_traj;
  b.cand = h->d_cand;
  b.dbg = h->d_dbg;
  b.host_res = nullptr;
  b.seq = h->d_seq;
  b.host_seq = h->h_seq;
  b.ksplit = h->ksplit;
  b.n_ranks = h->n_ranks;
  b.rank = h->rank;
  b.peers = h->p2p ? h->d_peers : nullptr;
  for (int g = 0; g < PEER_MAX_RANKS; ++g) b.peer_w[g] = h->p2p ? h->peer_ptr[g] : nullptr;
  b.seq_base = h->seq_base;
  b.trace = h->profiling == 2 ? 1 : 0;
  b.halt = nullptr;
  b.eps_stage = 0;
  b.ref = static_cast<const R*>(h->d_ref);
  b.cta_part = (R*)h->d_cta_part;
  b.n_cta = h->scan_ctas;
  b.fold_off = 0;
  b.ticket = h->d_ticket;
  b.tail_in_kernel = 0;
  return b;
}

inline size_t real_size(const mppi_handle* h) { return h->cfg.precision == MPPI_FP64 ? 8 : 4; }

}  // namespace mppi
