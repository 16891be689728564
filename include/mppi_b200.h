/*
 * mppi_b200.h — C ABI of libmppi_b200.so, a B200-native (sm_100a) MPPI engine.
 *
 * Drop-in boundary for the MPPI step of congkaishen/MotionPlanning
 * (OptimalControl/MPPI).  The reference has no FFI; its boundary is the Julia
 * function surface of OptimalControl/MPPI/src/setup.jl and MPPIUtils.jl.  Each
 * entry point below names the reference interface (file:line, relative to the
 * reference root) it replaces.  A Julia `ccall` shim over exactly these symbols
 * is shown in INTEGRATION.md (julia/MPPIB200.jl); a ctypes twin lives in
 * motionplanning_b200/_capi.py.
 *
 * Conventions
 *   - every function returns int32_t status (MPPI_OK == 0) unless stated;
 *   - never throws / aborts across the ABI; mppi_last_error() gives the text;
 *   - all host-facing arrays are `double`, column-major (== Julia Matrix{Float64});
 *     the one exception is the big noise tensor, whose element type follows
 *     cfg.precision and whose shape is Julia (m, K, N) == C [N][K][m];
 *   - the caller owns every buffer it passes, the library copies during the call
 *     and keeps no host pointer; the library owns the handle and device memory;
 *   - a handle is not thread-safe; distinct handles are independent.
 */
#ifndef MPPI_B200_H
#define MPPI_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPPI_B200_VERSION 200 /* 0.2.0 */
#define MPPI_MAX_GPUS 8
#define MPPI_MAX_NX 8
#define MPPI_MAX_CIRCLES 64
#define MPPI_MAX_BLOCKS 32
#define MPPI_MAX_REF 1024 /* reference-path points (mppi_set_reference_path) */
#define MPPI_MAX_N 256

/* status codes (the reference's only error path is error(...) in setup.jl:19-39) */
enum {
  MPPI_OK = 0,
  MPPI_EINVAL = 1, /* bad sizes/values: setup.jl:19-39 checks + K/N/device ranges */
  MPPI_ECUDA = 2,  /* CUDA runtime failure */
  MPPI_ENCCL = 3,  /* NCCL failure / NCCL not loadable */
  MPPI_ENOMEM = 4,
  MPPI_ESTATE = 5  /* call order violated (e.g. get_costs before a step) */
};

/* vehicle model = which vehicledynamics.jl the reference would `include`
 * (MPPIUtils.jl:3) */
enum {
  MPPI_MODEL_DYNAMIC7 = 0,  /* OptimalControl/MPPI/src/vehicledynamics.jl:1-54 */
  MPPI_MODEL_KINEMATIC5 = 1 /* OptimalControl/Parking/src/VehicleModel.jl:1-21 */
};
enum { MPPI_FP32 = 0, MPPI_FP64 = 1 };
enum {
  MPPI_NOISE_SUPPLIED = 0, /* deterministic verification: host passes eps */
  MPPI_NOISE_PHILOX = 1    /* in-kernel Philox4x32-10 + Box-Muller + chol(Sigma) */
};
/* planner = how the K rollouts are turned into a control:
 *   MPPI : softmin-weighted mean of the clamped sampled controls (MPPIUtils.jl:154-203)
 *   DWA  : constant control per candidate from a grid, arg-min selection
 *          (OptimalControl/DynamicWindow/src/DWAUtils.jl:141-163) */
enum { MPPI_PLANNER_MPPI = 0, MPPI_PLANNER_DWA = 1 };

/* per-sample flag bits returned by mppi_get_flags */
enum {
  MPPI_FLAG_FEASIBLE = 1, /* all(constraint_his .== 1), MPPIUtils.jl:55 */
  MPPI_FLAG_COLLIDED = 2, /* any ObstacleEvaluation hit, MPPIUtils.jl:125-126 / SAT */
  MPPI_FLAG_OOB = 4       /* any BoundEvaluation hit, MPPIUtils.jl:139-148 */
};

/*
 * POD configuration.  Fields mirror MPPISetting (types.jl:10-31) and the
 * positional arguments of defineMPPI (setup.jl:3-17); the constants the
 * reference hard-codes inside VehicleDynamics (vehicledynamics.jl:2-18,52),
 * ObstacleEvaluation (MPPIUtils.jl:127) and TrajectoryRollout (MPPIUtils.jl:54)
 * are explicit fields whose defaults (mppi_config_default) are those literals.
 */
typedef struct mppi_config {
  uint32_t struct_size; /* = sizeof(mppi_config_t); ABI versioning */
  int32_t model;        /* MPPI_MODEL_* */
  int32_t n_x;          /* numStates: 7 (dynamic) or 5 (kinematic) */
  int32_t m;            /* numControls: 2 (sr, ax) */
  int32_t K;            /* SamplingNumber, global over all ranks */
  int32_t N;            /* horizon steps */
  int32_t B;            /* independent instances (batched mode); 1 otherwise */
  int32_t precision;    /* MPPI_FP32 | MPPI_FP64 */
  int32_t noise_mode;   /* MPPI_NOISE_* */
  int32_t planner;      /* MPPI_PLANNER_* */
  int32_t feasibility_count; /* types.jl:24 (1300); >= K disables the cut-off */
  int32_t use_cuda_graph;    /* 1: one cudaGraphLaunch per step */
  int32_t device;            /* CUDA device ordinal of this handle */
  int32_t k_offset;          /* first global sample id owned by this handle */
  int32_t K_local;           /* samples owned by this handle (0 => K) */
  int32_t rollout_variant;   /* 0 auto | 1 one sample per thread | 2 two samples per thread (packed FP32x2; FP32 and
                              * even K_local only) | 3, 4 = 1, 2 with the Philox noise always drawn inside the
                              * rollout | 5 scan-form fused step | 6 two-stage warp pipeline (kinematic model, packed; what
                              * auto picks wherever it packs and the scene has circles only) | 7 = 6 with the softmin update
                              * done by the rollout's own blocks (no update kernel; no cut-off; the device-timed step is
                              * shorter, the result reaches the host later: for device-chained loops).  0-2, 5-7 pre-generate the
                              * next iteration's noise beside finalize whenever the step is a single-instance Philox MPPI
                              * step.
                              * Environment MPPI_ROLLOUT_VARIANT overrides; MPPI_PREGEN=0 disables pre-generation. */
  int32_t strict_math;       /* FP32 only.  1: the K-sample rollout runs the reference's operation sequence in IEEE
                              * round-to-nearest arithmetic without fused multiply-add and with the elementary
                              * functions of include/mppi_strict_math.h, so that per-sample costs, flags and the
                              * argmin are BIT-IDENTICAL to a float restatement of the reference using the same
                              * functions (tests/test_gpu_strict.py).  Verification flavour: ~10x slower. */
  int32_t noise_policy;      /* Philox steps, B == 1.  0 auto | 1 "latency": the NEXT iteration's noise tensor is drawn
                              * by noise_kernel beside finalize and the rollout reads it back (the result arrives
                              * earlier; +1 HBM pass over eps) | 2 "throughput": noise drawn inside the rollout
                              * (no extra pass).  auto = latency up to 131,072 samples per GPU, throughput beyond. */
  int32_t n_gpus;            /* 0 / 1: this handle drives cfg.device only.  2..8: ONE handle, ONE host thread, samples
                              * sharded over device_ids[0..n_gpus) (the reference's caller is a single Julia task,
                              * main.jl:55-83); results are read from device_ids[0].  Needs peer access between the
                              * devices (NVLink / NVSwitch).  B = 1, MPPI planner.  Every entry point below works on such
                              * a handle as on a one-GPU handle (costs / flags / noise / SUPPLIED eps are the full K-sample
                              * arrays) except the closed loop, the batched step and the split-phase / comm / peer calls,
                              * which return MPPI_EINVAL.  An ordinal may be listed more than once in device_ids: those
                              * shards then share that GPU (logical shards; what the 1-GPU tests run). */
  int32_t device_ids[MPPI_MAX_GPUS];
  int32_t reserved0;
  uint64_t seed;             /* Philox key */
  double T;                  /* horizon seconds; dt = T/N (setup.jl:53) */
  double lambda;             /* types.jl:28 */
  double Sigma[4];           /* m x m covariance, column-major (setup.jl:56-57) */
  double X0[MPPI_MAX_NX];
  double goal[2];
  double XL[MPPI_MAX_NX], XU[MPPI_MAX_NX];
  double CL[2], CU[2];
  double slack_penalty;    /* types.jl:30 = 1e5 */
  double obstacle_penalty; /* MPPIUtils.jl:127 = 100*712.5 (DWA: 10000*712.5) */
  double terminal_scale;   /* MPPIUtils.jl:54 = 1e4 */
  /* vehicledynamics.jl:2-11 */
  double la, lb, mass, Izz, g, mu, KFZF, KFZR, KFZX;
  double tire[6];
  double stage_w[6]; /* weights of v^2, r^2, ax^2, sr^2, sa^2, y^2 (:52) */
  /* vehicle rectangle for SAT (hybrid_astar_utils.jl:193, main_hybrid_astar.jl:21) */
  double veh_L, veh_W, veh_center_offset;
} mppi_config_t;

/* result of one step == MPPIResult (types.jl:36-44) + diagnostics.
 * Control / Traj point into caller-allocated storage (may be NULL to skip). */
typedef struct mppi_result {
  double* Control;        /* N x m col-major  = r.Control  (MPPIUtils.jl:194) */
  double* Traj;           /* (N+1) x n_x col-major = r.Traj (MPPIUtils.jl:193) */
  double cost;            /* r.cost of the nominal rollout (:196) */
  int32_t feasible;       /* r.Feasibility == :Feasible (:195) */
  int32_t rollout_count;  /* r.RolloutCount = R + 1 (off-by-one kept, :171,182,197) */
  int32_t feasible_count; /* r.FeasibleTrajCount (:198) */
  int32_t argmin_index;   /* 1-based global index of the min-cost sample (:155) */
  double rho;             /* min cost (:156) */
  double eta;             /* sum exp(-(c-rho)/lambda) (:159-161) */
  double time_s;          /* host wall time of the call = r.time (:173,199-200) */
} mppi_result_t;

/* One rank's contribution to the softmin combine (multi-GPU sample sharding):
 * local min cost and its global index, local eta and local weighted control sum,
 * both taken against the LOCAL rho; see DESIGN.md "multi-GPU". */
typedef struct mppi_partial {
  double rho;
  double eta;
  int64_t argmin_global; /* 0-based global sample id */
  int32_t rollouts;      /* samples rolled out on this rank */
  int32_t feasible;      /* feasible samples on this rank */
  double S[2 * MPPI_MAX_N]; /* sum_k w_k * u_k, N x m col-major, unnormalised */
} mppi_partial_t;

/* fill cfg with the reference literals (main.jl:7-27, vehicledynamics.jl, types.jl) */
int32_t mppi_config_default(mppi_config_t* cfg);
int32_t mppi_version(void);

/* defineMPPI (setup.jl:3-59): validates like setup.jl:19-39, allocates device state */
typedef struct mppi_handle mppi_handle_t;
int32_t mppi_create(const mppi_config_t* cfg, mppi_handle_t** out);
int32_t mppi_destroy(mppi_handle_t* h);
/* error text of the last failing call on h (h == NULL: last create failure) */
const char* mppi_last_error(const mppi_handle_t* h);

/* ShiftInitialCondition (MPPIUtils.jl:24-27); x0 is n x B col-major */
int32_t mppi_set_x0(mppi_handle_t* h, const double* x0, int32_t n);
/* defineMPPINominalControl!(mppi, U) (setup.jl:70-73); U is N x m (x B) col-major */
int32_t mppi_set_nominal(mppi_handle_t* h, const double* U, int32_t N, int32_t m);
/* defineMPPINominalControl!(mppi, U, Sigma) second form (setup.jl:74-78) + lambda */
int32_t mppi_set_sigma_lambda(mppi_handle_t* h, const double* Sigma, double lambda);
/* defineMPPIobs! (setup.jl:61-64); xyr is 3 x M_c col-major: [x, y, r] per circle */
int32_t mppi_set_circles(mppi_handle_t* h, const double* xyr, int32_t M_c);
/* polygon walls, block = [cx, cy, psi, half_len, half_wid] (CollisionDetection
 * utils.jl:14-25); 5 x M_p col-major */
int32_t mppi_set_blocks(mppi_handle_t* h, const double* blocks, int32_t M_p);
/* set goal (MPPISetting.goal) without re-creating the handle */
int32_t mppi_set_goal(mppi_handle_t* h, const double* goal2);
/* DWA candidate grid (DynamicWindow/src/setup.jl:59-85): K x m col-major constant
 * controls; only for planner == MPPI_PLANNER_DWA */
int32_t mppi_set_candidates(mppi_handle_t* h, const double* ctrl, int32_t K, int32_t m);

/* MPPIPlan (MPPIUtils.jl:169-203).  noise == NULL in PHILOX mode; in SUPPLIED
 * mode it is the host eps tensor [N][K_local][m] of float (FP32) or double (FP64).
 * Synchronous: on return `out` and the caller's Control/Traj are filled. */
int32_t mppi_step(mppi_handle_t* h, const void* noise, mppi_result_t* out);

/* main.jl:60-62 without the host round trip: U_nom <- last U_new, on device */
int32_t mppi_adopt_nominal(mppi_handle_t* h);
/* device-resident iteration for closed loops / benchmarking: launches one
 * iteration (adopting the previous result as nominal when adopt != 0) and does
 * NOT synchronise or copy results; pair with mppi_sync / mppi_fetch_result */
int32_t mppi_step_async(mppi_handle_t* h, int32_t adopt);
int32_t mppi_sync(mppi_handle_t* h);
int32_t mppi_fetch_result(mppi_handle_t* h, mppi_result_t* out);

/* per-sample outputs of the last step (TrajectoryCollection[i].cost / .Feasibility) */
int32_t mppi_get_costs(mppi_handle_t* h, double* costs /* K_local */);
int32_t mppi_get_flags(mppi_handle_t* h, uint8_t* flags /* K_local */);
/* dump the eps actually used by the last step ([N][K_local][m], element type by
 * precision) so the oracle can replay device-generated Philox noise */
int32_t mppi_get_noise(mppi_handle_t* h, void* eps);
/* plot support (MPPIUtils.jl:221-223): xy of every `stride`-th sampled trajectory,
 * float [n_samples][N+1][2]; re-rolls those samples on device */
int32_t mppi_get_sample_xy(mppi_handle_t* h, int32_t stride, float* xy, int32_t* n_samples);

/* ---- reference-path tracking term (an extension) -------
 * The reference's MPPI tracks the implicit lane y = 0 (`10*y^2`, vehicledynamics.jl:52, stage_w[5]).
 * This adds weight * (squared distance to the nearest point of a polyline) to every stage cost,
 * the nearest point searched in a window [idx, idx + window] that moves forward with the vehicle,
 * as the reference's tracker does (HybridAstar/src/tracker_utils.jl:36-41, main_Tracker.jl:82-83).
 * xy: 2 x n col-major, n <= MPPI_MAX_REF; n = 0 removes the term.  MPPI planner: PHILOX noise only. */
int32_t mppi_set_reference_path(mppi_handle_t* h, const double* xy, int32_t n, double weight, int32_t window);

/* ---- closed loop on the device ------------------------
 * The driver loop of OptimalControl/MPPI/main.jl:45-47,58-83 without host round trips: every MPC
 * iteration (ShiftInitialCondition + defineMPPINominalControl! + MPPIPlan, :59-62) is followed on
 * the same stream by a plant kernel that holds U_new with zero order (Constant{Previous}
 * interpolation of U_new over time_serial = range(0, T, N) sampled at range(0, update_time,
 * plant_steps), :45-46,65-67) and integrates VehicleDynamics with explicit Euler for plant_steps
 * steps of plant_dt in FP64 (:76-78), appends [t; states] to the history (:79), stops the episode
 * inside goal_radius of the goal (:80-82) and hands the new state to the next iteration as X0. */
typedef struct mppi_closed_loop {
  uint32_t struct_size;
  int32_t max_iters;     /* MPC iterations; main.jl:58 runs floor(15/dt_plant)/plant_steps = 150 */
  int32_t plant_steps;   /* update_idx = floor(update_time / plant_dt) = 100 (main.jl:20) */
  int32_t shift_nominal; /* 0: NominalControls = mppi.r.Control as the reference does (:62);
                            k > 0: conventional warm start, U_nom[j] = U_new[min(j + k, N - 1)] */
  double plant_dt;       /* dt of the plant, 1e-3 (main.jl:19) */
  double update_time;    /* 0.1 (main.jl:18) */
  double goal_radius;    /* 6 (main.jl:80) */
} mppi_closed_loop_t;
int32_t mppi_closed_loop_default(mppi_closed_loop_t* opt);
/* Runs from the handle's current X0 / nominal.  states_his: (1 + n_x) x (1 + max_iters * plant_steps)
 * col-major, column = [t; states] = one row of MPPITrajectory.csv (main.jl:87), first column
 * [0; X0] (:48); *n_cols = columns written.  iter_log (may be NULL): 4 x max_iters col-major
 * {nominal cost, feasible, RolloutCount, rho} per iteration; *n_iters = iterations that ran. */
int32_t mppi_run_closed_loop(mppi_handle_t* h, const mppi_closed_loop_t* opt, double* states_his,
                             int32_t* n_cols, double* iter_log, int32_t* n_iters);

/* batched multi-instance step (config 5): x0s n_x x B, Us N x m x B in,
 * Us_out N x m x B, costs_out B (nominal rollout cost), feasible_out B */
int32_t mppi_step_batched(mppi_handle_t* h, const double* x0s, const double* Us,
                          double* Us_out, double* costs_out, int32_t* feasible_out);

/* ---- multi-GPU sample sharding (one process per GPU) ------------------------
 * Split-phase form: step_local runs rollout + local softmin partials on this
 * rank's shard; the caller exchanges mppi_partial_t between ranks (any
 * transport); step_combine folds all ranks' partials in rank order and runs the
 * nominal rollout.  With mppi_comm_init the exchange is an in-library NCCL
 * all-gather captured in the step graph and mppi_step does all of it. */
int32_t mppi_step_local(mppi_handle_t* h, const void* noise, mppi_partial_t* partial_out);
int32_t mppi_step_combine(mppi_handle_t* h, const mppi_partial_t* partials, int32_t n_ranks,
                          mppi_result_t* out);
#define MPPI_COMM_ID_BYTES 128
int32_t mppi_comm_unique_id(uint8_t id[MPPI_COMM_ID_BYTES]);
int32_t mppi_comm_init(mppi_handle_t* h, const uint8_t id[MPPI_COMM_ID_BYTES], int32_t rank,
                       int32_t n_ranks);

/* Peer-memory exchange (preferred on NVLink boxes): every rank exports a CUDA-IPC handle of its
 * exchange window, the handles are all-gathered by any transport, and mppi_peer_init maps the
 * peers' windows.  mppi_step then does the softmin all-gather INSIDE the finalize kernel with
 * NVLink stores + flags (one store latency; no collective library on the data path).  Ranks
 * must call mppi_step in lockstep, as with any collective; a rank that waits > 2 s fails the
 * step with MPPI_ENCCL.  The sequential feasibility cut-off (MPPIUtils.jl:175,185; feasibility_count < K)
 * works across ranks on this transport and on n_gpus handles: the per-rank feasible counts are all-gathered
 * through the same windows between rollout and update, and every rank derives its local cut index. */
#define MPPI_PEER_HANDLE_BYTES 64
int32_t mppi_peer_export(mppi_handle_t* h, uint8_t handle[MPPI_PEER_HANDLE_BYTES]);
int32_t mppi_peer_init(mppi_handle_t* h, const uint8_t* handles /* n_ranks x 64, rank order */,
                       int32_t rank, int32_t n_ranks);

/* timings of the last step in microseconds (CUDA events on the handle's stream):
 * [0] rollout, [1] softmin+update, [2] combine+nominal rollout, [3] whole device step.
 * [0..2] need mppi_set_profiling(h, 1).  [3] under graph replay: the two timestamp operations that bracket a replay are
 * recorded only from the first mppi_get_timings / mppi_set_profiling call on (they cost ~2 us of every step), so the first
 * call returns 0 in [3] and arms them for the steps that follow. */
int32_t mppi_get_timings(mppi_handle_t* h, float us[4]);
/* the CUDA stream (cudaStream_t) every step of this handle is enqueued on, so that callers can
 * order their own work or record their own events against it */
int32_t mppi_get_stream(mppi_handle_t* h, void** stream);
/* SM-clock stamps (clock64) at the phase boundaries of the last finalize kernel, instance 0:
 * [0] entry, [1] partials reduced, [2] ranks combined, [3] U_new ready, [4] nominal state
 * recurrence done, [5] states costed, [6] result written.  [7] (profiling mode 2, scenes with walls): number of
 * (warp-step, wall) pairs of the last rollout that ran the EXACT rectangle test, i.e. that the warp-voted broad
 * phase did not cull; divided by warps x N x M_p it is the share of the SAT work actually executed.  Diagnostics only. */
int32_t mppi_get_phase_clocks(mppi_handle_t* h, uint64_t clocks[16]);
/* per-kernel timing: 0 off; 1 eager launches bracketed by CUDA events (no graph replay);
 * 2 graph replay in which every kernel stamps %globaltimer (ns) at its first CTA's start and last
 * CTA's end into the phase-clock block: clocks[8+2i], clocks[9+2i] for i = rollout, update,
 * finalize.  Re-arm with mppi_set_profiling(h, 2) before each traced step. */
int32_t mppi_set_profiling(mppi_handle_t* h, int32_t on);
/* which kernels the next single-instance step of this handle runs (diagnostics, tests, bench):
 * [0] 1 = scan-form fused step (rollout + softmin + update + tail in one kernel), 0 = rollout / update / finalize kernels
 * [1] 1 = the next iteration's Philox noise is pre-generated behind the step ("latency" policy)
 * [2] 1 = packed FP32x2 sequential rollout, 2 = packed two-stage warp pipeline (mppi_pipe.cuh), 3 = that pipeline with the softmin update in its blocks   [3] CTAs of the scan kernel   [4] update ksplit   [5] update jsplit
 * [6] strict_math   [7] GPUs driven by this handle */
int32_t mppi_get_plan_info(mppi_handle_t* h, int32_t info[8]);
/* number of kernels the library launched on this handle so far */
int64_t mppi_launch_count(const mppi_handle_t* h);

/* batched pose-vs-obstacle query on device: the block_collision_check of
 * PathPlanning/HybridAstar/src/hybrid_astar_utils.jl:193-201 for n poses
 * (x, y, psi) 3 x n col-major against the handle's circles and blocks;
 * hit[i] = number of obstacles pose i collides with */
int32_t mppi_check_poses(mppi_handle_t* h, const double* poses, int32_t n, int32_t* hit);

/* mppi_check_poses without any allocation (no cudaMalloc / cudaFree / allocating library call inside).
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The device staging of
 * poses / hit counts lives in one file-scope static __device__ array (448 KiB, held by every GPU that loads the
 * library until exit), so n is limited to MPPI_STATIC_POSES_MAX: larger n is rejected with MPPI_EINVAL and a
 * message naming the limit.  All calls share that array: calls that overlap on the GPU (different handles /
 * streams / threads) race; sequential calls are safe (the hit region is cleared per call).  Same results as
 * mppi_check_poses. */
#define MPPI_STATIC_POSES_MAX 16384
int32_t mppi_check_poses_static(mppi_handle_t* h, const double* poses, int32_t n, int32_t* hit);

/* FP32 FMA-chain micro-benchmark on this handle's device; returns TFLOP/s
 * (denominator of the rollout kernel's roofline) */
int32_t mppi_measure_fp32_peak(mppi_handle_t* h, double* tflops);

/* Measurement aid (no counterpart in the reference): the per-iteration call sequence of OptimalControl/MPPI/main.jl:59-61 —
 * mppi_set_x0 (ShiftInitialCondition), mppi_set_nominal (defineMPPINominalControl!), mppi_step (MPPIPlan) — made n times from
 * HOST buffers x0 [n_x] and U [N x 2 col-major], each iteration timed with the host's steady clock and NOTHING but those three
 * entry points inside the timed span: what a compiled caller (Julia's ccall) pays, without the interpreter overhead of the
 * Python mirror.  result_us[i] = call start -> mppi_step returned (the plan is in host memory); idle_us[i] = ... -> the
 * handle's stream is idle (mppi_sync; includes the next iteration's pre-generated noise).  If flush_dev is non-null, flush_bytes
 * of device memory at flush_dev are overwritten (and the stream drained) before every timed span, so that the L2 is cold.
 * Single-instance handles only (B = 1). */
int32_t mppi_time_reference_calls(mppi_handle_t* h, const double* x0, const double* U, int32_t n, void* flush_dev, size_t flush_bytes,
                                  double* result_us, double* idle_us);

#ifdef __cplusplus
}
#endif
#endif /* MPPI_B200_H */
