// mppi_pipeline.cuh — enqueue one MPPI iteration on the handle's stream, either
// eagerly (profiling) or under stream capture into a CUDA graph that is
// instantiated once per variant and replayed with a single cudaGraphLaunch.
#pragma once
#include <type_traits>

#include "mppi_handle.cuh"

namespace mppi {

// what a step variant contains (bitmask)
enum {
  V_H2D_IN = 1,      // pinned x0/U_nom -> device
  V_ADOPT = 2,       // U_nom <- previous U_new, device to device (main.jl:60-62)
  V_LOCAL = 4,       // scratch reset, rollout, [cut-off], update
  V_PARTIAL = 8,     // finalize<PH_PARTIAL>
  V_D2H_PARTIAL = 16,
  V_H2D_GATHERED = 32,
  V_NCCL = 64,       // in-library all-gather of the partials
  V_COMBINE = 128,   // finalize<PH_COMBINE> (+ nominal rollout)
  V_D2H_RES = 256,
  V_FUSED_FINAL = 512,  // finalize<PH_PARTIAL|PH_COMBINE> in one kernel (single rank)
  V_H2D_X0 = 1024,      // pinned x0 only -> device
  V_H2D_U = 2048,       // pinned U_nom only -> device
  V_HOST_RES = 4096,    // B == 1: finalize writes the result block to mapped host memory (no D2H node)
  V_XFINAL = 8192,      // finalize<PARTIAL|XCHG|COMBINE>: softmin all-gather in-kernel over NVLink peer memory
  V_PLANT = 16384,      // closed loop: plant kernel behind finalize advances X0 on the device (mppi_plant.cuh)
  V_SHIFT = 32768,      // closed loop: U_nom <- U_new shifted by shift_k knots (instead of V_ADOPT's plain copy)
  // bits 16..19 carry n_ranks (variant_key), bit 24 the tracing flag
  V_PREGEN = 1 << 20,   // rollout reads pre-generated Philox noise; noise_kernel for the NEXT iteration runs beside finalize
  V_NOISE_NOW = 1 << 21,   // ... and the buffer does not hold this iteration's noise yet: draw it ahead of the rollout
  V_SCAN = 1 << 22,        // scan-form fused step (mppi_scan.cuh) instead of rollout + update + finalize
  V_ENQ_MASK = 0xFFFF | V_PREGEN | V_NOISE_NOW | V_SCAN
};

template <class R> const StaticParams<R>& params_of(const mppi_handle* h);
template <> inline const StaticParams<float>& params_of<float>(const mppi_handle* h) { return h->pf; }
template <> inline const StaticParams<double>& params_of<double>(const mppi_handle* h) { return h->pd; }

// Launch `kernel`, optionally as a programmatic dependent of the kernel before it on the stream
// (pdl_wait / pdl_trigger in mppi_device.cuh).
template <class P, class Bf>
cudaError_t launch_step_kernel(void (*kernel)(P, Bf), dim3 grid, dim3 block, size_t smem, cudaStream_t s, bool dependent,
                               const P& p, const Bf& buf) {
  cudaLaunchConfig_t lc = {};
  lc.gridDim = grid; lc.blockDim = block; lc.dynamicSmemBytes = smem; lc.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  lc.attrs = at; lc.numAttrs = dependent ? 1 : 0;
  return cudaLaunchKernelEx(&lc, kernel, p, buf);
}

// a kernel whose blocks each stage tens of KB wants the SM's shared-memory carve-out at its maximum (once per instantiation)
template <class K>
void prefer_shared_carveout(K kernel) { cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared); }
// true the first time it is called for this call site's `mask` on the current device (function attributes are per device)
inline bool first_on_device(unsigned long long& mask) {
  int dev = 0; cudaGetDevice(&dev);
  const unsigned long long bit = 1ull << (dev & 63);
  if (mask & bit) return false;
  mask |= bit;
  return true;
}

// noise_kernel behind the scan kernel on the SAME stream, as its programmatic dependent
template <class R>
cudaError_t launch_noise_dependent(dim3 grid, cudaStream_t s, bool dependent, const StaticParams<R>& p, const Buffers<R>& buf, const unsigned int* iter_ptr) {
  cudaLaunchConfig_t lc = {};
  lc.gridDim = grid; lc.blockDim = dim3(NOISE_BS); lc.dynamicSmemBytes = 0; lc.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  lc.attrs = at; lc.numAttrs = dependent ? 1 : 0;
  return cudaLaunchKernelEx(&lc, noise_kernel<R>, p, buf, iter_ptr, 0);
}

// long horizons need more than the default 48 KB of dynamic shared memory: opt in once per instantiation
template <class R, class Model, int SRC, int PHASES, int BS>
cudaError_t launch_finalize_bs(int B, size_t smem, cudaStream_t s, bool dependent, const StaticParams<R>& p, const Buffers<R>& buf) {
  static size_t allowed_dev[64];   // function attributes are per device: one process may hold handles on several
  int dev = 0; cudaGetDevice(&dev);
  size_t& allowed = allowed_dev[dev & 63];
  if (allowed == 0) allowed = 48 * 1024;
  if (smem > allowed) {
    cudaError_t e = cudaFuncSetAttribute(finalize_kernel<R, Model, SRC, PHASES, BS>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e != cudaSuccess) return e;
    allowed = smem;
  }
  return launch_step_kernel(finalize_kernel<R, Model, SRC, PHASES, BS>, dim3(B), dim3(BS), smem, s, dependent, p, buf);
}
template <class R, class Model, int SRC, int PHASES>
cudaError_t launch_finalize(int B, size_t smem, cudaStream_t s, bool dependent, const StaticParams<R>& p, const Buffers<R>& buf) {
  if ((PHASES & PH_XCHG) == 0 && B >= FINAL_BATCH_MIN_B) return launch_finalize_bs<R, Model, SRC, PHASES, FINAL_BS_BATCH>(B, smem, s, dependent, p, buf);
  return launch_finalize_bs<R, Model, SRC, PHASES, FINAL_BS>(B, smem, s, dependent, p, buf);
}

// The scan-form fused step (mppi_scan.cuh): chunk length C covers the horizon with SCAN_L lanes per sample, NC = circles held
// in the unrolled test.  `tail`: the last CTA runs fold / exchange / combine / nominal rollout inside the kernel.
template <int C, int NC>
cudaError_t launch_scan_cn(dim3 grid, size_t smem, cudaStream_t s, bool dependent, const StaticParams<float>& p, const Buffers<float>& buf) {
  auto kernel = scan_rollout_kernel<C, NC, PH_FOLD | PH_XCHG | PH_COMBINE>;
  static size_t allowed_dev[64];
  int dev = 0; cudaGetDevice(&dev);
  size_t& allowed = allowed_dev[dev & 63];
  if (allowed == 0) allowed = 48 * 1024;
  if (smem > allowed) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e != cudaSuccess) return e;
    allowed = smem;
  }
  return launch_step_kernel(kernel, grid, dim3(SCAN_BS), smem, s, dependent, p, buf);
}
inline int scan_chunk(int N) { return N <= 7 * SCAN_L ? 7 : (N <= 13 * SCAN_L ? 13 : 0); }
inline cudaError_t launch_scan(int N, int n_circles, dim3 grid, size_t smem, cudaStream_t s, bool dependent, const StaticParams<float>& p, const Buffers<float>& buf) {
  const int C = scan_chunk(N);
  if (C == 7) return n_circles <= 3 ? launch_scan_cn<7, 3>(grid, smem, s, dependent, p, buf) : launch_scan_cn<7, 8>(grid, smem, s, dependent, p, buf);
  return n_circles <= 3 ? launch_scan_cn<13, 3>(grid, smem, s, dependent, p, buf) : launch_scan_cn<13, 8>(grid, smem, s, dependent, p, buf);
}

// the two-stage warp pipeline (mppi_pipe.cuh) runs this step's rollout when the handle asked for it and the scene has circles only
inline bool pipe_eligible(const mppi_handle* h) {
  return h->pipe && h->ref_path.empty() && h->blocks.empty() && h->circles.size() / 3 <= size_t(HOT_CIRCLES);
}
// ... and its block also does its share of the softmin update (rollout_pipe_kernel<., true>), so the step has no update_kernel: no
// cut-off (the cut needs every cost before any weight), one column of cta_part per block
inline bool pipe_folds(const mppi_handle* h) {
  return pipe_eligible(h) && h->pipe_fold && !h->cutoff && (h->K_local + PIPE_SAMPLES - 1) / PIPE_SAMPLES <= FOLD_MAX_CTAS;
}
template <int NQ, bool FOLD>
cudaError_t launch_pipe_nq(dim3 grid, size_t smem, cudaStream_t s, bool dependent, const StaticParams<float>& p, const Buffers<float>& buf) {
  static unsigned long long prepared = 0;
  if (first_on_device(prepared)) {
    prefer_shared_carveout(rollout_pipe_kernel<NQ, FOLD>);
    cudaFuncSetAttribute(rollout_pipe_kernel<NQ, FOLD>, cudaFuncAttributeMaxDynamicSharedMemorySize, PIPE_SLAB_MAX + 8192);
  }
  return launch_step_kernel(rollout_pipe_kernel<NQ, FOLD>, grid, dim3(PIPE_BS), smem, s, dependent, p, buf);
}
inline cudaError_t launch_pipe(int n_circles, bool fold, dim3 grid, size_t smem, cudaStream_t s, bool dependent, const StaticParams<float>& p, const Buffers<float>& buf) {
  if (fold) return n_circles <= 3 ? launch_pipe_nq<3, true>(grid, smem, s, dependent, p, buf) : launch_pipe_nq<8, true>(grid, smem, s, dependent, p, buf);
  return n_circles <= 3 ? launch_pipe_nq<3, false>(grid, smem, s, dependent, p, buf) : launch_pipe_nq<8, false>(grid, smem, s, dependent, p, buf);
}

template <class R, class Model, int SRC>
int enqueue_variant(mppi_handle* h, int v, int timed) {   // timed: 1 = all events, 2 = inner boundaries only (graph nodes)
  const StaticParams<R>& p = params_of<R>(h);
  Buffers<R> buf = make_buffers<R>(h);
  if (v & V_FUSED_FINAL) { buf.gathered = buf.partial; buf.n_ranks = 1; }   // single rank: combine reads its own partial
  const int B = h->cfg.B, N = h->cfg.N, K = h->K_local, nx = h->cfg.n_x;
  cudaStream_t s = h->stream;
  if ((v & V_HOST_RES) && B == 1) buf.host_res = reinterpret_cast<uint4*>(h->h_ll);
  if (v & V_PLANT) buf.halt = h->d_pstate + 1;
  const size_t in_bytes = size_t(B) * (MPPI_MAX_NX + 2 * N) * sizeof(double);
  const size_t res_bytes = size_t(B) * (RES_SCALARS + 2 * N) * sizeof(double);
  const size_t smemU = (size_t(N) * 4 + size_t(8) * p.n_blocks + size_t(2) * p.n_ref) * sizeof(R);   // nominal + wall table + reference polyline
  const bool track = p.n_ref > 0;
  // programmatic dependent launches between consecutive kernels of the step (not when events sit between them)
  const bool pdl = h->pdl && !timed && (v & V_LOCAL);
  const bool pdl_final = pdl && SRC != SRC_DWA;   // DWA: finalize follows the rollout directly and reads its scratch before any wait
  if (track && SRC == SRC_SUPPLIED) { h->err = "a reference path needs PHILOX noise (the tracking rollout is not built for supplied noise)"; return MPPI_EINVAL; }

  if (timed == 1) MPPI_CUDA(h, cudaEventRecord(h->ev[0], s));
  if (v & (V_H2D_IN | V_H2D_X0 | V_H2D_U)) {
    const int nx0 = B * MPPI_MAX_NX, nU = B * 2 * N;
    const int first = (v & V_H2D_U) ? nx0 : 0;
    const int count = (v & V_H2D_IN) ? nx0 + nU : ((v & V_H2D_X0) ? nx0 : nU);
    if (B == 1) {   // one small kernel reading the mapped host block: cheaper than a memcpy node
      begin_kernel<<<1, 128, 0, s>>>(h->h_in, h->d_in, first, count);
      h->launches++;
    } else {
      MPPI_CUDA(h, cudaMemcpyAsync(h->d_in + first, h->h_in + first, size_t(count) * sizeof(double), cudaMemcpyHostToDevice, s));
    }
  }
  if (v & V_ADOPT)
    MPPI_CUDA(h, cudaMemcpyAsync(h->d_in + size_t(B) * MPPI_MAX_NX, h->d_res + size_t(B) * RES_SCALARS,
                                 size_t(B) * 2 * N * sizeof(double), cudaMemcpyDeviceToDevice, s));
  if (v & V_SHIFT) {
    shift_nominal_kernel<<<1, 256, 0, s>>>(h->d_in + size_t(B) * MPPI_MAX_NX, h->d_res + size_t(B) * RES_SCALARS, N, h->shift_k);
    h->launches++;
  }
  // the rollout as a programmatic dependent of begin_kernel (nothing else between them on the stream)
  const bool pdl_in = h->pdl && !timed && B == 1 && (v & (V_H2D_IN | V_H2D_X0 | V_H2D_U)) && !(v & (V_ADOPT | V_SHIFT | V_NOISE_NOW)) && (v & V_LOCAL);
  const bool pregen = SRC == SRC_PHILOX && (v & V_PREGEN) && (v & V_LOCAL);
  // A rollout that READS eps (host-supplied or pre-generated) stages its block's slab in shared memory by bulk
  // copies when the slab is small enough to keep the occupancy: N rows of ROLLOUT_BS * LANES samples.
  size_t smemE = 0;
  if ((SRC == SRC_SUPPLIED || pregen) && (v & V_LOCAL)) {
    const int lanes = (h->packed_seq && sizeof(R) == 4) ? 2 : 1;
    const size_t slab = size_t(N) * ROLLOUT_BS * lanes * 2 * sizeof(R);
    const bool aligned = sizeof(R) == 8 || (K % 2) == 0;   // rows start and end on 16-byte boundaries
    if (slab <= size_t(lanes == 2 ? EPS_SLAB_MAX : 32 * 1024) && aligned && N <= EPS_GROUP * EPS_MAX_GROUPS && !getenv("MPPI_NO_EPS_STAGE")) {
      buf.eps_stage = 1;
      smemE = ((smemU + 127) & ~size_t(127)) - smemU + slab;
    }
  }
  const dim3 noise_grid((K + NOISE_BS - 1) / NOISE_BS, (N + 2 * NOISE_PAIRS - 1) / (2 * NOISE_PAIRS), B);
  if (pregen && (v & V_NOISE_NOW)) {
    noise_kernel<R><<<noise_grid, NOISE_BS, 0, s>>>(p, buf, h->d_iter, 0);
    h->launches++;
  }
  if (timed == 1) MPPI_CUDA(h, cudaEventRecord(h->ev[5], s));   // the rollout's interval starts after the inputs are staged
  // ---- scan-form fused step: rollout + cost + online softmin + weighted sum in ONE kernel whose last CTA also runs the tail ----
  bool scan_done = false, scan_tail = false, folded = false;   // folded: the pipeline kernel left per-block softmin partials, no update_kernel
  if constexpr (std::is_same<R, float>::value && std::is_same<Model, Kin5<float>>::value && SRC != SRC_DWA) {
    if ((v & V_LOCAL) && (v & V_SCAN)) {
      const int C = scan_chunk(N);
      const size_t plen_ = PART_HDR + 2 * size_t(N);
      const size_t smemF_ = (RES_SCALARS + 2 * size_t(N) + size_t(N + 1) * 5 + plen_ + size_t(h->n_ranks) * plen_) * sizeof(double) +
                            (size_t(N) * 6 + size_t(N + 1) * (5 + 2 + 3 + 1)) * sizeof(float);
      const size_t smemS = size_t(SCAN_L) * C * sizeof(float4) + size_t(SCAN_NW) * (2 * N + 2) * sizeof(float);
      // the tail runs in-kernel for whole steps (fused / peer-exchange finalize); eager profiling and split-phase steps launch it separately
      scan_tail = !timed && (v & (V_FUSED_FINAL | V_XFINAL)) != 0;
      buf.tail_in_kernel = scan_tail ? 1 : 0;
      buf.n_cta = h->scan_ctas;
      if (v & V_FUSED_FINAL) { buf.n_ranks = 1; buf.peers = nullptr; }
      MPPI_CUDA(h, launch_scan(N, p.n_circles, dim3(h->scan_ctas), std::max(smemS, smemF_), s, pdl_in, p, buf));
      h->launches++;
      if (timed) MPPI_CUDA(h, cudaEventRecord(h->ev[1], s));
      if (timed) MPPI_CUDA(h, cudaEventRecord(h->ev[2], s));
      scan_done = true;
    }
  }
  if ((v & V_LOCAL) && !scan_done) {
    bool launched = pregen, piped = false;
    folded = false;
    if constexpr (std::is_same<R, float>::value && SRC != SRC_DWA) {
      if (h->cfg.strict_math) {   // verification flavour: the reference's operation sequence in unfused IEEE FP32 (mppi_strict.cuh); reads eps
        if (SRC == SRC_PHILOX) {
          noise_kernel<R><<<noise_grid, NOISE_BS, 0, s>>>(p, buf, h->d_iter, 0);
          h->launches++;
        }
        const size_t smemS = (size_t(2) * N + size_t(10) * p.n_blocks) * sizeof(float);
        const dim3 sg((K + strict::BS - 1) / strict::BS);
        if (Model::NX == 7) strict::rollout_strict_kernel<MPPI_MODEL_DYNAMIC7><<<sg, strict::BS, smemS, s>>>(h->sp, buf);
        else strict::rollout_strict_kernel<MPPI_MODEL_KINEMATIC5><<<sg, strict::BS, smemS, s>>>(h->sp, buf);
        launched = true;
      }
    }
    if constexpr (std::is_same<R, float>::value && std::is_same<Model, Kin5<float>>::value && SRC != SRC_DWA) {
      if (!h->cfg.strict_math && (SRC == SRC_SUPPLIED || pregen) && pipe_eligible(h)) {   // two-stage warp pipeline: stage A / stage B warps per sample pair
        const size_t smemP = ((size_t(N) * 16 + 127) & ~size_t(127)) + size_t(N) * PIPE_SAMPLES * 8;
        const dim3 pgrid((K + PIPE_SAMPLES - 1) / PIPE_SAMPLES, B);
        folded = pipe_folds(h);
        if (folded) buf.n_cta = int(pgrid.x);
        MPPI_CUDA(h, launch_pipe(p.n_circles, folded, pgrid, smemP, s, pdl_in, p, buf));
        launched = piped = true;
      }
    }
    if (pregen && !piped) {   // eps read back (the host keeps V_PREGEN off for tracking handles)
      bool done = false;
      if constexpr (std::is_same<R, float>::value) {
        if (h->packed_seq) {
          using PK = typename Rebind<Model, f32x2>::type;
          dim3 grid((K + 2 * ROLLOUT_BS - 1) / (2 * ROLLOUT_BS), B);
          static unsigned long long carved = 0;
          if (first_on_device(carved)) {
            prefer_shared_carveout(rollout_kernel<f32x2, PK, SRC_SUPPLIED>);
            cudaFuncSetAttribute(rollout_kernel<f32x2, PK, SRC_SUPPLIED>, cudaFuncAttributeMaxDynamicSharedMemorySize, EPS_SLAB_MAX + 8192);
          }
          MPPI_CUDA(h, launch_step_kernel(rollout_kernel<f32x2, PK, SRC_SUPPLIED>, grid, dim3(ROLLOUT_BS), smemU + smemE, s, pdl_in, p, buf));
          done = true;
        }
      }
      if (!done) {
        dim3 grid((K + ROLLOUT_BS - 1) / ROLLOUT_BS, B);
        static unsigned long long carved = 0;
        if (first_on_device(carved)) prefer_shared_carveout(rollout_kernel<R, Model, SRC_SUPPLIED>);
        MPPI_CUDA(h, launch_step_kernel(rollout_kernel<R, Model, SRC_SUPPLIED>, grid, dim3(ROLLOUT_BS), smemU + smemE, s, pdl_in, p, buf));
      }
    }
    if constexpr (std::is_same<R, float>::value) {
      if (h->packed_seq && !pregen && !launched) {   // two samples per thread on the packed FP32 pipe (mppi_pack.cuh)
        dim3 grid((K + 2 * ROLLOUT_BS - 1) / (2 * ROLLOUT_BS), B);
        if constexpr (SRC != SRC_SUPPLIED) {
          if (track) rollout_kernel<f32x2, typename Rebind<Model, f32x2>::type, SRC, true><<<grid, ROLLOUT_BS, smemU, s>>>(p, buf);
        }
        if (!track) {
          using PK = typename Rebind<Model, f32x2>::type;
          static unsigned long long carved = 0;
          if (SRC == SRC_SUPPLIED && first_on_device(carved)) {   // host-supplied eps, slab staged in shared memory (up to EPS_SLAB_MAX > 48 KB)
            prefer_shared_carveout(rollout_kernel<f32x2, PK, SRC>);
            cudaFuncSetAttribute(rollout_kernel<f32x2, PK, SRC>, cudaFuncAttributeMaxDynamicSharedMemorySize, EPS_SLAB_MAX + 8192);
          }
          MPPI_CUDA(h, launch_step_kernel(rollout_kernel<f32x2, PK, SRC>, grid, dim3(ROLLOUT_BS), smemU + smemE, s, pdl_in, p, buf));
        }
        launched = true;
      }
    }
    if (!launched) {
      dim3 grid((K + ROLLOUT_BS - 1) / ROLLOUT_BS, B);
      if constexpr (SRC != SRC_SUPPLIED) {
        if (track) rollout_kernel<R, Model, SRC, true><<<grid, ROLLOUT_BS, smemU, s>>>(p, buf);
      }
      if (!track) {
        static unsigned long long carved = 0;
        if (SRC == SRC_SUPPLIED && first_on_device(carved)) prefer_shared_carveout(rollout_kernel<R, Model, SRC>);
        MPPI_CUDA(h, launch_step_kernel(rollout_kernel<R, Model, SRC>, grid, dim3(ROLLOUT_BS), smemU + smemE, s, pdl_in, p, buf));
      }
    }
    h->launches++;
    if (h->cutoff && SRC != SRC_DWA) {
      cutoff_kernel<R><<<B, CUTOFF_BS, 0, s>>>(p, buf);
      h->launches++;
    }
    if (timed) MPPI_CUDA(h, cudaEventRecord(h->ev[1], s));
    if (SRC != SRC_DWA && !folded) {
      const int ngroups = (N + UPDATE_JG - 1) / UPDATE_JG;
      dim3 ug(h->ksplit, h->jsplit, B);
      const size_t smemR = size_t((ngroups + h->jsplit - 1) / h->jsplit) * (UPDATE_BS / 32) * 2 * UPDATE_JG * sizeof(R);
      MPPI_CUDA(h, launch_step_kernel(update_kernel<R>, ug, dim3(UPDATE_BS), smemR, s, pdl, p, buf));   // rollout (or cut-off) -> update
      h->launches++;
    }
    if (pregen) {   // fork: the next iteration's noise may overwrite eps as soon as the update has streamed it
      MPPI_CUDA(h, cudaEventRecord(h->ev_fork, s));
      MPPI_CUDA(h, cudaStreamWaitEvent(h->stream2, h->ev_fork, 0));
      noise_kernel<R><<<noise_grid, NOISE_BS, 0, h->stream2>>>(p, buf, h->d_iter + 1, 0);
      MPPI_CUDA(h, cudaEventRecord(h->ev_join, h->stream2));
      h->launches++;
    }
    if (timed) MPPI_CUDA(h, cudaEventRecord(h->ev[2], s));
  }
  // finalize's shared memory: result block + own partial (+ G gathered partials for the in-kernel
  // exchange) in double, then nominal controls, states and per-step costs in R
  const size_t plen = PART_HDR + 2 * size_t(N);
  const size_t smemF = (RES_SCALARS + 2 * size_t(N) + size_t(N + 1) * Model::NX + plen) * sizeof(double) +
                       (size_t(N) * 6 + size_t(N + 1) * (Model::NX + 2 + 3 + 1) + size_t(2) * p.n_ref) * sizeof(R);
  const size_t smemX = smemF + size_t(h->n_ranks) * plen * sizeof(double);
  if (scan_done || folded) {
    if constexpr (std::is_same<R, float>::value && std::is_same<Model, Kin5<float>>::value && SRC != SRC_DWA) {
      const bool dep = folded && pdl_final;   // behind the pipeline kernel finalize is its programmatic dependent, as it is behind update_kernel
      // the per-block partial table staged in finalize's shared memory by one bulk copy, when it fits behind the block's own data
      const size_t tab = ((size_t(2 * N + 2) * size_t(buf.n_cta) * sizeof(R)) + 15) & ~size_t(15);
      const bool xf = (v & V_XFINAL) != 0;
      const size_t base = ((xf ? smemX : smemF) + 15) & ~size_t(15);
      const bool stage = folded && B == 1 && base + tab <= size_t(200) * 1024;
      if (stage) buf.fold_off = int(base);
      const size_t smF = stage ? base + tab : smemF, smX = stage ? base + tab : smemX;
      if (!scan_tail && (v & V_FUSED_FINAL)) {
        MPPI_CUDA(h, (launch_finalize<R, Model, SRC, PH_FOLD | PH_COMBINE>(B, smF, s, dep, p, buf)));
        h->launches++;
      }
      if (!scan_tail && xf) {
        MPPI_CUDA(h, (launch_finalize<R, Model, SRC, PH_FOLD | PH_XCHG | PH_COMBINE>(B, smX, s, dep, p, buf)));
        h->launches++;
      }
      if (v & V_PARTIAL) {
        MPPI_CUDA(h, (launch_finalize<R, Model, SRC, PH_FOLD>(B, smF, s, dep, p, buf)));
        h->launches++;
      }
      buf.fold_off = 0;
    }
  } else {
  if (v & V_FUSED_FINAL) {
    MPPI_CUDA(h, (launch_finalize<R, Model, SRC, PH_PARTIAL | PH_COMBINE>(B, smemF, s, pdl_final, p, buf)));
    h->launches++;
  }
  if (v & V_XFINAL) {
    MPPI_CUDA(h, (launch_finalize<R, Model, SRC, PH_PARTIAL | PH_XCHG | PH_COMBINE>(B, smemX, s, pdl_final, p, buf)));
    h->launches++;
  }
  if (v & V_PARTIAL) {
    MPPI_CUDA(h, (launch_finalize<R, Model, SRC, PH_PARTIAL>(B, smemF, s, pdl_final, p, buf)));
    h->launches++;
  }
  }
  if (v & V_D2H_PARTIAL)
    MPPI_CUDA(h, cudaMemcpyAsync(h->h_partial, h->d_partial, size_t(B) * h->plen * sizeof(double), cudaMemcpyDeviceToHost, s));
  if (v & V_H2D_GATHERED)
    MPPI_CUDA(h, cudaMemcpyAsync(h->d_gathered, h->h_gathered, size_t(h->n_ranks) * B * h->plen * sizeof(double),
                                 cudaMemcpyHostToDevice, s));
  if (v & V_NCCL) {
    NcclApi& api = nccl_api();
    ncclResult_t r = api.AllGather(h->d_partial, h->d_gathered, size_t(B) * h->plen, ncclFloat64, h->comm, s);
    if (r != ncclSuccess) { h->err = std::string("ncclAllGather: ") + api.GetErrorString(r); return MPPI_ENCCL; }
  }
  if (v & V_COMBINE) {
    MPPI_CUDA(h, (launch_finalize<R, Model, SRC, PH_COMBINE>(B, smemF, s, false, p, buf)));
    h->launches++;
  }
  if (timed) MPPI_CUDA(h, cudaEventRecord(h->ev[3], s));
  if (pregen && scan_done) {
    // next iteration's noise behind the scan kernel: with the tail in-kernel it is that kernel's programmatic dependent, so
    // its blocks take the SMs as the scan CTAs drain and run beside the one-CTA tail
    MPPI_CUDA(h, launch_noise_dependent<R>(noise_grid, s, scan_tail && h->pdl, p, buf, h->d_iter + 1));
    h->launches++;
  }
  if (pregen && !scan_done) MPPI_CUDA(h, cudaStreamWaitEvent(s, h->ev_join, 0));   // join
  if (v & V_PLANT) {
    plant_kernel<typename Rebind<Model, double>::type><<<1, 32, 0, s>>>(h->pd, h->plant);
    h->launches++;
  }
  if (v & V_D2H_RES) {
    const size_t traj_bytes = (B == 1) ? size_t(N + 1) * nx * sizeof(double) : 0;   // contiguous with the scalars
    MPPI_CUDA(h, cudaMemcpyAsync(h->h_res, h->d_res, res_bytes + traj_bytes, cudaMemcpyDeviceToHost, s));
  }
  if (timed == 1) MPPI_CUDA(h, cudaEventRecord(h->ev[4], s));
  MPPI_CUDA(h, cudaGetLastError());
  return MPPI_OK;
}

inline int source_of(const mppi_handle* h) {
  if (h->cfg.planner == MPPI_PLANNER_DWA) return SRC_DWA;
  return h->cfg.noise_mode == MPPI_NOISE_PHILOX ? SRC_PHILOX : SRC_SUPPLIED;
}

template <class R, class Model>
int enqueue_src(mppi_handle* h, int v, int timed) {
  switch (source_of(h)) {
    case SRC_PHILOX: return enqueue_variant<R, Model, SRC_PHILOX>(h, v, timed);
    case SRC_DWA: return enqueue_variant<R, Model, SRC_DWA>(h, v, timed);
    default: return enqueue_variant<R, Model, SRC_SUPPLIED>(h, v, timed);
  }
}
template <class R>
int enqueue_model(mppi_handle* h, int v, int timed) {
#ifndef MPPI_DEV_FAST   // development builds (make FAST=1): FP32 kinematic instantiations only, a quarter of the compile time
  if (h->cfg.model == MPPI_MODEL_DYNAMIC7) return enqueue_src<R, Dyn7<R>>(h, v, timed);
#endif
  return enqueue_src<R, Kin5<R>>(h, v, timed);
}
inline int enqueue(mppi_handle* h, int v, int timed) {
#ifndef MPPI_DEV_FAST
  if (h->cfg.precision == MPPI_FP64) return enqueue_model<double>(h, v, timed);
#endif
  return enqueue_model<float>(h, v, timed);
}

// number of kernels a variant launches (for mppi_launch_count under graph replay)
inline int kernels_in(const mppi_handle* h, int v) {
  int n = 0;
  const bool dwa = h->cfg.planner == MPPI_PLANNER_DWA;
  if ((v & (V_H2D_IN | V_H2D_X0 | V_H2D_U)) && h->cfg.B == 1) n++;
  if ((v & V_LOCAL) && (v & V_SCAN)) {   // scan kernel (tail inside for whole steps) [+ partial-only finalize] [+ next noise]
    n += 1 + ((v & V_PARTIAL) ? 1 : 0) + ((v & V_PREGEN) ? 1 + ((v & V_NOISE_NOW) ? 1 : 0) : 0) + ((v & V_COMBINE) ? 1 : 0) + ((v & V_PLANT) ? 1 : 0) + ((v & V_SHIFT) ? 1 : 0);
    return n;
  }
  if ((v & V_LOCAL) && h->cfg.strict_math && h->cfg.noise_mode == MPPI_NOISE_PHILOX && !dwa) n++;   // strict flavour draws its noise ahead of the rollout
  const bool reads_eps = h->cfg.noise_mode == MPPI_NOISE_SUPPLIED || (v & V_PREGEN);
  const bool folded = (v & V_LOCAL) && !dwa && !h->cfg.strict_math && reads_eps && pipe_folds(h);
  if (v & V_LOCAL) n += 1 + ((h->cutoff && !dwa) ? 1 : 0) + ((dwa || folded) ? 0 : 1);
  if ((v & V_LOCAL) && (v & V_PREGEN)) n += 1 + ((v & V_NOISE_NOW) ? 1 : 0);
  if (v & V_FUSED_FINAL) n++;
  if (v & V_XFINAL) n++;
  if (v & V_PARTIAL) n++;
  if (v & V_COMBINE) n++;
  if (v & V_PLANT) n++;
  if (v & V_SHIFT) n++;
  return n;
}

// Pre-generated noise applies to single-instance Philox MPPI steps with one-sample rollout threads: the
// latency-bound regime, where the rollout's issue slots are the step's critical path.  Large sample
// counts (packed rollout) are throughput-bound: the same instructions cost the same wherever they run.
inline bool scan_eligible(const mppi_handle* h) {   // the scan kernel reads eps from memory: host-supplied, or pre-generated
  if (!h->scan_ok || !h->ref_path.empty() || !h->blocks.empty() || h->circles.size() / 3 > 8) return false;
  return h->cfg.noise_mode == MPPI_NOISE_SUPPLIED || h->pregen_ok;
}
inline int noise_bits(const mppi_handle* h) {
  int v = scan_eligible(h) ? V_SCAN : 0;
  if (!h->pregen_ok || !h->ref_path.empty()) return v;
  return v | V_PREGEN | (h->noise_ready ? 0 : V_NOISE_NOW);
}

// pending host-side input changes -> the H2D node(s) the next step must carry
inline int take_dirty(mppi_handle* h) {
  const int v = (h->x0_dirty && h->u_dirty) ? V_H2D_IN : (h->x0_dirty ? V_H2D_X0 : (h->u_dirty ? V_H2D_U : 0));
  h->x0_dirty = h->u_dirty = false;
  return v;
}

// Wait for the step whose finalize carries sequence number seq_launched.  B == 1: the result block
// arrives in mapped host memory as LL lines, each 8-byte half {value bits | flag} written atomically
// with flag = the step's sequence number, so "every line carries this step's flag" IS completion:
// one PCIe one-way trip after the last store, no fence and no second flag write on the GPU.  The
// stream is checked now and then so that a faulted launch cannot hang the caller.  Otherwise: stream sync.
inline int wait_step(mppi_handle* h, bool flag_ok) {
  if (flag_ok && h->cfg.B == 1) {
    const unsigned long long f = (unsigned long long)(uint32_t(h->seq_launched)) << 32;
    const volatile unsigned long long* L = h->h_ll;
    const int n = h->n_ll;
    unsigned spins = 0;
    bool faulted = false;
    for (int i = n - 1, done = 0; done < n;) {   // last line first (it is stored last), then the rest in order
      unsigned long long w0, w1;
      for (;;) {
        w0 = L[2 * i]; w1 = L[2 * i + 1];
        if ((w0 & 0xFFFFFFFF00000000ull) == f && (w1 & 0xFFFFFFFF00000000ull) == f) break;
        if ((++spins & 0x3FFF) == 0) {
          const cudaError_t e = cudaStreamQuery(h->stream);
          if (e != cudaErrorNotReady && e != cudaSuccess) { h->err = std::string("step failed: ") + cudaGetErrorString(e); return MPPI_ECUDA; }
          if (e == cudaSuccess && (spins & 0x3FFFF) == 0) { faulted = true; break; }   // retired long ago, line never came
        }
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
      }
      if (faulted) break;
      const unsigned long long bits = (w0 & 0xFFFFFFFFull) | (w1 << 32);
      memcpy(&h->h_res[i], &bits, sizeof bits);
      ++done;
      i = (done == 1) ? 0 : i + 1;
    }
    if (!faulted) { h->timings_pending = true; h->inputs_busy = false; return MPPI_OK; }
    h->err = "step retired without delivering its result block"; return MPPI_ECUDA;
  }
  MPPI_CUDA(h, cudaStreamSynchronize(h->stream));
  h->in_flight = h->inputs_busy = false;
  h->timings_pending = true;
  return MPPI_OK;
}

// Capture + instantiate + upload the graph of variant `v` if this handle does not hold it yet.  Split from run_variant because
// a single-process multi-GPU step must do ALL host-side preparation for ALL members before the first member is launched:
// a member's step kernel spins for its peers' partials, and capture / instantiation / first-launch upload of a later member can
// block on device activity (seen: 2 s exchange time-outs when a new variant was captured with earlier members already spinning).
inline bool uses_graph(const mppi_handle* h) { return h->cfg.use_cuda_graph && h->profiling != 1; }
inline int graph_key(const mppi_handle* h, int v) { return h->profiling == 2 ? (v | (1 << 24)) : v; }   // graphs whose kernels trace %globaltimer are cached separately
inline int ensure_graph(mppi_handle* h, int v) {
  if (!uses_graph(h)) return MPPI_OK;
  v = graph_key(h, v);
  if (h->graphs.find(v) != h->graphs.end()) return MPPI_OK;
  cudaGraph_t g = nullptr;
  const int64_t before = h->launches;
  MPPI_CUDA(h, cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
  int rc = enqueue(h, v & V_ENQ_MASK, 0);
  cudaError_t e = cudaStreamEndCapture(h->stream, &g);
  h->launches = before;
  if (rc != MPPI_OK) { if (g) cudaGraphDestroy(g); return rc; }
  if (e != cudaSuccess) { h->err = std::string("cudaStreamEndCapture: ") + cudaGetErrorString(e); return MPPI_ECUDA; }
  cudaGraphExec_t ex = nullptr;
  e = cudaGraphInstantiate(&ex, g, 0);
  cudaGraphDestroy(g);
  if (e != cudaSuccess) { h->err = std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e); return MPPI_ECUDA; }
  h->graphs.emplace(v, ex);
  if (h->spmg_member) {   // pay the first-launch upload now, with no member spinning yet
    MPPI_CUDA(h, cudaGraphUpload(ex, h->stream));
    MPPI_CUDA(h, cudaStreamSynchronize(h->stream));
  }
  return MPPI_OK;
}

// run a variant: graph replay (captured on first use) or eager when profiling / graphs are off
inline int run_variant(mppi_handle* h, int v) {
  if (!uses_graph(h)) {
    h->in_flight = h->inputs_busy = true;
    if (v & (V_COMBINE | V_FUSED_FINAL | V_XFINAL)) h->seq_launched++;
    h->noise_ready = (v & V_PREGEN) != 0;
    return enqueue(h, v & V_ENQ_MASK, h->profiling ? 1 : 0);
  }
  { const int rc = ensure_graph(h, v); if (rc != MPPI_OK) return rc; }
  auto it = h->graphs.find(graph_key(h, v));
  // bracket the replay with events on the launching stream so callers can read the
  // device time of the whole step (mppi_get_timings()[3]) without a profiler
  // The bracket costs two timestamp operations on the stream around every replay, so it is recorded only once a caller has
  // asked for timings (mppi_get_timings / mppi_set_profiling arm it; MPPI_STEP_EVENTS=1 forces it from the first step).
  if (h->step_events) MPPI_CUDA(h, cudaEventRecord(h->ev[0], h->stream));
  MPPI_CUDA(h, cudaGraphLaunch(it->second, h->stream));
  if (h->step_events) MPPI_CUDA(h, cudaEventRecord(h->ev[4], h->stream));
  h->step_timed = h->step_events;
  h->launches += kernels_in(h, v & V_ENQ_MASK);
  h->noise_ready = (v & V_PREGEN) != 0;   // any other step leaves this iteration's own noise (or none) in eps
  if (v & (V_COMBINE | V_FUSED_FINAL | V_XFINAL)) h->seq_launched++;
  h->in_flight = h->inputs_busy = true;
  return MPPI_OK;
}

}  // namespace mppi
