#!/usr/bin/env python
"""bench.py — MPPI sample-steps/s and per-iteration latency on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA, through the C ABI)
  python bench.py --impl reference --gpus N --steps K ...  reference arm: the CPU path (oracle
                                                           port of the Julia code) on host cores

One "step" = one MPPI iteration (MPPIPlan, MPPIUtils.jl:169-203): K-sample rollout + cost,
softmin weights, weighted control update, nominal rollout.  Default workload = BASELINE.json
configs[1]: kinematic bicycle, 3 circles, K = 65,536, N = 50, in-kernel Philox noise, one
CUDA-graph launch per iteration.  For --gpus N > 1 (torchrun, one rank per GPU) samples are
sharded, 65,536 per GPU (weak scaling), and combined by one in-library NCCL all-gather.
Prints ONE JSON line on rank 0.
"""
import argparse
import gc
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "mppi_sample_steps_per_s"
UNIT = "sample-steps/s"
WORKLOADS = {
    1: "cfg2: kinematic bicycle, 3 circles, K=65536/GPU, N=50, Philox, CUDA graph",
    0: "cfg1: kinematic bicycle, 3 circles, K=1024, N=50",
    2: "cfg3: parking layout polygon walls (SAT), kinematic, K=262144, N=80",
    3: "cfg4: dynamic bicycle + polygon walls, K=1048576 total, N=100",
    4: "cfg5: 4096 instances x K=1024 x N=50 (batched)",
}


# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu captures
# (profiles/r01_ncu_*.md), keyed by --config: (rollout kernel, update kernel); N = 1 only.
NCU_TRAFFIC = {
    1: (26_285_824 + 0, 26_491_136 + 0),                        # rollout reads the pre-generated eps (26.2 MB algorithmic)
    2: (167_845_632 + 4_443_136, 168_838_656 + 3_827_200),
    3: (839_461_376 + 9_430_784, 843_085_056 + 3_552_768),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=1, help="index into BASELINE.json configs (default 1)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N > 1: softmin all-gather in-kernel over NVLink peer memory (default) or in-library NCCL")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons sampled with NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.max_mhz, self.stop_flag = [], set(), None, False
        try:
            import pynvml as nv

            nv.nvmlInit()
            self.nv, self.h = nv, nv.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def workload(cfg_idx, world, rank):
    """defineMPPI arguments for this rank.  Sample-sharded configs keep the per-GPU shard fixed
    (weak scaling); cfg4's K = 1,048,576 is the TOTAL (strong scaling, as BASELINE.json states it);
    cfg5 shards instances (512 per GPU at 8 GPUs)."""
    from motionplanning_b200 import scenarios as sc

    if cfg_idx == 3:
        args, kw, circles, blocks = sc.config(3)
        K = args[5]
        scaling = "strong"
    elif cfg_idx == 4:
        args, kw, circles, blocks = sc.config(4)
        kw["B"] = 4096 // world if world > 1 else 4096
        return args, kw, circles, blocks, "strong", args[5], 0, args[5]
    else:
        args, kw, circles, blocks = sc.config(cfg_idx)
        K = args[5] * world
        scaling = "weak"
    from motionplanning_b200.dist import shard_range

    off, kl = shard_range(K, rank, world)
    args = args[:5] + (K,) + args[6:]
    kw.update(k_offset=off, K_local=kl, FeasibilityCount=K)
    return args, kw, circles, blocks, scaling, K, off, kl


def cpu_baseline(cfg_idx, budget_s=12.0, threads=None):
    """Time the CPU path (C++ oracle port of the reference's Julia MPPIPlan, FP64) on this box's
    host cores, on a bounded sample of the workload: at most 65,536 samples per repetition,
    repeated until ~budget_s of wall time."""
    import oracle_lib as orc
    from motionplanning_b200 import _capi as capi
    from motionplanning_b200 import scenarios as sc

    threads = threads or host_cores()
    args, kw, circles, blocks = sc.config(cfg_idx)
    K = min(args[5], 65536)
    N = args[7]
    args = args[:5] + (K,) + args[6:]
    kw = dict(kw, precision=capi.FP64, noise_mode=capi.NOISE_SUPPLIED, FeasibilityCount=K)
    kw.pop("B", None)
    cfg = orc.make_config(args, kw)
    eps = orc.generate_noise(args[6], sc.SEED, 0, K, N, capi.FP64)
    U = np.zeros((N, 2))
    secs, reps = [], 0
    t_end = time.perf_counter() + budget_s
    while reps < 2 or (time.perf_counter() < t_end and reps < 200):
        r = orc.plan(cfg, circles, blocks, args[3], U, eps, capi.FP64, threads=threads)
        U = r["Control"]
        secs.append(r["seconds"])
        reps += 1
    best = float(np.median(secs))
    return {"value": K * N / best, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{K} of the workload's samples x N={N}, full MPPIPlan (FP64, supplied noise), "
                      f"median of {reps} repetitions, OpenMP over samples",
            "ms_per_iteration": best * 1e3}, K, N


def run_reference(a, rank, world):
    """Reference arm: the reference's own CPU implementation of the path.  Julia is not
    installed on either box and the reference has no compiled code, so this times the oracle
    port (kind "port") with all host threads.  Each step = one MPPIPlan on a bounded sample."""
    if rank != 0:
        return
    import oracle_lib as orc
    from motionplanning_b200 import _capi as capi
    from motionplanning_b200 import scenarios as sc

    threads = host_cores()
    args, kw, circles, blocks = sc.config(a.config)
    K = min(args[5], 65536)
    N = args[7]
    args = args[:5] + (K,) + args[6:]
    kw = dict(kw, precision=capi.FP64, noise_mode=capi.NOISE_SUPPLIED, FeasibilityCount=K)
    kw.pop("B", None)
    cfg = orc.make_config(args, kw)
    eps = orc.generate_noise(args[6], sc.SEED, 0, K, N, capi.FP64)
    U = np.zeros((N, 2))
    steps = min(a.steps, 50)   # bounded: ~0.1-1 s per CPU iteration
    warm = min(a.warmup, 3)
    for _ in range(warm):
        U = orc.plan(cfg, circles, blocks, args[3], U, eps, capi.FP64, threads=threads)["Control"]
    t0 = time.perf_counter()
    for _ in range(steps):
        U = orc.plan(cfg, circles, blocks, args[3], U, eps, capi.FP64, threads=threads)["Control"]
    dt = time.perf_counter() - t0
    v = K * N * steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOADS[a.config], "samples_per_step": K, "horizon": N},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"{K} samples x N={N} per step, {steps} steps, OpenMP over samples; "
                                       "reference Julia path: unavailable on this box (no julia)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_ours(a, rank, world, local_rank):
    import torch

    import motionplanning_b200 as m
    from motionplanning_b200 import dist as mdist
    from motionplanning_b200 import scenarios as sc

    capi = m.capi
    torch.cuda.set_device(local_rank)
    args, kw, circles, blocks, scaling, K, off, kl = workload(a.config, world, rank)
    N, nx = args[7], args[0]
    B = kw.get("B", 1)
    kw.update(device=local_rank, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
    batched = B > 1
    if batched:
        x0s = sc.jittered_x0(B, seed=7 + rank)
        args = args[:3] + (x0s,) + args[4:]
    mppi = m.defineMPPI(*args, **kw)
    if circles:
        m.defineMPPIobs(mppi, circles)
    if blocks:
        m.defineMPPIblocks(mppi, blocks)
    exchange = "none"
    if world > 1 and not batched:
        exchange = a.exchange
        if exchange == "peer":
            try:
                mdist.init_peer_exchange(mppi)
            except Exception as e:  # no peer access between these devices: NCCL carries the exchange
                print(f"[rank {rank}] peer exchange unavailable ({e}); using NCCL", file=sys.stderr)
                exchange = "nccl"
        if exchange == "nccl":
            mdist.init_library_comm(mppi)

    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def l2_flush():
        flush.zero_()
        torch.cuda.synchronize()

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    # ---- nominal: 5 closed-loop iterations from zeros, then frozen ----
    if batched:
        mppi.step_batched()
    else:
        m.MPPIPlan(mppi)
    for _ in range(4):
        mppi.step_async(adopt=True)
    mppi.sync()
    mppi.adopt_nominal()

    # ---- device-resident timing: one graph launch per iteration, CUDA events on the LIBRARY's
    #      stream, L2 flushed (256 MiB write on the same stream) before every timed iteration.
    #      Everything is enqueued asynchronously, so host jitter never enters the timed spans and
    #      the ranks of a sharded run stay in lockstep through the step's own exchange. ----
    lib_stream = torch.cuda.ExternalStream(mppi.stream(), device=torch.device("cuda", local_rank))
    for _ in range(max(a.warmup, 3)):
        mppi.step_async(adopt=False)
    mppi.sync()
    clocks = ClockSampler(local_rank)
    launches0 = mppi.launch_count()
    ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps)]
    ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps)]
    barrier()
    clocks.start()
    for i in range(a.steps):
        with torch.cuda.stream(lib_stream):
            flush.zero_()
        ev0[i].record(lib_stream)
        mppi.step_async(adopt=False)
        ev1[i].record(lib_stream)
    mppi.sync()
    barrier()
    clocks.stop_flag = True
    dev_us = [ev0[i].elapsed_time(ev1[i]) * 1e3 for i in range(a.steps)]
    launches = mppi.launch_count() - launches0
    total_s = float(np.sum(dev_us)) * 1e-6
    if world > 1:
        t = torch.tensor([total_s], dtype=torch.float64, device="cuda")
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        total_s = float(t.item())
    units_per_step = (K if not batched else B * world * K) * N
    value = units_per_step * a.steps / total_s

    # ---- end to end through the reference-facing API with HOST buffers, every step:
    #      ShiftInitialCondition + defineMPPINominalControl! + MPPIPlan (main.jl:59-62) ----
    U_host = np.array(mppi.s.NominalControl if batched else mppi.r.Control, copy=True)
    x_host = np.array(mppi.s.X0, dtype=np.float64, copy=True)
    e2e_s = []
    if batched:   # host buffers in the ABI's own layout (what a Julia caller holds): no reshuffling in the timed call
        xb = np.ascontiguousarray(x_host.T)                                  # (B, n_x)
        Ub = np.ascontiguousarray(np.transpose(U_host, (2, 1, 0))) if U_host.ndim == 3 else np.zeros((B, 2, N))
        Uo, cb, fb = np.empty_like(Ub), np.empty(B), np.empty(B, dtype=np.int32)
    gc.disable()   # the timed calls are ~60 us each: a collector pause would be the largest term in the sum
    for i in range(a.steps + 3):
        l2_flush()
        t0 = time.perf_counter()
        if batched:
            mppi.step_batched_abi(xb, Ub, Uo, cb, fb)
        else:
            m.ShiftInitialCondition(mppi, x_host)
            m.defineMPPINominalControl(mppi, U_host)
            m.MPPIPlan(mppi)
            _ = mppi.r.Control[0, 0]
        t1 = time.perf_counter()
        if i >= 3:
            e2e_s.append(t1 - t0)
    gc.enable()
    e2e_total = float(np.sum(e2e_s))
    if world > 1:
        t = torch.tensor([e2e_total], dtype=torch.float64, device="cuda")
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        e2e_total = float(t.item())
    e2e_value = units_per_step * a.steps / e2e_total
    h2d = B * (nx + 2 * N) * 8
    d2h = B * (8 + 2 * N) * 8 + (0 if batched else (N + 1) * nx * 8)

    # ---- per-kernel times (eager launches bracketed by events on the library stream) ----
    mppi.set_profiling(True)
    k_us = []
    for _ in range(12):
        l2_flush()
        mppi.step_async(adopt=False)
        mppi.sync()
        k_us.append(mppi.timings_us())
    mppi.set_profiling(False)
    k_us = np.median(np.array(k_us[2:]), axis=0)
    # ---- the same kernels inside the replayed graph: %globaltimer first-block-in / last-block-out spans
    #      (mppi_set_profiling(2)); free of the launch and event gaps an eager bracket includes ----
    g_us = None
    if not batched:
        spans = []
        for i in range(14):
            l2_flush()
            mppi.set_profiling(2)
            mppi.step_async(adopt=False)
            mppi.sync()
            c = mppi.phase_clocks()
            if i >= 2:
                noise = c[15] > 0 and c[14] < 2 ** 63   # noise_kernel for the next iteration ran beside finalize (V_PREGEN)
                spans.append([(c[9 + 2 * j] - c[8 + 2 * j]) * 1e-3 for j in range(3)] + [(c[13] - c[8]) * 1e-3] +
                             ([(c[15] - c[14]) * 1e-3, (max(c[13], c[15]) - c[8]) * 1e-3] if noise else [0.0, (c[13] - c[8]) * 1e-3]))
        mppi.set_profiling(0)
        g_us = np.median(np.array(spans), axis=0)
    fp32_peak = mppi.measure_fp32_peak()
    peaks, peak_src = measured_peaks()
    F = sc.flop_per_sample_step(kw["model"], len(circles), len(blocks))
    local_units = kl * N * B
    roll_tflops = local_units * F / (k_us[0] * 1e-6) / 1e12
    upd_bytes = local_units * 2 * 4 + kl * B * 4 + N * 2 * 4 * B
    upd_gbs = upd_bytes / (k_us[1] * 1e-6) / 1e9

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": total_s / a.steps * 1e3, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOADS[a.config], "K_total": K if not batched else K, "K_per_gpu": kl, "N": N,
                       "instances": B * (world if batched else 1), "noise": ("philox4x32-10 on the device, one tensor drawn per step inside the timed region: noise_kernel draws the next "
                                 "iteration's tensor on a graph branch beside finalize, the rollout reads it back (TMA-staged)"
                                 if not batched else "philox4x32-10 inside the rollout kernel"),
                       "parallelism": (f"samples sharded x{world}, softmin all-gather in the finalize kernel over NVLink peer memory"
                                       if exchange == "peer" else f"samples sharded x{world}, one NCCL all-gather per step") if world > 1 and not batched
                       else ("instances sharded, no collective" if batched and world > 1 else "1 GPU"),
                       "l2": "flushed between timed iterations (256 MiB write)",
                       "timing": "CUDA events on the library stream around one graph launch per iteration (flush + step enqueued asynchronously), summed, max over ranks"},
            "latency_us": {"device_median": float(np.median(dev_us)), "device_p99": float(np.percentile(dev_us, 99)),
                           "e2e_median": float(np.median(e2e_s) * 1e6), "e2e_p99": float(np.percentile(e2e_s, 99) * 1e6)},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": ("mppi_step_batched (host x0s/Us in, Us_out/costs/feasible out, synchronous)" if batched else
                            "ShiftInitialCondition + defineMPPINominalControl + MPPIPlan (host buffers, synchronous)")},
            "gpu_launches": int(launches),
            "kernels_us": {"rollout_cost": float(k_us[0]), "softmin_update": float(k_us[1]), "finalize_nominal": float(k_us[2]),
                           "step_eager": float(k_us[3]),
                           "source": "CUDA events around eager launches on the library stream, L2 flushed, median of 10"},
            "kernels_us_in_graph": None if g_us is None else {
                "rollout_cost": float(g_us[0]), "softmin_update": float(g_us[1]), "finalize_nominal": float(g_us[2]),
                "rollout_start_to_finalize_end": float(g_us[3]),
                "noise_next_iteration": float(g_us[4]), "rollout_start_to_last_kernel_end": float(g_us[5]),
                "source": "%globaltimer spans written by the kernels themselves inside the replayed graph (mppi_set_profiling(2))"},
            "roofline": {"kernel": "rollout_kernel", "bound": "fp32", "achieved": roll_tflops, "peak": fp32_peak,
                         "unit": "TFLOP/s", "frac": roll_tflops / fp32_peak if fp32_peak else None,
                         "traffic": NCU_TRAFFIC.get(a.config, (None, None))[0] if world == 1 else None,
                         "traffic_source": "profiles/r01_ncu_*.md (ncu --set full, per launch)",
                         "flop_per_sample_step": F,
                         "frac_in_graph": None if g_us is None or not fp32_peak else local_units * F / (g_us[0] * 1e-6) / 1e12 / fp32_peak,
                         "peak_source": "FFMA-chain probe measured in this run (mppi_measure_fp32_peak)"},
            "roofline_update": {"kernel": "update_kernel", "bound": "hbm", "achieved": upd_gbs, "peak": peaks.get("hbm_gbs"),
                                "unit": "GB/s", "frac": upd_gbs / peaks["hbm_gbs"],
                                "traffic": NCU_TRAFFIC.get(a.config, (None, None))[1] if world == 1 else None,
                                "bytes_per_launch": upd_bytes, "peak_source": f"MEASURED_PEAKS.json ({peak_src})",
                                "frac_in_graph": None if g_us is None else upd_bytes / (g_us[1] * 1e-6) / 1e9 / peaks["hbm_gbs"],
                                "note": "eps was read (or produced) by the rollout kernel earlier in the same step; when it fits the "
                                        "126 MB L2 the stream is served from L2 and can exceed the HBM peak"},
            "clocks": clocks.summary(),
        }
        if not a.no_cpu_baseline:
            cb, _, _ = cpu_baseline(a.config if a.config != 4 else 0)
            line["cpu_baseline"] = cb
    return line


def main():
    a = parse()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference":
        run_reference(a, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist

        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    line = run_ours(a, rank, world, local_rank)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
