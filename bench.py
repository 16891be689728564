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
# (profiles/r02_ncu_*.md): (rollout / scan kernel, update kernel); N = 1 only.
NCU_TRAFFIC = {   # key: (--config, --walls); the start pose does not change the traffic
    (1, 0): (26245376 + 0, 26491392 + 0),                          # profiles/r02_ncu_cfg2_K65536_N50.md
    (2, 0): (545536 + 112067072, 168837120 + 3288832),             # profiles/r02_ncu_cfg3_K262144_N80.md (eps written by the rollout: Philox inside)
    (3, 0): (838968576 + 8367104, 843080192 + 3372800),            # profiles/r02_ncu_cfg4_K1M_N100_walls_reached.md (eps read back: pre-generated)
    (4, 0): (4166000 + 1685743000, 1697900000 + 3697000),          # profiles/r02_ncu_cfg5_B4096.md
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=1, help="index into BASELINE.json configs (default 1)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra.cfg4 strong-scaling block that rides along with the default line")
    ap.add_argument("--walls", type=int, default=0, help="cfg4 only: number of wall blocks M_p (default 3; SURVEY 8(d) asks for 3 and 16)")
    ap.add_argument("--x0", default="", help="override the start state, comma separated (cfg4: a pose from which the walls are reached)")
    ap.add_argument("--noise-policy", type=int, default=0, help="0 auto | 1 latency (pre-generated noise) | 2 throughput (Philox inside the rollout)")
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


def load_scenarios():
    """motionplanning_b200/scenarios.py loaded BY PATH: pure host data, and the CPU legs (cpu_baseline, --impl reference)
    must not import the product package or map libmppi_b200.so."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("mppi_scenarios", os.path.join(ROOT, "motionplanning_b200", "scenarios.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def config_keys(cfg_idx, world):
    """`config` of the JSON line: the workload and nothing else, identical in both arms."""
    sc = load_scenarios()
    args, kw, _, _ = sc.config(cfg_idx)
    K, N, B = args[5], args[7], kw.get("B", 1)
    if cfg_idx in (0, 1, 2):
        K *= world          # sample-sharded, weak scaling: 65,536 (cfg2) per GPU
    return {"workload": WORKLOADS[cfg_idx], "K_total": K, "N": N, "instances": B}


def cpu_problem(cfg_idx):
    """The bounded CPU sample of a workload: at most 65,536 samples of ONE instance, supplied noise, FP64."""
    import oracle_lib as orc

    sc = load_scenarios()
    args, kw, circles, blocks = sc.config(cfg_idx)
    K = min(args[5], 65536)
    N = args[7]
    args = args[:5] + (K,) + args[6:]
    kw = dict(kw, precision=orc.FP64, noise_mode=orc.NOISE_SUPPLIED, FeasibilityCount=K)
    kw.pop("B", None)
    cfg = orc.make_config(args, kw)
    eps = orc.generate_noise(args[6], sc.SEED, 0, K, N, orc.FP64)
    return orc, cfg, circles, blocks, args, eps, K, N


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
    repeated until ~budget_s of wall time.  Also: the same plan on ONE thread (the reference has no
    threading; 4,096-sample slice) and the number of samples of the slice that sit inside the FP32
    guard band (closest threshold comparison within 1e-4 relative of flipping)."""
    orc, cfg, circles, blocks, args, eps, K, N = cpu_problem(cfg_idx)
    threads = threads or host_cores()
    U = np.zeros((N, 2))
    secs, reps = [], 0
    t_end = time.perf_counter() + budget_s
    while reps < 2 or (time.perf_counter() < t_end and reps < 200):
        r = orc.plan(cfg, circles, blocks, args[3], U, eps, orc.FP64, threads=threads)
        U = r["Control"]
        secs.append(r["seconds"])
        reps += 1
    best = float(np.median(secs))
    # single thread, faithful to the reference's structure (serial K loop): a 4,096-sample slice, scaled per sample-step
    K1 = min(K, 4096)
    cfg1 = orc.make_config(args[:5] + (K1,) + args[6:], dict(model=cfg.model, precision=orc.FP64, noise_mode=orc.NOISE_SUPPLIED,
                                                             FeasibilityCount=K1, veh_L=cfg.veh_L, veh_W=cfg.veh_W,
                                                             veh_center_offset=cfg.veh_center_offset))
    eps1 = np.ascontiguousarray(eps[:, :K1, :])
    one = [orc.plan(cfg1, circles, blocks, args[3], U, eps1, orc.FP64, threads=1, want_margin=(i == 0)) for i in range(3)]
    t1 = float(np.median([o["seconds"] for o in one]))
    guard = int(np.sum(one[0]["margins"] < 1e-4))
    return {"value": K * N / best, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{K} of the workload's samples x N={N}, full MPPIPlan (FP64, supplied noise), "
                      f"median of {reps} repetitions, OpenMP over samples",
            "ms_per_iteration": best * 1e3,
            "single_thread": {"value": K1 * N / t1, "unit": UNIT, "cores": 1,
                              "sample": f"{K1} samples x N={N}, serial K loop as in MPPIUtils.jl:175, median of 3"},
            "guard_band_samples": guard, "guard_band_of": K1,
            "guard_band_note": "samples whose closest obstacle/bound comparison is within 1e-4 relative of flipping in the FP64 oracle "
                               "(the FP32 fast path may legitimately flip these; strict_math = 1 matches a float oracle bit for bit)"}, K, N


def run_reference(a, rank, world):
    """Reference arm: the reference's own CPU implementation of the path.  Julia is not
    installed on either box and the reference has no compiled code, so this times the oracle
    port (kind "port") with all host threads.  Each step = one MPPIPlan on a bounded sample.
    Imports nothing of the product (tests/test_oracle_config.py::test_oracle_wrapper_never_loads_the_product)."""
    if rank != 0:
        return
    orc, cfg, circles, blocks, args, eps, K, N = cpu_problem(a.config)
    threads = host_cores()
    U = np.zeros((N, 2))
    steps, warm = a.steps, max(a.warmup, 3)
    for _ in range(warm):
        U = orc.plan(cfg, circles, blocks, args[3], U, eps, orc.FP64, threads=threads)["Control"]
    t0 = time.perf_counter()
    for _ in range(steps):
        U = orc.plan(cfg, circles, blocks, args[3], U, eps, orc.FP64, threads=threads)["Control"]
    dt = time.perf_counter() - t0
    v = K * N * steps / dt
    scaling = "strong" if a.config in (3, 4) else "weak"
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_keys(a.config, world),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"{K} samples x N={N} per step, {steps} steps, OpenMP over samples; "
                                       "reference Julia path: unavailable on this box (no julia)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def parity_across_ranks(m, mdist, args, kw, circles, blocks, exchange, rank, world, local_rank, iters=3):
    """N > 1, outside the timed region: the sharded step (this rank's shard + the exchange under test) against ONE
    unsharded handle with K = K_total on this rank's GPU, same seed, same iteration, `iters` chained iterations.
    Costs come from the same kernel on both sides (the unsharded handle is forced onto the sharded handle's rollout
    variant), so what is checked is exactly the exchange + combine: argmin, RolloutCount, FeasibleTrajCount identical,
    |U_new - U_ref| <= 2e-6 max|U_ref|, and U_new bit-identical on all ranks.  Raises on mismatch (rc != 0)."""
    import torch
    import torch.distributed as dist

    capi = m.capi
    sh = m.defineMPPI(*args, **kw)
    if circles:
        m.defineMPPIobs(sh, circles)
    if blocks:
        m.defineMPPIblocks(sh, blocks)
    if exchange == "peer":
        mdist.init_peer_exchange(sh)
    else:
        mdist.init_library_comm(sh)
    info = sh.plan_info()
    kw_ref = dict(kw, k_offset=0, K_local=0)
    kw_ref["rollout_variant"] = 5 if info["scan"] else ({3: 6, 2: 6, 1: 2}.get(info["packed"], 1))   # 3 (in-block update) -> 6: same rollout, and K_total may exceed variant 7's block limit   # 6 = two-stage warp pipeline, 2 = packed sequential
    kw_ref["noise_policy"] = 1 if info["pregen"] else 2
    ref = m.defineMPPI(*args, **kw_ref)
    if circles:
        m.defineMPPIobs(ref, circles)
    if blocks:
        m.defineMPPIblocks(ref, blocks)
    N = args[7]
    U = np.zeros((N, 2))
    max_du, ok_arg, ok_cnt = 0.0, True, True
    for _ in range(iters):
        m.defineMPPINominalControl(sh, U)
        m.defineMPPINominalControl(ref, U)
        m.MPPIPlan(sh)
        m.MPPIPlan(ref)
        scale = max(float(np.abs(ref.r.Control).max()), 1e-12)
        max_du = max(max_du, float(np.abs(sh.r.Control - ref.r.Control).max()) / scale)
        ok_arg &= sh.r.argmin == ref.r.argmin
        ok_cnt &= sh.r.RolloutCount == ref.r.RolloutCount and sh.r.FeasibleTrajCount == ref.r.FeasibleTrajCount
        U = sh.r.Control.copy()
    mine = torch.from_numpy(np.ascontiguousarray(U).view(np.uint8).reshape(-1).copy()).cuda()
    allU = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allU, mine)
    identical = all(bool(torch.equal(allU[0], t)) for t in allU)
    flags = torch.tensor([max_du, 0.0 if ok_arg else 1.0, 0.0 if ok_cnt else 1.0, 0.0 if identical else 1.0], dtype=torch.float64, device="cuda")
    dist.all_reduce(flags, op=dist.ReduceOp.MAX)
    out = {"checked": True, "iterations": iters, "exchange": exchange, "max_du": float(flags[0]), "tol_du": 2e-6,
           "argmin_equal": flags[1].item() == 0.0, "counts_equal": flags[2].item() == 0.0,
           "ranks_bit_identical": flags[3].item() == 0.0,
           "against": "one unsharded handle (K = K_total) per rank, same seed / iteration / rollout kernel"}
    sh.close()
    ref.close()
    if not (out["max_du"] <= out["tol_du"] and out["argmin_equal"] and out["counts_equal"] and out["ranks_bit_identical"]):
        raise SystemExit(f"[rank {rank}] sharded step differs from the unsharded one: {json.dumps(out)}")
    return out


def strong_scaling_cfg4(m, mdist, exchange, rank, world, local_rank, steps=20, warm=5):
    """BASELINE.json config 4 (dynamic bicycle + polygon walls, K = 1,048,576 TOTAL, N = 100) on `world` GPUs, strong scaling:
    device time per step (CUDA events on the library stream, L2 flushed, max over ranks) and the exchange's parity check.
    Reported as `extra.cfg4` of the headline line so that the driver's 1/2/4/8-GPU record carries the north-star target."""
    import torch

    args, kw, circles, blocks, scaling, K, off, kl = workload(3, world, rank)
    N = args[7]
    kw.update(device=local_rank, precision=m.capi.FP32, noise_mode=m.capi.NOISE_PHILOX)
    parity = None
    if world > 1:
        parity = parity_across_ranks(m, mdist, args, kw, circles, blocks, exchange, rank, world, local_rank, iters=2)
    mp = m.defineMPPI(*args, **kw)
    m.defineMPPIblocks(mp, blocks)
    if world > 1:
        (mdist.init_peer_exchange if exchange == "peer" else mdist.init_library_comm)(mp)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    st = torch.cuda.ExternalStream(mp.stream(), device=torch.device("cuda", local_rank))
    sync_t = torch.zeros(1, device="cuda")
    m.MPPIPlan(mp)
    for _ in range(warm):
        mp.step_async(adopt=True)
    mp.sync()
    mp.adopt_nominal()
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    with torch.cuda.stream(st):
        torch.cuda._sleep(16_000_000)   # the host enqueues the timed iterations while the stream spins (see run_ours)
    for i in range(steps):
        with torch.cuda.stream(st):
            flush.zero_()
            if world > 1:
                torch.distributed.all_reduce(sync_t)
        e0[i].record(st)
        mp.step_async(adopt=False)
        e1[i].record(st)
    mp.sync()
    total = float(np.sum([e0[i].elapsed_time(e1[i]) for i in range(steps)])) * 1e-3
    if world > 1:
        t = torch.tensor([total], dtype=torch.float64, device="cuda")
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        total = float(t.item())
    info = mp.plan_info()
    mp.close()
    del flush
    return {"workload": WORKLOADS[3], "scaling": "strong", "K_total": K, "K_per_gpu": kl, "N": N, "steps": steps, "ms_per_step": total / steps * 1e3,
            "value": K * N * steps / total, "unit": UNIT, "plan_info": info, "parity_n": parity}


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
    if a.walls:
        blocks = sc.wall_blocks(a.walls)
    if a.x0:
        args = args[:3] + ([float(v) for v in a.x0.split(",")],) + args[4:]
    if a.noise_policy:
        kw["noise_policy"] = a.noise_policy
    batched = B > 1
    if batched:
        x0s = sc.jittered_x0(B, seed=7 + rank)
        args = args[:3] + (x0s,) + args[4:]
    exchange = "none"
    parity_n = None
    if world > 1 and not batched:
        exchange = a.exchange
        # correctness of the sharded step, outside the timed region, recorded in the JSON line (fails the run on mismatch)
        parity_n = parity_across_ranks(m, mdist, args, kw, circles, blocks, exchange, rank, world, local_rank)
    mppi = m.defineMPPI(*args, **kw)
    if circles:
        m.defineMPPIobs(mppi, circles)
    if blocks:
        m.defineMPPIblocks(mppi, blocks)
    if exchange == "peer":
        mdist.init_peer_exchange(mppi)   # no silent fallback: without peer access between the GPUs this raises; ask for --exchange nccl
    elif exchange == "nccl":
        mdist.init_library_comm(mppi)
    info = mppi.plan_info()

    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    lib_stream = torch.cuda.ExternalStream(mppi.stream(), device=torch.device("cuda", local_rank))
    sync_t = torch.zeros(1, device="cuda")

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
    #      Everything is enqueued asynchronously, so host jitter never enters the timed spans.  N > 1: the ranks
    #      meet in a one-element all-reduce on the same stream between the flush and the first event, so that the
    #      differing durations of the ranks' flushes do not count as step time; the step's own exchange is inside. ----
    warm = max(a.warmup, 3)
    for _ in range(warm):
        mppi.step_async(adopt=False)
    mppi.sync()
    clocks = ClockSampler(local_rank)
    launches0 = mppi.launch_count()
    ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps)]
    ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps)]
    barrier()
    clocks.start()
    # Host jitter must not reach the device-timed spans: the stream first spins for ~8 ms (torch.cuda._sleep), during which the host
    # enqueues the timed iterations, so every rank's GPU then runs them back to back out of a full queue.  (Seen at 8 GPUs without
    # it: one 1.8 ms step in a hundred — a rank whose host thread fell behind its stream — which the summed time then carries.)
    with torch.cuda.stream(lib_stream):
        torch.cuda._sleep(16_000_000)
    for i in range(a.steps):
        with torch.cuda.stream(lib_stream):
            flush.zero_()
            if world > 1 and not batched:
                torch.distributed.all_reduce(sync_t)
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
    #      ShiftInitialCondition + defineMPPINominalControl! + MPPIPlan (main.jl:59-62).  The timed span of a step
    #      ends when the handle's stream is IDLE (mppi.sync() after the call returns), so work the step leaves behind
    #      the delivery of its result (the next iteration's noise under the "latency" policy) is inside it; the time
    #      to the result alone is reported as latency_us.e2e_result_*. ----
    U_host = np.array(mppi.s.NominalControl if batched else mppi.r.Control, copy=True)
    x_host = np.array(mppi.s.X0, dtype=np.float64, copy=True)
    e2e_s, e2e_res_s = [], []
    if batched:   # host buffers in the ABI's own layout (what a Julia caller holds): no reshuffling in the timed call
        xb = np.ascontiguousarray(x_host.T)                                  # (B, n_x)
        Ub = np.ascontiguousarray(np.transpose(U_host, (2, 1, 0))) if U_host.ndim == 3 else np.zeros((B, 2, N))
        Uo, cb_, fb = np.empty_like(Ub), np.empty(B), np.empty(B, dtype=np.int32)
    gc.disable()   # the timed calls are tens of us each: a collector pause would be the largest term in the sum
    for i in range(a.steps + 3):
        l2_flush()
        if world > 1:
            torch.distributed.barrier()
        t0 = time.perf_counter()
        if batched:
            mppi.step_batched_abi(xb, Ub, Uo, cb_, fb)
        else:
            m.ShiftInitialCondition(mppi, x_host)
            m.defineMPPINominalControl(mppi, U_host)
            m.MPPIPlan(mppi)
            _ = mppi.r.Control[0, 0]
        t1 = time.perf_counter()
        mppi.sync()
        t2 = time.perf_counter()
        if i >= 3:
            e2e_res_s.append(t1 - t0)
            e2e_s.append(t2 - t0)
    gc.enable()
    e2e_total = float(np.sum(e2e_s))
    if world > 1:
        t = torch.tensor([e2e_total], dtype=torch.float64, device="cuda")
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        e2e_total = float(t.item())
    e2e_value = units_per_step * a.steps / e2e_total
    h2d = B * (nx + 2 * N) * 8
    d2h = B * (8 + 2 * N) * 8 + (0 if batched else (N + 1) * nx * 8)
    # ---- the same three calls made by a COMPILED caller (what Julia's ccall pays): timed inside the library with the host's steady
    #      clock, nothing but mppi_set_x0 + mppi_set_nominal + mppi_step in the span, L2 flushed before every span
    #      (mppi_time_reference_calls).  The e2e key above stays the Python mirror's number; this one isolates the interpreter's share. ----
    c_abi = None
    if not batched and world == 1:
        torch.cuda.synchronize()
        res_us, idle_us = mppi.time_reference_calls(x_host, U_host, a.steps + 3, flush.data_ptr(), flush.numel())
        c_abi = {"result_median_us": float(np.median(res_us[3:])), "result_p99_us": float(np.percentile(res_us[3:], 99)),
                 "idle_median_us": float(np.median(idle_us[3:])),
                 "value": units_per_step * a.steps / float(np.sum(idle_us[3:]) * 1e-6), "unit": UNIT,
                 "source": "mppi_time_reference_calls: mppi_set_x0 + mppi_set_nominal + mppi_step from host buffers, host steady clock inside the library, "
                           "L2 flushed before every span; `value` counts every span to stream idle like the e2e key"}
    # ---- per-kernel times: eager launches bracketed by CUDA events on the library stream (mppi_set_profiling(1)); the L2
    #      flush is enqueued on the same stream right before, without a host synchronise in between, so the events do not
    #      also time an idle GPU's launch latency.  The scan-form step launches its tail as a separate kernel in this mode. ----
    mppi.set_profiling(1)
    k_us = []
    for _ in range(14):
        with torch.cuda.stream(lib_stream):
            flush.zero_()
        mppi.step_async(adopt=False)
        mppi.sync()
        k_us.append(mppi.timings_us())
    mppi.set_profiling(0)
    k_us = np.median(np.array(k_us[2:]), axis=0)
    # ---- the same kernels inside the replayed graph: %globaltimer first-block-in / last-block-out spans
    #      (mppi_set_profiling(2)); free of the launch and event gaps an eager bracket includes ----
    g_us = None
    wall_tests = None
    if not batched:
        spans = []
        for i in range(14):
            l2_flush()
            mppi.set_profiling(2)
            mppi.step_async(adopt=False)
            mppi.sync()
            c = mppi.phase_clocks()
            wall_tests = c[7]
            if i >= 2:
                def span(j):
                    return (c[9 + 2 * j] - c[8 + 2 * j]) * 1e-3 if (c[9 + 2 * j] > 0 and c[8 + 2 * j] < 2 ** 63) else 0.0
                noise = c[15] > 0 and c[14] < 2 ** 63   # noise_kernel for the next iteration ran behind the step
                last = max(c[13], c[15]) if noise else c[13]
                spans.append([span(0), span(1), span(2), (c[13] - c[8]) * 1e-3, span(3), (last - c[8]) * 1e-3])
        mppi.set_profiling(0)
        g_us = np.median(np.array(spans), axis=0)
    fp32_peak = mppi.measure_fp32_peak()
    peaks, peak_src = measured_peaks()
    F = sc.flop_per_sample_step(kw["model"], len(circles), len(blocks))
    local_units = kl * N * B
    fused = bool(info["scan"])
    folded = info["packed"] == 3   # warp pipeline whose blocks also do the softmin update (rollout_variant 7): no update_kernel
    roll_tflops = local_units * F / (k_us[0] * 1e-6) / 1e12
    eps_bytes = local_units * 2 * 4
    upd_bytes = eps_bytes + kl * B * 4 + N * 2 * 4 * B
    traffic = NCU_TRAFFIC.get((a.config, a.walls or 0), (None, None)) if world == 1 else (None, None)
    # share of the rectangle (SAT) tests the warp-voted broad phase left standing, counted by the kernel itself in the traced
    # steps (mppi_get_phase_clocks()[7]); the algorithmic flop count F credits every wall at every step, F_exec only those run
    sat_frac = F_exec = None
    if blocks and wall_tests is not None:
        warps = -(-kl // (32 * (2 if info["packed"] else 1)))
        sat_frac = wall_tests / float(warps * N * len(blocks))
        F_exec = F - 102 * len(blocks) * (1.0 - sat_frac)

    # the north star's multi-GPU target (cfg4, strong scaling) rides along with the headline line at every N
    extra_cfg4 = None
    if a.config == 1 and not a.no_extra:
        mppi.close()
        del flush
        torch.cuda.empty_cache()
        extra_cfg4 = strong_scaling_cfg4(m, mdist, exchange if world > 1 else "none", rank, world, local_rank)

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": warm,
            "ms_per_step": total_s / a.steps * 1e3, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": config_keys(a.config, world),
            "method": {
                "K_per_gpu": kl, "walls": len(blocks), "circles": len(circles),
                "step": ("scan-form fused kernel: horizon-parallel rollout + cost + online softmin + weighted update, tail (fold, exchange, "
                         "combine, nominal rollout) in its last CTA" if fused else
                         ("rollout_pipe_kernel (two-stage warp pipeline, softmin update in its blocks) -> finalize_kernel (folds one partial per block)" if folded else
                          ("rollout_pipe_kernel (two-stage warp pipeline: producer / consumer warps per sample pair) -> update_kernel -> finalize_kernel"
                           if info["packed"] == 2 else "rollout_kernel -> update_kernel -> finalize_kernel"))),
                "noise": ("philox4x32-10 on the device, one tensor per step inside the timed region: noise_kernel draws the NEXT iteration's tensor "
                          "behind the step, the rollout reads it back" if info["pregen"] else "philox4x32-10 inside the rollout kernel"),
                "parallelism": ((f"samples sharded x{world}, softmin all-gather inside the step kernel over NVLink peer memory"
                                 if exchange == "peer" else f"samples sharded x{world}, one NCCL all-gather per step") if world > 1 and not batched
                                else ("instances sharded, no collective" if batched and world > 1 else "1 GPU")),
                "l2": "flushed between timed iterations (256 MiB write)",
                "timing": "CUDA events on the library stream around one graph launch per iteration (flush + step enqueued asynchronously behind "
                          "an 8 ms device-side spin, so the queue is full when the timed iterations run), summed, max over ranks" + ("; ranks aligned by a one-element all-reduce between flush and first event" if world > 1 and not batched else ""),
                "plan_info": info},
            "latency_us": {"device_median": float(np.median(dev_us)), "device_p99": float(np.percentile(dev_us, 99)),
                           "device_max": float(np.max(dev_us)),
                           "e2e_result_median": float(np.median(e2e_res_s) * 1e6), "e2e_result_p99": float(np.percentile(e2e_res_s, 99) * 1e6),
                           "e2e_idle_median": float(np.median(e2e_s) * 1e6), "e2e_idle_p99": float(np.percentile(e2e_s, 99) * 1e6)},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "span": "call start -> result in host memory -> handle's stream idle (the tail a step leaves behind its result is inside)",
                    "api": ("mppi_step_batched (host x0s/Us in, Us_out/costs/feasible out, synchronous)" if batched else
                            "ShiftInitialCondition + defineMPPINominalControl + MPPIPlan (host buffers, synchronous)")},
            "e2e_c_abi": c_abi,
            "gpu_launches": int(launches),
            "kernels_us": {"rollout_cost": float(k_us[0]), "softmin_update": float(k_us[1]), "finalize_nominal": float(k_us[2]),
                           "step_eager": float(k_us[3]),
                           "source": "CUDA events around eager launches on the library stream, L2 flushed on the same stream, median of 12"
                                     + ("; rollout_cost = the fused scan kernel (softmin + update inside), finalize_nominal = its tail launched separately" if fused else "")},
            "kernels_us_in_graph": None if g_us is None else {
                "rollout_cost": float(g_us[0]), "softmin_update": float(g_us[1]), "finalize_nominal": float(g_us[2]),
                "rollout_start_to_finalize_end": float(g_us[3]),
                "noise_next_iteration": float(g_us[4]), "rollout_start_to_last_kernel_end": float(g_us[5]),
                "source": "%globaltimer spans written by the kernels themselves inside the replayed graph (mppi_set_profiling(2))"},
            "roofline": {"kernel": "scan_rollout_kernel (rollout + cost + softmin + weighted update)" if fused else ("rollout_pipe_kernel<., true> (rollout + cost + softmin weights + weighted update of the block's samples)" if folded else
                                                                                                                ("rollout_pipe_kernel" if info["packed"] == 2 else "rollout_kernel")),
                         "bound": "fp32", "achieved": roll_tflops, "peak": fp32_peak,
                         "unit": "TFLOP/s",
                         # wall scenes: `frac` counts only the rectangle tests the kernel actually ran (sat_exact_fraction, counted by the
                         # kernel); the figure that credits every wall at every step is kept as frac_algorithmic_unculled
                         "frac": (roll_tflops / fp32_peak if F_exec is None else local_units * F_exec / (k_us[0] * 1e-6) / 1e12 / fp32_peak) if fp32_peak else None,
                         "frac_algorithmic_unculled": roll_tflops / fp32_peak if fp32_peak else None,
                         "traffic": traffic[0], "traffic_source": "profiles/r02_ncu_*.md (ncu --set full, per launch)",
                         "flop_per_sample_step": F, "algorithmic_flop_per_launch": local_units * F,
                         "sat_exact_fraction": sat_frac, "executed_flop_per_sample_step": F_exec,

                         "frac_in_graph": None if g_us is None or not fp32_peak or g_us[0] <= 0 else local_units * (F if F_exec is None else F_exec) / (g_us[0] * 1e-6) / 1e12 / fp32_peak,
                         "peak_source": "FFMA-chain probe measured in this run (mppi_measure_fp32_peak)",
                         "hbm_view": {"algorithmic_bytes": eps_bytes + kl * B * 5, "gbs": (eps_bytes + kl * B * 5) / (k_us[0] * 1e-6) / 1e9,
                                      "frac_of_measured_peak": (eps_bytes + kl * B * 5) / (k_us[0] * 1e-6) / 1e9 / peaks["hbm_gbs"]}
                         if (fused or info["pregen"]) else None},
            "clocks": clocks.summary(),
        }
        if folded:
            line["roofline_update"] = {"kernel": None, "fused": True,
                                       "note": "no update kernel: every block of rollout_pipe_kernel weights its own 256 samples and sums their clamped controls "
                                               "from a second fetch of its eps slab (L2 hits), finalize folds one partial per block"}
        elif fused:
            line["roofline_update"] = {"kernel": None, "fused": True,
                                       "note": "no separate update kernel at this size: the weighted control sum is accumulated by the scan kernel from "
                                               "the eps values it already holds in registers, so eps is read ONCE per step (update_kernel's second pass, "
                                               f"{upd_bytes} B here, is gone); update_kernel's HBM roofline is reported on the configurations that run it "
                                               "(profiles/r02_bench_cfg3.json, cfg4, cfg5)"}
        else:
            upd_gbs = upd_bytes / (k_us[1] * 1e-6) / 1e9
            in_l2 = eps_bytes <= 100e6
            line["roofline_update"] = {"kernel": "update_kernel", "bound": "hbm", "achieved": upd_gbs, "peak": peaks.get("hbm_gbs"),
                                       "unit": "GB/s", "frac": upd_gbs / peaks["hbm_gbs"], "frac_of_nominal_8TBs": upd_gbs / 8000.0,
                                       "traffic": traffic[1], "bytes_per_launch": upd_bytes, "peak_source": f"MEASURED_PEAKS.json ({peak_src})",
                                       ("l2_frac_in_graph" if in_l2 else "frac_in_graph"):
                                           None if g_us is None or g_us[1] <= 0 else upd_bytes / (g_us[1] * 1e-6) / 1e9 / peaks["hbm_gbs"],
                                       "note": ("in the graph eps (smaller than the 126 MB L2) was just touched by the rollout: that figure is an L2 rate, not HBM evidence"
                                                if in_l2 else "eps exceeds the L2: the in-graph figure is an HBM rate")}
        if parity_n is not None:
            line["parity_n"] = parity_n
        if extra_cfg4 is not None:
            line["extra"] = {"cfg4": extra_cfg4}
        if not a.no_cpu_baseline:
            cb, _, _ = cpu_baseline(a.config if a.config != 4 else 0)
            line["cpu_baseline"] = cb
            line["guard_band_samples"] = {"count": cb["guard_band_samples"], "of": cb["guard_band_of"], "band": 1e-4}
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
