"""Step time of ONE handle driving G GPUs from one host thread (cfg.n_gpus; headline config, 65,536 samples per GPU):
device-side span per step measured with CUDA events on rank 0's stream around mppi_step_async, and the wall time of the
synchronous reference-facing calls (ShiftInitialCondition + defineMPPINominalControl + MPPIPlan)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc

capi = m.capi
out = []
for G in [int(a) for a in (sys.argv[1:] or [1, 2, 4, 8])]:
    if G > torch.cuda.device_count():
        continue
    args, kw, circles, blocks = sc.config(1, K=65536 * G)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
    if G > 1:
        kw.update(n_gpus=G, device_ids=list(range(G)))
    mp = m.defineMPPI(*args, **kw)
    m.defineMPPIobs(mp, circles)
    m.MPPIPlan(mp)
    for _ in range(10):
        mp.step_async(adopt=True)
    mp.sync()
    mp.adopt_nominal()
    st = torch.cuda.ExternalStream(mp.stream(), device=torch.device("cuda", 0))
    n = 200
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
    for i in range(n):
        e0[i].record(st)
        mp.step_async(adopt=False)
        e1[i].record(st)
        mp.sync()
    dev = np.array([e0[i].elapsed_time(e1[i]) * 1e3 for i in range(n)])[20:]
    U = mp.r.Control.copy()
    x0 = np.asarray(args[3], dtype=np.float64)
    wall = []
    for i in range(n):
        t0 = time.perf_counter()
        m.ShiftInitialCondition(mp, x0)
        m.defineMPPINominalControl(mp, U)
        m.MPPIPlan(mp)
        wall.append(time.perf_counter() - t0)
    wall = np.array(wall[20:]) * 1e6
    row = {"n_gpus": G, "K_total": 65536 * G, "N": 50, "device_us_rank0_median": float(np.median(dev)), "e2e_us_median": float(np.median(wall)),
           "e2e_us_p99": float(np.percentile(wall, 99)), "sample_steps_per_s_e2e": 65536 * G * 50 / (np.median(wall) * 1e-6)}
    print(json.dumps(row), flush=True)
    out.append(row)
    mp.close()
