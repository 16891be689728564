import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi = m.capi
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
def run(K, mode):
    args, kw, circles, blocks = sc.config(1, K=K)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
    mp = m.defineMPPI(*args, **kw); m.defineMPPIobs(mp, circles)
    st = torch.cuda.ExternalStream(mp.stream())
    for _ in range(5): mp.step_async(adopt=False)
    mp.sync()
    n = 60
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]; e1 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
    for i in range(n):
        with torch.cuda.stream(st):
            if mode == "flush": flush.zero_()
        e0[i].record(st); mp.step_async(adopt=False); e1[i].record(st)
    mp.sync()
    us = np.array([e0[i].elapsed_time(e1[i]) * 1e3 for i in range(n)])[5:]
    # amortised: 50 steps between one pair of events
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for i in range(50): mp.step_async(adopt=False)
    b.record(st); mp.sync()
    print(f"K={K:6d} mode={mode:6s} events={os.environ.get('MPPI_STEP_EVENTS', 'off')}: per-step median {np.median(us):6.1f} us  min {us.min():6.1f} | 50 back-to-back steps: {a.elapsed_time(b) * 1e3 / 50:6.1f} us per step", flush=True)
    mp.close()
for K in (64, 65536):
    for mode in ("flush", "none"):
        run(K, mode)
