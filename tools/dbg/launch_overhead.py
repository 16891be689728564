"""Per-step device time minus the kernels' own span (first CTA of the rollout -> end of the last kernel), i.e. what graph launch,
completion and the bench's two events cost, with and without the forked noise branch (MPPI_PREGEN=0/1 in the environment)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi = m.capi
for K in (64, 65536):
    args, kw, circles, blocks = sc.config(1, K=K)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
    mp = m.defineMPPI(*args, **kw); m.defineMPPIobs(mp, circles)
    st = torch.cuda.ExternalStream(mp.stream())
    for _ in range(5): mp.step_async(adopt=False)
    mp.sync()
    n = 60
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]; e1 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
    for i in range(n):
        e0[i].record(st); mp.step_async(adopt=False); e1[i].record(st); mp.sync()
    us = np.median([e0[i].elapsed_time(e1[i]) * 1e3 for i in range(10, n)])
    sp = []
    for i in range(20):
        mp.set_profiling(2); mp.step_async(adopt=False); mp.sync()
        c = mp.phase_clocks()
        last = max(c[13], c[15]) if (c[15] > 0 and c[14] < 2 ** 63) else c[13]
        sp.append((last - c[8]) * 1e-3)
    mp.set_profiling(0)
    print(f"K={K:6d} pregen={mp.plan_info()['pregen']} step {us:5.1f} us, kernels first->last {np.median(sp[3:]):5.1f} us, outside {us - np.median(sp[3:]):4.1f} us", flush=True)
    mp.close()
