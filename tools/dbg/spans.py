import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi = m.capi
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
def run(K, variant):
    args, kw, circles, blocks = sc.config(1, K=K)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=variant)
    try:
        mp = m.defineMPPI(*args, **kw)
    except Exception as e:
        print(K, variant, "n/a"); return
    m.defineMPPIobs(mp, circles)
    for _ in range(5): mp.step_async(adopt=False)
    mp.sync()
    sp = []
    for i in range(12):
        flush.zero_(); torch.cuda.synchronize()
        mp.set_profiling(2); mp.step_async(adopt=False); mp.sync()
        c = mp.phase_clocks()
        f = lambda j: (c[9 + 2 * j] - c[8 + 2 * j]) * 1e-3
        sp.append([f(0), (c[10] - c[9]) * 1e-3, f(1), (c[12] - c[11]) * 1e-3, f(2), (c[13] - c[8]) * 1e-3, f(3)])
    mp.set_profiling(0)
    s = np.median(np.array(sp[2:]), axis=0)
    # finalize phase clocks (cycles)
    ph = [c[i + 1] - c[i] for i in range(6)]
    print(f"K={K:6d} v={variant} info={mp.plan_info()['packed']} rollout {s[0]:5.1f} gap {s[1]:4.1f} update {s[2]:4.1f} gap {s[3]:4.1f} finalize {s[4]:4.1f} | first->last {s[5]:5.1f} noise {s[6]:4.1f} | finalize phases (cycles) {ph}", flush=True)
    mp.close()
for K in (64, 4096, 16384, 65536):
    for v in (1, 2, 6):
        run(K, v)
