"""Rollout kernel time (eager, CUDA events) for one-sample vs packed two-sample threads."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc

capi = m.capi
cases = [("kin+circles", 1, {}), ("kin+walls", 2, {}), ("dyn+walls3", 3, {}), ("dyn+walls16", 3, {"M_p": 16})]
Ks = [int(k) for k in (sys.argv[1:] or [65536, 131072, 262144, 1048576])]
for name, idx, extra in cases:
    for K in Ks:
        row = []
        for variant in (1, 2):
            args, kw, circles, blocks = sc.config(idx, K=K, **extra)
            kw.update(precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=variant)
            mp = m.defineMPPI(*args, **kw)
            if circles:
                m.defineMPPIobs(mp, circles)
            if blocks:
                m.defineMPPIblocks(mp, blocks)
            m.MPPIPlan(mp)
            for _ in range(3):
                mp.step_async(adopt=True)
            mp.sync(); mp.adopt_nominal()
            mp.set_profiling(1)
            ts = []
            for _ in range(12):
                mp.step_async(); mp.sync(); ts.append(mp.timings_us())
            t = np.median(np.array(ts[3:]), axis=0)
            row.append(t[0])
            mp.close()
        N = args[7]
        print(f"{name:12s} K={K:8d} N={N:3d} rollout us: single {row[0]:9.1f}  packed {row[1]:9.1f}  ratio {row[0]/row[1]:.2f}  "
              f"ns/sample-step {row[0]*1e3/(K*N):.4f} / {row[1]*1e3/(K*N):.4f}", flush=True)
