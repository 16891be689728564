"""Whole-step device time (graph replay, L2 flushed before every step) for the rollout variants:
3 = Philox inside the rollout (one sample per thread), 1 = pre-generated noise + one sample per thread,
2 = pre-generated noise + packed two-sample threads, 4 = packed with Philox inside the rollout."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc

capi = m.capi
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
cases = [("kin+circles", 1, {}), ("dyn+walls3", 3, {})]
Ks = [int(k) for k in (sys.argv[1:] or [1024, 4096, 16384, 32768, 65536, 131072, 262144])]
for name, idx, extra in cases:
    for K in Ks:
        row = []
        for variant, env in ((3, "1"), (1, "1"), (2, "1"), (4, "1")):
            os.environ["MPPI_PREGEN"] = env
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
            ts = []
            for _ in range(14):
                flush.zero_(); torch.cuda.synchronize()
                mp.step_async(); mp.sync(); ts.append(mp.timings_us()[3])
            row.append(float(np.median(ts[3:])))
            mp.close()
        print(f"{name:12s} K={K:8d} N={args[7]:3d} step us: philox-in-rollout {row[0]:8.1f} | pregen scalar {row[1]:8.1f} | pregen packed {row[2]:8.1f} | packed, philox in rollout {row[3]:8.1f}", flush=True)
