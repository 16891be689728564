"""In-graph spans and finalize phase clocks of the headline step with / without the pipeline kernel's in-block softmin update."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi = m.capi
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
def run(K, fold):
    os.environ["MPPI_PIPE_FOLD"] = str(fold)
    args, kw, circles, blocks = sc.config(1, K=K)
    mp = m.defineMPPI(*args, **dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX))
    m.defineMPPIobs(mp, circles)
    for _ in range(5): mp.step_async(adopt=False)
    mp.sync()
    rows, phs = [], []
    for i in range(14):
        flush.zero_(); torch.cuda.synchronize()
        mp.set_profiling(2); mp.step_async(adopt=False); mp.sync()
        c = mp.phase_clocks()
        rows.append([(c[9] - c[8]) * 1e-3, (c[10] - c[8]) * 1e-3, (c[11] - c[8]) * 1e-3, (c[12] - c[8]) * 1e-3, (c[13] - c[8]) * 1e-3, (c[15] - c[8]) * 1e-3])
        phs.append([c[i + 1] - c[i] for i in range(6)])
    mp.set_profiling(0)
    s = np.median(np.array(rows[2:]), axis=0); ph = np.median(np.array(phs[2:]), axis=0)
    print(f"K={K} fold={fold} launches/step={mp.launch_count()}: rollout phase ends {s[0]:5.1f} | slot1 (update / tail) {s[1]:5.1f}..{s[2]:5.1f} | finalize {s[3]:5.1f}..{s[4]:5.1f} | noise ends {s[5]:5.1f} | finalize phases (cycles) {ph.astype(int).tolist()}", flush=True)
    mp.close()
for K in (65536, 16384):
    for fold in (0, 1, 0, 1):
        run(K, fold)
