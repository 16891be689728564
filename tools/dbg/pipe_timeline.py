import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi = m.capi
args, kw, circles, blocks = sc.config(1)
kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
mp = m.defineMPPI(*args, **kw); m.defineMPPIobs(mp, circles)
print(mp.plan_info())
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
for i in range(3):
    flush.zero_(); torch.cuda.synchronize()
    print("--- step", i, "flushed", flush=True)
    mp.step_async(adopt=(i > 0)); mp.sync()
for i in range(2):
    print("--- step warm L2", i, flush=True)
    mp.step_async(adopt=True); mp.sync()
