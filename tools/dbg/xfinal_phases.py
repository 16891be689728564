import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi = m.capi
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
for G, fl in ((1, 1), (2, 1), (2, 0), (1, 0)):
    args, kw, circles, blocks = sc.config(1, K=65536 * G)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
    if G > 1: kw.update(n_gpus=G, device_ids=[0] * G)
    mp = m.defineMPPI(*args, **kw); m.defineMPPIobs(mp, circles)
    for _ in range(10): mp.step_async(adopt=False)
    mp.sync()
    ph = []
    for i in range(40):
        if fl: flush.zero_(); torch.cuda.synchronize()
        mp.step_async(adopt=False); mp.sync()
        c = mp.phase_clocks(); ph.append([c[j + 1] - c[j] for j in range(6)])
    ph = np.median(np.array(ph[5:]), axis=0)
    print(f"G={G} flush={fl}: partial {ph[0]:.0f} xchg+combine {ph[1]:.0f} U_new {ph[2]:.0f} states {ph[3]:.0f} costing {ph[4]:.0f} sum/write {ph[5]:.0f}", flush=True)
    mp.close()
