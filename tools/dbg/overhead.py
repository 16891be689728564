import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi = m.capi
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
small = torch.empty(1024, dtype=torch.uint8, device="cuda")
def run(K, graph, mode, policy=0):
    args, kw, circles, blocks = sc.config(1, K=K)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, use_cuda_graph=graph, noise_policy=policy)
    mp = m.defineMPPI(*args, **kw); m.defineMPPIobs(mp, circles)
    st = torch.cuda.ExternalStream(mp.stream())
    for _ in range(5): mp.step_async(adopt=False)
    mp.sync()
    n = 60
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]; e1 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
    for i in range(n):
        with torch.cuda.stream(st):
            if mode == "flush": flush.zero_()
            elif mode == "small": small.zero_()
        e0[i].record(st); mp.step_async(adopt=False); e1[i].record(st)
        if mode == "sync": mp.sync()
    mp.sync()
    us = np.array([e0[i].elapsed_time(e1[i]) * 1e3 for i in range(n)])[5:]
    print(f"K={K:6d} graph={graph} mode={mode:6s} policy={policy}: median {np.median(us):6.1f} us  min {us.min():6.1f}  p90 {np.percentile(us, 90):6.1f}", flush=True)
    mp.close()
for K in (64, 65536):
    for graph in (1, 0):
        for mode in ("flush", "small", "none", "sync"):
            run(K, graph, mode)
run(65536, 1, "flush", 2)
run(65536, 1, "none", 2)
