import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi = m.capi
G = int(sys.argv[1]); sync_between = int(sys.argv[2]); pre = int(sys.argv[3])
K, N = 8192, 50
args, kw, circles, blocks = sc.config(0, K=K)
kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=1)
def mk(**e):
    mp = m.defineMPPI(*args, **dict(kw, **e)); m.defineMPPIobs(mp, circles); return mp
mp = mk(n_gpus=G, device_ids=[0]*G); one = mk()
for it in range(pre):
    m.MPPIPlan(mp); m.MPPIPlan(one)
    m.defineMPPINominalControl(mp, mp.r.Control); m.defineMPPINominalControl(one, one.r.Control)
if pre: print("pre", mp.r.argmin, one.r.argmin, np.abs(mp.r.Control-one.r.Control).max())
for i in range(4):
    mp.step_async(adopt=(i > 0 or pre > 0)); one.step_async(adopt=(i > 0 or pre > 0))
    if sync_between:
        mp.sync(); one.sync()
        mp.fetch_result(); one.fetch_result()
        print(i, mp.r.argmin, one.r.argmin, mp.r.FeasibleTrajCount, one.r.FeasibleTrajCount, np.abs(mp.r.Control-one.r.Control).max(), np.array_equal(mp.noise(), one.noise()))
mp.sync(); one.sync(); mp.fetch_result(); one.fetch_result()
print("end", mp.r.argmin, one.r.argmin, mp.r.FeasibleTrajCount, one.r.FeasibleTrajCount, np.abs(mp.r.Control-one.r.Control).max())
