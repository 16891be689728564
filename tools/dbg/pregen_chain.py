import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi = m.capi
K = int(sys.argv[1]); pre = int(sys.argv[2]); G = int(sys.argv[3]) if len(sys.argv) > 3 else 0
N = 50
args, kw, circles, blocks = sc.config(0, K=K)
kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, use_cuda_graph=int(os.environ.get('UG', '1')))
def mk(**e):
    mp = m.defineMPPI(*args, **dict(kw, **e)); m.defineMPPIobs(mp, circles); return mp
a = mk(rollout_variant=1, **(dict(n_gpus=G, device_ids=[0]*G) if G else {})); b = mk(rollout_variant=3)
print(a.plan_info(), b.plan_info())
for it in range(pre):
    m.MPPIPlan(a); m.MPPIPlan(b)
    m.defineMPPINominalControl(a, a.r.Control); m.defineMPPINominalControl(b, b.r.Control)
for i in range(4):
    a.step_async(adopt=(i > 0 or pre > 0)); b.step_async(adopt=(i > 0 or pre > 0))
    a.sync(); b.sync(); a.fetch_result(); b.fetch_result()
    ca, cb = a.p.costs, b.p.costs
    Gs = max(G, 1)
    per = [float(np.abs(ca[g*K//Gs:(g+1)*K//Gs] - cb[g*K//Gs:(g+1)*K//Gs]).max() / np.abs(cb).max()) for g in range(Gs)]
    print(i, a.r.argmin, b.r.argmin, a.r.FeasibleTrajCount, b.r.FeasibleTrajCount, np.abs(a.r.Control-b.r.Control).max(), "per-shard cost diff", ["%.1e" % x for x in per])
