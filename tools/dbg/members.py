import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi = m.capi
K = int(sys.argv[1]); pre = int(sys.argv[2]); G = int(sys.argv[3])
N = 50
args, kw, circles, blocks = sc.config(0, K=K)
kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
def mk(**e):
    mp = m.defineMPPI(*args, **dict(kw, **e)); m.defineMPPIobs(mp, circles); return mp
a = mk(rollout_variant=1, n_gpus=G, device_ids=[0]*G)
b = mk(rollout_variant=int(os.environ.get('BV', '3')))
for it in range(pre):
    m.MPPIPlan(a); m.MPPIPlan(b)
    m.defineMPPINominalControl(a, a.r.Control); m.defineMPPINominalControl(b, b.r.Control)
for i in range(4):
    a.step_async(adopt=(i > 0 or pre > 0)); b.step_async(adopt=(i > 0 or pre > 0))
    a.sync(); b.sync(); b.fetch_result()
    outs = []
    for g in range(G):
        os.environ["MPPI_DEBUG_MEMBER"] = str(g)
        a.fetch_result()
        outs.append((a.r.Control.copy(), a.r.argmin, a.r.rho, a.r.eta, a.r.RolloutCount))
    os.environ["MPPI_DEBUG_MEMBER"] = "0"
    print(i, "b:", b.r.argmin, [(o[1], "%.6f" % o[2], "%.6f" % o[3], o[4], "%.2e" % np.abs(o[0] - outs[0][0]).max()) for o in outs])
