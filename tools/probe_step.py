import sys, numpy as np
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi=m.capi
args,kw,circles,blocks=sc.config(1)
kw.update(precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
mp=m.defineMPPI(*args,**kw); m.defineMPPIobs(mp,circles)
m.MPPIPlan(mp)
for _ in range(4): mp.step_async(adopt=True)
mp.sync(); mp.adopt_nominal()
for mode in (0,1):
    mp.set_profiling(mode)
    ts=[]
    for _ in range(30):
        mp.step_async(); mp.sync(); ts.append(mp.timings_us())
    print('mode',mode,np.median(np.array(ts[5:]),axis=0))
ck=mp.phase_clocks(); print('finalize phase cycles', [ck[i+1]-ck[i] for i in range(6)], 'total', ck[6]-ck[0])
