import sys, os, numpy as np
sys.path.insert(0, os.getcwd())
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi=m.capi
def run(noise_mode):
    args,kw,circles,blocks=sc.config(1)
    kw.update(precision=capi.FP32, noise_mode=noise_mode)
    mp=m.defineMPPI(*args,**kw); m.defineMPPIobs(mp,circles)
    K,N=args[5],args[7]
    eps=None
    if noise_mode==capi.NOISE_SUPPLIED:
        eps=(np.random.default_rng(0).standard_normal((N,K,2))*np.sqrt([0.05,0.1])).astype(np.float32)
    m.MPPIPlan(mp,eps)
    for mode in (1,2):
        ts=[]
        for _ in range(12):
            mp.set_profiling(mode)
            m.MPPIPlan(mp,eps)
            if mode==1: ts.append(mp.timings_us())
            else:
                c=mp.phase_clocks(); t0=c[8]; ts.append([c[9+2*i]-c[8+2*i] for i in range(3)])
        print('noise_mode',noise_mode,'mode',mode,np.median(np.array(ts[3:]),axis=0))
run(capi.NOISE_PHILOX); run(capi.NOISE_SUPPLIED)
