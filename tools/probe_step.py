import sys, numpy as np
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc
capi=m.capi
args,kw,circles,blocks=sc.config(1)
kw.update(precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, use_cuda_graph=int(os.environ.get('PROBE_GRAPH', '1')))
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
mp.set_profiling(2); mp.step_async(); mp.sync()   # first traced replay captures the graph
spans=[]
for _ in range(20):
    mp.set_profiling(2); mp.step_async(); mp.sync(); c=mp.phase_clocks()
    t0=c[8]; spans.append([(c[8+2*i]-t0, c[9+2*i]-t0) for i in range(3)])
print('in-graph spans ns (start,end) rollout/update/finalize, median:', np.median(np.array(spans),axis=0).tolist())
mp.set_profiling(0)
ck=mp.phase_clocks()
print('finalize phase cycles', [ck[i+1]-ck[i] for i in range(6)], 'total', ck[6]-ck[0])

# ---- host-side cost of the three reference-facing calls (main.jl:59-61) ----
import time
mp.set_profiling(0)
x = np.array(mp.s.X0, dtype=np.float64); U = np.array(mp.r.Control, copy=True)
T = {k: [] for k in ("shift", "nominal", "plan", "raw_step")}
for i in range(300):
    t0 = time.perf_counter(); m.ShiftInitialCondition(mp, x)
    t1 = time.perf_counter(); m.defineMPPINominalControl(mp, U)
    t2 = time.perf_counter(); m.MPPIPlan(mp)
    t3 = time.perf_counter()
    T["shift"].append(t1 - t0); T["nominal"].append(t2 - t1); T["plan"].append(t3 - t2)
import ctypes as C
res = capi.Result()
for i in range(300):
    t0 = time.perf_counter(); mp._lib.mppi_step(mp._h, None, C.byref(res)); T["raw_step"].append(time.perf_counter() - t0)
print({k: round(float(np.median(v[50:])) * 1e6, 2) for k, v in T.items()}, "device us", mp.timings_us()[3])
