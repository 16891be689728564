"""The reference's own driver scenario (OptimalControl/MPPI/main.jl:7-30: dynamic bicycle, three circles,
N = 20, T = 3 s, replanning every 0.1 s, 1 ms plant) as ONE device-resident episode, at the reference's
sample count (1,500) and at the headline's (65,536).  Prints iterations, final pose, wall time."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import motionplanning_b200 as m  # noqa: E402

capi = m.capi
X0 = [0, 0, 0, 0, 0, 5, 0]                                    # main.jl:7
XL = [-10, -20, -2, -np.pi / 2, -np.pi / 2, 1, -np.pi / 9]    # :8
XU = [130, 20, 2, np.pi / 2, np.pi / 2, 10, np.pi / 9]        # :9
for K in (1500, 65536):
    mppi = m.defineMPPI(7, 2, 25.0, X0, [110, 0], K, [[0.05, 0], [0, 0.1]], 20, 3.0, XL, XU, [-0.5, -2.5], [0.5, 2.5],
                        model=capi.MODEL_DYNAMIC7, noise_mode=capi.NOISE_PHILOX, FeasibilityCount=max(K, 1300))
    m.defineMPPIobs(mppi, [[50.0, 1, 2.5], [70.0, -1, 2.5], [90.0, 1, 2.5]])   # :30
    mppi.run_closed_loop(max_iters=2)                                           # graph capture, warm-up
    m.ShiftInitialCondition(mppi, X0)
    m.defineMPPINominalControl(mppi, np.zeros((20, 2)))
    t0 = time.perf_counter()
    his, log = mppi.run_closed_loop()
    dt = time.perf_counter() - t0
    d = np.hypot(his[:, 1][:, None] - np.array([50.0, 70.0, 90.0]), his[:, 2][:, None] - np.array([1.0, -1.0, 1.0])).min()
    print(f"K={K}: {len(log)} MPC iterations, {len(his) - 1} plant steps, simulated {his[-1, 0]:.3f} s, "
          f"final (x, y) = ({his[-1, 1]:.2f}, {his[-1, 2]:.2f}), closest approach to an obstacle centre {d:.2f} m (r = 2.5), "
          f"feasible plans {int(log[:, 1].sum())}/{len(log)}; wall {dt * 1e3:.2f} ms = {dt / len(log) * 1e6:.1f} us per iteration")
    if K == 1500:
        m.write_trajectory_csv("gpurun_out/MPPITrajectory_b200.csv", his)
