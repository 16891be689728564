"""Circle-obstacle episodes on the reference's own scenario file:
PathPlanning/Scenarios/obstacle_field.mat holds 100 random fields of 5-12 circles on a 120 m x 120 m
square, used by the reference's A*/RRT/Hybrid-A* demos with start (0, 0) and goal (120, 120)
(Astar/main.jl:11-22).  Here each field is one device-resident MPPI episode (mppi_run_closed_loop)
of the kinematic bicycle from (0, 0) heading 45 degrees towards (120, 120).  The lane term 10*y^2 of
vehicledynamics.jl:52 belongs to the straight-road scenario and is switched off (stage_w[5] = 0).

usage: python tools/obstacle_field_bench.py [file.mat] [K]   (default: the fixture subset, K = 65536)"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import motionplanning_b200 as m  # noqa: E402
from motionplanning_b200.scenario_io import read_obstacle_field  # noqa: E402

capi = m.capi
path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "obstacle_field_subset.mat")
K = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
trials = read_obstacle_field(path)
N, T = 50, 5.0
X0 = [0.0, 0.0, np.pi / 4, 5.0, 0.0]
XL, XU = [-5, -5, -4 * np.pi, 1.0, -0.5], [125, 125, 4 * np.pi, 10.0, 0.5]
mppi = m.defineMPPI(5, 2, 25.0, X0, [120.0, 120.0], K, [[0.05, 0], [0, 0.1]], N, T, XL, XU, [-0.5, -2.5], [0.5, 2.5],
                    model=capi.MODEL_KINEMATIC5, noise_mode=capi.NOISE_PHILOX, FeasibilityCount=K,
                    stage_w=[1.0, 1.0, 5.0, 3.0, 2.0, 0.0])
ok = 0
for i, circles in enumerate(trials):
    m.defineMPPIobs(mppi, [list(c) for c in circles])
    if i == 0:
        mppi.run_closed_loop(max_iters=2)   # graph capture
    m.ShiftInitialCondition(mppi, X0)
    m.defineMPPINominalControl(mppi, np.zeros((N, 2)))
    t0 = time.perf_counter()
    his, log = mppi.run_closed_loop(max_iters=400)
    dt = time.perf_counter() - t0
    d = np.hypot(his[:, 1][:, None] - circles[:, 0], his[:, 2][:, None] - circles[:, 1]) - circles[:, 2]
    reached = (his[-1, 1] - 120.0) ** 2 + (his[-1, 2] - 120.0) ** 2 <= 36.0
    clear = d.min() > 0.0
    ok += int(reached and clear)
    print(f"trial {i + 1}: {len(circles)} circles, {len(log)} iterations, simulated {his[-1, 0]:.1f} s, reached={bool(reached)}, "
          f"min clearance {d.min():.2f} m, wall {dt * 1e3:.1f} ms")
print(f"{ok}/{len(trials)} episodes reached the goal without touching an obstacle (K = {K}, N = {N})")
