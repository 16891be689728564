"""The five BASELINE.json configurations as synthetic scenes.

Pure host-side data: positional arguments for defineMPPI plus obstacle lists.
Used by tests/ and bench.py so both run the same workloads.  Deliberately import-free (no
`_capi`, so no product library): bench.py's reference arm loads this file by path.
"""
import math

import numpy as np

MODEL_DYNAMIC7, MODEL_KINEMATIC5 = 0, 1   # include/mppi_b200.h

SEED = 0x5EED0001
CIRCLES = [[50.0, 1, 2.5], [70.0, -1, 2.5], [90.0, 1, 2.5]]  # MPPI/main.jl:30
PARKING_BLOCKS = [  # HybridAstar/main_hybrid_astar.jl:17
    [0.0, -1.0, 0.0, 5.5 / 2 + 1.0, 1.0],
    [-5.5 / 2, 2.7432 / 2, 0.0, 1.0, 2.7432 / 2],
    [5.5 / 2, 2.7432 / 2, 0.0, 1.0, 2.7432 / 2],
]
CL, CU = [-0.5, -2.5], [0.5, 2.5]  # MPPI/main.jl:10-11
SIGMA = [[0.05, 0.0], [0.0, 0.1]]  # MPPI/main.jl:27
LAMBDA = 25.0  # MPPI/main.jl:24
GOAL = [110.0, 0.0]  # MPPI/main.jl:23
PI = math.pi
# MPPI/main.jl:7-9 (7-state) and the matching 5-state subset
X0_DYN = [0, 0, 0, 0, 0, 5, 0]
XL_DYN = [-10, -20, -2, -PI / 2, -PI / 2, 1, -PI / 9]
XU_DYN = [130, 20, 2, PI / 2, PI / 2, 10, PI / 9]
X0_KIN = [0, 0, 0, 5, 0]
XL_KIN = [-10, -20, -PI / 2, 1, -PI / 9]
XU_KIN = [130, 20, PI / 2, 10, PI / 9]


def wall_blocks(M_p, seed=1234):
    """cfg 4 walls: the three circles of cfg 1 as blocks, plus M_p-3 random blocks that
    leave the y = 0 corridor passable (SURVEY.md 8(d) cfg 4)."""
    blocks = [[50.0, 1, 0, 2.5, 1.25], [70.0, -1, 0, 2.5, 1.25], [90.0, 1, 0, 2.5, 1.25]]
    rng = np.random.RandomState(seed)
    while len(blocks) < M_p:
        x, y = rng.uniform(20, 100), rng.uniform(-8, 8)
        psi = rng.uniform(-PI, PI)
        hl, hw = rng.uniform(0.5, 2.5), rng.uniform(0.5, 2.5)
        if abs(y) - math.hypot(hl, hw) < 3.0:  # keep the corridor |y| < 3 free
            continue
        blocks.append([x, y, psi, hl, hw])
    return blocks[:M_p]


def config(idx, K=None, N=None, **kw):
    """Return (args, kwargs, circles, blocks) for defineMPPI reproducing BASELINE.json configs[idx]."""
    if idx in (0, 1, 4):  # kinematic + circles
        k = {0: 1024, 1: 65536, 4: 1024}[idx]
        n = 50
        K, N = K or k, N or n
        args = (5, 2, LAMBDA, X0_KIN, GOAL, K, SIGMA, N, 0.1 * N, XL_KIN, XU_KIN, CL, CU)
        kwargs = dict(model=MODEL_KINEMATIC5, FeasibilityCount=K, seed=SEED)
        if idx == 4:
            kwargs["B"] = 4096
        kwargs.update(kw)
        return args, kwargs, CIRCLES, []
    if idx == 2:  # parking layout, kinematic + polygon walls
        K, N = K or 262144, N or 80
        x0 = [7, 0, PI / 2, 0.5, 0]  # main_hybrid_astar.jl:15
        xl = [-5, 0, -PI, -2, -PI / 6]  # :29, :22
        xu = [10, 10, PI, 2, PI / 6]
        args = (5, 2, LAMBDA, x0, [0.0, 0.5], K, SIGMA, N, 0.1 * N, xl, xu, CL, CU)
        kwargs = dict(model=MODEL_KINEMATIC5, FeasibilityCount=K, seed=SEED, veh_L=3.0, veh_W=2.0,
                      veh_center_offset=1.5)
        kwargs.update(kw)
        return args, kwargs, [], PARKING_BLOCKS
    if idx == 3:  # dynamic + polygon walls
        K, N = K or 1048576, N or 100
        M_p = kw.pop("M_p", 3)
        args = (7, 2, LAMBDA, X0_DYN, GOAL, K, SIGMA, N, 0.05 * N, XL_DYN, XU_DYN, CL, CU)
        kwargs = dict(model=MODEL_DYNAMIC7, FeasibilityCount=K, seed=SEED)
        kwargs.update(kw)
        return args, kwargs, [], wall_blocks(M_p)
    raise ValueError(idx)


def jittered_x0(B, seed=7):
    """cfg 5 initial states: x~U[0,40], y~U[-2,2], psi~U[-0.2,0.2], ux~U[3,8], sa=0; shape (5, B)."""
    rng = np.random.RandomState(seed)
    return np.stack([rng.uniform(0, 40, B), rng.uniform(-2, 2, B), rng.uniform(-0.2, 0.2, B),
                     rng.uniform(3, 8, B), np.zeros(B)])


# algorithmic flop per sample-step, SURVEY.md 8(a) closed forms
def flop_per_sample_step(model, n_circles, n_blocks):
    if model == MODEL_KINEMATIC5:
        base = 83 if n_blocks == 0 else 121
    else:
        base = 201 if n_blocks == 0 else 239
    return base + 7 * n_circles + 102 * n_blocks
