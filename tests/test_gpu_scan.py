"""The scan-form fused step (csrc/mppi_scan.cuh: horizon-parallel kinematic rollout + online softmin +
weighted control sum + in-kernel tail) against the oracle and against the sequential kernels.

Reference semantics reproduced: OptimalControl/Parking/src/VehicleModel.jl:14-18 (model),
OptimalControl/MPPI/src/MPPIUtils.jl:31-57 (rollout and cost), :76-90 (Heun), :120-151 (tests),
:154-167 and :186-190 (softmin weights and weighted update).  The scan form changes only the ORDER of the
running sums, so costs agree with the FP64 oracle to the FP32 tolerance of north_star (1e-4 relative)."""
import numpy as np
import pytest

import motionplanning_b200 as m
import oracle_lib as orc
from motionplanning_b200 import scenarios as sc
from test_gpu_parity import GUARD, check, run_pair, warm_nominal

capi = m.capi
pytestmark = pytest.mark.gpu


def _args(K, N, x0=None):
    args, kw, circles, blocks = sc.config(0, K=K, N=N)
    if x0 is not None:
        args = args[:3] + (x0,) + args[4:]
    return args, kw, circles, blocks


@pytest.mark.parametrize("K,N", [(1024, 50), (1021, 50), (4096, 23), (2048, 56), (2050, 80), (1024, 100), (515, 7), (64, 1)])
def test_scan_step_matches_oracle(built, K, N):
    """Whole MPPIPlan on supplied noise: ragged K (not a multiple of the 4-sample warp tile), horizons that fill, underfill
    and overflow the 7-step chunk (N = 56 is exactly 8 x 7; N > 56 uses the 13-step chunk), N = 1."""
    args, kw, circles, blocks = _args(K, N)
    mppi, ref = run_pair(args, dict(kw, rollout_variant=5), circles, blocks, capi.FP32, Unom=warm_nominal(N))
    assert mppi.plan_info()["scan"] == 1
    check(mppi, ref, capi.FP32)


@pytest.mark.parametrize("n_circles", [0, 1, 3, 5, 8])
def test_scan_circle_counts_and_collisions(built, n_circles):
    """Start next to the first obstacle so that hits, out-of-bounds samples and feasible samples all occur."""
    rng = np.random.RandomState(5)
    circles = [[50.0, 1, 2.5], [70.0, -1, 2.5], [90.0, 1, 2.5]] + [[float(rng.uniform(40, 100)), float(rng.uniform(-6, 6)), float(rng.uniform(0.5, 2))]
                                                                  for _ in range(5)]
    args, kw, _, blocks = _args(4096, 50, x0=[38.0, -1.0, 0.05, 8.0, 0.02])
    mppi, ref = run_pair(args, dict(kw, rollout_variant=5), circles[:n_circles], blocks, capi.FP32, Unom=warm_nominal(50))
    if n_circles >= 3:
        assert (ref["flags"] & capi.FLAG_COLLIDED).any() and (ref["flags"] & capi.FLAG_FEASIBLE).any() and (ref["flags"] & capi.FLAG_OOB).any()
    check(mppi, ref, capi.FP32)


def test_scan_equals_sequential_kernels(built):
    """Same supplied noise through the scan kernel and through rollout_kernel + update_kernel + finalize_kernel."""
    args, kw, circles, blocks = _args(8192, 50)
    U = warm_nominal(50)
    a, ref = run_pair(args, dict(kw, rollout_variant=5), circles, blocks, capi.FP32, Unom=U)
    b, _ = run_pair(args, dict(kw, rollout_variant=1), circles, blocks, capi.FP32, Unom=U)
    assert a.plan_info()["scan"] == 1 and b.plan_info()["scan"] == 0
    ca, cb = a.p.costs, b.p.costs
    assert (np.abs(ca - cb) / np.abs(cb)).max() < 2e-5
    clear = ref["margins"] > GUARD
    assert np.array_equal(a.p.flags[clear], b.p.flags[clear])
    assert a.r.argmin == b.r.argmin == ref["argmin"]
    assert np.abs(a.r.Control - b.r.Control).max() < 2e-6 * max(1.0, np.abs(b.r.Control).max())
    assert a.r.RolloutCount == b.r.RolloutCount and a.r.FeasibleTrajCount == b.r.FeasibleTrajCount


def test_scan_is_deterministic_and_graph_equals_eager(built):
    args, kw, circles, blocks = _args(16384, 50)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=5)
    outs = []
    for graph in (True, True, False):
        mp = m.defineMPPI(*args, **dict(kw, use_cuda_graph=graph))
        m.defineMPPIobs(mp, circles)
        for _ in range(3):
            m.MPPIPlan(mp)
            m.defineMPPINominalControl(mp, mp.r.Control)
        outs.append((mp.r.Control.copy(), mp.r.cost, mp.r.argmin, mp.p.costs.copy()))
    for o in outs[1:]:
        assert np.array_equal(o[0], outs[0][0]) and o[1] == outs[0][1] and o[2] == outs[0][2] and np.array_equal(o[3], outs[0][3])


@pytest.mark.parametrize("K", [4096, 65536])
def test_scan_philox_chain_matches_oracle_replay(built, K):
    """Device Philox noise (pre-generated behind the previous step), four chained iterations; every iteration's dumped
    noise replayed through the FP64 oracle.  One kernel per step + the noise kernel."""
    args, kw, circles, blocks = _args(K, 50)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=5)
    mp = m.defineMPPI(*args, **kw)
    m.defineMPPIobs(mp, circles)
    info = mp.plan_info()
    assert info["scan"] == 1 and info["pregen"] == 1
    U = np.zeros((50, 2))
    cfg = orc.make_config(args, dict(kw, precision=capi.FP64, noise_mode=capi.NOISE_SUPPLIED))
    for it in range(4):
        m.defineMPPINominalControl(mp, U)
        n0 = mp.launch_count()
        m.MPPIPlan(mp)
        n1 = mp.launch_count()
        assert n1 - n0 <= 4   # begin + scan (tail inside) + next noise + this iteration's noise (redrawn: mp.noise() consumed the buffer)
        eps = mp.noise().astype(np.float64)
        sub = slice(0, 4096)
        ref = orc.plan(orc.make_config(args[:5] + (4096,) + args[6:], dict(kw, precision=capi.FP64, noise_mode=capi.NOISE_SUPPLIED, FeasibilityCount=4096)),
                       circles, blocks, args[3], U, np.ascontiguousarray(eps[:, sub, :]), capi.FP64, threads=8, want_margin=True)
        costs = mp.p.costs[sub]
        ok = ref["margins"] > GUARD
        assert (~ok).sum() <= 0.02 * 4096
        assert (np.abs(costs - ref["costs"]) / np.abs(ref["costs"]))[ok].max() < 1e-4
        assert np.array_equal(mp.p.flags[sub][ok], ref["flags"][ok])
        if K == 4096:
            full = orc.plan(cfg, circles, blocks, args[3], U, eps, capi.FP64, threads=8)
            assert np.abs(mp.r.Control - full["Control"]).max() <= 1e-4 * max(1.0, np.abs(full["Control"]).max())
            assert mp.r.argmin == full["argmin"]
        U = mp.r.Control.copy()


def test_scan_profiling_split_tail_is_bit_identical(built):
    """Profiling mode 1 launches the tail as its own kernel (finalize_kernel<PH_FOLD | PH_COMBINE>) so that events can
    bracket the scan kernel alone; the arithmetic is the same block function."""
    args, kw, circles, blocks = _args(8192, 50)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=5)
    res = []
    for prof in (0, 1):
        mp = m.defineMPPI(*args, **kw)
        m.defineMPPIobs(mp, circles)
        mp.set_profiling(prof)
        for _ in range(2):
            mp.step_async(adopt=False)
            mp.sync()
        mp.fetch_result()
        res.append((mp.r.Control.copy(), mp.r.cost))
        if prof:
            us = mp.timings_us()
            assert us[0] > 0 and us[2] > 0
    assert np.array_equal(res[0][0], res[1][0]) and res[0][1] == res[1][1]
