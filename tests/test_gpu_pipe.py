"""The two-stage warp pipeline rollout (csrc/mppi_pipe.cuh: a producer warp walks controls / (sa, ux) / tan(sa)/L, a consumer
warp walks psi / x / y / tests a few steps behind, values handed over in place of the consumed eps in shared memory) against
the oracle and against the sequential kernels.

Reference semantics reproduced: OptimalControl/Parking/src/VehicleModel.jl:14-18 (model), OptimalControl/MPPI/src/
MPPIUtils.jl:13-20 (PushInBounds), :31-57 (rollout and cost), :76-90 (Heun), :120-151 (tests).  The pipeline changes only the
ORDER in which the cost terms are summed (two partial sums), so costs agree with the FP64 oracle to the FP32 tolerance of
north_star (1e-4 relative) and with the sequential kernel to FP32 rounding; flags are compared outside the guard band."""
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


@pytest.mark.parametrize("K,N", [(1024, 50), (1022, 50), (4096, 23), (2048, 56), (770, 8), (514, 7), (64, 1), (65536, 50)])
def test_pipe_rollout_matches_oracle(built, K, N):
    """Whole MPPIPlan on supplied noise: ragged K (the last CTA holds fewer than 256 samples, a last row copy of 16 bytes),
    horizons that end inside, on and one past a group of 8 steps, N = 1, and the headline size."""
    args, kw, circles, blocks = _args(K, N)
    mppi, ref = run_pair(args, dict(kw, rollout_variant=6), circles, blocks, capi.FP32, Unom=warm_nominal(N))
    assert mppi.plan_info()["packed"] == 2
    check(mppi, ref, capi.FP32)


@pytest.mark.parametrize("n_circles", [0, 1, 3, 5, 8])
def test_pipe_circle_counts_and_collisions(built, n_circles):
    """Start next to the first obstacle so that hits, out-of-bounds samples and feasible samples all occur."""
    rng = np.random.RandomState(5)
    circles = [[50.0, 1, 2.5], [70.0, -1, 2.5], [90.0, 1, 2.5]] + [[float(rng.uniform(40, 100)), float(rng.uniform(-6, 6)), float(rng.uniform(0.5, 2))]
                                                                  for _ in range(5)]
    args, kw, _, blocks = _args(4096, 50, x0=[38.0, -1.0, 0.05, 8.0, 0.02])
    mppi, ref = run_pair(args, dict(kw, rollout_variant=6), circles[:n_circles], blocks, capi.FP32, Unom=warm_nominal(50))
    assert mppi.plan_info()["packed"] == 2
    if n_circles >= 3:
        assert (ref["flags"] & capi.FLAG_COLLIDED).any() and (ref["flags"] & capi.FLAG_FEASIBLE).any() and (ref["flags"] & capi.FLAG_OOB).any()
    check(mppi, ref, capi.FP32)


def test_pipe_equals_sequential_kernels(built):
    """Same supplied noise through the pipeline and through the packed and the one-sample sequential rollout."""
    args, kw, circles, blocks = _args(8192, 50, x0=[38.0, -1.0, 0.05, 8.0, 0.02])
    U = warm_nominal(50)
    a, ref = run_pair(args, dict(kw, rollout_variant=6), circles, blocks, capi.FP32, Unom=U)
    for v in (1, 2):
        b, _ = run_pair(args, dict(kw, rollout_variant=v), circles, blocks, capi.FP32, Unom=U)
        assert a.plan_info()["packed"] == 2 and b.plan_info()["packed"] == v - 1
        ca, cb = a.p.costs, b.p.costs
        assert (np.abs(ca - cb) / np.abs(cb)).max() < 2e-6   # summation order only
        clear = ref["margins"] > GUARD
        assert np.array_equal(a.p.flags[clear], b.p.flags[clear])
        assert a.r.argmin == b.r.argmin == ref["argmin"]
        assert np.abs(a.r.Control - b.r.Control).max() < 2e-6 * max(1.0, np.abs(b.r.Control).max())
        assert a.r.RolloutCount == b.r.RolloutCount and a.r.FeasibleTrajCount == b.r.FeasibleTrajCount


def test_pipe_is_deterministic_and_graph_equals_eager(built):
    args, kw, circles, blocks = _args(16384, 50)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=6)
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
def test_pipe_philox_chain_matches_oracle_replay(built, K):
    """Device Philox noise (pre-generated behind the previous step), four chained iterations through the auto policy at the
    headline size; every iteration's dumped noise replayed through the FP64 oracle."""
    args, kw, circles, blocks = _args(K, 50)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=(0 if K >= 65536 else 6))
    mp = m.defineMPPI(*args, **kw)
    m.defineMPPIobs(mp, circles)
    info = mp.plan_info()
    assert info["packed"] == 2 and info["pregen"] == 1   # auto picks the pipeline at the headline size
    U = np.zeros((50, 2))
    cfg = orc.make_config(args, dict(kw, precision=capi.FP64, noise_mode=capi.NOISE_SUPPLIED))
    for it in range(4):
        m.defineMPPINominalControl(mp, U)
        m.MPPIPlan(mp)
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


def test_pipe_falls_back_where_it_does_not_apply(built):
    """Walls, a reference path, long horizons or the dynamic model keep the sequential kernels; forcing variant 6 there fails."""
    args, kw, circles, blocks = sc.config(2, K=4096)   # parking walls
    mp = m.defineMPPI(*args, **dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=2))
    m.defineMPPIblocks(mp, blocks)
    assert mp.plan_info()["packed"] == 1
    mp.close()
    args, kw, circles, blocks = _args(4096, 50)
    mp = m.defineMPPI(*args, **dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=6))
    assert mp.plan_info()["packed"] == 2
    m.defineMPPIblocks(mp, sc.PARKING_BLOCKS)           # a wall arrives later: the next step takes the sequential packed kernel
    assert mp.plan_info()["packed"] == 1
    m.MPPIPlan(mp)
    mp.close()
    # auto: the pipeline at every (even) sample count; below the packing threshold a step that cannot take it (walls) falls back to
    # one sample per thread, at or above it to the packed sequential kernel
    for K, fallback in ((1024, 0), (65536, 1)):
        args, kw, circles, blocks = _args(K, 50)
        mp = m.defineMPPI(*args, **dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX))
        m.defineMPPIobs(mp, circles)
        assert mp.plan_info()["packed"] == 2
        m.defineMPPIblocks(mp, sc.PARKING_BLOCKS)
        assert mp.plan_info()["packed"] == fallback
        m.MPPIPlan(mp)
        mp.close()
    args, kw, circles, blocks = _args(1023, 50)           # odd K cannot pack
    mp = m.defineMPPI(*args, **dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX))
    assert mp.plan_info()["packed"] == 0
    mp.close()
    with pytest.raises(m.MPPIError, match="two-stage warp pipeline"):
        m.defineMPPI(*_args(4096, 80)[0], **dict(_args(4096, 80)[1], precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=6))
    a7, k7, c7, b7 = sc.config(3, K=4096)
    with pytest.raises(m.MPPIError, match="two-stage warp pipeline"):
        m.defineMPPI(*a7, **dict(k7, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=6))


# ---- rollout_variant 7: the pipeline's blocks also do the softmin update (no update_kernel), finalize folds one partial per block ----
@pytest.mark.parametrize("K,N", [(1024, 50), (1022, 50), (4096, 23), (770, 8), (64, 1), (65536, 50)])
def test_pipe_with_in_block_update_matches_oracle(built, K, N):
    """Whole MPPIPlan on supplied noise (MPPIUtils.jl:154-167 weights, :186-190 weighted sum of the clamped controls): per-block
    weights against the block's own minimum, rescaled against the global one in finalize; ragged last block, N = 1."""
    args, kw, circles, blocks = _args(K, N)
    mppi, ref = run_pair(args, dict(kw, rollout_variant=7), circles, blocks, capi.FP32, Unom=warm_nominal(N))
    assert mppi.plan_info()["packed"] == 3
    check(mppi, ref, capi.FP32)


def test_pipe_in_block_update_equals_update_kernel(built):
    """Same noise through variant 7 and variant 6 (pipeline + update_kernel): identical costs and flags (same rollout code), U_new to
    FP32 rounding (different partition of the weighted sum), one kernel fewer per step; Philox chain of four iterations."""
    args, kw, circles, blocks = _args(16384, 50, x0=[38.0, -1.0, 0.05, 8.0, 0.02])
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
    a = m.defineMPPI(*args, **dict(kw, rollout_variant=7))
    b = m.defineMPPI(*args, **dict(kw, rollout_variant=6))
    for mp in (a, b):
        m.defineMPPIobs(mp, circles)
    Ua = Ub = np.zeros((50, 2))
    for it in range(4):
        m.defineMPPINominalControl(a, Ua); m.MPPIPlan(a)
        m.defineMPPINominalControl(b, Ub); m.MPPIPlan(b)
        if it == 0:
            assert np.array_equal(a.p.costs, b.p.costs) and np.array_equal(a.p.flags, b.p.flags)
        assert a.r.argmin == b.r.argmin and a.r.RolloutCount == b.r.RolloutCount and a.r.FeasibleTrajCount == b.r.FeasibleTrajCount
        assert np.abs(a.r.Control - b.r.Control).max() < 2e-6 * max(1.0, np.abs(b.r.Control).max())
        assert abs(a.r.cost - b.r.cost) <= 1e-5 * abs(b.r.cost)
        Ua, Ub = a.r.Control.copy(), b.r.Control.copy()
    assert np.array_equal(a.noise(), b.noise())
    n0 = a.launch_count(); m.MPPIPlan(a); na = a.launch_count() - n0
    n0 = b.launch_count(); m.MPPIPlan(b); nb = b.launch_count() - n0
    assert na == nb - 1
    a.close(); b.close()


@pytest.mark.parametrize("G", [2, 4])
def test_pipe_in_block_update_across_shards(built, G):
    """One handle, G logical shards (and real GPUs when the box has them): every shard folds its own blocks' partials, the peer
    exchange combines the shards (PH_FOLD | PH_XCHG | PH_COMBINE)."""
    import torch
    args, kw, circles, blocks = _args(8192, 50)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
    one = m.defineMPPI(*args, **dict(kw, rollout_variant=6))
    m.defineMPPIobs(one, circles)
    m.MPPIPlan(one)
    for ids in [[0] * G] + ([list(range(G))] if torch.cuda.device_count() >= G else []):
        mp = m.defineMPPI(*args, **dict(kw, rollout_variant=7, n_gpus=G, device_ids=ids))
        m.defineMPPIobs(mp, circles)
        m.MPPIPlan(mp)
        assert np.array_equal(mp.p.costs, one.p.costs) and mp.r.argmin == one.r.argmin
        assert np.abs(mp.r.Control - one.r.Control).max() < 2e-6 * max(1.0, np.abs(one.r.Control).max())
        mp.close()
    one.close()


def test_pipe_in_block_update_needs_no_cutoff(built):
    args, kw, circles, blocks = _args(4096, 50)
    with pytest.raises(m.MPPIError, match="no feasibility cut-off"):
        m.defineMPPI(*args, **dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=7, FeasibilityCount=100))
    mp = m.defineMPPI(*args, **dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=7))
    m.defineMPPIobs(mp, circles)
    assert mp.plan_info()["packed"] == 3
    m.defineMPPIblocks(mp, sc.PARKING_BLOCKS)   # a wall arrives: sequential packed rollout + update_kernel from the next step on
    assert mp.plan_info()["packed"] == 1
    m.MPPIPlan(mp)
    m.defineMPPIblocks(mp, [])
    assert mp.plan_info()["packed"] == 3
    m.MPPIPlan(mp)
    mp.close()


def test_pipe_in_block_update_split_phase_and_device_loops(built):
    """The other callers of the step with variant 7: split-phase shards on supplied noise (PH_FOLD alone -> exchange format ->
    combine), the device-resident chain (mppi_step_async with adopt) and the closed loop on the device (main.jl:58-83)."""
    from motionplanning_b200 import dist as mdist
    K, N, G = 4096, 50, 2
    args, kw, circles, blocks = _args(K, N)
    eps = orc.generate_noise(args[6], sc.SEED, 0, K, N, capi.FP64).astype(np.float32)
    U = warm_nominal(N)
    kws = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_SUPPLIED)
    whole = m.defineMPPI(*args, **dict(kws, rollout_variant=6))
    m.defineMPPIobs(whole, circles)
    m.defineMPPINominalControl(whole, U)
    m.MPPIPlan(whole, eps)
    shards, partials = [], []
    for g in range(G):
        off, kl = mdist.shard_range(K, g, G)
        s = m.defineMPPI(*args, **dict(kws, rollout_variant=7, k_offset=off, K_local=kl))
        m.defineMPPIobs(s, circles)
        m.defineMPPINominalControl(s, U)
        assert s.plan_info()["packed"] == 3
        partials.append(s.step_local(np.ascontiguousarray(eps[:, off:off + kl, :])))
        shards.append(s)
    assert np.array_equal(np.concatenate([s.p.costs for s in shards]), whole.p.costs)
    for s in shards:
        s.step_combine(partials)
        assert s.r.argmin == whole.r.argmin and s.r.RolloutCount == whole.r.RolloutCount and s.r.rho == whole.r.rho
        assert np.abs(s.r.Control - whole.r.Control).max() <= 2e-6 * max(1.0, np.abs(whole.r.Control).max())
        s.close()
    whole.close()
    # device-resident loops, Philox noise
    kwp = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
    a = m.defineMPPI(*args, **dict(kwp, rollout_variant=7))
    b = m.defineMPPI(*args, **dict(kwp, rollout_variant=6))
    for mp in (a, b):
        m.defineMPPIobs(mp, circles)
        m.MPPIPlan(mp)
        for _ in range(4):
            mp.step_async(adopt=True)
        mp.sync()
        mp.fetch_result()
    assert a.r.argmin == b.r.argmin
    assert np.abs(a.r.Control - b.r.Control).max() <= 1e-4 * max(1.0, np.abs(b.r.Control).max())   # five chained FP32 updates
    ha, la = a.run_closed_loop(max_iters=4, plant_steps=20)
    hb, lb = b.run_closed_loop(max_iters=4, plant_steps=20)
    assert ha.shape == hb.shape and np.isfinite(ha).all()
    assert np.abs(ha - hb).max() <= 1e-3 * max(1.0, np.abs(hb).max())
    a.close(); b.close()
