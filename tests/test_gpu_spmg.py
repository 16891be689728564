"""Single-process multi-GPU handle (cfg.n_gpus, cfg.device_ids) and the cross-GPU feasibility cut-off.

The reference's caller is ONE Julia task (OptimalControl/MPPI/main.jl:55-83: ShiftInitialCondition,
defineMPPINominalControl!, MPPIPlan at :59-61), so the drop-in must reach several GPUs from one handle and
one host thread.  What is checked: a handle with n_gpus = G gives the result of the unsharded handle
(MPPIUtils.jl:154-167,186-198 — softmin weights, weighted update, nominal rollout) and of the oracle, and the
sequential cut-off of MPPIUtils.jl:175,185 (stop after the (FeasibilityCount+1)-th feasible sample IN GLOBAL
SAMPLE ORDER) is reproduced when the samples are sharded.  On a 1-GPU box the G shards share the GPU (the same
ordinal repeated in device_ids) — same kernels, same peer-window exchange, plain device pointers; with >= 2 GPUs
visible the same tests also run across real devices."""
import ctypes as C

import numpy as np
import pytest

import motionplanning_b200 as m
import oracle_lib as orc
from motionplanning_b200 import scenarios as sc
from test_gpu_parity import check, warm_nominal

capi = m.capi
pytestmark = pytest.mark.gpu


def _n_devices():
    import torch

    return torch.cuda.device_count()


def _device_sets(G):
    """device_ids to test for G shards: logical shards on GPU 0, and real devices when the box has them."""
    sets = [[0] * G]
    if _n_devices() >= G:
        sets.append(list(range(G)))
    return sets


def _make(args, kw, circles, blocks, **extra):
    mp = m.defineMPPI(*args, **dict(kw, **extra))
    if circles:
        m.defineMPPIobs(mp, circles)
    if blocks:
        m.defineMPPIblocks(mp, blocks)
    return mp


@pytest.mark.parametrize("precision", [capi.FP64, capi.FP32])
@pytest.mark.parametrize("G", [2, 3, 8])
def test_one_handle_many_gpus_matches_unsharded_and_oracle(built, precision, G):
    """Supplied noise: the caller passes the FULL eps tensor to the one handle, which splits it over its shards."""
    K, N = 4099, 50   # ragged: the shards differ in size
    args, kw, circles, blocks = sc.config(0, K=K)
    kw = dict(kw, precision=precision, noise_mode=capi.NOISE_SUPPLIED, rollout_variant=1)
    eps = orc.generate_noise(args[6], sc.SEED, 0, K, N, capi.FP64)
    U = warm_nominal(N)
    one = _make(args, kw, circles, blocks)
    m.defineMPPINominalControl(one, U)
    m.MPPIPlan(one, eps)
    eps_ref = eps.astype(np.float32).astype(np.float64) if precision == capi.FP32 else eps
    ref = orc.plan(orc.make_config(args, kw), circles, blocks, args[3], U, eps_ref, capi.FP64, threads=8, want_margin=True)
    for ids in _device_sets(G):
        mp = _make(args, kw, circles, blocks, n_gpus=G, device_ids=ids)
        assert mp.plan_info()["n_gpus"] == G
        m.defineMPPINominalControl(mp, U)
        m.MPPIPlan(mp, eps)
        assert np.array_equal(mp.p.costs, one.p.costs) and np.array_equal(mp.p.flags, one.p.flags)   # same kernel, same samples
        assert mp.r.argmin == one.r.argmin and mp.r.RolloutCount == one.r.RolloutCount and mp.r.FeasibleTrajCount == one.r.FeasibleTrajCount
        tol = 1e-12 if precision == capi.FP64 else 2e-6
        assert np.abs(mp.r.Control - one.r.Control).max() <= tol * max(1.0, np.abs(one.r.Control).max())
        check(mp, ref, precision)
        mp.close()
    one.close()


@pytest.mark.parametrize("G", [2, 4])
def test_one_handle_philox_chain_and_device_loop(built, G):
    """Philox counters use global sample ids, so the G-GPU handle draws the noise of the unsharded one; three chained
    iterations through the reference-facing calls, then the device-resident form (step_async / adopt)."""
    K, N = 8192, 50
    args, kw, circles, blocks = sc.config(0, K=K)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=1)
    for ids in _device_sets(G):
        mp = _make(args, kw, circles, blocks, n_gpus=G, device_ids=ids)
        one2 = _make(args, kw, circles, blocks)
        U = np.zeros((N, 2))
        V = np.zeros((N, 2))
        for it in range(3):
            m.defineMPPINominalControl(mp, U)
            m.ShiftInitialCondition(mp, np.asarray(args[3], dtype=np.float64))
            m.MPPIPlan(mp)
            m.defineMPPINominalControl(one2, V)
            m.MPPIPlan(one2)
            if it == 0:   # same nominal on both sides: same kernel, same samples, same bits
                assert np.array_equal(mp.p.costs, one2.p.costs) and np.array_equal(mp.p.flags, one2.p.flags)
            U, V = mp.r.Control.copy(), one2.r.Control.copy()
            # the combine folds G partials instead of one: U_new agrees to FP32 rounding, and so do the next iteration's costs
            assert np.abs(U - V).max() <= 2e-6 * max(1.0, np.abs(V).max())
        assert np.array_equal(mp.noise(), one2.noise())   # the noise does not depend on the nominal: identical bits
        assert np.abs(mp.p.costs - one2.p.costs).max() <= 1e-5 * np.abs(one2.p.costs).max()
        assert mp.r.argmin == one2.r.argmin
        # device-resident chain
        for _ in range(3):
            mp.step_async(adopt=True)
            one2.step_async(adopt=True)
        mp.sync(); one2.sync()
        mp.fetch_result(); one2.fetch_result()
        assert mp.r.argmin == one2.r.argmin
        # six chained iterations: the FP32 rounding difference of the G-way combine is fed back through the nominal each time
        assert np.abs(mp.r.Control - one2.r.Control).max() <= 1e-3 * max(1.0, np.abs(one2.r.Control).max())
        xy = mp.p.sample_xy(stride=64)
        assert xy.shape == (sum((kl + 63) // 64 for kl in _shard_sizes(K, G)), N + 1, 2) and np.isfinite(xy).all()
        mp.close(); one2.close()


def _shard_sizes(K, G):
    return [K // G + (1 if g < K % G else 0) for g in range(G)]


@pytest.mark.parametrize("precision", [capi.FP64, capi.FP32])
@pytest.mark.parametrize("G", [2, 4, 8])
def test_feasibility_cutoff_across_shards(built, precision, G):
    """MPPIUtils.jl:175,185 with sharded samples: the cut falls on the first shard, on a middle one, on the last one, or
    nowhere (fewer feasible samples than the limit); in every case RolloutCount, FeasibleTrajCount, argmin and U_new equal
    those of the unsharded handle and of the oracle."""
    K, N = 2048, 20
    X0 = [40.0, -1.5, 0.0, 0.0, -0.05, 4.0, 0.0]   # next to the first obstacle: a mix of feasible and infeasible samples
    eps = orc.generate_noise(sc.SIGMA, sc.SEED, 0, K, N, capi.FP64)
    eps_ref = eps.astype(np.float32).astype(np.float64) if precision == capi.FP32 else eps
    U = warm_nominal(N)
    base_kw = dict(precision=precision, noise_mode=capi.NOISE_SUPPLIED, rollout_variant=1)
    probe_args = (7, 2, 25.0, X0, [110, 0], K, sc.SIGMA, N, 3.0, sc.XL_DYN, sc.XU_DYN, sc.CL, sc.CU)
    probe = orc.plan(orc.make_config(probe_args, dict(base_kw, FeasibilityCount=K)), sc.CIRCLES, [], X0, U, eps_ref, capi.FP64, threads=8)
    n_feas = int((probe["flags"] & capi.FLAG_FEASIBLE).sum())
    assert 200 < n_feas < K - 200
    seen = set()
    for fc in (5, n_feas // 2, n_feas - 3, n_feas + 10):
        kw = dict(base_kw, FeasibilityCount=fc)
        ref = orc.plan(orc.make_config(probe_args, kw), sc.CIRCLES, [], X0, U, eps_ref, capi.FP64, threads=8, want_margin=True)
        one = _make(probe_args, kw, sc.CIRCLES, [])
        m.defineMPPINominalControl(one, U)
        m.MPPIPlan(one, eps)
        for ids in _device_sets(G):
            mp = _make(probe_args, kw, sc.CIRCLES, [], n_gpus=G, device_ids=ids)
            m.defineMPPINominalControl(mp, U)
            m.MPPIPlan(mp, eps)
            assert mp.r.RolloutCount == one.r.RolloutCount and mp.r.FeasibleTrajCount == one.r.FeasibleTrajCount
            assert mp.r.argmin == one.r.argmin
            tol = 1e-12 if precision == capi.FP64 else 2e-6
            assert np.abs(mp.r.Control - one.r.Control).max() <= tol * max(1.0, np.abs(one.r.Control).max())
            check(mp, ref, precision)
            # a second step on the same handle: the exchange slots alternate parity and must not see stale counts
            m.MPPIPlan(mp, eps)
            assert mp.r.RolloutCount == one.r.RolloutCount and mp.r.argmin == one.r.argmin
            mp.close()
        cut = one.r.RolloutCount - 1
        seen.add(min(G - 1, next((g for g in range(G) if cut <= sum(_shard_sizes(K, G)[:g + 1])), G - 1)))
        one.close()
    assert 0 in seen and (G - 1) in seen   # the cut landed on the first and on the last shard at least once


def test_multi_gpu_handle_error_behaviour(built):
    args, kw, circles, blocks = sc.config(0, K=1024)
    with pytest.raises(m.MPPIError, match="shards the samples of ONE MPPI instance"):
        m.defineMPPI(*args, **dict(kw, n_gpus=2, device_ids=[0, 0], B=2))
    with pytest.raises(m.MPPIError, match="owns all K samples"):
        m.defineMPPI(*args, **dict(kw, n_gpus=2, device_ids=[0, 0], K_local=512))
    with pytest.raises(m.MPPIError, match="out of range"):
        m.defineMPPI(*args, **dict(kw, n_gpus=2, device_ids=[0, 63]))
    mp = m.defineMPPI(*args, **dict(kw, n_gpus=2, device_ids=[0, 0], precision=capi.FP32, noise_mode=capi.NOISE_PHILOX))
    with pytest.raises(m.MPPIError, match="not available on a multi-GPU handle"):
        mp.step_local()
    with pytest.raises(m.MPPIError, match="not available on a multi-GPU handle"):
        mp.run_closed_loop(max_iters=2)
    with pytest.raises(m.MPPIError, match="not available on a multi-GPU handle"):
        mp.peer_export()
    m.MPPIPlan(mp)   # still usable after the rejected calls
    assert mp.r.RolloutCount == 1025
    mp.close()
    # split-phase handles cannot apply the cut-off across shards
    sh = m.defineMPPI(*args, **dict(kw, k_offset=0, K_local=512, FeasibilityCount=100, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX))
    with pytest.raises(m.MPPIError, match="feasibility cut-off"):
        sh.step_local()
    sh.close()
