"""GPU tests beyond supplied-noise parity: Philox mode (device noise replayed through the oracle),
sample sharding (logical shards on one GPU), cut-off, DWA planner, batched instances, pose
queries, plot export, error behaviour, and size-independent properties at BASELINE.json sizes."""
import numpy as np
import pytest

import motionplanning_b200 as m
import oracle_lib as orc
from motionplanning_b200 import dist as mdist
from motionplanning_b200 import scenarios as sc
from test_gpu_parity import FP32_TOL, GUARD, check, warm_nominal

capi = m.capi
pytestmark = pytest.mark.gpu


def make(args, kw, circles, blocks, Unom=None):
    mppi = m.defineMPPI(*args, **kw)
    if circles:
        m.defineMPPIobs(mppi, circles)
    if blocks:
        m.defineMPPIblocks(mppi, blocks)
    if Unom is not None:
        m.defineMPPINominalControl(mppi, Unom)
    return mppi


@pytest.mark.parametrize("precision", [capi.FP64, capi.FP32])
def test_philox_noise_replayed_through_oracle(built, precision):
    """Config 2's mode at a size the oracle finishes quickly: in-kernel Philox noise is dumped
    (mppi_get_noise) and replayed through the oracle's supplied-noise path."""
    args, kw, circles, blocks = sc.config(1, K=8192)
    kw = dict(kw, precision=precision, noise_mode=capi.NOISE_PHILOX)
    U = warm_nominal(50)
    mppi = make(args, kw, circles, blocks, U)
    m.MPPIPlan(mppi)
    eps = mppi.noise()
    ref = orc.plan(orc.make_config(args, kw), circles, blocks, args[3], U, eps.astype(np.float64), capi.FP64,
                   threads=8, want_margin=True)
    check(mppi, ref, precision)
    # distribution: eps ~ N(0, Sigma), Sigma a covariance (setup.jl:57)
    flat = eps.reshape(-1, 2).astype(np.float64)
    assert abs(flat.mean(0)).max() < 2e-3
    np.testing.assert_allclose(np.cov(flat.T), sc.SIGMA, atol=1.5e-3)
    # the device generator follows the oracle's Philox/Box-Muller statement up to libm rounding
    host = orc.generate_noise(sc.SIGMA, sc.SEED, 0, 8192, 50, precision)
    assert np.abs(host.astype(np.float64) - eps).max() < 5e-5


def test_philox_determinism_iteration_and_seed(built):
    args, kw, circles, blocks = sc.config(1, K=4096)
    a = make(args, kw, circles, blocks)
    b = make(args, kw, circles, blocks)
    m.MPPIPlan(a), m.MPPIPlan(b)
    assert np.array_equal(a.noise(), b.noise()) and np.array_equal(a.r.Control, b.r.Control)
    e0 = a.noise()
    m.MPPIPlan(a)
    assert not np.array_equal(e0, a.noise())          # iteration counter advances the stream
    host1 = orc.generate_noise(sc.SIGMA, sc.SEED, 1, 4096, 50, capi.FP32)
    assert np.abs(host1 - a.noise()).max() < 5e-5
    c = make(args, dict(kw, seed=12345), circles, blocks)
    m.MPPIPlan(c)
    assert not np.array_equal(e0, c.noise())


@pytest.mark.parametrize("precision", [capi.FP64, capi.FP32])
@pytest.mark.parametrize("G", [2, 4, 8])
def test_logical_shards_match_unsharded(built, precision, G):
    """G logical shards on one GPU (split-phase local / combine) == the unsharded step:
    argmin / counts identical, U_new equal up to the rounding of the rescale."""
    K = 8192
    args, kw, circles, blocks = sc.config(1, K=K)
    # one rollout flavour on both sides (the split-phase step keeps the in-rollout Philox generator, where auto would give the
    # unsharded handle the warp pipeline, whose cost sum has another order): what is compared here is the sharding arithmetic
    kw = dict(kw, precision=precision, rollout_variant=3)
    U = warm_nominal(50)
    whole = make(args, kw, circles, blocks, U)
    m.MPPIPlan(whole)
    shards, partials = [], []
    for g in range(G):
        off, kl = mdist.shard_range(K, g, G)
        s = make(args, dict(kw, k_offset=off, K_local=kl), circles, blocks, U)
        partials.append(s.step_local())
        shards.append(s)
    # Philox counters use the global sample id: the shards drew exactly the unsharded noise
    assert np.array_equal(np.concatenate([s.noise() for s in shards], axis=1), whole.noise())
    assert np.array_equal(np.concatenate([s.p.costs for s in shards]), whole.p.costs)
    for s in shards:
        s.step_combine(partials)
    tol = 1e-12 if precision == capi.FP64 else 2e-6
    for s in shards:
        assert s.r.argmin == whole.r.argmin and s.r.RolloutCount == whole.r.RolloutCount
        assert s.r.FeasibleTrajCount == whole.r.FeasibleTrajCount
        assert s.r.rho == whole.r.rho
        assert s.r.eta == pytest.approx(whole.r.eta, rel=tol)
        assert np.abs(s.r.Control - whole.r.Control).max() <= tol * max(np.abs(whole.r.Control).max(), 1)
    # every rank folds the same partials in the same order: bitwise identical results
    for s in shards[1:]:
        assert np.array_equal(s.r.Control, shards[0].r.Control) and s.r.cost == shards[0].r.cost


def test_partial_roundtrip_through_exchange_format(built):
    args, kw, circles, blocks = sc.config(0)
    s = make(args, kw, circles, blocks)
    p = s.step_local()
    q = mdist.array_to_partial(mdist.partial_to_array(p, 50), 50)
    assert (q.rho, q.eta, q.argmin_global, q.rollouts, q.feasible) == (p.rho, p.eta, p.argmin_global, p.rollouts, p.feasible)
    assert list(q.S[:100]) == list(p.S[:100])


@pytest.mark.parametrize("precision", [capi.FP64, capi.FP32])
def test_dwa_planner(built, precision):
    """DWA mode (DynamicWindow/src/DWAUtils.jl:141-163, setup.jl:59-85): 31 x 41 constant-control grid,
    no lambda term, obstacle penalty 10000*712.5, arg-min selection."""
    CLd, CUd = [-0.3, -2.5], [0.3, 2.5]
    g0, g1 = np.linspace(CLd[0], CUd[0], 31), np.linspace(CLd[1], CUd[1], 41)
    cand = np.array([[a, b] for a in g0 for b in g1])
    K = len(cand)
    x0 = [38.0, -1.2, 0.0, 0.0, 0.0, 5.0, 0.0]
    args = (7, 2, 25.0, x0, sc.GOAL, K, sc.SIGMA, 20, 3.0, sc.XL_DYN, sc.XU_DYN, CLd, CUd)
    kw = dict(planner=capi.PLANNER_DWA, obstacle_penalty=10000.0 * 712.5, FeasibilityCount=K, precision=precision)
    mppi = make(args, kw, sc.CIRCLES, [])
    cf = np.asfortranarray(cand)
    mppi._check(mppi._lib.mppi_set_candidates(mppi._h, cf.ctypes.data_as(m.mppi.C.POINTER(m.mppi.C.c_double)), K, 2))
    m.MPPIPlan(mppi)
    ref = orc.plan(orc.make_config(args, kw), sc.CIRCLES, [], x0, None, cand, capi.FP64, threads=8)
    tol = 1e-9 if precision == capi.FP64 else FP32_TOL
    rel = np.abs(mppi.p.costs - ref["costs"]) / np.abs(ref["costs"])
    assert rel.max() <= tol
    assert np.array_equal(mppi.p.flags, ref["flags"])
    assert mppi.r.argmin == ref["argmin"]
    assert np.array_equal(mppi.r.Control, ref["Control"])
    assert np.abs(mppi.r.Traj - ref["Traj"]).max() <= (1e-9 if precision == capi.FP64 else 1e-3)
    assert mppi.r.cost == pytest.approx(ref["cost"], rel=tol)


def test_check_poses_matches_block_collision_check(built):
    """mppi_check_poses == block_collision_check (hybrid_astar_utils.jl:193-201) per pose, incl. the
    ConvexCollision quirk and headings outside [-pi, pi] (modπ)."""
    cfg = capi.default_config()
    rng = np.random.RandomState(11)
    blocks = sc.PARKING_BLOCKS + sc.wall_blocks(16)
    poses = np.stack([rng.uniform(-8, 100, 50000), rng.uniform(-10, 10, 50000), rng.uniform(-9, 9, 50000)], axis=1)
    mppi = make(*sc.config(0)[:2], sc.CIRCLES, blocks)
    hit = mppi.check_poses(poses)
    ref, margin = orc.check_poses(cfg, sc.CIRCLES, blocks, poses)
    ok = margin > 1e-9
    assert ok.mean() > 0.999 and np.array_equal(hit[ok], ref[ok])
    assert (ref > 0).any() and (ref == 0).any()
    # the derived known-answer cases of CollisionDetection/main.jl:7-12 (free) and the quirk
    one = make(*sc.config(0)[:2], [], [[0.0, -1.0, 0.0, 3.75, 1.0]])
    assert one.check_poses([[5.5 - 1.5, 1.0, 0.0]])[0] == 0          # vehicle block [5.5, 1, 0, 1.5, 1]
    unit = make(sc.config(0)[0], dict(sc.config(0)[1], veh_L=2.0, veh_W=2.0, veh_center_offset=0.0), [], [[0, 0, 0, 1, 1]])
    assert unit.check_poses([[2.45, 0.0, np.pi / 4]])[0] == 1         # disjoint, yet ConvexCollision = false
    assert unit.check_poses([[2.0, 0.0, 0.0]])[0] == 0                # touching counts as separated
    assert unit.check_poses([[3.0, 0.0, np.pi / 4]])[0] == 1                # still the quirk: 3/sqrt2 < 1 + sqrt2
    assert unit.check_poses([[3.5, 0.0, np.pi / 4]])[0] == 0


def test_batched_instances_match_single_instance_oracle(built):
    """Config 5 shape at a small size: B instances with their own x0 / U_nom, Philox noise per
    (instance, sample); every instance replayed through the oracle."""
    B, K, N = 6, 512, 50
    args, kw, circles, blocks = sc.config(4, K=K)
    kw = dict(kw, B=B, precision=capi.FP64)
    x0s = sc.jittered_x0(B)
    Us = np.stack([warm_nominal(N, seed=b) for b in range(B)], axis=2)
    mppi = make(args[:3] + (x0s,) + args[4:], kw, circles, blocks)
    Uout, costs, feas = mppi.step_batched(x0s, Us)
    dt = np.float64
    eps = np.empty((B, N, K, 2), dtype=dt)
    mppi._check(mppi._lib.mppi_get_noise(mppi._h, eps.ctypes.data_as(m.mppi.C.c_void_p)))
    allc = mppi.p.costs.reshape(B, K)
    for b in range(B):
        ref = orc.plan(orc.make_config(args, dict(kw, B=1)), circles, blocks, x0s[:, b], Us[:, :, b], eps[b], capi.FP64, threads=8)
        assert (np.abs(allc[b] - ref["costs"]) / np.abs(ref["costs"])).max() < 1e-9
        assert np.abs(Uout[:, :, b] - ref["Control"]).max() < 1e-9
        assert costs[b] == pytest.approx(ref["cost"], rel=1e-9) and bool(feas[b]) == ref["feasible"]
    assert not np.array_equal(eps[0], eps[1])     # instances draw independent noise


def test_batched_many_instances_small_blocks_path(built, monkeypatch):
    """B >= 64 switches finalize to 128-thread blocks and the update kernel to one block per instance
    walking all horizon groups; both must give what the single-instance kernels give (FP32, bitwise:
    same operations in the same order for an instance whatever B is)."""
    B, K, N = 80, 1024, 50
    args, kw, circles, blocks = sc.config(4, K=K)
    x0s = sc.jittered_x0(B)
    Us = np.stack([warm_nominal(N, seed=b) for b in range(B)], axis=2)
    big = make(args[:3] + (x0s,) + args[4:], dict(kw, B=B), circles, blocks)
    Uout, costs, feas = big.step_batched(x0s, Us)
    small_B = 4
    small = make(args[:3] + (x0s[:, :small_B],) + args[4:], dict(kw, B=small_B), circles, blocks)
    Uo2, c2, f2 = small.step_batched(x0s[:, :small_B], Us[:, :, :small_B])
    # instance b draws noise from counter (k, j, iter, b): the first instances see identical noise in both handles
    assert np.allclose(Uout[:, :, :small_B], Uo2, rtol=2e-6, atol=1e-7)
    assert np.allclose(costs[:small_B], c2, rtol=1e-6) and np.array_equal(feas[:small_B], f2)
    # update-kernel grid shapes: forced one-group-per-block and one-block-per-instance agree
    for js in ("1", "13"):
        monkeypatch.setenv("MPPI_UPDATE_JSPLIT", js)
        alt = make(args[:3] + (x0s[:, :small_B],) + args[4:], dict(kw, B=small_B), circles, blocks)
        Uo3, c3, _ = alt.step_batched(x0s[:, :small_B], Us[:, :, :small_B])
        assert np.allclose(Uo3, Uo2, rtol=2e-6, atol=1e-7) and np.allclose(c3, c2, rtol=1e-6)
    monkeypatch.delenv("MPPI_UPDATE_JSPLIT")


def test_batched_full_size_cfg5_spot_parity(built):
    """BASELINE.json config 5 at its full size (4,096 instances x K = 1,024 x N = 50, FP32, Philox): 16 instances spread over
    the batch are replayed through the FP64 oracle on the device's own noise (costs of every sample, U_new, nominal cost), and
    size-independent properties hold for ALL instances: U_new inside [CL, CU], finite, and an instance's result does not
    depend on its position in the batch (instances 0..3 equal those of a 4-instance handle)."""
    B, K, N = 4096, 1024, 50
    args, kw, circles, blocks = sc.config(4, K=K)
    x0s = sc.jittered_x0(B)
    Us = np.stack([warm_nominal(N, seed=b % 7) for b in range(B)], axis=2)
    big = make(args[:3] + (x0s,) + args[4:], dict(kw, B=B), circles, blocks)
    Uout, costs, feas = big.step_batched(x0s, Us)
    assert np.isfinite(Uout).all() and np.isfinite(costs).all()
    assert (Uout >= np.array(sc.CL)[None, :, None] - 1e-6).all() and (Uout <= np.array(sc.CU)[None, :, None] + 1e-6).all()
    eps = np.empty((B, N, K, 2), dtype=np.float32)
    big._check(big._lib.mppi_get_noise(big._h, eps.ctypes.data_as(m.mppi.C.c_void_p)))
    allc = big.p.costs.reshape(B, K)
    allf = big.p.flags.reshape(B, K)
    n_guard = 0
    for b in np.linspace(0, B - 1, 16).astype(int):
        ref = orc.plan(orc.make_config(args, dict(kw, B=1, precision=capi.FP64)), circles, blocks, x0s[:, b], Us[:, :, b],
                       eps[b].astype(np.float64), capi.FP64, threads=8, want_margin=True)
        ok = ref["margins"] > GUARD
        n_guard += int((~ok).sum())
        assert (np.abs(allc[b] - ref["costs"]) / np.abs(ref["costs"]))[ok].max() <= FP32_TOL
        assert np.array_equal(allf[b][ok], ref["flags"][ok])
        assert np.abs(Uout[:, :, b] - ref["Control"]).max() <= FP32_TOL * max(1.0, np.abs(ref["Control"]).max())
        if ref["nominal_margin"] > GUARD:
            assert costs[b] == pytest.approx(ref["cost"], rel=FP32_TOL) and bool(feas[b]) == ref["feasible"]
    print(f"[guard band] {n_guard} of {16 * K} samples excluded")
    assert n_guard <= 0.02 * 16 * K
    small = make(args[:3] + (x0s[:, :4],) + args[4:], dict(kw, B=4), circles, blocks)
    Uo2, c2, f2 = small.step_batched(x0s[:, :4], Us[:, :, :4])
    assert np.allclose(Uout[:, :, :4], Uo2, rtol=2e-6, atol=1e-7) and np.allclose(costs[:4], c2, rtol=1e-6) and np.array_equal(feas[:4], f2)


def test_sample_xy_export(built):
    args, kw, circles, blocks = sc.config(0)
    kw = dict(kw, precision=capi.FP64)
    mppi = make(args, kw, circles, blocks)
    m.MPPIPlan(mppi)
    xy = mppi.p.sample_xy(stride=64)
    assert xy.shape == (16, 51, 2)
    eps = mppi.noise()
    cfg = orc.make_config(args, kw)
    for i in (0, 7, 15):
        u = np.clip(eps[:, 64 * i, :], sc.CL, sc.CU)
        ref = orc.rollout(cfg, circles, blocks, args[3], np.zeros((50, 2)), u)
        assert np.abs(xy[i] - ref["Traj"][:, :2]).max() < 1e-4


def test_error_behaviour(built):
    """setup.jl:19-39 errors become MPPIError; state errors are reported, never crash."""
    args, kw, circles, blocks = sc.config(0)
    with pytest.raises(m.MPPIError, match="Length of XL"):
        m.defineMPPI(*(args[:9] + ([0.0] * 4,) + args[10:]), **kw)
    with pytest.raises(m.MPPIError, match="Controls"):
        m.defineMPPI(*(args[:1] + (0,) + args[2:]), **kw)
    mppi = make(args, dict(kw, noise_mode=capi.NOISE_SUPPLIED), circles, blocks)
    with pytest.raises(m.MPPIError, match="needs the eps tensor"):
        m.MPPIPlan(mppi)
    with pytest.raises(m.MPPIError):
        mppi.p.costs   # no step has run
    with pytest.raises(m.MPPIError, match="must match number of states"):
        m.ShiftInitialCondition(mppi, [0.0] * 7)
    with pytest.raises(m.MPPIError):
        m.defineMPPIobs(mppi, [[0.0, 0.0, 1.0]] * 65)
    with pytest.raises(m.MPPIError, match="positive definite"):
        m.defineMPPINominalControl(mppi, np.zeros((50, 2)), [[1.0, 2.0], [2.0, 1.0]])


def test_closed_loop_device_resident_equals_host_loop(built):
    """main.jl:59-62 (U_nom <- r.Control each iteration): the device-resident adopt path gives the
    same controls as the host round trip."""
    args, kw, circles, blocks = sc.config(1, K=4096)
    a = make(args, kw, circles, blocks)
    b = make(args, kw, circles, blocks)
    m.MPPIPlan(a)
    m.MPPIPlan(b)
    for _ in range(4):
        m.defineMPPINominalControl(a, a.r.Control)
        m.MPPIPlan(a)
        b.step_async(adopt=True)
    b.sync()
    b.fetch_result()
    assert np.array_equal(a.r.Control, b.r.Control) and a.r.cost == b.r.cost
    assert a.r.cost < 1e4 and a.r.Feasibility == "Feasible"


def _wavy_path(n=240):
    x = np.linspace(-5.0, 115.0, n)
    return np.column_stack([x, 1.5 * np.sin(x / 12.0)])


@pytest.mark.parametrize("model_cfg,precision,variant", [(1, capi.FP64, 0), (1, capi.FP32, 1), (1, capi.FP32, 2),
                                                          (3, capi.FP64, 0), (3, capi.FP32, 2)])
def test_reference_path_tracking_term_matches_oracle(built, model_cfg, precision, variant):
    """the windowed nearest-point tracking term, device (scalar and packed rollout,
    finalize's nominal costing) against the oracle's restatement on the device's own Philox noise."""
    K, N = 4096, 40
    args, kw, circles, blocks = sc.config(model_cfg, K=K, N=N)
    kw = dict(kw, precision=precision, noise_mode=capi.NOISE_PHILOX, rollout_variant=variant, stage_w=[1, 1, 5, 3, 2, 0.0])
    U = warm_nominal(N)
    path = _wavy_path()
    mppi = make(args, kw, circles, blocks, U)
    m.MPPIPlan(mppi)
    plain_costs = mppi.p.costs.copy()
    m.defineMPPIrefpath(mppi, path, weight=4.0, window=12)
    m.defineMPPINominalControl(mppi, U)
    m.MPPIPlan(mppi)
    eps = mppi.noise()
    assert (mppi.p.costs > plain_costs.min()).all() and not np.allclose(mppi.p.costs[:64], plain_costs[:64])
    orc.set_reference_path(path, weight=4.0, window=12)
    try:
        ref = orc.plan(orc.make_config(args, kw), circles, blocks, args[3], U, eps.astype(np.float64), capi.FP64,
                       threads=8, want_margin=True)
    finally:
        orc.set_reference_path(None)
    check(mppi, ref, precision)
    # removing the path restores the plain cost exactly
    m.defineMPPIrefpath(mppi, None)
    m.defineMPPINominalControl(mppi, U)
    m.MPPIPlan(mppi)
    assert mppi.r.rho > 0 and np.isfinite(mppi.p.costs).all()


def test_reference_path_pulls_the_plan_and_rejects_supplied_noise(built):
    """A path offset to y = +3 m with the lane term off: a few iterations move the nominal plan's
    lateral position towards it; supplied-noise handles refuse the term."""
    args, kw, circles, blocks = sc.config(1, K=16384, N=40)
    kw = dict(kw, noise_mode=capi.NOISE_PHILOX, stage_w=[1, 1, 5, 3, 2, 0.0])
    a = make(args, kw, [], [])
    x = np.linspace(-5, 120, 200)
    m.defineMPPIrefpath(a, np.column_stack([x, np.full_like(x, 3.0)]), weight=20.0, window=20)
    for _ in range(12):
        m.MPPIPlan(a)
        m.defineMPPINominalControl(a, a.r.Control)
    assert a.r.Traj[-1, 1] > 1.0            # started at y = 0, plain MPPI stays there
    b = make(args, dict(kw, noise_mode=capi.NOISE_SUPPLIED), [], [])
    with pytest.raises(m.MPPIError, match="PHILOX"):
        m.defineMPPIrefpath(b, np.column_stack([x, x * 0]))
    with pytest.raises(m.MPPIError):
        m.defineMPPIrefpath(a, np.zeros((2000, 2)))


@pytest.mark.parametrize("model_cfg,precision", [(0, capi.FP64), (0, capi.FP32), (3, capi.FP64)])
def test_closed_loop_on_device_matches_host_loop_with_oracle_plant(built, model_cfg, precision):
    """mppi_run_closed_loop (main.jl:58-83 as one chain of graph launches, plant kernel behind finalize)
    against the same loop driven from the host: MPPIPlan per iteration + the ORACLE's plant
    (oracle_plant_advance, pinned to the reference's MPPITrajectory.csv in test_oracle_golden.py)."""
    iters, steps = 6, 100
    args, kw, circles, blocks = sc.config(model_cfg, K=2048, N=30)
    kw = dict(kw, precision=precision, noise_mode=capi.NOISE_PHILOX)
    dev = make(args, kw, circles, blocks)
    # the closed-loop chain draws its noise inside the (one-sample) rollout; the host loop is pinned to that flavour, since auto would
    # give it the warp pipeline, whose cost sum has another order (1e-7 relative in FP32, more than this test's 1e-9)
    host = make(args, dict(kw, rollout_variant=3), circles, blocks)
    his, log = dev.run_closed_loop(max_iters=iters, plant_steps=steps)
    cfg = host._cfg
    x = np.array(args[3], dtype=np.float64)
    rows = [np.concatenate([[0.0], x])]
    U = np.zeros((cfg.N, 2))
    costs = []
    for it in range(iters):
        m.ShiftInitialCondition(host, x)
        m.defineMPPINominalControl(host, U)
        m.MPPIPlan(host)
        U = np.array(host.r.Control, copy=True)
        costs.append(host.r.cost)
        x, h, done = orc.plant_advance(cfg, x, U, steps=steps, time_idx0=it * steps)
        rows.extend(h)
        if done:
            break
    ref = np.array(rows)
    assert his.shape == ref.shape == (1 + iters * steps, 1 + cfg.n_x)
    assert np.array_equal(his[:, 0], ref[:, 0])
    scale = np.maximum(1.0, np.abs(ref[:, 1:]).max(axis=0))
    assert (np.abs(his[:, 1:] - ref[:, 1:]) / scale).max() <= 1e-9
    assert np.allclose(log[:, 0], costs, rtol=1e-12) and (log[:, 2] == 2048 + 1).all()
    # the handle carries on from the episode's end state, host mirrors included
    assert np.abs(np.asarray(dev.s.X0) - ref[-1, 1:]).max() <= 1e-9 * scale.max()
    m.MPPIPlan(dev)
    m.ShiftInitialCondition(host, x)
    m.defineMPPINominalControl(host, U)
    m.MPPIPlan(host)
    assert np.abs(dev.r.Control - host.r.Control).max() <= 1e-6


def test_closed_loop_stops_at_goal_and_shift_option(built, tmp_path):
    """Goal test (main.jl:80-82) ends the episode on the device; the history is a valid
    MPPITrajectory.csv; the optional warm-start shift changes the plan but not the contract."""
    args, kw, circles, blocks = sc.config(0, K=4096, N=20)
    kw = dict(kw, noise_mode=capi.NOISE_PHILOX)
    start = np.array(args[3], dtype=np.float64)
    start[0] = 95.0                     # 15 m from the goal (110, 0) at 5 m/s: ~1.8 s
    a = make((args[0], args[1], args[2], start) + tuple(args[4:]), kw, circles, blocks)
    his, log = a.run_closed_loop(max_iters=40)
    assert 5 <= len(log) < 40 and len(his) <= 1 + len(log) * 100
    gx, gy = his[-1, 1] - 110.0, his[-1, 2]
    assert gx * gx + gy * gy <= 36.0 and ((his[:-1, 1] - 110.0) ** 2 + his[:-1, 2] ** 2 > 36.0).all()
    assert np.allclose(np.diff(his[:, 0]), 1e-3, atol=1e-12)
    path = tmp_path / "MPPITrajectory.csv"
    m.write_trajectory_csv(path, his)
    back = np.loadtxt(path, delimiter=",")
    assert np.array_equal(back, his)
    b = make((args[0], args[1], args[2], start) + tuple(args[4:]), kw, circles, blocks)
    his_s, log_s = b.run_closed_loop(max_iters=40, shift_nominal=1)
    assert len(log_s) >= 5 and not np.array_equal(his_s[:len(his)], his[:len(his_s)])
    g2 = (his_s[-1, 1] - 110.0) ** 2 + his_s[-1, 2] ** 2
    assert g2 <= 36.0


@pytest.mark.parametrize("cfg_idx,K,N", [(1, 65536, 50), (2, 262144, 80), (3, 131072, 100), (3, 1048576, 100)])
def test_full_size_properties(built, cfg_idx, K, N):
    """BASELINE.json sizes (cfg4 at the 8-GPU per-rank shard), checked through size-independent
    properties: rho is the cost of the argmin sample and the minimum of all costs; eta equals the
    softmin sum recomputed in float64 from the device costs; U_new equals the weighted mean recomputed
    in float64 from the dumped noise; U_new stays inside [CL, CU]; counts match the flags."""
    args, kw, circles, blocks = sc.config(cfg_idx, K=K, N=N)
    U = warm_nominal(N)
    mppi = make(args, kw, circles, blocks, U)
    m.MPPIPlan(mppi)
    costs, flags = mppi.p.costs, mppi.p.flags
    assert mppi.r.rho == costs.min() and costs[mppi.r.argmin - 1] == mppi.r.rho
    assert mppi.r.argmin - 1 == int(np.argmin(costs))
    w = np.exp(-(costs - costs.min()) / sc.LAMBDA)
    assert mppi.r.eta == pytest.approx(w.sum(), rel=2e-5)
    assert mppi.r.FeasibleTrajCount == int((flags & capi.FLAG_FEASIBLE).sum())
    assert mppi.r.RolloutCount == K + 1
    eps = mppi.noise()
    Uref = np.zeros((N, 2))
    wn = w / w.sum()
    for c0 in range(0, K, 65536):   # chunked: the full cfg4 tensor is 0.8 GB in FP32
        u = np.clip(U[:, None, :] + eps[:, c0:c0 + 65536, :].astype(np.float64), sc.CL, sc.CU)
        Uref += np.einsum("k,jkc->jc", wn[c0:c0 + 65536], u)
    assert np.abs(mppi.r.Control - Uref).max() <= FP32_TOL * np.abs(Uref).max()
    assert (mppi.r.Control >= np.array(sc.CL) - 1e-6).all() and (mppi.r.Control <= np.array(sc.CU) + 1e-6).all()
    # spot parity: 256 samples spread over the shard replayed through the oracle
    idx = np.linspace(0, K - 1, 256).astype(int)
    sub = orc.plan(orc.make_config(args[:5] + (256,) + args[6:], dict(kw, FeasibilityCount=256)), circles, blocks,
                   args[3], U, np.ascontiguousarray(eps[:, idx, :].astype(np.float64)), capi.FP64, threads=8, want_margin=True)
    ok = sub["margins"] > GUARD
    rel = np.abs(costs[idx] - sub["costs"]) / np.abs(sub["costs"])
    assert rel[ok].max() <= FP32_TOL and np.array_equal(flags[idx][ok], sub["flags"][ok])


@pytest.mark.parametrize("precision,variants", [(capi.FP64, (3, 1)), (capi.FP32, (3, 1)), (capi.FP32, (4, 2))])
@pytest.mark.parametrize("model_cfg", [0, 3])
def test_pregenerated_noise_is_bit_identical_to_in_rollout_philox(built, precision, variants, model_cfg):
    """Step variant V_PREGEN (noise_kernel draws iteration i+1's eps beside finalize of iteration i, the
    rollout reads it back) against rollout_variant = 3 / 4 (Philox inside the one-sample / packed rollout): same counters, same
    arithmetic, so costs, flags, U_new and the dumped eps agree bit for bit over a chain of iterations,
    across mppi_get_noise (which redraws the consumed eps and forces the next step to draw its own) and
    across a change of Sigma (which invalidates the pre-generated buffer)."""
    args, kw, circles, blocks = sc.config(model_cfg, K=4096 + 64, N=31)     # odd horizon: last Philox pair half used
    kw = dict(kw, precision=precision, noise_mode=capi.NOISE_PHILOX)
    ref = make(args, dict(kw, rollout_variant=variants[0]), circles, blocks)      # Philox inside the rollout (3 scalar, 4 packed)
    pre = make(args, dict(kw, rollout_variant=variants[1]), circles, blocks)      # pre-generated noise
    assert pre.launch_count() == ref.launch_count() == 0

    def both(f):
        f(ref), f(pre)

    def same():
        assert np.array_equal(ref.p.costs, pre.p.costs) and np.array_equal(ref.p.flags, pre.p.flags)
        assert np.array_equal(ref.r.Control, pre.r.Control) and np.array_equal(ref.r.Traj, pre.r.Traj)
        assert (ref.r.argmin, ref.r.cost, ref.r.rho, ref.r.eta) == (pre.r.argmin, pre.r.cost, pre.r.rho, pre.r.eta)

    for it in range(5):
        both(m.MPPIPlan)
        same()
        if it in (1, 2):
            assert np.array_equal(ref.noise(), pre.noise())
        if it == 3:
            xy_r, xy_p = ref.p.sample_xy(64), pre.p.sample_xy(64)
            assert np.array_equal(xy_r, xy_p)
        both(lambda s: m.defineMPPINominalControl(s, s.r.Control.copy()))
    Sigma2 = np.array([[0.08, 0.01], [0.01, 0.05]])
    both(lambda s: m.defineMPPINominalControl(s, s.r.Control.copy(), Sigma2))
    both(m.MPPIPlan)
    same()
    assert np.array_equal(ref.noise(), pre.noise())
    # asynchronous chain (mppi_step_async with adopt): U_nom <- U_new on the device
    for _ in range(3):
        both(lambda s: s.step_async(adopt=True))
    both(lambda s: s.sync())
    both(lambda s: s.fetch_result())
    same()
    # the pre-generating handle launched one extra kernel per step (+ one when it had to draw ahead of the rollout)
    assert pre.launch_count() > ref.launch_count()


def test_time_reference_calls_measures_the_three_entry_points(built):
    """mppi_time_reference_calls (measurement aid): n iterations of mppi_set_x0 + mppi_set_nominal + mppi_step from host buffers
    (main.jl:59-61), timed inside the library; the handle is left as after n ordinary plans."""
    args, kw, circles, blocks = sc.config(0)
    a = make(args, kw, circles, blocks)
    b = make(args, kw, circles, blocks)
    U = warm_nominal(50)
    x0 = np.asarray(args[3], dtype=np.float64)
    res, idle = a.time_reference_calls(x0, U, 4)
    assert res.shape == (4,) and (res > 0).all() and (idle >= res).all() and (res < 1e5).all()
    for _ in range(4):
        m.ShiftInitialCondition(b, x0)
        m.defineMPPINominalControl(b, U)
        m.MPPIPlan(b)
    a.fetch_result()
    assert np.array_equal(a.r.Control, b.r.Control) and a.r.cost == b.r.cost and a.r.argmin == b.r.argmin
    args5, kw5, c5, b5 = sc.config(4, K=64)
    many = m.defineMPPI(*args5, **dict(kw5, B=3))
    with pytest.raises(m.MPPIError, match="single-instance"):
        many.time_reference_calls(np.zeros(5), np.zeros((50, 2)), 1)
