"""CPU tests of the oracle's MPPI restatement (MPPIUtils.jl) against by-hand numpy evaluations,
plus Philox known answers.  No GPU needed."""
import math

import numpy as np
import pytest

import oracle_lib as orc
from motionplanning_b200 import _capi as capi
from motionplanning_b200 import scenarios as sc


def _kin_f(x, u, L=3.2):
    return np.array([x[3] * math.cos(x[2]), x[3] * math.sin(x[2]), x[3] / L * math.tan(x[4]), u[1], u[0]])


def _kin_cost(x, u):
    return 5 * u[1] ** 2 + 3 * u[0] ** 2 + 2 * x[4] ** 2 + 10 * x[1] ** 2


def _heun(x, u, dt):
    k1 = _kin_f(x, u)
    k2 = _kin_f(x + k1 * dt, u)
    return x + dt * (k1 + k2) / 2


def _rollout_by_hand(x0, ctrl, nominal, circles, XL, XU, CL, CU, goal, dt, lam, Sigma):
    """Literal numpy transcription of TrajectoryRollout (MPPIUtils.jl:31-57) for the kinematic model."""
    N = ctrl.shape[0]
    x = np.array(x0, dtype=float)
    cost_his = np.zeros(N + 1)
    cons_his = np.zeros(N + 1)

    def obs(s):
        c, v = 1, 0.0
        for ox, oy, r in circles:
            if (s[0] - ox) ** 2 + (s[1] - oy) ** 2 <= r ** 2:
                c, v = 0, v + 100.0 * 712.5
        return c, v

    def bnd(s):
        c, v = 1, 0.0
        for i in range(5):
            if s[i] < XL[i]:
                c, v = 0, v + 1e5 * abs(s[i] - XL[i])
            if s[i] > XU[i]:
                c, v = 0, v + 1e5 * abs(s[i] - XU[i])
        return c, v

    Sinv = np.linalg.inv(np.asarray(Sigma))
    for j in range(N):
        ci = obs(x) if j > 0 else (1, 0.0)
        bi = bnd(x) if j > 0 else (1, 0.0)
        p = _kin_cost(x, ctrl[j])
        xn = _heun(x, ctrl[j], dt)
        cons_his[j] = 1 * ci[0] * bi[0]
        cost_his[j] = p + bi[1] + ci[1] + lam * nominal[j] @ Sinv @ (ctrl[j] - nominal[j])
        x = xn
    ci, bi = obs(x), bnd(x)
    cons_his[N] = ci[0] * bi[0]
    cost_his[N] = _kin_cost(x, [0, 0]) + bi[1] + ci[1]
    d0 = (x0[0] - goal[0]) ** 2 + (x0[1] - goal[1]) ** 2
    total = cost_his.sum() + ((x[0] - goal[0]) ** 2 + (x[1] - goal[1]) ** 2) / d0 * 10000.0
    return total, bool((cons_his == 1).all())


def test_tiny_mppi_by_hand(built):
    """K = 2, N = 1 (SURVEY 8(c) MPPI-tiny) and a K = 16, N = 6 variant, against the literal transcription."""
    for K, N in [(2, 1), (16, 6)]:
        x0 = [48.0, 0.5, 0.1, 5.0, 0.02]
        args = (5, 2, 25.0, x0, sc.GOAL, K, sc.SIGMA, N, 0.5 * N, sc.XL_KIN, sc.XU_KIN, sc.CL, sc.CU)
        kw = dict(FeasibilityCount=K, model=capi.MODEL_KINEMATIC5)
        cfg = orc.make_config(args, kw)
        rng = np.random.RandomState(1)
        eps = rng.randn(N, K, 2) * [0.3, 1.5]
        nominal = rng.uniform(-0.2, 0.2, (N, 2))
        ref = orc.plan(cfg, sc.CIRCLES, [], x0, nominal, eps, capi.FP64)
        costs = np.zeros(K)
        ctrls = np.zeros((K, N, 2))
        for k in range(K):
            ctrls[k] = np.clip(nominal + eps[:, k, :], sc.CL, sc.CU)
            costs[k], feas = _rollout_by_hand(x0, ctrls[k], nominal, sc.CIRCLES, sc.XL_KIN, sc.XU_KIN, sc.CL, sc.CU,
                                               sc.GOAL, 0.5, 25.0, sc.SIGMA)
            assert bool(ref["flags"][k] & capi.FLAG_FEASIBLE) == feas
        np.testing.assert_allclose(ref["costs"], costs, rtol=1e-12)
        rho = costs.min()
        w = np.exp(-(costs - rho) / 25.0)
        eta = w.sum()
        U = np.tensordot(w / eta, ctrls, axes=1)
        assert ref["argmin"] == int(np.argmin(costs)) + 1
        assert ref["rho"] == pytest.approx(rho, rel=1e-12) and ref["eta"] == pytest.approx(eta, rel=1e-12)
        np.testing.assert_allclose(ref["Control"], U, rtol=1e-12, atol=1e-15)
        c_nom, f_nom = _rollout_by_hand(x0, U, nominal, sc.CIRCLES, sc.XL_KIN, sc.XU_KIN, sc.CL, sc.CU, sc.GOAL, 0.5,
                                        25.0, sc.SIGMA)
        assert ref["cost"] == pytest.approx(c_nom, rel=1e-12) and ref["feasible"] == f_nom
        assert ref["rollout_count"] == K + 1   # off-by-one of MPPIUtils.jl:171,182,197


def test_weights_argmin_first_index_and_nan_last(built):
    # types.jl:9 isless: first minimal index; NaN sorts last
    w, rho, eta, am = orc.weights([5.0, 3.0, 3.0, float("nan"), 9.0][:3] + [9.0], 2.0)
    assert am == 1 and rho == 3.0
    assert w.sum() == pytest.approx(1.0)
    w, rho, eta, am = orc.weights([float("nan"), 4.0, 2.0], 1.0)
    assert am == 2 and rho == 2.0


def test_feasibility_cutoff_semantics(built):
    """MPPIUtils.jl:175: stop right after the (FeasibilityCount+1)-th feasible sample."""
    args, kw, circles, blocks = sc.config(0, K=256, N=10)
    kw = dict(kw, FeasibilityCount=20)
    cfg = orc.make_config(args, kw)
    eps = orc.generate_noise(sc.SIGMA, 1, 0, 256, 10)
    full = orc.plan(orc.make_config(args, dict(kw, FeasibilityCount=256)), circles, blocks, args[3], None, eps)
    cut = orc.plan(cfg, circles, blocks, args[3], None, eps)
    feas = (full["flags"] & 1).astype(int)
    R = int(np.searchsorted(np.cumsum(feas), 21)) + 1   # index after the 21st feasible
    assert cut["rollout_count"] == R + 1 and cut["feasible_count"] == 21
    par = orc.plan(cfg, circles, blocks, args[3], None, eps, threads=4)   # parallel path == serial path
    assert par["rollout_count"] == cut["rollout_count"]
    np.testing.assert_array_equal(par["Control"], cut["Control"])
    w = np.exp(-(full["costs"][:R] - full["costs"][:R].min()) / 25.0)
    assert cut["eta"] == pytest.approx(w.sum(), rel=1e-12)


def test_philox4x32_10_known_answers(built):
    # Random123 kat_vectors for philox4x32-10
    kats = [
        ([0, 0, 0, 0], [0, 0], [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]),
        ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]),
        ([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0],
         [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]),
    ]
    for ctr, key, want in kats:
        assert [int(v) for v in orc.philox(ctr, key)] == want


def test_noise_distribution_and_sharding(built):
    """eps ~ N(0, Sigma) with Sigma a covariance (setup.jl:57); global sample ids make the stream
    independent of the shard layout."""
    Sigma = [[0.05, 0.02], [0.02, 0.1]]
    e = orc.generate_noise(Sigma, 42, 0, 20000, 10)
    flat = e.reshape(-1, 2)
    assert abs(flat.mean(0)).max() < 3e-3
    np.testing.assert_allclose(np.cov(flat.T), Sigma, atol=2e-3)
    a = orc.generate_noise(Sigma, 42, 3, 64, 7, k_offset=0)
    b = orc.generate_noise(Sigma, 42, 3, 32, 7, k_offset=32)
    np.testing.assert_array_equal(a[:, 32:, :], b)
    c = orc.generate_noise(Sigma, 42, 4, 64, 7)
    assert not np.array_equal(a, c)   # iteration counter changes the stream


def test_dwa_plan_argmin(built):
    """DWAPlan (DWAUtils.jl:141-163): constant control per candidate, arg-min selection, no lambda term."""
    nS = [5, 7]
    CLd, CUd = [-0.3, -2.5], [0.3, 2.5]
    g0, g1 = np.linspace(CLd[0], CUd[0], nS[0]), np.linspace(CLd[1], CUd[1], nS[1])
    cand = np.array([[a, b] for a in g0 for b in g1])   # setup.jl:59-85 ordering: first control slowest
    K = len(cand)
    args = (7, 2, 25.0, sc.X0_DYN, sc.GOAL, K, sc.SIGMA, 20, 3.0, sc.XL_DYN, sc.XU_DYN, CLd, CUd)
    kw = dict(planner=capi.PLANNER_DWA, obstacle_penalty=10000.0 * 712.5, FeasibilityCount=K)
    cfg = orc.make_config(args, kw)
    ref = orc.plan(cfg, sc.CIRCLES, [], sc.X0_DYN, None, cand)
    k = ref["argmin"] - 1
    assert ref["costs"][k] == ref["costs"].min() == ref["rho"]
    np.testing.assert_array_equal(ref["Control"], np.tile(cand[k], (20, 1)))
    one = orc.rollout(cfg, sc.CIRCLES, [], sc.X0_DYN, None, np.tile(cand[3], (20, 1)))
    assert one["cost"] == pytest.approx(ref["costs"][3], rel=1e-14)


def test_tracking_term_against_numpy_restatement(built):
    """Oracle's windowed nearest-point term (mppi_oracle.hpp track_scan / track_start) against a direct
    numpy evaluation along the oracle's own nominal trajectory."""
    from motionplanning_b200 import scenarios as sc
    args, kw, circles, blocks = sc.config(0, K=8, N=25)
    kw = dict(kw, stage_w=[1, 1, 5, 3, 2, 0.0])
    cfg = orc.make_config(args, kw)
    rng = np.random.default_rng(3)
    ctrl = np.column_stack([rng.uniform(-0.3, 0.3, 25), rng.uniform(-1, 1, 25)])
    base = orc.rollout(cfg, [], [], args[3], None, ctrl)
    x = np.linspace(-2, 40, 90)
    path = np.column_stack([x, 0.8 * np.cos(x / 5.0)])
    W, w = 6, 3.0
    orc.set_reference_path(path, weight=w, window=W)
    try:
        got = orc.rollout(cfg, [], [], args[3], None, ctrl)
    finally:
        orc.set_reference_path(None)
    traj = base["Traj"]
    idx = int(np.argmin(((path - traj[0, :2]) ** 2).sum(1)))
    extra = 0.0
    for j in range(26):
        hi = min(idx + W, len(path) - 1)
        d = ((path[idx:hi + 1] - traj[j, :2]) ** 2).sum(1)
        idx += int(np.argmin(d))
        extra += w * d.min()
    assert np.array_equal(got["Traj"], traj)
    assert got["cost"] - base["cost"] == pytest.approx(extra, rel=1e-12)
