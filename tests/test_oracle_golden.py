"""Pin the oracle's VehicleDynamics (vehicledynamics.jl:1-54) against the reference's own
committed closed-loop logs (tests/golden/*.csv, made by tests/golden/make_golden.py from
OptimalControl/MPPI/MPPITrajectory.csv and OptimalControl/DynamicWindow/DWATrajectory.csv).

main.jl:76-78 integrates the plant with explicit Euler at 1 ms:
    row[k+1] = row[k] + VehicleDynamics(row[k], u)[1] * 1e-3
so each consecutive row pair is a known-answer vector for the derivative output.
"""
import os

import numpy as np
import pytest

import oracle_lib as orc
from motionplanning_b200 import _capi as capi

GOLD = os.path.join(os.path.dirname(__file__), "golden")
DT = 1e-3


def _pairs(name):
    a = np.loadtxt(os.path.join(GOLD, name), delimiter=",", comments="#")
    return a[:, 0].astype(int), a[:, 1:8], a[:, 8:15]


@pytest.mark.parametrize("name", ["mppi_plant_pairs.csv", "dwa_plant_pairs.csv"])
def test_vehicle_dynamics_matches_reference_logs(built, name):
    cfg = capi.default_config()
    ks, xk, xk1 = _pairs(name)
    assert len(ks) > 300
    worst = 0.0
    for x, xn in zip(xk, xk1):
        # controls recovered from the log: dsa = sr, dux = ax  (vehicledynamics.jl:48-49)
        sr = (xn[6] - x[6]) / DT
        ax = (xn[5] - x[5]) / DT
        dx, cons, cost = orc.vehicle_dynamics(cfg, x, [sr, ax])
        pred = x + dx * DT
        worst = max(worst, np.abs(pred - xn).max())
        assert cons == 0.0
    assert worst <= 1e-12, worst


def test_stage_cost_formula(built):
    # vehicledynamics.jl:52 evaluated by hand
    cfg = capi.default_config()
    x = np.array([3.0, 0.7, -0.3, 0.2, 0.1, 6.0, 0.05])
    u = np.array([0.25, -1.5])
    _, _, cost = orc.vehicle_dynamics(cfg, x, u)
    want = x[2] ** 2 * 1 + 1 * x[3] ** 2 + 5 * u[1] ** 2 + 3 * u[0] ** 2 + 2 * x[6] ** 2 + 10 * x[1] ** 2
    assert cost == pytest.approx(want, rel=1e-15)


def test_ux_clamp_only_affects_position_and_dv(built):
    # vehicledynamics.jl:40-42: ux <- max(ux, 0) AFTER the slip angles
    cfg = capi.default_config()
    x = np.array([0.0, 0.0, 0.4, 0.1, 0.3, -2.0, 0.02])
    dx, _, _ = orc.vehicle_dynamics(cfg, x, [0.0, 0.0])
    assert dx[0] == pytest.approx(-0.4 * np.sin(0.3), rel=1e-14)   # ux term vanished
    assert dx[1] == pytest.approx(0.4 * np.cos(0.3), rel=1e-14)
    # slips still use ux = -2: compare with a by-hand evaluation
    mu, B, Cc, E = 0.8, -10.4 / 0.8, 1.3, 0.1556
    den = -2.0 + 0.01
    af = np.arctan((0.4 + 1.56 * 0.1) / den) - 0.02
    ar = np.arctan((0.4 - 1.64 * 0.1) / den)
    rv = 0.1 * 0.4
    FZF = 2 * (1018.28 / 2 * 9.81 - (0 - rv) * 186.22)
    FZR = 2 * (963.34 / 2 * 9.81 + (0 - rv) * 186.22)
    fy = lambda FZ, a: mu * FZ * 1.0 * np.sin(Cc * np.arctan(B * a - E * (B * a - np.arctan(B * a))))
    FY1, FY2 = fy(FZF, af), fy(FZR, ar)
    assert dx[2] == pytest.approx((FY1 + FY2) / 2020 - 0.1 * 0.0, rel=1e-13)
    assert dx[3] == pytest.approx((FY1 * 1.56 - FY2 * 1.64) / 4095, rel=1e-13)


def test_kinematic_model(built):
    # Parking/src/VehicleModel.jl:14-18
    cfg = capi.default_config()
    cfg.model, cfg.n_x = capi.MODEL_KINEMATIC5, 5
    x = np.array([1.0, -0.5, 0.3, 4.0, 0.1])
    dx, cons, cost = orc.vehicle_dynamics(cfg, x, [0.2, -1.0])
    L = 1.56 + 1.64
    np.testing.assert_allclose(dx, [4 * np.cos(0.3), 4 * np.sin(0.3), 4 / L * np.tan(0.1), -1.0, 0.2], rtol=1e-15)
    assert cons == 0.0
    assert cost == pytest.approx(5 * 1.0 + 3 * 0.04 + 2 * 0.01 + 10 * 0.25, rel=1e-15)


def test_heun_step(built):
    # MPPIUtils.jl:76-90
    cfg = capi.default_config()
    x = np.array([0.0, 0.1, 0.2, 0.05, 0.1, 5.0, 0.01])
    u = [0.1, 1.0]
    dt = cfg.T / cfg.N
    k1, _, c1 = orc.vehicle_dynamics(cfg, x, u)
    k2, _, _ = orc.vehicle_dynamics(cfg, x + k1 * dt, u)
    xn, cons, pc = orc.rungekutta2(cfg, x, u)
    np.testing.assert_allclose(xn, x + dt * (k1 + k2) / 2, rtol=0, atol=1e-15)
    assert cons == 1 and pc == c1


def test_plant_advance_reproduces_reference_log_rows(built):
    """oracle_plant_advance (main.jl:65-67,76-79) against the reference's MPPITrajectory.csv rows: one
    plant step from row k with the controls the log implies gives row k+1, stamped (k+1) * 1 ms."""
    cfg = capi.default_config()
    ks, xk, xk1 = _pairs("mppi_plant_pairs.csv")
    for k, x, xn in list(zip(ks, xk, xk1))[:120]:
        sr, ax = (xn[6] - x[6]) / DT, (xn[5] - x[5]) / DT
        U = np.tile([sr, ax], (cfg.N, 1))
        x1, his, done = orc.plant_advance(cfg, x, U, steps=1, time_idx0=int(k))
        assert np.abs(x1 - xn).max() <= 1e-12
        assert his.shape == (1, 8) and his[0, 0] == (int(k) + 1) * DT and np.array_equal(his[0, 1:], x1)
        assert not done


def test_plant_zero_order_hold_and_goal_stop(built):
    """Constant{Previous} interpolation over time_serial = range(0, T, N) (main.jl:45,65-67) and the
    6 m goal test (:80-82), against a direct numpy restatement."""
    cfg = capi.default_config()
    N, T = cfg.N, cfg.T
    rng = np.random.default_rng(7)
    U = np.column_stack([rng.uniform(-0.3, 0.3, N), rng.uniform(-2, 2, N)])
    x0 = np.array([0.0, 0.0, 0.0, 0.0, 0.0, 5.0, 0.0])
    steps, upd = 400, 1.0     # long enough to cross several knots of time_serial
    x1, his, done = orc.plant_advance(cfg, x0, U, steps=steps, update_time=upd)
    knots = np.linspace(0.0, T, N)
    fine = np.linspace(0.0, upd, steps)
    x = x0.copy()
    for i in range(steps):
        j = int(np.searchsorted(knots, fine[i], side="right")) - 1
        dx, _, _ = orc.vehicle_dynamics(cfg, x, U[j])
        x = x + dx * DT
        assert np.abs(his[i, 1:] - x).max() <= 1e-12
    assert len(np.unique([int(np.searchsorted(knots, t, side="right")) - 1 for t in fine])) > 3
    assert not done and np.abs(x1 - x).max() <= 1e-12
    # start 6.5 m short of the goal at 5 m/s: the goal test fires after ~0.1 s and cuts the block short
    near = np.array([cfg.goal[0] - 6.5, cfg.goal[1], 0.0, 0.0, 0.0, 5.0, 0.0])
    x2, his2, done2 = orc.plant_advance(cfg, near, np.zeros((N, 2)), steps=400)
    assert done2 and 90 <= len(his2) <= 110
    assert (his2[-1, 1] - cfg.goal[0]) ** 2 + (his2[-1, 2] - cfg.goal[1]) ** 2 <= 36.0
    assert (his2[-2, 1] - cfg.goal[0]) ** 2 + (his2[-2, 2] - cfg.goal[1]) ** 2 > 36.0


# ---------------------------------------------------------------------------------------------------
# The whole plan, pinned on the reference's own output.  DynamicWindow/main.jl has no random sampling:
# every DWAPlan of the committed run can be replayed, and the control it selected can be read back from
# the log (tests/golden/make_golden.py: dwa_plan_cases).  One case exercises the candidate grid
# (setup.jl:59-95), TrajectoryRollout (DWAUtils.jl:16-42), Rungekutta2 (:61-75), VehicleDynamics,
# ObstacleEvaluation (:105-118), BoundEvaluation (:121-137), the terminal term (:39,98-101) and the
# arg-min selection (:152) -- rows a3..a10 of SURVEY.md 8(a), which MPPIPlan shares line for line.
# ---------------------------------------------------------------------------------------------------
def _dwa_cases():
    a = np.loadtxt(os.path.join(GOLD, "dwa_plan_cases.csv"), delimiter=",", comments="#")
    return a[:-1], a[-1]


def _dwa_scene(**over):
    from motionplanning_b200 import defineDWAcontrols
    from motionplanning_b200 import scenarios as sc
    CLd, CUd = [-0.3, -2.5], [0.3, 2.5]                       # DynamicWindow/main.jl:10-11
    cand = defineDWAcontrols([31, 41], CLd, CUd)              # main.jl:25
    K = len(cand)
    T = over.pop("T", 3.0)
    args = (7, 2, 25.0, sc.X0_DYN, [110, 0], K, sc.SIGMA, 20, T, sc.XL_DYN, sc.XU_DYN, CLd, CUd)   # main.jl:7-9,14,20,26
    kw = dict(planner=capi.PLANNER_DWA, obstacle_penalty=10000.0 * 712.5, FeasibilityCount=K)
    kw.update(over)
    return orc.make_config(args, kw), cand


DWA_CIRCLES = [[50.0, 1, 2.5], [70.0, -1, 2.5], [90.0, 1, 2.5]]   # DynamicWindow/main.jl:27


def _dwa_mismatches(cfg, cand, circles, cases):
    bad = 0
    for row in cases:
        got = orc.plan(cfg, circles, [], row[2:9], None, cand)["Control"][0]
        bad += not np.allclose(got, row[9:11], rtol=0, atol=1e-9)
    return bad


def test_dwa_plan_reproduces_every_selection_of_the_reference_log(built):
    cases, last = _dwa_cases()
    assert len(cases) == 127 and int(last[0]) == 12632
    cfg, cand = _dwa_scene()
    assert cand.shape == (31 * 41, 2)
    assert _dwa_mismatches(cfg, cand, DWA_CIRCLES, cases) == 0


def test_dwa_pin_has_teeth(built):
    """The same replay with one ingredient changed must disagree with the log, i.e. the 127 selections
    really constrain obstacles, horizon, dynamics and the terminal term."""
    cases, _ = _dwa_cases()
    cfg, cand = _dwa_scene()
    assert _dwa_mismatches(cfg, cand, [], cases) >= 40                                    # no obstacles
    assert _dwa_mismatches(cfg, cand, [[x, y, 2.4] for x, y, _ in DWA_CIRCLES], cases) >= 10
    assert _dwa_mismatches(*_dwa_scene(T=2.9), DWA_CIRCLES, cases) >= 40                  # dt = T / N
    assert _dwa_mismatches(*_dwa_scene(la=1.6), DWA_CIRCLES, cases) >= 10                 # tyre model geometry
    assert _dwa_mismatches(*_dwa_scene(terminal_scale=5000.0), DWA_CIRCLES, cases) >= 10  # DWAUtils.jl:39


def test_dwa_closed_loop_reproduces_the_reference_log(built):
    """DynamicWindow/main.jl:46-67 replayed end to end with the oracle (plan, then 100 Euler plant steps
    under the held control, goal radius 7.2): every 100th row and the final row of DWATrajectory.csv."""
    cases, last = _dwa_cases()
    cfg, cand = _dwa_scene()
    x = np.array(cases[0, 2:9])
    worst, k = 0.0, 0
    for i in range(200):
        assert np.abs(x - cases[i, 2:9]).max() <= 1e-9 and k == int(cases[i, 0])
        u = orc.plan(cfg, DWA_CIRCLES, [], x, None, cand)["Control"]
        x, his, done = orc.plant_advance(cfg, x, u, steps=100, goal_radius=7.2, time_idx0=k)
        k += len(his)
        worst = max(worst, np.abs(x - (cases[i + 1, 2:9] if i + 1 < len(cases) else last[2:9])).max()) if not done else worst
        if done:
            break
    assert done and k == int(last[0])
    assert np.abs(x - last[2:9]).max() <= 1e-9 and his[-1, 0] == pytest.approx(last[1], abs=1e-12)
    assert worst <= 1e-9
