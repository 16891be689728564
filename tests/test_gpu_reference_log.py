"""The CUDA path against the REFERENCE'S OWN OUTPUT (not only against the oracle).

OptimalControl/DynamicWindow/main.jl is deterministic; its committed log DWATrajectory.csv records,
through the sa / ux columns, the control every DWAPlan selected (tests/golden/make_golden.py:
dwa_plan_cases -> tests/golden/dwa_plan_cases.csv).  The DWA planner runs the same rollout kernel as
MPPI (rollout_kernel<.., SRC_DWA>: Heun step, VehicleDynamics, stage / bound / obstacle / terminal cost,
arg-min epilogue), so these are known answers from the reference for rows a3..a10 of SURVEY.md 8(a),
driven through the reference's own DWA interface (defineDWA / defineDWAobs! / ShiftInitialCondition /
DWAPlan, DynamicWindow/src/setup.jl:3-57, DWAUtils.jl:9-12,141-163).
"""
import os

import numpy as np
import pytest

import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc

capi = m.capi
pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")

CLd, CUd = [-0.3, -2.5], [0.3, 2.5]                               # DynamicWindow/main.jl:10-11
DWA_CIRCLES = [[50.0, 1, 2.5], [70.0, -1, 2.5], [90.0, 1, 2.5]]   # main.jl:27


def _cases():
    a = np.loadtxt(os.path.join(GOLD, "dwa_plan_cases.csv"), delimiter=",", comments="#")
    return a[:-1], a[-1]


def _dwa(precision):
    dwa = m.defineDWA(7, 2, sc.X0_DYN, [110, 0], [31, 41], 3.0, sc.XL_DYN, sc.XU_DYN, CLd, CUd, precision=precision)   # main.jl:26
    m.defineDWAobs(dwa, DWA_CIRCLES)
    return dwa


@pytest.mark.parametrize("precision", [capi.FP64, capi.FP32])
def test_dwa_plan_selects_what_the_reference_selected(built, precision):
    """127 plans of the reference run: the GPU's selection is the logged one.  In FP32 mode a plan whose two
    best candidates tie within the FP32 cost tolerance (1e-4 relative) may pick the other one."""
    cases, _ = _cases()
    dwa = _dwa(precision)
    near_ties = 0
    for row in cases:
        m.ShiftInitialCondition(dwa, row[2:9])
        m.DWAPlan(dwa)
        if np.allclose(dwa.r.Control, row[9:11], rtol=0, atol=1e-9):
            continue
        assert precision == capi.FP32, f"FP64 selection differs from the reference at log row {int(row[0])}"
        costs = dwa.p.costs
        logged = int(np.argmin(np.abs(dwa.s.ControlSamples - row[9:11]).sum(1)))
        assert abs(costs[logged] - costs.min()) <= 1e-4 * abs(costs.min())
        near_ties += 1
    assert near_ties <= 3
    assert dwa.r.Traj.shape == (21, 7) and dwa.r.Control.shape == (2,)


@pytest.mark.parametrize("precision", [capi.FP64, capi.FP32])
def test_dwa_closed_loop_on_the_device_reproduces_the_reference_log(built, precision):
    """main.jl:46-67 as one device-resident episode (mppi_run_closed_loop: plan -> 100 Euler plant steps
    under the held control -> 7.2 m goal test): same number of rows as DWATrajectory.csv, every 100th row and
    the last row equal to the log."""
    cases, last = _cases()
    dwa = _dwa(precision)
    his, log = dwa.run_closed_loop(max_iters=200, plant_steps=100, plant_dt=1e-3, update_time=0.1, goal_radius=7.2)
    assert his.shape == (int(last[0]) + 1, 8) and len(log) == len(cases)
    tol = 1e-9 if precision == capi.FP64 else 1e-9   # the plant is FP64 in both modes; only the selection could differ
    for row in cases:
        k = int(row[0])
        assert np.abs(his[k] - row[1:9]).max() <= tol, f"row {k}"
    assert np.abs(his[-1] - last[1:9]).max() <= tol
    # the controls held over each block, read back the way make_golden.py reads the reference's log
    sr = (his[1::100, 7][:len(cases)] - his[0::100, 7][:len(cases)]) / 1e-3
    ax = (his[1::100, 6][:len(cases)] - his[0::100, 6][:len(cases)]) / 1e-3
    assert np.abs(sr - cases[:, 9]).max() <= 1e-9 and np.abs(ax - cases[:, 10]).max() <= 1e-9
