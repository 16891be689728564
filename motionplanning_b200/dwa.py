"""Host-side mirror of the reference's DynamicWindow (DWA) interface, over the same C ABI.

The DWA planner is the MPPI rollout's nearest relative: one constant
control per candidate from a SampleNumber[1] x SampleNumber[2] grid, no control-cost term, selection by
arg-min.  Same names and positional arguments as OptimalControl/DynamicWindow/src/setup.jl and
DWAUtils.jl, so a test reads like the reference's DynamicWindow/main.jl:

    dwa = defineDWA(numStates, numControls, X0, goal, SampleNumber, T, XL, XU, CL, CU)
    defineDWAobs(dwa, [[50.0, 1, 2.5], ...])
    ShiftInitialCondition(dwa, states)
    DWAPlan(dwa)
    dwa.r.Control, dwa.r.Traj, dwa.r.Feasibility, dwa.r.cost

Everything numerical happens in libmppi_b200.so (rollout_kernel<.., SRC_DWA>).  No CPU path.
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from .mppi import MPPIError, _dptr, defineMPPI, defineMPPIobs

DWA_N = 20                         # types.jl:24   N::Int64 = 20
DWA_SLACK_PENALTY = 1e5            # types.jl:27
DWA_OBSTACLE_PENALTY = 10000.0 * 712.5   # DWAUtils.jl:112
DWA_TERMINAL_SCALE = 10000.0       # DWAUtils.jl:39


def defineDWAcontrols(SampleNumber, CL, CU):
    """defineDWAcontrols! (setup.jl:59-85) with DecodeControl (:87-95): row i of ControlSamples is
    (CtlVector[1][i1], CtlVector[2][i2]) with the FIRST control the slowest index; CtlVector[c] =
    LinRange(CL[c], CU[c], SampleNumber[c])."""
    grids = [np.linspace(float(CL[c]), float(CU[c]), int(SampleNumber[c])) for c in range(len(SampleNumber))]
    mesh = np.meshgrid(*grids, indexing="ij")
    return np.stack([g.reshape(-1) for g in mesh], axis=1)


def defineDWA(numStates=0, numControls=0, X0=None, goal=None, SampleNumber=(10, 10), T=3.0,
              XL=None, XU=None, CL=None, CU=None, *, precision=capi.FP32, device=0, **overrides):
    """defineDWA (setup.jl:3-52): same checks, same order, same wording (they are defineMPPI's)."""
    if numControls != 2 and numControls > 0:
        raise MPPIError(f"Controls ({numControls}) must be 2: the vehicle models take (sr, ax)")
    SampleNumber = [int(v) for v in SampleNumber]
    K = int(np.prod(SampleNumber)) if numControls > 0 else 1
    kw = dict(planner=capi.PLANNER_DWA, obstacle_penalty=DWA_OBSTACLE_PENALTY, slack_penalty=DWA_SLACK_PENALTY,
              terminal_scale=DWA_TERMINAL_SCALE, FeasibilityCount=K, precision=precision, device=device)
    kw.update(overrides)
    dwa = defineMPPI(numStates, numControls, 1.0, X0, goal, K, np.eye(max(numControls, 1)), DWA_N, T,
                     XL, XU, CL, CU, **kw)
    s = dwa.s
    s.SampleNumber = SampleNumber
    s.SlackPenalty = DWA_SLACK_PENALTY
    s.ControlSamples = defineDWAcontrols(SampleNumber, s.CL, s.CU)          # setup.jl:50
    cf = np.asfortranarray(s.ControlSamples)
    dwa._check(dwa._lib.mppi_set_candidates(dwa._h, _dptr(cf), K, numControls))
    return dwa


def defineDWAobs(dwa, obstacle_list):
    """defineDWAobs! (setup.jl:54-57)."""
    defineMPPIobs(dwa, obstacle_list)


def DWAPlan(dwa):
    """DWAPlan (DWAUtils.jl:141-163): every candidate rolled out, BestCand = minimum by cost.
    r.Control is the selected candidate's control VECTOR (BestCand.Control = ctrl_list, :156)."""
    res = dwa._res
    res.Control, res.Traj = dwa._ctrl_ptr, dwa._traj_ptr           # staging arrays of the handle (addresses taken once)
    if dwa._f_step(dwa._hv, None, dwa._res_ptr):
        dwa._check(1)
    r = dwa.r
    r.Traj = dwa._traj_buf.copy(order="F")                          # :155
    r.Control = dwa._ctrl_buf[0].copy()                             # :156  (constant over the horizon)
    r.Feasibility = "Feasible" if res.feasible else "InFeasible"    # :157
    r.cost = res.cost                                               # :158
    r.time = res.time_s                                             # :160
    r.RolloutCount, r.FeasibleTrajCount = res.rollout_count, res.feasible_count
    r.argmin = res.argmin_index
    return None
