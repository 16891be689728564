"""Generate tests/golden/*.csv from the reference's own committed artefacts.

Run in the authoring container only (needs /root/reference, which does not exist
on the GPU box).  The reference ships no tests; the only numeric artefacts on
the MPPI hot path are two closed-loop logs written by
OptimalControl/MPPI/main.jl:87 and OptimalControl/DynamicWindow/main.jl:
columns t, x, y, v, r, psi, ux, sa at a 1 ms explicit-Euler plant step
(main.jl:76-78: states += VehicleDynamics(states, u)[1] * 1e-3).
Consecutive rows therefore pin the derivative output of VehicleDynamics
(vehicledynamics.jl:1-54): row[k+1] = row[k] + f(row[k], u) * 1e-3, where
u = (sr, ax) is recovered from the sa / ux columns (dsa = sr, dux = ax).

We keep every STRIDE-th consecutive row pair (plus the first 50 pairs, where the
transient is richest) as known-answer vectors: x_k (7), x_{k+1} (7).
"""
import sys
import numpy as np

REF = "/root/reference/OptimalControl"
SRC = {
    "mppi_plant_pairs.csv": f"{REF}/MPPI/MPPITrajectory.csv",
    "dwa_plant_pairs.csv": f"{REF}/DynamicWindow/DWATrajectory.csv",
}
STRIDE = 37
DWA_LOG = f"{REF}/DynamicWindow/DWATrajectory.csv"
UPDATE_IDX = 100     # DynamicWindow/main.jl:17-19: one DWAPlan every update_time / dt = 100 plant steps


def main():
    for out, src in SRC.items():
        a = np.loadtxt(src, delimiter=",")
        idx = sorted(set(range(0, 50)) | set(range(0, a.shape[0] - 1, STRIDE)))
        rows = [np.concatenate([[k], a[k, 1:], a[k + 1, 1:]]) for k in idx]
        with open(f"{sys.path[0]}/{out}", "w") as f:
            f.write("# k, x_k[7] = x,y,v,r,psi,ux,sa, x_{k+1}[7]; plant step 1e-3; from " + src.replace("/root/reference/", "") + "\n")
            for r in rows:
                f.write(",".join([str(int(r[0]))] + [repr(float(v)) for v in r[1:]]) + "\n")
        print(out, len(rows), "pairs of", a.shape[0], "rows")


def dwa_plan_cases():
    """DynamicWindow/main.jl is fully deterministic (no sampling): DWAPlan runs at time_idx = 1, 101, ...
    on the state of log row 100*i (main.jl:48-50) and its result dwa.r.Control = (sr, ax) is held for the
    next 100 plant rows (main.jl:53,61-62), where it can be read back exactly from the sa / ux columns
    (dsa = sr, dux = ax).  Each (state, selected control) pair is therefore a known answer for the WHOLE
    plan: candidate grid, rollout, Heun step, stage / bound / obstacle / terminal cost, arg-min selection.
    The last log row (goal test fired, main.jl:64-66) closes the file with nan controls."""
    a = np.loadtxt(DWA_LOG, delimiter=",")
    rows = []
    for k in range(0, a.shape[0] - 1, UPDATE_IDX):
        sr = (a[k + 1, 7] - a[k, 7]) / 1e-3
        ax = (a[k + 1, 6] - a[k, 6]) / 1e-3
        sr, ax = round(sr / 0.02) * 0.02, round(ax / 0.125) * 0.125      # snap to the 31 x 41 grid of main.jl:25
        assert abs(sr - (a[k + 1, 7] - a[k, 7]) / 1e-3) < 1e-9 and abs(ax - (a[k + 1, 6] - a[k, 6]) / 1e-3) < 1e-9
        # the control is held over the whole block
        blk = a[k:min(k + UPDATE_IDX, a.shape[0] - 1) + 1]
        assert np.abs(np.diff(blk[:, 7]) / 1e-3 - sr).max() < 1e-9 and np.abs(np.diff(blk[:, 6]) / 1e-3 - ax).max() < 1e-9
        rows.append(np.concatenate([[k], a[k], [sr, ax]]))
    rows.append(np.concatenate([[a.shape[0] - 1], a[-1], [np.nan, np.nan]]))
    with open(f"{sys.path[0]}/dwa_plan_cases.csv", "w") as f:
        f.write("# log row k, t, x,y,v,r,psi,ux,sa at row k, control (sr, ax) DWAPlan selected there; last line = final row; from "
                "OptimalControl/DynamicWindow/DWATrajectory.csv\n")
        for r in rows:
            f.write(",".join([str(int(r[0]))] + [repr(float(v)) for v in r[1:]]) + "\n")
    print("dwa_plan_cases.csv", len(rows) - 1, "plans,", a.shape[0], "log rows")


if __name__ == "__main__":
    main()
    dwa_plan_cases()
