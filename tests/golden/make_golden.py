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


if __name__ == "__main__":
    main()
