"""Fixture for the scenario reader (motionplanning_b200/scenario_io.py), made from the reference's
PathPlanning/Scenarios/obstacle_field.mat (100 trials of n x 3 [x y r]).  Run in the authoring
container only (needs /root/reference and scipy).

1. checks the repo's own MAT-5 reader against scipy.io.loadmat on the WHOLE reference file;
2. writes the trials Astar/main.jl uses (Trial_num = 53) plus the first four as
   obstacle_field_subset.mat (compressed, as MATLAB writes it) and obstacle_field_subset_raw.mat
   (uncompressed), and their values as obstacle_field_subset.csv (trial (1-based), x, y, r).
"""
import os
import sys

import numpy as np
import scipy.io as sio

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from motionplanning_b200.scenario_io import read_obstacle_field  # noqa: E402

SRC = "/root/reference/PathPlanning/Scenarios/obstacle_field.mat"
TRIALS = [1, 2, 3, 4, 53]   # 1-based, as Julia indexes them


def main():
    ref = sio.loadmat(SRC)["obstacle_field"].ravel()
    mine = read_obstacle_field(SRC)
    assert len(mine) == len(ref) == 100 and all(np.array_equal(a, b) for a, b in zip(mine, ref))
    print("reader == scipy.io.loadmat on all", len(ref), "trials,", sum(len(a) for a in ref), "obstacles")
    cell = np.empty((len(TRIALS), 1), dtype=object)
    for i, t in enumerate(TRIALS):
        cell[i, 0] = np.asarray(ref[t - 1], dtype=np.float64)
    sio.savemat(os.path.join(HERE, "obstacle_field_subset.mat"), {"obstacle_field": cell}, do_compression=True)
    sio.savemat(os.path.join(HERE, "obstacle_field_subset_raw.mat"), {"obstacle_field": cell}, do_compression=False)
    with open(os.path.join(HERE, "obstacle_field_subset.csv"), "w") as f:
        f.write("# trial (1-based, PathPlanning/Scenarios/obstacle_field.mat), x, y, r\n")
        for t in TRIALS:
            for row in ref[t - 1]:
                f.write(f"{t}," + ",".join(repr(float(v)) for v in row) + "\n")


if __name__ == "__main__":
    main()
