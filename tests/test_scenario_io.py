"""Scenario I/O: the MAT-5 reader against the fixture cut from the reference's
obstacle_field.mat (tests/golden/make_obstacle_field.py, which also checked the reader against
scipy.io.loadmat on the whole reference file), and the MPPITrajectory.csv writer."""
import os

import numpy as np
import pytest

from motionplanning_b200 import scenario_io as sio_b200
from motionplanning_b200.mppi import write_trajectory_csv

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _expected():
    a = np.loadtxt(os.path.join(GOLD, "obstacle_field_subset.csv"), delimiter=",", comments="#")
    trials = []
    for t in dict.fromkeys(a[:, 0].astype(int)):   # file order
        trials.append(a[a[:, 0] == t, 1:])
    return trials


@pytest.mark.parametrize("name", ["obstacle_field_subset.mat", "obstacle_field_subset_raw.mat"])
def test_obstacle_field_reader_matches_reference_values(name):
    got = sio_b200.read_obstacle_field(os.path.join(GOLD, name))
    want = _expected()
    assert len(got) == len(want) == 5
    for g, w in zip(got, want):
        assert g.dtype == np.float64 and g.shape == w.shape and np.array_equal(g, w)
    assert got[-1].shape[1] == 3 and (got[-1][:, 2] > 0).all()   # Trial 53 of Astar/main.jl:18: radii


def test_reader_rejects_other_files(tmp_path):
    p = tmp_path / "x.mat"
    p.write_bytes(b"not a mat file" * 20)
    with pytest.raises(sio_b200.MatFormatError):
        sio_b200.read_mat5(p)
    with pytest.raises(KeyError):
        sio_b200.read_obstacle_field(os.path.join(GOLD, "obstacle_field_subset.mat"), name="nope")


def test_trajectory_csv_round_trip_and_reference_layout(tmp_path):
    """main.jl:87 writes states_his' without a header: t, x, y, v, r, psi, ux, sa per row; the
    reference's own log starts `0.0,0.0,...,5.0,0.0` (tests/golden/mppi_plant_pairs.csv row 0)."""
    pairs = np.loadtxt(os.path.join(GOLD, "mppi_plant_pairs.csv"), delimiter=",", comments="#")
    his = np.column_stack([pairs[:5, 0] * 1e-3, pairs[:5, 1:8]])
    path = tmp_path / "MPPITrajectory.csv"
    write_trajectory_csv(path, his)
    text = open(path).read().splitlines()
    assert len(text) == 5 and text[0] == "0.0,0.0,0.0,0.0,0.0,0.0,5.0,0.0"
    assert np.array_equal(np.loadtxt(path, delimiter=","), his)
