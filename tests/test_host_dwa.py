"""Host-side mirror of the reference's DWA interface (motionplanning_b200/dwa.py): what can be checked
without a GPU -- the candidate grid of defineDWAcontrols! / DecodeControl and the argument checks."""
import math

import numpy as np
import pytest

import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc


def _decode_control(number, CtlDecode):
    """DecodeControl (DynamicWindow/src/setup.jl:87-95), 1-based as in the reference."""
    idx = []
    for d in CtlDecode:
        i = int(math.ceil(number / d))
        idx.append(i)
        number = number - (i - 1) * d
    return idx


@pytest.mark.parametrize("SampleNumber", [[31, 41], [3, 2], [1, 5], [4, 1]])
def test_control_samples_follow_decode_control(SampleNumber):
    """defineDWAcontrols! (setup.jl:59-85): ControlSamples[i, j] = LinRange(CL[j], CU[j], n_j)[DecodeControl(i)[j]],
    CtlDecode = [n_2, 1]."""
    CL, CU = [-0.3, -2.5], [0.3, 2.5]
    got = m.defineDWAcontrols(SampleNumber, CL, CU)
    grids = [np.linspace(CL[c], CU[c], SampleNumber[c]) for c in range(2)]
    CtlDecode = [SampleNumber[1], 1]
    want = np.array([[grids[c][_decode_control(i, CtlDecode)[c] - 1] for c in range(2)]
                     for i in range(1, SampleNumber[0] * SampleNumber[1] + 1)])
    assert got.shape == want.shape and np.array_equal(got, want)


def test_define_dwa_checks_are_the_references(built):
    """setup.jl:16-36: same seven checks, same order, same wording -- raised before any device work."""
    ok = dict(numStates=7, numControls=2, X0=sc.X0_DYN, goal=[110, 0], SampleNumber=[3, 3], T=3.0,
              XL=sc.XL_DYN, XU=sc.XU_DYN, CL=[-0.3, -2.5], CU=[0.3, 2.5])
    for change, text in [(dict(numControls=0), "Controls (0) must be > 0"),
                         (dict(numStates=0), "States (0) must be > 0"),
                         (dict(X0=[0, 0, 0]), "Length of X0 (3) must match number of states (7)"),
                         (dict(XL=[0]), "Length of XL (1) must match number of states (7)"),
                         (dict(XU=[0, 1]), "Length of XU (2) must match number of states (7)"),
                         (dict(CL=[0]), "Length of CL (1) must match number of controls (2)"),
                         (dict(CU=[0, 1, 2]), "Length of CU (3) must match number of controls (2)")]:
        with pytest.raises(m.MPPIError) as e:
            m.defineDWA(**dict(ok, **change))
        assert text in str(e.value)
