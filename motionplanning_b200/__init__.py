"""motionplanning_b200 — B200-native MPPI engine behind the reference's MPPI interface.

Product = motionplanning_b200/csrc/libmppi_b200.so (hand-written sm_100a CUDA, C ABI in
include/mppi_b200.h).  This package is the thin host-side mirror of the reference's
Julia API (OptimalControl/MPPI/src/setup.jl, MPPIUtils.jl); see julia/MPPIB200.jl for
the `ccall` twin.
"""
from . import _capi as capi  # noqa: F401
from .dwa import DWAPlan, defineDWA, defineDWAcontrols, defineDWAobs  # noqa: F401
from .mppi import (  # noqa: F401
    MPPIError,
    MPPIPlan,
    MPPISearcher,
    ShiftInitialCondition,
    defineMPPI,
    defineMPPIblocks,
    defineMPPINominalControl,
    defineMPPIobs,
    defineMPPIrefpath,
    write_trajectory_csv,
)
