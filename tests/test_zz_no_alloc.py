"""mppi_check_poses_static: the allocation-free variant of mppi_check_poses (its device scratch is a file-scope static
__device__ array, no counterpart in the reference).  Called through the
C ABI and checked against the existing entry point on two calls in a row (the shared array keeps the previous call's
contents), at the size limit, and beyond it.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import motionplanning_b200 as m
from motionplanning_b200 import scenarios as sc

capi = m.capi
pytestmark = pytest.mark.gpu


def _both(mppi, poses):
    lib = capi.load()
    p = np.ascontiguousarray(np.asarray(poses, dtype=np.float64).reshape(-1, 3))
    n = p.shape[0]
    a = np.full(n, -7, dtype=np.int32)
    b = np.full(n, -9, dtype=np.int32)
    dp = p.ctypes.data_as(C.POINTER(C.c_double))
    assert lib.mppi_check_poses(mppi._h, dp, n, a.ctypes.data_as(C.POINTER(C.c_int32))) == capi.MPPI_OK
    rc = lib.mppi_check_poses_static(mppi._h, dp, n, b.ctypes.data_as(C.POINTER(C.c_int32)))
    assert rc == capi.MPPI_OK, lib.mppi_last_error(mppi._h)
    return a, b


def test_static_scratch_entry_point_matches_check_poses_twice_in_a_row(built):
    args, kw, circles, _ = sc.config(0)
    mppi = m.defineMPPI(*args, **dict(kw, precision=capi.FP64))
    m.defineMPPIobs(mppi, circles)
    m.defineMPPIblocks(mppi, sc.wall_blocks(16))
    rng = np.random.RandomState(11)

    def poses(n):
        return np.column_stack([rng.uniform(15, 105, n), rng.uniform(-9, 9, n), rng.uniform(-np.pi, np.pi, n)])

    # first call: many poses, many hits; second call: FEWER poses far from every obstacle -> stale hits of the first
    # call must not leak through the shared array
    a, b = _both(mppi, poses(5000))
    assert np.array_equal(a, b) and (a > 0).any() and (a == 0).any()      # integer work: exact
    far = np.column_stack([np.full(300, -500.0), np.linspace(-5, 5, 300), np.zeros(300)])
    a, b = _both(mppi, far)
    assert np.array_equal(a, b) and not b.any()
    a, b = _both(mppi, poses(capi.MPPI_STATIC_POSES_MAX))                  # exactly the limit
    assert np.array_equal(a, b)
    # beyond the limit: rejected with a clear error, nothing written
    lib = capi.load()
    big = np.ascontiguousarray(poses(capi.MPPI_STATIC_POSES_MAX + 1))
    out = np.full(big.shape[0], -3, dtype=np.int32)
    rc = lib.mppi_check_poses_static(mppi._h, big.ctypes.data_as(C.POINTER(C.c_double)), big.shape[0],
                                     out.ctypes.data_as(C.POINTER(C.c_int32)))
    assert rc == capi.MPPI_EINVAL and b"MPPI_STATIC_POSES_MAX" in lib.mppi_last_error(mppi._h)
    assert (out == -3).all()
    # a second handle (its own stream) shares the same array: sequential use is safe
    other = m.defineMPPI(*args, **dict(kw, precision=capi.FP32))
    m.defineMPPIobs(other, circles)
    a, b = _both(other, poses(1000))
    assert np.array_equal(a, b)
