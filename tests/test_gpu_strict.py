"""FP32 `strict_math` flavour: per-sample costs, feasible / collided / out-of-bounds flags and the argmin are
BIT-IDENTICAL to a float restatement of the reference (north_star: "argmin sample index and collision flags must be
bit-exact"), with ZERO exclusions.

Both sides execute the reference's evaluation order (MPPIUtils.jl:31-57,76-90,113-151; vehicledynamics.jl:30-52;
Parking/src/VehicleModel.jl:14-18; CollisionDetection/src/utils.jl:14-74) in unfused IEEE single precision with the
elementary functions of include/mppi_strict_math.h: the GPU in csrc/mppi_strict.cuh, the checker as the oracle's
float instantiation with `oracle_set_strict_f32(1)`.  The comparisons that must not flip are MPPIUtils.jl:55,125,139,144,155."""
import numpy as np
import pytest

import motionplanning_b200 as m
import oracle_lib as orc
from motionplanning_b200 import scenarios as sc
from test_gpu_parity import warm_nominal

capi = m.capi
pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _strict_oracle():
    orc.set_strict_f32(True)
    yield
    orc.set_strict_f32(False)


def strict_pair(args, kw, circles, blocks, Unom=None, x0=None):
    K, N = args[5], args[7]
    if x0 is not None:
        args = args[:3] + (x0,) + args[4:]
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_SUPPLIED, strict_math=1)
    eps = orc.generate_noise(args[6], kw.get("seed", sc.SEED), 0, K, N, capi.FP32)
    mppi = m.defineMPPI(*args, **kw)
    assert mppi.plan_info()["strict"] == 1
    if circles:
        m.defineMPPIobs(mppi, circles)
    if blocks:
        m.defineMPPIblocks(mppi, blocks)
    U = np.zeros((N, 2)) if Unom is None else Unom
    m.defineMPPINominalControl(mppi, U)
    m.MPPIPlan(mppi, eps)
    ref = orc.plan(orc.make_config(args, kw), circles, blocks, args[3], U, eps, capi.FP32, threads=8)
    return mppi, ref


def check_bit_exact(mppi, ref):
    R = ref["rollout_count"] - 1
    costs, flags = mppi.p.costs, mppi.p.flags
    same = costs[:R].astype(np.float32).view(np.uint32) == ref["costs"][:R].astype(np.float32).view(np.uint32)
    assert same.all(), f"{(~same).sum()} of {R} per-sample costs differ in their bits (first at {np.argmin(same)})"
    assert np.array_equal(flags[:R], ref["flags"][:R]), "flags differ"
    assert mppi.r.argmin == ref["argmin"], (mppi.r.argmin, ref["argmin"])
    assert mppi.r.RolloutCount == ref["rollout_count"] and mppi.r.FeasibleTrajCount == ref["feasible_count"]
    assert np.float32(mppi.r.rho) == np.float32(ref["rho"])
    # the weighted update sums K terms in a different order: tolerance of the FP32 mode
    scale = max(np.abs(ref["Control"]).max(), 1e-12)
    assert np.abs(mppi.r.Control - ref["Control"]).max() / scale <= 1e-4


def test_strict_cfg1_kinematic_circles(built):
    args, kw, circles, blocks = sc.config(0)
    check_bit_exact(*strict_pair(args, kw, circles, blocks))
    check_bit_exact(*strict_pair(args, kw, circles, blocks, Unom=warm_nominal(args[7])))


def test_strict_cfg2_shape_with_collisions(built):
    """K = 16,384 of cfg 2's scene started next to the first obstacle: hits, bound violations and feasible samples occur."""
    args, kw, circles, blocks = sc.config(1, K=16384)
    mppi, ref = strict_pair(args, kw, circles, blocks, Unom=warm_nominal(50), x0=[38.0, -1.0, 0.05, 8.0, 0.02])
    f = ref["flags"]
    assert (f & capi.FLAG_COLLIDED).any() and (f & capi.FLAG_OOB).any() and (f & capi.FLAG_FEASIBLE).any()
    check_bit_exact(mppi, ref)


def test_strict_cfg3_parking_walls_sat(built):
    """The literal 4-edge x 10-point x 2-direction SeparatingAxisTheorem against the parking layout."""
    args, kw, circles, blocks = sc.config(2, K=4096, N=80)
    mppi, ref = strict_pair(args, kw, circles, blocks, Unom=warm_nominal(80))
    assert (ref["flags"] & capi.FLAG_COLLIDED).any()
    check_bit_exact(mppi, ref)


@pytest.mark.parametrize("M_p", [3, 16])
def test_strict_cfg4_dynamic_walls(built, M_p):
    args, kw, circles, blocks = sc.config(3, K=4096, N=100, M_p=M_p)
    check_bit_exact(*strict_pair(args, kw, circles, blocks, Unom=warm_nominal(100)))
    # from a pose inside the wall field so that the exact SAT decides many samples
    mppi, ref = strict_pair(args, kw, circles, blocks, Unom=warm_nominal(100), x0=[42.0, -0.5, 0.0, 0.0, 0.05, 8.0, 0.0])
    assert (ref["flags"] & capi.FLAG_COLLIDED).any()
    check_bit_exact(mppi, ref)


def test_strict_reference_demo_with_cutoff(built):
    """The shipped scenario (main.jl:7-30, K = 1500, N = 20, FeasibilityCount = 1300): the cut-off index is data
    dependent on the feasibility flags, so it is bit-exact only if every flag is."""
    args = (7, 2, 25.0, sc.X0_DYN, [110, 0], 2048, sc.SIGMA, 20, 3.0, sc.XL_DYN, sc.XU_DYN, sc.CL, sc.CU)
    mppi, ref = strict_pair(args, dict(FeasibilityCount=1300), sc.CIRCLES, [])
    assert ref["rollout_count"] - 1 < 2048
    check_bit_exact(mppi, ref)


def test_strict_philox_noise_replayed(built):
    """PHILOX mode in the strict flavour: the device draws the noise, the checker replays the dumped tensor."""
    args, kw, circles, blocks = sc.config(0, K=4096)
    kw = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, strict_math=1)
    mppi = m.defineMPPI(*args, **kw)
    m.defineMPPIobs(mppi, circles)
    U = np.zeros((50, 2))
    for _ in range(3):
        m.defineMPPINominalControl(mppi, U)
        m.MPPIPlan(mppi)
        eps = mppi.noise()
        ref = orc.plan(orc.make_config(args, dict(kw, noise_mode=capi.NOISE_SUPPLIED)), circles, blocks, args[3], U, eps, capi.FP32, threads=8)
        check_bit_exact(mppi, ref)
        U = mppi.r.Control.copy()


def test_strict_math_functions_are_accurate():
    """The portable functions are libm-class: the strict float oracle agrees with the FP64 oracle to the FP32 tolerance."""
    args, kw, circles, blocks = sc.config(0, K=512)
    kw32 = dict(kw, precision=capi.FP32, noise_mode=capi.NOISE_SUPPLIED)
    eps = orc.generate_noise(args[6], sc.SEED, 0, 512, 50, capi.FP32)
    U = warm_nominal(50)
    a = orc.plan(orc.make_config(args, kw32), circles, blocks, args[3], U, eps, capi.FP32)
    b = orc.plan(orc.make_config(args, kw32), circles, blocks, args[3], U, eps.astype(np.float64), capi.FP64)
    assert (np.abs(a["costs"] - b["costs"]) / np.abs(b["costs"])).max() < 1e-4
