"""Small invocations of every kernel flavour of the MPPI step through the reference-facing calls (2 chained iterations each,
results checked for finiteness only — parity is the tests' job).  A quick tour for a fresh GPU box: `python tools/flavour_tour.py
[case ...]`, about 3 s for all 21 cases.  It was written as the workload for compute-sanitizer (memcheck / racecheck / synccheck);
that tool is closed on this GPU pool, so the kernels' own guards stand in for it: bounds are asserted on the host before every
launch (mppi_api.cu validate), exchanges time out into MPPI_ENCCL, and every flavour is compared with the CPU oracle in tests/.
Product path only: nothing from oracle/ is loaded here."""
import sys

import numpy as np

sys.path.insert(0, ".")
import motionplanning_b200 as m   # noqa: E402
from motionplanning_b200 import dwa, scenarios as sc   # noqa: E402

capi = m.capi
ONLY = set(sys.argv[1:])


def case(name):
    def deco(fn):
        if ONLY and name not in ONLY:
            return fn
        try:
            fn()
            print(f"case {name}: ok", flush=True)
        except Exception as e:   # keep going, report at the end
            print(f"case {name}: FAILED {type(e).__name__}: {e}", flush=True)
        return fn
    return deco


def chain(mp, steps=2, noise=None):
    U = np.zeros((mp.s.N, 2))
    for _ in range(steps):
        m.defineMPPINominalControl(mp, U)
        m.MPPIPlan(mp, noise)
        U = mp.r.Control.copy()
    assert np.isfinite(mp.r.Control).all() and np.isfinite(mp.p.costs).all()
    mp.p.flags
    mp.p.sample_xy(stride=16)


def make(idx, K, N=None, circles=True, blocks=True, **kw):
    args, k, c, b = sc.config(idx, K=K, N=N)
    mp = m.defineMPPI(*args, **dict(k, **kw))
    if c and circles:
        m.defineMPPIobs(mp, c)
    if b and blocks:
        m.defineMPPIblocks(mp, b)
    return mp


for v, K in ((6, 1022), (6, 64), (2, 1022), (1, 1023), (3, 1023), (4, 1022)):
    @case(f"kin_circles_fp32_variant{v}_K{K}")
    def _(v=v, K=K):
        mp = make(0, K, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=v)
        chain(mp)
        mp.close()


@case("kin_circles_fp32_graph_off")
def _():
    mp = make(0, 1022, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=6, use_cuda_graph=False)
    chain(mp)
    mp.close()


@case("kin_circles_fp32_scan_form")
def _():
    mp = make(0, 1024, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=5)
    chain(mp)
    mp.close()


@case("kin_circles_fp64_supplied")
def _():
    mp = make(0, 777, precision=capi.FP64, noise_mode=capi.NOISE_SUPPLIED)
    eps = np.random.RandomState(1).standard_normal((50, 777, 2)) * 0.3
    chain(mp, noise=eps)
    mp.close()


@case("kin_circles_fp32_strict")
def _():
    mp = make(0, 1024, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, strict_math=1)
    chain(mp)
    mp.close()


@case("parking_walls_fp32")
def _():
    mp = make(2, 2048, N=80, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
    chain(mp)
    mp.close()


for M_p in (3, 16):
    @case(f"dynamic_walls_fp32_Mp{M_p}")
    def _(M_p=M_p):
        args, k, c, b = sc.config(3, K=2048, N=100, M_p=M_p)
        args = args[:3] + ([38.0, 0, 0, 0, 0, 8.0, 0],) + args[4:]
        mp = m.defineMPPI(*args, **dict(k, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX))
        m.defineMPPIblocks(mp, b)
        chain(mp)
        mp.close()


@case("dynamic_fp64_cutoff")
def _():
    args, k, c, b = sc.config(3, K=1024, N=40)
    mp = m.defineMPPI(*args, **dict(k, precision=capi.FP64, noise_mode=capi.NOISE_PHILOX, FeasibilityCount=100))
    m.defineMPPIobs(mp, sc.CIRCLES)
    chain(mp)
    mp.close()


@case("refpath_tracking")
def _():
    mp = make(0, 1024, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
    xs = np.linspace(0, 110, 64)
    m.defineMPPIrefpath(mp, np.stack([xs, 0.5 * np.sin(xs / 20)], 1))
    chain(mp)
    mp.close()


@case("batched_B6")
def _():
    args, k, c, b = sc.config(4, K=256)
    mp = m.defineMPPI(*args, **dict(k, B=6, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX))
    m.defineMPPIobs(mp, c)
    out, costs, feas = mp.step_batched(sc.jittered_x0(6), np.zeros((50, 2, 6)))
    out, costs, feas = mp.step_batched(None, out)
    assert np.isfinite(out).all()
    mp.close()


@case("device_chain_and_closed_loop")
def _():
    mp = make(0, 1024, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
    m.MPPIPlan(mp)   # adopt needs a previous result
    for _ in range(3):
        mp.step_async(adopt=True)
    mp.sync()
    mp.fetch_result()
    his, log = mp.run_closed_loop(max_iters=3, plant_steps=20)
    assert np.isfinite(his).all()
    mp.close()


@case("check_poses")
def _():
    mp = make(2, 256, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX)
    poses = np.random.RandomState(2).uniform(-5, 10, (1000, 3))
    mp.check_poses(poses)
    mp.close()


@case("dwa_plan")
def _():
    d = dwa.defineDWA(7, 2, sc.X0_DYN, sc.GOAL, (10, 10), 3.0, sc.XL_DYN, sc.XU_DYN, sc.CL, sc.CU)
    dwa.defineDWAobs(d, sc.CIRCLES)
    dwa.DWAPlan(d)
    d.close()


@case("split_phase_shards")
def _():
    args, k, c, b = sc.config(0, K=1024)
    hs = [m.defineMPPI(*args, **dict(k, k_offset=512 * r, K_local=512, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX,
                                     rollout_variant=3)) for r in range(2)]
    for h in hs:
        m.defineMPPIobs(h, c)
        m.defineMPPINominalControl(h, np.zeros((50, 2)))
    parts = [h.step_local() for h in hs]
    for h in hs:
        h.step_combine(parts)
    for h in hs:
        h.close()


@case("one_handle_two_logical_shards")
def _():
    mp = make(0, 1024, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, rollout_variant=1, n_gpus=2, device_ids=[0, 0])
    chain(mp)
    mp.close()


print("flavour_tour: done", flush=True)
