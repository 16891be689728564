"""ctypes binding of include/mppi_b200.h (the C ABI of libmppi_b200.so).

This is the ctypes twin of the Julia `ccall` shim in julia/MPPIB200.jl: both bind
exactly the symbols the header declares.  There is no CPU fallback: if the
shared library is missing, importing the product API raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libmppi_b200.so")
if os.environ.get("MPPI_B200_LIB"):   # development aid: A/B runs of experimental kernel builds (tools/build_exp.sh); never a CPU path
    LIB_PATH = os.path.abspath(os.environ["MPPI_B200_LIB"])

MPPI_MAX_NX = 8
MPPI_MAX_N = 256
MPPI_MAX_GPUS = 8
MPPI_STATIC_POSES_MAX = 16384
MPPI_COMM_ID_BYTES = 128
MPPI_PEER_HANDLE_BYTES = 64

MPPI_OK, MPPI_EINVAL, MPPI_ECUDA, MPPI_ENCCL, MPPI_ENOMEM, MPPI_ESTATE = range(6)
MODEL_DYNAMIC7, MODEL_KINEMATIC5 = 0, 1
FP32, FP64 = 0, 1
NOISE_SUPPLIED, NOISE_PHILOX = 0, 1
PLANNER_MPPI, PLANNER_DWA = 0, 1
FLAG_FEASIBLE, FLAG_COLLIDED, FLAG_OOB = 1, 2, 4

_d = C.c_double
_i = C.c_int32


class Config(C.Structure):
    """mppi_config_t"""

    _fields_ = [
        ("struct_size", C.c_uint32),
        ("model", _i), ("n_x", _i), ("m", _i), ("K", _i), ("N", _i), ("B", _i),
        ("precision", _i), ("noise_mode", _i), ("planner", _i),
        ("feasibility_count", _i), ("use_cuda_graph", _i), ("device", _i),
        ("k_offset", _i), ("K_local", _i), ("rollout_variant", _i),
        ("strict_math", _i), ("noise_policy", _i), ("n_gpus", _i), ("device_ids", _i * MPPI_MAX_GPUS), ("reserved0", _i),
        ("seed", C.c_uint64),
        ("T", _d), ("lambda_", _d),
        ("Sigma", _d * 4),
        ("X0", _d * MPPI_MAX_NX),
        ("goal", _d * 2),
        ("XL", _d * MPPI_MAX_NX), ("XU", _d * MPPI_MAX_NX),
        ("CL", _d * 2), ("CU", _d * 2),
        ("slack_penalty", _d), ("obstacle_penalty", _d), ("terminal_scale", _d),
        ("la", _d), ("lb", _d), ("mass", _d), ("Izz", _d), ("g", _d), ("mu", _d),
        ("KFZF", _d), ("KFZR", _d), ("KFZX", _d),
        ("tire", _d * 6),
        ("stage_w", _d * 6),
        ("veh_L", _d), ("veh_W", _d), ("veh_center_offset", _d),
    ]


class Result(C.Structure):
    """mppi_result_t"""

    _fields_ = [
        ("Control", C.c_void_p), ("Traj", C.c_void_p),   # double*: plain addresses, so the per-step path assigns ints
        ("cost", _d),
        ("feasible", _i), ("rollout_count", _i), ("feasible_count", _i), ("argmin_index", _i),
        ("rho", _d), ("eta", _d), ("time_s", _d),
    ]


class Partial(C.Structure):
    """mppi_partial_t"""

    _fields_ = [
        ("rho", _d), ("eta", _d),
        ("argmin_global", C.c_int64),
        ("rollouts", _i), ("feasible", _i),
        ("S", _d * (2 * MPPI_MAX_N)),
    ]


class ClosedLoop(C.Structure):
    """mppi_closed_loop_t"""

    _fields_ = [
        ("struct_size", C.c_uint32), ("max_iters", _i), ("plant_steps", _i), ("shift_nominal", _i),
        ("plant_dt", _d), ("update_time", _d), ("goal_radius", _d),
    ]


_P = C.POINTER
_H = C.c_void_p

# name -> (restype, argtypes); every symbol include/mppi_b200.h declares
SIGNATURES = {
    "mppi_config_default": (_i, [_P(Config)]),
    "mppi_version": (_i, []),
    "mppi_create": (_i, [_P(Config), _P(_H)]),
    "mppi_destroy": (_i, [_H]),
    "mppi_last_error": (C.c_char_p, [_H]),
    "mppi_set_x0": (_i, [_H, _P(_d), _i]),
    "mppi_set_nominal": (_i, [_H, _P(_d), _i, _i]),
    "mppi_set_sigma_lambda": (_i, [_H, _P(_d), _d]),
    "mppi_set_circles": (_i, [_H, _P(_d), _i]),
    "mppi_set_blocks": (_i, [_H, _P(_d), _i]),
    "mppi_set_goal": (_i, [_H, _P(_d)]),
    "mppi_set_reference_path": (_i, [_H, _P(_d), _i, _d, _i]),
    "mppi_set_candidates": (_i, [_H, _P(_d), _i, _i]),
    "mppi_step": (_i, [_H, C.c_void_p, _P(Result)]),
    "mppi_adopt_nominal": (_i, [_H]),
    "mppi_step_async": (_i, [_H, _i]),
    "mppi_sync": (_i, [_H]),
    "mppi_closed_loop_default": (_i, [_P(ClosedLoop)]),
    "mppi_run_closed_loop": (_i, [_H, _P(ClosedLoop), _P(_d), _P(_i), _P(_d), _P(_i)]),
    "mppi_fetch_result": (_i, [_H, _P(Result)]),
    "mppi_get_costs": (_i, [_H, _P(_d)]),
    "mppi_get_flags": (_i, [_H, _P(C.c_uint8)]),
    "mppi_get_noise": (_i, [_H, C.c_void_p]),
    "mppi_get_sample_xy": (_i, [_H, _i, _P(C.c_float), _P(_i)]),
    "mppi_step_batched": (_i, [_H, _P(_d), _P(_d), _P(_d), _P(_d), _P(_i)]),
    "mppi_step_local": (_i, [_H, C.c_void_p, _P(Partial)]),
    "mppi_step_combine": (_i, [_H, _P(Partial), _i, _P(Result)]),
    "mppi_comm_unique_id": (_i, [_P(C.c_uint8)]),
    "mppi_comm_init": (_i, [_H, _P(C.c_uint8), _i, _i]),
    "mppi_peer_export": (_i, [_H, _P(C.c_uint8)]),
    "mppi_peer_init": (_i, [_H, _P(C.c_uint8), _i, _i]),
    "mppi_get_timings": (_i, [_H, _P(C.c_float)]),
    "mppi_get_stream": (_i, [_H, _P(C.c_void_p)]),
    "mppi_get_phase_clocks": (_i, [_H, _P(C.c_uint64)]),
    "mppi_set_profiling": (_i, [_H, _i]),
    "mppi_get_plan_info": (_i, [_H, _P(_i)]),
    "mppi_launch_count": (C.c_int64, [_H]),
    "mppi_check_poses": (_i, [_H, _P(_d), _i, _P(_i)]),
    "mppi_check_poses_static": (_i, [_H, _P(_d), _i, _P(_i)]),
    "mppi_measure_fp32_peak": (_i, [_H, _P(_d)]),
    "mppi_time_reference_calls": (_i, [_H, _P(_d), _P(_d), _i, C.c_void_p, C.c_size_t, _P(_d), _P(_d)]),
}

_lib = None


def load():
    """Load libmppi_b200.so (built in-tree by __graft_entry__.build()).  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C motionplanning_b200/csrc`.  There is no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the ABI is incomplete
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def default_config():
    cfg = Config()
    rc = load().mppi_config_default(C.byref(cfg))
    if rc != MPPI_OK:
        raise RuntimeError(f"mppi_config_default failed: {rc}")
    return cfg
