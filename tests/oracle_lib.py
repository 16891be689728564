"""ctypes wrapper of the CPU oracle (oracle/libmppi_oracle.so).

Test infrastructure: imported only by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  Never by the product.

Independent of the product: this module does not import motionplanning_b200 and never loads
libmppi_b200.so.  It carries its own ctypes twin of the POD `mppi_config_t` (the layout of
include/mppi_b200.h is the only thing shared) and takes every default from the oracle's own
`oracle_config_default` (reference literals typed from the reference sources);
tests/test_oracle_config.py checks both twins and both default tables against each other.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(_ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libmppi_oracle.so")

# enums of include/mppi_b200.h
MPPI_MAX_NX, MPPI_MAX_N, MPPI_MAX_GPUS = 8, 256, 8
MODEL_DYNAMIC7, MODEL_KINEMATIC5 = 0, 1
FP32, FP64 = 0, 1
NOISE_SUPPLIED, NOISE_PHILOX = 0, 1
PLANNER_MPPI, PLANNER_DWA = 0, 1
FLAG_FEASIBLE, FLAG_COLLIDED, FLAG_OOB = 1, 2, 4

_dd = C.c_double
_ii = C.c_int32


class Config(C.Structure):
    """mppi_config_t (include/mppi_b200.h)"""

    _fields_ = [
        ("struct_size", C.c_uint32),
        ("model", _ii), ("n_x", _ii), ("m", _ii), ("K", _ii), ("N", _ii), ("B", _ii),
        ("precision", _ii), ("noise_mode", _ii), ("planner", _ii),
        ("feasibility_count", _ii), ("use_cuda_graph", _ii), ("device", _ii),
        ("k_offset", _ii), ("K_local", _ii), ("rollout_variant", _ii),
        ("strict_math", _ii), ("noise_policy", _ii), ("n_gpus", _ii), ("device_ids", _ii * MPPI_MAX_GPUS),
        ("reserved0", _ii),
        ("seed", C.c_uint64),
        ("T", _dd), ("lambda_", _dd),
        ("Sigma", _dd * 4),
        ("X0", _dd * MPPI_MAX_NX),
        ("goal", _dd * 2),
        ("XL", _dd * MPPI_MAX_NX), ("XU", _dd * MPPI_MAX_NX),
        ("CL", _dd * 2), ("CU", _dd * 2),
        ("slack_penalty", _dd), ("obstacle_penalty", _dd), ("terminal_scale", _dd),
        ("la", _dd), ("lb", _dd), ("mass", _dd), ("Izz", _dd), ("g", _dd), ("mu", _dd),
        ("KFZF", _dd), ("KFZR", _dd), ("KFZX", _dd),
        ("tire", _dd * 6),
        ("stage_w", _dd * 6),
        ("veh_L", _dd), ("veh_W", _dd), ("veh_center_offset", _dd),
    ]


_lib = None
_dp = C.POINTER(C.c_double)


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _fill(arr, vals):
    for i, v in enumerate(vals):
        arr[i] = float(v)


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            subprocess.check_call(["make", "-C", ORACLE_DIR])
        _lib = C.CDLL(ORACLE_SO)
        _lib.oracle_mod_pi.restype = C.c_double
        _lib.oracle_mod_pi.argtypes = [C.c_double]
    return _lib


def default_config():
    """The reference's literals from the oracle's own table (oracle_config_default)."""
    cfg = Config()
    load().oracle_config_default(C.byref(cfg))
    assert cfg.struct_size == C.sizeof(Config), (cfg.struct_size, C.sizeof(Config))
    return cfg


def set_strict_f32(on):
    """Strict FP32 flavour: the float instantiation uses include/mppi_strict_math.h (process-wide)."""
    load().oracle_set_strict_f32(C.c_int(1 if on else 0))


def make_config(args, kwargs):
    """Build the mppi_config_t that defineMPPI(*args, **kwargs) would build, without a GPU."""
    (numStates, numControls, lambda_, X0, goal, K, Sigma, N, T, XL, XU, CL, CU) = args
    kw = dict(kwargs)
    cfg = default_config()
    cfg.model = kw.pop("model", MODEL_DYNAMIC7 if numStates == 7 else MODEL_KINEMATIC5)
    cfg.n_x, cfg.m, cfg.K, cfg.N, cfg.B = numStates, numControls, K, N, kw.pop("B", 1)
    cfg.precision = kw.pop("precision", FP32)
    cfg.noise_mode = kw.pop("noise_mode", NOISE_PHILOX)
    cfg.planner = kw.pop("planner", PLANNER_MPPI)
    cfg.feasibility_count = kw.pop("FeasibilityCount", 1300)
    cfg.seed = kw.pop("seed", 0x5EED0001)
    cfg.k_offset, cfg.K_local = kw.pop("k_offset", 0), kw.pop("K_local", 0)
    for k in ("device", "use_cuda_graph", "n_gpus", "device_ids", "noise_policy", "rollout_variant"):
        kw.pop(k, None)
    cfg.strict_math = kw.pop("strict_math", 0)
    cfg.T, cfg.lambda_ = float(T), float(lambda_)
    _fill(cfg.Sigma, np.asarray(Sigma, dtype=np.float64).reshape(-1, order="F"))
    x0 = np.asarray(X0, dtype=np.float64)
    _fill(cfg.X0, x0 if x0.ndim == 1 else x0[:, 0])
    _fill(cfg.goal, goal), _fill(cfg.XL, XL), _fill(cfg.XU, XU)
    _fill(cfg.CL, CL), _fill(cfg.CU, CU)
    for k, v in kw.items():
        cur = getattr(cfg, k)
        if hasattr(cur, "__len__"):
            _fill(cur, v)
        else:
            setattr(cfg, k, v)
    return cfg


def _obs(circles, blocks):
    c = np.ascontiguousarray(np.asarray(circles, dtype=np.float64).reshape(-1, 3))
    b = np.ascontiguousarray(np.asarray(blocks, dtype=np.float64).reshape(-1, 5))
    return c, b


def plan(cfg, circles, blocks, x0, Unom, eps_or_cand, precision=FP64, threads=1, want_margin=False):
    """MPPIPlan / DWAPlan on supplied noise.  Returns a dict of numpy outputs."""
    lib = load()
    c, b = _obs(circles, blocks)
    N, nx, K = cfg.N, cfg.n_x, cfg.K
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    Un = None if Unom is None else np.asfortranarray(np.asarray(Unom, dtype=np.float64))
    if cfg.planner == PLANNER_DWA:
        src = np.asfortranarray(np.asarray(eps_or_cand, dtype=np.float64))
    else:
        src = np.ascontiguousarray(eps_or_cand, dtype=np.float64 if precision == FP64 else np.float32)
        assert src.shape == (N, K, 2), src.shape
    costs = np.empty(K)
    flags = np.empty(K, dtype=np.uint8)
    margins = np.full(K, np.inf)
    U_new = np.empty((N, 2), order="F")
    traj = np.empty((N + 1, nx), order="F")
    sc = np.zeros(10)
    lib.oracle_plan(C.byref(cfg), _d(c), c.shape[0], _d(b), b.shape[0], _d(x0), _d(Un),
                    src.ctypes.data_as(C.c_void_p), precision, threads, 1 if want_margin else 0,
                    _d(costs), flags.ctypes.data_as(C.POINTER(C.c_uint8)), _d(margins), _d(U_new), _d(traj), _d(sc))
    return dict(costs=costs, flags=flags, margins=margins, Control=U_new, Traj=traj, cost=sc[0],
                feasible=bool(sc[1]), rollout_count=int(sc[2]), feasible_count=int(sc[3]), argmin=int(sc[4]),
                rho=sc[5], eta=sc[6], seconds=sc[7], nominal_margin=sc[8], nominal_flags=int(sc[9]))


def rollout(cfg, circles, blocks, x0, Unom, ctrl, precision=FP64):
    lib = load()
    c, b = _obs(circles, blocks)
    N, nx = cfg.N, cfg.n_x
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    Un = None if Unom is None else np.asfortranarray(np.asarray(Unom, dtype=np.float64))
    u = np.asfortranarray(np.asarray(ctrl, dtype=np.float64))
    traj = np.empty((N + 1, nx), order="F")
    sc = np.zeros(5)
    lib.oracle_rollout(C.byref(cfg), _d(c), c.shape[0], _d(b), b.shape[0], _d(x0), _d(Un), _d(u), precision,
                       _d(traj), _d(sc))
    return dict(Traj=traj, cost=sc[0], feasible=bool(sc[1]), collided=bool(sc[2]), oob=bool(sc[3]), margin=sc[4])


def vehicle_dynamics(cfg, x, u):
    lib = load()
    x = np.ascontiguousarray(x, dtype=np.float64)
    u = np.ascontiguousarray(u, dtype=np.float64)
    dx = np.zeros(cfg.n_x)
    cc = np.zeros(2)
    lib.oracle_vehicle_dynamics(C.byref(cfg), _d(x), _d(u), _d(dx), _d(cc))
    return dx, cc[0], cc[1]


def rungekutta2(cfg, x, u):
    lib = load()
    x = np.ascontiguousarray(x, dtype=np.float64)
    u = np.ascontiguousarray(u, dtype=np.float64)
    xn = np.zeros(cfg.n_x)
    cc = np.zeros(2)
    lib.oracle_rungekutta2(C.byref(cfg), _d(x), _d(u), _d(xn), _d(cc))
    return xn, int(cc[0]), cc[1]


def plant_advance(cfg, x, U_new, steps=100, plant_dt=1e-3, update_time=0.1, goal_radius=6.0, time_idx0=0):
    """main.jl:65-67,76-82 for one MPC iteration: returns (x_next, history rows [t, states], done)."""
    lib = load()
    x = np.array(x, dtype=np.float64, copy=True)
    U = np.asfortranarray(np.asarray(U_new, dtype=np.float64))
    his = np.zeros((steps, 1 + cfg.n_x))
    done = C.c_int(0)
    lib.oracle_plant_advance.restype = C.c_int
    n = lib.oracle_plant_advance(C.byref(cfg), _d(x), _d(U), C.c_int(steps), C.c_double(plant_dt), C.c_double(update_time),
                                 C.c_double(goal_radius), C.c_int(time_idx0), _d(his), C.byref(done))
    return x, his[:n].copy(), bool(done.value)


def set_reference_path(pts=None, weight=10.0, window=16):
    """Process-wide reference path of the oracle's tracking term; None clears it."""
    lib = load()
    a = np.zeros((0, 2)) if pts is None else np.ascontiguousarray(np.asarray(pts, dtype=np.float64).reshape(-1, 2))
    lib.oracle_set_reference_path.restype = None
    lib.oracle_set_reference_path(_d(a), C.c_int(a.shape[0]), C.c_double(weight), C.c_int(window))


def get_rectangle_pts(block):
    lib = load()
    b = np.ascontiguousarray(block, dtype=np.float64)
    pts = np.zeros((5, 2))
    lib.oracle_get_rectangle_pts(_d(b), _d(pts))
    return pts.T.copy()  # 2 x 5 like Julia


def _cols(p):
    return np.ascontiguousarray(np.asarray(p, dtype=np.float64).T)  # (n, 2) rows == 2 x n col-major


def sat(base, other):
    lib = load()
    a, b = _cols(base), _cols(other)
    return bool(lib.oracle_sat(_d(a), a.shape[0], _d(b), b.shape[0]))


def convex_collision(p1, p2):
    lib = load()
    a, b = _cols(p1), _cols(p2)
    return bool(lib.oracle_convex_collision(_d(a), a.shape[0], _d(b), b.shape[0]))


def mod_pi(a):
    return load().oracle_mod_pi(float(a))


def check_poses(cfg, circles, blocks, poses):
    lib = load()
    c, b = _obs(circles, blocks)
    p = np.ascontiguousarray(np.asarray(poses, dtype=np.float64).reshape(-1, 3))
    hit = np.zeros(p.shape[0], dtype=np.int32)
    mg = np.zeros(p.shape[0])
    lib.oracle_check_poses(C.byref(cfg), _d(c), c.shape[0], _d(b), b.shape[0], _d(p), p.shape[0],
                           hit.ctypes.data_as(C.POINTER(C.c_int32)), _d(mg))
    return hit, mg


def philox(ctr, key):
    lib = load()
    c = np.asarray(ctr, dtype=np.uint32)
    k = np.asarray(key, dtype=np.uint32)
    out = np.zeros(4, dtype=np.uint32)
    u32 = C.POINTER(C.c_uint32)
    lib.oracle_philox4x32_10(c.ctypes.data_as(u32), k.ctypes.data_as(u32), out.ctypes.data_as(u32))
    return out


def generate_noise(Sigma, seed, iteration, K, N, precision=FP64, k_offset=0, instance=0):
    lib = load()
    S = np.asarray(Sigma, dtype=np.float64).reshape(-1, order="F").copy()
    out = np.empty((N, K, 2), dtype=np.float64 if precision == FP64 else np.float32)
    lib.oracle_generate_noise(_d(S), C.c_uint64(seed), C.c_uint32(iteration), C.c_uint32(instance), k_offset, K, N,
                              precision, out.ctypes.data_as(C.c_void_p))
    return out


def weights(costs, lambda_):
    lib = load()
    c = np.ascontiguousarray(costs, dtype=np.float64)
    w = np.empty_like(c)
    re = np.zeros(2)
    am = C.c_int(0)
    lib.oracle_weights(_d(c), c.shape[0], C.c_double(lambda_), _d(w), _d(re), C.byref(am))
    return w, re[0], re[1], am.value
