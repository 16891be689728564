"""Host-side mirror of the reference's Julia MPPI interface, over the C ABI.

Same names, positional arguments and error behaviour as
OptimalControl/MPPI/src/setup.jl and MPPIUtils.jl (Julia's `!` suffix is dropped:
``defineMPPIobs!`` -> ``defineMPPIobs``), so the parity tests read like the
reference's own ``main.jl``:

    mppi = defineMPPI(numStates, numControls, lambda_, X0, goal, SamplingNumber, Sigma, N, T, XL, XU, CL, CU)
    defineMPPIobs(mppi, [[50.0, 1, 2.5], ...])
    ShiftInitialCondition(mppi, states)
    defineMPPINominalControl(mppi, NominalControls)
    MPPIPlan(mppi)
    mppi.r.Control, mppi.r.Traj, mppi.r.Feasibility, mppi.r.cost, mppi.r.RolloutCount

Everything numerical happens in libmppi_b200.so (CUDA, sm_100a).  No CPU path.
julia/MPPIB200.jl is the same shim written with `ccall`.
"""
import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _capi as capi


class MPPIError(RuntimeError):
    """Raised where the reference calls error(...) (setup.jl:19-39) or the library fails."""


def _dptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


@dataclass
class MPPISetting:  # types.jl:10-31
    numStates: int = 0
    numControls: int = 0
    X0: np.ndarray = None
    XL: np.ndarray = None
    XU: np.ndarray = None
    CL: np.ndarray = None
    CU: np.ndarray = None
    dt: float = 0.0
    T: float = 0.0
    N: int = 0
    goal: np.ndarray = None
    obstacle_list: list = field(default_factory=list)
    block_list: list = field(default_factory=list)
    SamplingNumber: int = 0
    FeasibilityCount: int = 1300
    NominalControl: np.ndarray = None
    Σ: np.ndarray = None
    Sigma: np.ndarray = None
    lambda_: float = 0.02
    SlackPenalty: float = 1e5


class MPPIPlanner:  # types.jl:32-34
    """``p.TrajectoryCollection`` is materialised lazily from device memory."""

    def __init__(self, searcher):
        self._m = searcher

    @property
    def costs(self):
        return self._m._costs()

    @property
    def flags(self):
        return self._m._flags()

    def sample_xy(self, stride=1):
        """xy of every stride-th sampled trajectory (plotRes, MPPIUtils.jl:221-223)."""
        return self._m._sample_xy(stride)


class MPPIResult:  # types.jl:36-44
    """mppi.r of the reference.  `Control` and `Traj` are fresh arrays per plan, like the reference's; on the per-step path they are
    copied out of the handle's staging buffers when first READ (a 20 ms control loop reads Control[1, :] and rarely the
    trajectory: two unconditional copies cost 1 us of every step's host time)."""

    def __init__(self):
        self._Traj = self._Control = self._pending = None
        self.Feasibility = "InFeasible"
        self.cost = 1e6
        self.time = 0.0
        self.FeasibleTrajCount = 0
        self.RolloutCount = 0
        self.argmin = 0
        self.rho = 0.0
        self.eta = 0.0

    @property
    def Control(self):
        if self._Control is None and self._pending is not None:
            self._Control = self._pending[0].copy(order="F")
        return self._Control

    @Control.setter
    def Control(self, v):
        self._Control = v

    @property
    def Traj(self):
        if self._Traj is None and self._pending is not None:
            self._Traj = self._pending[1].copy(order="F")
        return self._Traj

    @Traj.setter
    def Traj(self, v):
        self._Traj = v


class MPPISearcher:  # types.jl:46-50
    def __init__(self, cfg):
        self._lib = capi.load()
        self._cfg = cfg
        self._h = capi._H()
        rc = self._lib.mppi_create(C.byref(cfg), C.byref(self._h))
        if rc != capi.MPPI_OK:
            raise MPPIError(self._lib.mppi_last_error(None).decode())
        self.s = MPPISetting()
        self.p = MPPIPlanner(self)
        self.r = MPPIResult()
        # Per-step fast path (main.jl:59-61 runs these three calls every 20 ms): staging buffers whose
        # addresses are taken once, prototypes that take plain addresses, one reusable result struct.
        # Building a ctypes pointer object per call costs more than the copy it describes.
        lib, vp = self._lib, C.c_void_p
        self._hv = self._h.value
        self._x0_buf = np.zeros(cfg.n_x, dtype=np.float64)
        self._u_buf = np.zeros((cfg.N, 2), dtype=np.float64, order="F")
        self._x0_ptr, self._u_ptr = self._x0_buf.ctypes.data, self._u_buf.ctypes.data
        self._f_set_x0 = C.CFUNCTYPE(C.c_int32, vp, vp, C.c_int32)(("mppi_set_x0", lib))
        self._f_set_nominal = C.CFUNCTYPE(C.c_int32, vp, vp, C.c_int32, C.c_int32)(("mppi_set_nominal", lib))
        self._f_step = C.CFUNCTYPE(C.c_int32, vp, vp, vp)(("mppi_step", lib))
        self._res = capi.Result()
        self._res_ptr = C.addressof(self._res)
        # result staging: the library fills these two arrays (addresses taken once: `ndarray.ctypes.data` costs
        # 2 us a time, and it sits in FRONT of the step); every plan hands out fresh copies, like the reference
        self._ctrl_buf = np.zeros((cfg.N, 2), dtype=np.float64, order="F")
        self._traj_buf = np.zeros((cfg.N + 1, cfg.n_x), dtype=np.float64, order="F")
        self._ctrl_ptr, self._traj_ptr = self._ctrl_buf.ctypes.data, self._traj_buf.ctypes.data
        self._pending = (self._ctrl_buf, self._traj_buf)
        self._res_step = capi.Result()
        self._res_step.Control, self._res_step.Traj = self._ctrl_ptr, self._traj_ptr
        self._res_step_ptr = C.addressof(self._res_step)

    # -- plumbing ---------------------------------------------------------
    def _check(self, rc):
        if rc != capi.MPPI_OK:
            raise MPPIError(self._lib.mppi_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.mppi_destroy(self._h)
            self._h = capi._H()
            self._hv = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def K_local(self):
        """Samples this handle exposes (costs, flags, noise): its shard, or all K for a single-process multi-GPU handle."""
        if self._cfg.n_gpus > 1:
            return self._cfg.K
        return self._cfg.K_local if self._cfg.K_local > 0 else self._cfg.K

    def _costs(self):
        out = np.empty(self._cfg.B * self.K_local, dtype=np.float64)
        self._check(self._lib.mppi_get_costs(self._h, _dptr(out)))
        return out

    def _flags(self):
        out = np.empty(self._cfg.B * self.K_local, dtype=np.uint8)
        self._check(self._lib.mppi_get_flags(self._h, out.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out

    def noise(self):
        """eps used by the last step, shape (N, K_local, m) in C order == Julia (m, K, N)."""
        dt = np.float64 if self._cfg.precision == capi.FP64 else np.float32
        out = np.empty((self._cfg.N, self.K_local, 2), dtype=dt)
        self._check(self._lib.mppi_get_noise(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def _sample_xy(self, stride):
        n = C.c_int32(0)
        self._check(self._lib.mppi_get_sample_xy(self._h, stride, None, C.byref(n)))
        out = np.empty((n.value, self._cfg.N + 1, 2), dtype=np.float32)
        self._check(self._lib.mppi_get_sample_xy(self._h, stride, out.ctypes.data_as(C.POINTER(C.c_float)), C.byref(n)))
        return out

    def timings_us(self):
        us = (C.c_float * 4)()
        self._check(self._lib.mppi_get_timings(self._h, us))
        return list(us)

    def stream(self):
        """cudaStream_t (as int) the handle's steps are enqueued on."""
        st = C.c_void_p()
        self._check(self._lib.mppi_get_stream(self._h, C.byref(st)))
        return st.value or 0

    def phase_clocks(self):
        """clock64 stamps at the phase boundaries of the last finalize kernel (diagnostics)."""
        ck = (C.c_uint64 * 16)()
        self._check(self._lib.mppi_get_phase_clocks(self._h, ck))
        return list(ck)

    def plan_info(self):
        """Which kernels the next step runs: dict(scan, pregen, packed, scan_ctas, ksplit, jsplit, strict, n_gpus)."""
        a = (C.c_int32 * 8)()
        self._check(self._lib.mppi_get_plan_info(self._h, a))
        return dict(zip(("scan", "pregen", "packed", "scan_ctas", "ksplit", "jsplit", "strict", "n_gpus"), list(a)))

    def launch_count(self):
        return int(self._lib.mppi_launch_count(self._h))

    # -- device-resident loop (main.jl:59-62 without the host round trip) ---
    def step_async(self, adopt=False):
        self._check(self._lib.mppi_step_async(self._h, 1 if adopt else 0))

    def sync(self):
        self._check(self._lib.mppi_sync(self._h))

    def run_closed_loop(self, max_iters=None, plant_steps=None, shift_nominal=0, plant_dt=None, update_time=None,
                        goal_radius=None):
        """The driver loop of OptimalControl/MPPI/main.jl:58-83 on the device: MPPIPlan, then the 1 ms
        explicit-Euler plant under the zero-order-held new controls, until the goal radius or max_iters.
        Returns (states_his, iter_log): states_his rows = [t, states...] = the rows of MPPITrajectory.csv
        (main.jl:87), iter_log rows = [nominal cost, feasible, RolloutCount, rho] per MPC iteration."""
        o = capi.ClosedLoop()
        self._check(self._lib.mppi_closed_loop_default(C.byref(o)))
        for k, v in (("max_iters", max_iters), ("plant_steps", plant_steps), ("plant_dt", plant_dt),
                     ("update_time", update_time), ("goal_radius", goal_radius)):
            if v is not None:
                setattr(o, k, v)
        o.shift_nominal = int(shift_nominal)
        nx = self._cfg.n_x
        his = np.zeros((1 + o.max_iters * o.plant_steps, 1 + nx), dtype=np.float64)   # C rows == col-major columns
        log = np.zeros((o.max_iters, 4), dtype=np.float64)
        n_cols, n_iters = C.c_int32(0), C.c_int32(0)
        self._check(self._lib.mppi_run_closed_loop(self._h, C.byref(o), _dptr(his), C.byref(n_cols), _dptr(log),
                                                   C.byref(n_iters)))
        self.s.X0 = his[n_cols.value - 1, 1:].copy()
        return his[:n_cols.value].copy(), log[:n_iters.value].copy()

    def adopt_nominal(self):
        self._check(self._lib.mppi_adopt_nominal(self._h))

    def fetch_result(self):
        N, nx = self._cfg.N, self._cfg.n_x
        ctrl = np.empty((N, 2), dtype=np.float64, order="F")
        traj = np.empty((N + 1, nx), dtype=np.float64, order="F")
        res = capi.Result()
        res.Control, res.Traj = ctrl.ctypes.data, traj.ctypes.data
        self._check(self._lib.mppi_fetch_result(self._h, C.byref(res)))
        _store(self, res, ctrl, traj)

    def set_profiling(self, on):
        self._check(self._lib.mppi_set_profiling(self._h, int(on)))

    def measure_fp32_peak(self):
        v = C.c_double(0)
        self._check(self._lib.mppi_measure_fp32_peak(self._h, C.byref(v)))
        return v.value

    def time_reference_calls(self, x0, U, n, flush_ptr=None, flush_bytes=0):
        """n iterations of mppi_set_x0 + mppi_set_nominal + mppi_step from host buffers, timed inside the library with nothing else in
        the timed span (what a compiled caller pays); returns (result_us, idle_us) arrays."""
        xa = np.ascontiguousarray(np.asarray(x0, dtype=np.float64))
        Ua = np.asfortranarray(np.asarray(U, dtype=np.float64))
        res, idle = np.zeros(n), np.zeros(n)
        self._check(self._lib.mppi_time_reference_calls(self._h, _dptr(xa), _dptr(Ua), int(n), C.c_void_p(flush_ptr) if flush_ptr else None,
                                                        C.c_size_t(int(flush_bytes)), _dptr(res), _dptr(idle)))
        return res, idle

    def check_poses(self, poses):
        """hit count per pose (x, y, psi) against the handle's circles and blocks
        (block_collision_check, HybridAstar/src/hybrid_astar_utils.jl:193-201)."""
        p = np.ascontiguousarray(np.asarray(poses, dtype=np.float64).reshape(-1, 3))
        hit = np.zeros(p.shape[0], dtype=np.int32)
        self._check(self._lib.mppi_check_poses(self._h, _dptr(p), p.shape[0], hit.ctypes.data_as(C.POINTER(C.c_int32))))
        return hit

    # -- sample sharding across GPUs -----------------------------------------
    def step_local(self, noise=None):
        """Rollout + local softmin partial on this handle's shard; returns capi.Partial."""
        part = capi.Partial()
        nptr = None
        if noise is not None:
            want = np.float64 if self._cfg.precision == capi.FP64 else np.float32
            noise = np.ascontiguousarray(noise, dtype=want)
            nptr = noise.ctypes.data_as(C.c_void_p)
        self._check(self._lib.mppi_step_local(self._h, nptr, C.byref(part)))
        return part

    def step_combine(self, partials):
        """Fold all ranks' partials (rank order) and run the nominal rollout; fills self.r."""
        arr = (capi.Partial * len(partials))(*partials)
        N, nx = self._cfg.N, self._cfg.n_x
        ctrl = np.empty((N, 2), dtype=np.float64, order="F")
        traj = np.empty((N + 1, nx), dtype=np.float64, order="F")
        res = capi.Result()
        res.Control, res.Traj = ctrl.ctypes.data, traj.ctypes.data
        self._check(self._lib.mppi_step_combine(self._h, arr, len(partials), C.byref(res)))
        _store(self, res, ctrl, traj)

    def comm_init(self, unique_id, rank, n_ranks):
        """In-library NCCL: the softmin exchange becomes an all-gather inside the step graph."""
        buf = (C.c_uint8 * capi.MPPI_COMM_ID_BYTES).from_buffer_copy(bytes(unique_id))
        self._check(self._lib.mppi_comm_init(self._h, buf, rank, n_ranks))

    def step_batched_abi(self, x0s, Us, Us_out, costs, feas):
        """Config 5 without host-side reshuffling: all arrays already in the C ABI's layout
        (== the memory of Julia's Array{Float64}): x0s (B, n_x), Us / Us_out (B, m, N) C-ordered
        float64, costs (B,) float64, feas (B,) int32; outputs are written in place."""
        self._check(self._lib.mppi_step_batched(self._h, _dptr(x0s) if x0s is not None else None,
                                                _dptr(Us) if Us is not None else None, _dptr(Us_out), _dptr(costs),
                                                feas.ctypes.data_as(C.POINTER(C.c_int32))))

    def peer_export(self):
        """CUDA-IPC handle (bytes) of this rank's exchange window."""
        buf = (C.c_uint8 * capi.MPPI_PEER_HANDLE_BYTES)()
        self._check(self._lib.mppi_peer_export(self._h, buf))
        return bytes(buf)

    def peer_init(self, handles, rank, n_ranks):
        """Map all ranks' exchange windows: the softmin all-gather then runs inside the finalize
        kernel over NVLink peer memory."""
        blob = b"".join(handles)
        buf = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
        self._check(self._lib.mppi_peer_init(self._h, buf, rank, n_ranks))

    def step_batched(self, x0s=None, Us=None):
        """Config 5: B independent instances in one call. x0s (n_x, B), Us (N, m, B)."""
        B, N = self._cfg.B, self._cfg.N
        xp = up = None
        if x0s is not None:
            xa = np.ascontiguousarray(np.asarray(x0s, dtype=np.float64).T)
            xp = _dptr(xa)
        if Us is not None:
            ua = np.ascontiguousarray(np.transpose(np.asarray(Us, dtype=np.float64), (2, 1, 0)))
            up = _dptr(ua)
        out = np.empty((B, 2, N))
        costs = np.empty(B)
        feas = np.empty(B, dtype=np.int32)
        self._check(self._lib.mppi_step_batched(self._h, xp, up, _dptr(out), _dptr(costs),
                                                feas.ctypes.data_as(C.POINTER(C.c_int32))))
        return np.transpose(out, (2, 1, 0)), costs, feas


def comm_unique_id():
    """NCCL unique id (rank 0 creates it; ship it to the other ranks with any transport)."""
    buf = (C.c_uint8 * capi.MPPI_COMM_ID_BYTES)()
    lib = capi.load()
    if lib.mppi_comm_unique_id(buf) != capi.MPPI_OK:
        raise MPPIError(lib.mppi_last_error(None).decode())
    return bytes(buf)


def _fill(arr, vals):
    for i, v in enumerate(vals):
        arr[i] = float(v)


def defineMPPI(numStates=0, numControls=0, lambda_=0.25, X0=None, goal=None, SamplingNumber=500, Sigma=None,
               N=15, T=3.0, XL=None, XU=None, CL=None, CU=None, *, model=None, precision=capi.FP32,
               noise_mode=capi.NOISE_PHILOX, seed=0x5EED0001, device=0, FeasibilityCount=1300,
               use_cuda_graph=True, k_offset=0, K_local=0, B=1, planner=capi.PLANNER_MPPI, strict_math=0,
               noise_policy=0, rollout_variant=0, n_gpus=0, device_ids=None, **overrides):
    """defineMPPI (setup.jl:3-59).  Positional arguments as in the reference; keyword-only
    arguments select what the reference fixes by `include` or does not have (model,
    precision, noise source, device, sample shard)."""
    X0 = np.full(numStates, np.nan) if X0 is None else np.asarray(X0, dtype=np.float64)
    goal = np.full(2, np.nan) if goal is None else np.asarray(goal, dtype=np.float64)
    Sigma = np.eye(max(numControls, 0)) if Sigma is None else np.asarray(Sigma, dtype=np.float64)
    XL = np.full(numStates, np.nan) if XL is None else np.asarray(XL, dtype=np.float64)
    XU = np.full(numStates, np.nan) if XU is None else np.asarray(XU, dtype=np.float64)
    CL = np.full(numControls, np.nan) if CL is None else np.asarray(CL, dtype=np.float64)
    CU = np.full(numControls, np.nan) if CU is None else np.asarray(CU, dtype=np.float64)
    # setup.jl:19-39, same order and wording
    if numControls <= 0:
        raise MPPIError(f"Controls ({numControls}) must be > 0")
    if numStates <= 0:
        raise MPPIError(f"States ({numStates}) must be > 0")
    x0_len = X0.shape[0] if X0.ndim == 1 else X0.shape[0]
    if x0_len != numStates:
        raise MPPIError(f"Length of X0 ({x0_len}) must match number of states ({numStates})")
    if len(XL) != numStates:
        raise MPPIError(f"Length of XL ({len(XL)}) must match number of states ({numStates})")
    if len(XU) != numStates:
        raise MPPIError(f"Length of XU ({len(XU)}) must match number of states ({numStates})")
    if len(CL) != numControls:
        raise MPPIError(f"Length of CL ({len(CL)}) must match number of controls ({numControls})")
    if len(CU) != numControls:
        raise MPPIError(f"Length of CU ({len(CU)}) must match number of controls ({numControls})")

    cfg = capi.default_config()
    if model is None:
        model = capi.MODEL_DYNAMIC7 if numStates == 7 else capi.MODEL_KINEMATIC5
    cfg.model, cfg.n_x, cfg.m = model, numStates, numControls
    cfg.K, cfg.N, cfg.B = int(SamplingNumber), int(N), int(B)
    cfg.precision, cfg.noise_mode, cfg.planner = precision, noise_mode, planner
    cfg.feasibility_count = int(FeasibilityCount)
    cfg.use_cuda_graph = 1 if use_cuda_graph else 0
    cfg.device, cfg.k_offset, cfg.K_local = device, k_offset, K_local
    cfg.seed = seed
    cfg.strict_math, cfg.noise_policy, cfg.rollout_variant = int(strict_math), int(noise_policy), int(rollout_variant)
    cfg.n_gpus = int(n_gpus)
    for i, d in enumerate(device_ids if device_ids is not None else range(max(int(n_gpus), 0))):
        cfg.device_ids[i] = int(d)
    cfg.T, cfg.lambda_ = float(T), float(lambda_)
    _fill(cfg.Sigma, np.asarray(Sigma, dtype=np.float64).reshape(-1, order="F"))
    x0_first = X0 if X0.ndim == 1 else X0[:, 0]
    _fill(cfg.X0, x0_first)
    _fill(cfg.goal, goal)
    _fill(cfg.XL, XL)
    _fill(cfg.XU, XU)
    _fill(cfg.CL, CL)
    _fill(cfg.CU, CU)
    for k, v in overrides.items():
        cur = getattr(cfg, k)
        if hasattr(cur, "__len__"):
            _fill(cur, v)
        else:
            setattr(cfg, k, v)

    mppi = MPPISearcher(cfg)
    s = mppi.s
    s.goal, s.numStates, s.numControls, s.lambda_ = goal, numStates, numControls, lambda_
    s.X0, s.XL, s.XU, s.CL, s.CU = X0, XL, XU, CL, CU
    s.T, s.N, s.dt = T, N, T / N  # setup.jl:53
    s.SamplingNumber = SamplingNumber
    s.FeasibilityCount = FeasibilityCount
    shape = (N, numControls) if B == 1 else (N, numControls, B)
    s.NominalControl = np.zeros(shape)  # setup.jl:55
    s.Σ = s.Sigma = Sigma
    if X0.ndim == 2:
        ShiftInitialCondition(mppi, X0)
    return mppi


def defineMPPIobs(mppi, obstacle_list):
    """defineMPPIobs! (setup.jl:61-64): list of [x, y, r] circles."""
    mppi.s.obstacle_list = obstacle_list
    a = np.ascontiguousarray(np.asarray(obstacle_list, dtype=np.float64).reshape(-1, 3))
    mppi._check(mppi._lib.mppi_set_circles(mppi._h, _dptr(a), a.shape[0]))


def defineMPPIblocks(mppi, block_list):
    """Polygon walls, block = [cx, cy, psi, half_len, half_wid] (CollisionDetection utils.jl:14-25)."""
    mppi.s.block_list = block_list
    a = np.ascontiguousarray(np.asarray(block_list, dtype=np.float64).reshape(-1, 5))
    mppi._check(mppi._lib.mppi_set_blocks(mppi._h, _dptr(a), a.shape[0]))


def defineMPPIrefpath(mppi, pts, weight=10.0, window=16):
    """Reference-path tracking term: `pts` is (n, 2) [x, y] (None / empty removes it);
    every stage cost gains weight * squared distance to the nearest point of a forward-moving window."""
    a = np.zeros((0, 2)) if pts is None else np.ascontiguousarray(np.asarray(pts, dtype=np.float64).reshape(-1, 2))
    mppi.s.ref_path = a
    mppi._check(mppi._lib.mppi_set_reference_path(mppi._h, _dptr(a), a.shape[0], float(weight), int(window)))


def defineMPPINominalControl(mppi, *args):
    """defineMPPINominalControl! (setup.jl:70-80): (U) or (U, Sigma)."""
    U = np.asarray(args[0], dtype=np.float64)
    mppi.s.NominalControl = U
    if len(args) == 1 and U.shape == mppi._u_buf.shape:  # the per-step call (main.jl:60)
        mppi._u_buf[...] = U
        if mppi._f_set_nominal(mppi._hv, mppi._u_ptr, U.shape[0], 2):
            mppi._check(1)
        return
    Uf = np.asfortranarray(U)
    if U.ndim == 3:  # (N, m, B): instance-major blocks of col-major N x m
        Uf = np.ascontiguousarray(np.transpose(U, (2, 1, 0)))
    mppi._check(mppi._lib.mppi_set_nominal(mppi._h, _dptr(Uf), U.shape[0], U.shape[1]))
    if len(args) > 1:
        Sigma = np.asarray(args[1], dtype=np.float64)
        mppi.s.Σ = mppi.s.Sigma = Sigma
        Sf = np.asfortranarray(Sigma)
        mppi._check(mppi._lib.mppi_set_sigma_lambda(mppi._h, _dptr(Sf), float(mppi.s.lambda_)))


def ShiftInitialCondition(mppi, X0):
    """ShiftInitialCondition (MPPIUtils.jl:24-27)."""
    X0 = np.asarray(X0, dtype=np.float64)
    mppi.s.X0 = X0
    if X0.shape == mppi._x0_buf.shape:  # the per-step call (main.jl:59)
        mppi._x0_buf[...] = X0
        if mppi._f_set_x0(mppi._hv, mppi._x0_ptr, X0.shape[0]):
            mppi._check(1)
        return
    a = np.ascontiguousarray(X0 if X0.ndim == 1 else X0.T)  # (B, n_x) rows == n_x x B col-major
    mppi._check(mppi._lib.mppi_set_x0(mppi._h, _dptr(a), X0.shape[0]))


def MPPIPlan(mppi, noise=None):
    """MPPIPlan (MPPIUtils.jl:169-203).  `noise` (N, K, m) only in SUPPLIED-noise mode."""
    if noise is None:   # the per-step call (main.jl:61): its own result struct, whose two array addresses are set once
        res = mppi._res_step
        if mppi._f_step(mppi._hv, None, mppi._res_step_ptr):
            mppi._check(1)
        r = mppi.r   # the staging buffers stay untouched until the next per-step plan, which re-arms this
        r._Control = r._Traj = None
        r._pending = mppi._pending
        r.Feasibility = "Feasible" if res.feasible else "InFeasible"  # MPPIUtils.jl:195
        r.cost, r.RolloutCount, r.FeasibleTrajCount, r.time = res.cost, res.rollout_count, res.feasible_count, res.time_s  # :196-200
        r.argmin, r.rho, r.eta = res.argmin_index, res.rho, res.eta
        return None
    cfg = mppi._cfg
    N, nx = cfg.N, cfg.n_x
    res = mppi._res
    ctrl = np.empty((N, 2), dtype=np.float64, order="F")
    traj = np.empty((N + 1, nx), dtype=np.float64, order="F")
    res.Control, res.Traj = ctrl.ctypes.data, traj.ctypes.data
    nptr = None
    if noise is not None:
        want = np.float64 if cfg.precision == capi.FP64 else np.float32
        noise = np.ascontiguousarray(noise, dtype=want)
        if noise.shape != (N, mppi.K_local, 2):
            raise MPPIError(f"noise must have shape (N, K, m) = ({N}, {mppi.K_local}, 2), got {noise.shape}")
        nptr = noise.ctypes.data_as(C.c_void_p)
    mppi._check(mppi._lib.mppi_step(mppi._h, nptr, C.byref(res)))
    _store(mppi, res, ctrl, traj)
    return None


def _store(mppi, res, ctrl, traj):
    r = mppi.r
    r.Traj, r.Control = traj, ctrl  # MPPIUtils.jl:193-194
    r.Feasibility = "Feasible" if res.feasible else "InFeasible"  # :195
    r.cost = res.cost  # :196
    r.RolloutCount = res.rollout_count  # :197
    r.FeasibleTrajCount = res.feasible_count  # :198
    r.time = res.time_s  # :200
    r.argmin, r.rho, r.eta = res.argmin_index, res.rho, res.eta


def write_trajectory_csv(path, states_his):
    """CSV.write("MPPITrajectory.csv", Tables.table(states_his'), writeheader=false) (main.jl:87):
    one row per plant step, [t, states...], no header, shortest round-trip float formatting."""
    with open(path, "w") as f:
        for row in np.asarray(states_his, dtype=np.float64):
            f.write(",".join(repr(float(v)) for v in row) + "\n")
