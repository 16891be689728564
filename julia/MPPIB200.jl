# MPPIB200.jl — `ccall` shim over libmppi_b200.so (include/mppi_b200.h).
#
# Drop-in for the MPPI step of congkaishen/MotionPlanning: replace
#     include("src/MPPIUtils.jl")                       # OptimalControl/MPPI/main.jl:1
# by
#     include("/path/to/julia/MPPIB200.jl"); using .MPPIB200
#     include("src/vehicledynamics.jl")                 # the 1 ms plant of main.jl:76 still calls VehicleDynamics
#     include("src/plotting.jl")                        # circleShape / plotRes / plotEnv / vehicleShape = MPPIUtils.jl:205-260, moved
#                                                       # to a file of their own (they are host-side Plots.jl code, main.jl:32,68)
# and lines 29-30 and 59-68 of OptimalControl/MPPI/main.jl run unchanged:
#     mppi = defineMPPI(numStates, numControls, lambda, X0, goal, SamplingNumber, Σ, N, T, XL, XU, CL, CU)
#     defineMPPIobs!(mppi, [[50.0, 1, 2.5], [70.0, -1, 2.5], [90.0, 1, 2.5]])
#     ShiftInitialCondition(mppi, states); defineMPPINominalControl!(mppi, NominalControls); MPPIPlan(mppi)
#     mppi.r.Control, mppi.r.Feasibility, mppi.r.time, mppi.r.RolloutCount, mppi.r.Traj
#     plotRes(mppi, states_his)                         # reads mppi.s.obstacle_list, mppi.s.goal, mppi.r.Traj and
#                                                       # mppi.p.TrajectoryCollection[i].Trajectory[:, 1:2] (MPPIUtils.jl:221-223)
# Several GPUs from the one Julia task: defineMPPI(...; n_gpus = 8) (samples sharded inside the library, one handle).
#
# Julia is not installed in the build container, so this file is the syntax-reviewed twin of
# motionplanning_b200/_capi.py + mppi.py (ctypes), which IS exercised by tests/ against the same
# symbols.  The struct layouts below mirror include/mppi_b200.h field by field.
module MPPIB200

export defineMPPI, defineMPPIobs!, defineMPPIblocks!, defineMPPINominalControl!,
       ShiftInitialCondition, MPPIPlan, MPPISearcher, MPPIClosedLoop, defineMPPIrefpath!, MPPIPlanBatch,
       defineDWA, defineDWAobs!, DWAPlan

const LIB = get(ENV, "MPPI_B200_LIB", joinpath(@__DIR__, "..", "motionplanning_b200", "csrc", "libmppi_b200.so"))
const MAX_NX = 8

# mppi_config_t (include/mppi_b200.h) — isbits, same field order
struct Config
    struct_size::UInt32
    model::Int32; n_x::Int32; m::Int32; K::Int32; N::Int32; B::Int32
    precision::Int32; noise_mode::Int32; planner::Int32
    feasibility_count::Int32; use_cuda_graph::Int32; device::Int32
    k_offset::Int32; K_local::Int32; rollout_variant::Int32
    strict_math::Int32; noise_policy::Int32; n_gpus::Int32
    device_ids::NTuple{8,Int32}
    reserved0::Int32
    seed::UInt64
    T::Float64; lambda::Float64
    Sigma::NTuple{4,Float64}
    X0::NTuple{8,Float64}
    goal::NTuple{2,Float64}
    XL::NTuple{8,Float64}; XU::NTuple{8,Float64}
    CL::NTuple{2,Float64}; CU::NTuple{2,Float64}
    slack_penalty::Float64; obstacle_penalty::Float64; terminal_scale::Float64
    la::Float64; lb::Float64; mass::Float64; Izz::Float64; g::Float64; mu::Float64
    KFZF::Float64; KFZR::Float64; KFZX::Float64
    tire::NTuple{6,Float64}
    stage_w::NTuple{6,Float64}
    veh_L::Float64; veh_W::Float64; veh_center_offset::Float64
end

# mppi_result_t
mutable struct CResult
    Control::Ptr{Float64}; Traj::Ptr{Float64}
    cost::Float64
    feasible::Int32; rollout_count::Int32; feasible_count::Int32; argmin_index::Int32
    rho::Float64; eta::Float64; time_s::Float64
    CResult() = new(C_NULL, C_NULL, 0.0, 0, 0, 0, 0, 0.0, 0.0, 0.0)
end

# the (s, p, r) shape of MPPISearcher (types.jl:46-50)
mutable struct Setting                 # types.jl:10-31
    numStates::Int; numControls::Int
    X0::Vector{Float64}; XL::Vector{Float64}; XU::Vector{Float64}; CL::Vector{Float64}; CU::Vector{Float64}
    dt::Float64; T::Float64; N::Int; goal::Vector{Float64}
    obstacle_list::Vector{Vector{Float64}}
    SamplingNumber::Int; FeasibilityCount::Int
    NominalControl::Matrix{Float64}; Σ::Matrix{Float64}; lambda::Float64; SlackPenalty::Float64
end
mutable struct Result                  # types.jl:36-44
    Traj::Matrix{Float64}; Control::Matrix{Float64}; Feasibility::Symbol
    cost::Float64; time::Float64; FeasibleTrajCount::Int; RolloutCount::Int
end
# MPPIHolder (types.jl:3-8) as plotRes reads it: .Trajectory is (N+1) x 2 [x y]; .cost / .Feasibility of that sample
struct Holder
    Trajectory::Matrix{Float64}
    cost::Float64
    Feasibility::Bool
end
# MPPIPlanner (types.jl:32-34).  The K sampled trajectories never leave the GPU during MPPIPlan; reading
# `mppi.p.TrajectoryCollection` fetches every `stride`-th one (re-rolled on the device from the same noise) with its cost and
# feasibility flag.  stride = 1 gives the reference's full collection.
mutable struct Planner
    h::Ptr{Cvoid}
    N::Int; K::Int; n_gpus::Int
    stride::Int
end
function Base.getproperty(p::Planner, name::Symbol)
    name === :TrajectoryCollection ? trajectory_collection(p) : getfield(p, name)
end
# global sample ids exported by mppi_get_sample_xy: every stride-th sample of every GPU's contiguous shard, in shard order
function exported_ids(K, stride, G)
    G = max(G, 1); base = div(K, G); rem = K % G; ids = Int[]
    for g in 0:G-1
        off = g * base + min(g, rem); kl = base + (g < rem ? 1 : 0)
        append!(ids, off .+ (1:stride:kl))
    end
    return ids
end
function trajectory_collection(p::Planner)
    h = getfield(p, :h); N = getfield(p, :N); K = getfield(p, :K); stride = getfield(p, :stride)
    n = Ref{Int32}(0)
    check(h, ccall((:mppi_get_sample_xy, LIB), Int32, (Ptr{Cvoid}, Int32, Ptr{Float32}, Ref{Int32}), h, stride, C_NULL, n))
    xy = Array{Float32}(undef, 2, N + 1, n[])                                          # C [n][N+1][2]
    check(h, ccall((:mppi_get_sample_xy, LIB), Int32, (Ptr{Cvoid}, Int32, Ptr{Float32}, Ref{Int32}), h, stride, xy, n))
    costs = Vector{Float64}(undef, K); flags = Vector{UInt8}(undef, K)
    check(h, ccall((:mppi_get_costs, LIB), Int32, (Ptr{Cvoid}, Ptr{Float64}), h, costs))
    check(h, ccall((:mppi_get_flags, LIB), Int32, (Ptr{Cvoid}, Ptr{UInt8}), h, flags))
    ids = exported_ids(K, stride, getfield(p, :n_gpus))
    return [Holder(Float64.(permutedims(xy[:, :, i])), costs[ids[i]], (flags[ids[i]] & 0x01) != 0) for i in 1:n[]]
end

mutable struct MPPISearcher
    h::Ptr{Cvoid}
    s::Setting
    p::Planner
    r::Result
    cres::CResult
end

lasterr(h) = unsafe_string(ccall((:mppi_last_error, LIB), Cstring, (Ptr{Cvoid},), h))
check(h, rc) = rc == 0 ? nothing : error(lasterr(h))     # same exception behaviour as setup.jl:19-39

pad8(v) = ntuple(i -> i <= length(v) ? Float64(v[i]) : 0.0, 8)

function defineMPPI(numStates, numControls, lambda, X0, goal, SamplingNumber, Σ, N, T, XL, XU, CL, CU;
                    model = numStates == 7 ? 0 : 1, precision = 0, noise_mode = 1, seed = 0x5EED0001,
                    device = 0, FeasibilityCount = 1300, planner = 0, obstacle_penalty = nothing,
                    strict_math = 0, noise_policy = 0, rollout_variant = 0, n_gpus = 0, device_ids = 0:7, plot_stride = 16, B = 1)
    numControls > 0 || error("Controls ($numControls) must be > 0")                 # setup.jl:19-21
    numStates > 0 || error("States ($numStates) must be > 0")                       # :22-24
    length(X0) == numStates || error("Length of X0 ($(length(X0))) must match number of states ($numStates)")
    length(XL) == numStates || error("Length of XL ($(length(XL))) must match number of states ($numStates)")
    length(XU) == numStates || error("Length of XU ($(length(XU))) must match number of states ($numStates)")
    length(CL) == numControls || error("Length of CL ($(length(CL))) must match number of controls ($numControls)")
    length(CU) == numControls || error("Length of CU ($(length(CU))) must match number of controls ($numControls)")
    d = Ref{Config}()
    ccall((:mppi_config_default, LIB), Int32, (Ref{Config},), d)
    c = d[]
    cfg = Config(c.struct_size, model, numStates, numControls, SamplingNumber, N, B, precision, noise_mode, planner,
                 FeasibilityCount, 1, device, 0, 0, rollout_variant, strict_math, noise_policy, n_gpus,
                 ntuple(i -> i <= length(device_ids) ? Int32(device_ids[i]) : Int32(0), 8), 0, seed, T, lambda,
                 (Σ[1, 1], Σ[2, 1], Σ[1, 2], Σ[2, 2]), pad8(X0), (goal[1], goal[2]), pad8(XL), pad8(XU),
                 (CL[1], CL[2]), (CU[1], CU[2]), c.slack_penalty,
                 obstacle_penalty === nothing ? c.obstacle_penalty : Float64(obstacle_penalty), c.terminal_scale,
                 c.la, c.lb, c.mass, c.Izz, c.g, c.mu, c.KFZF, c.KFZR, c.KFZX, c.tire, c.stage_w,
                 c.veh_L, c.veh_W, c.veh_center_offset)
    hp = Ref{Ptr{Cvoid}}(C_NULL)
    rc = ccall((:mppi_create, LIB), Int32, (Ref{Config}, Ref{Ptr{Cvoid}}), cfg, hp)
    rc == 0 || error(lasterr(C_NULL))
    s = Setting(numStates, numControls, Float64.(X0), Float64.(XL), Float64.(XU), Float64.(CL), Float64.(CU),
                T / N, T, N, Float64.(goal), Vector{Vector{Float64}}(), SamplingNumber, FeasibilityCount,
                zeros(N, numControls), Float64.(Σ), lambda, 1e5)                      # setup.jl:53-55
    r = Result(zeros(N + 1, numStates), zeros(N, numControls), :InFeasible, 1e6, 0.0, 0, 0)
    mppi = MPPISearcher(hp[], s, Planner(hp[], N, SamplingNumber, n_gpus, plot_stride), r, CResult())
    finalizer(m -> ccall((:mppi_destroy, LIB), Int32, (Ptr{Cvoid},), m.h), mppi)
    return mppi
end

function defineMPPIobs!(mppi::MPPISearcher, obstacle_list)                           # setup.jl:61-64
    mppi.s.obstacle_list = obstacle_list
    xyr = Float64[o[i] for i in 1:3, o in obstacle_list]                              # 3 x M_c col-major
    check(mppi.h, ccall((:mppi_set_circles, LIB), Int32, (Ptr{Cvoid}, Ptr{Float64}, Int32), mppi.h, xyr, size(xyr, 2)))
end

# polygon walls [cx, cy, ψ, half_len, half_wid] (CollisionDetection/src/utils.jl:14-25)
function defineMPPIblocks!(mppi::MPPISearcher, block_list)
    b = Float64[o[i] for i in 1:5, o in block_list]
    check(mppi.h, ccall((:mppi_set_blocks, LIB), Int32, (Ptr{Cvoid}, Ptr{Float64}, Int32), mppi.h, b, size(b, 2)))
end

# reference polyline for the tracking term: pts is n x 2 [x y] (as ref_pts in main_Tracker.jl:47); empty removes it
function defineMPPIrefpath!(mppi::MPPISearcher, pts::Matrix{Float64}; weight = 10.0, window = 16)
    xy = Matrix(transpose(pts))                                                       # 2 x n col-major
    check(mppi.h, ccall((:mppi_set_reference_path, LIB), Int32, (Ptr{Cvoid}, Ptr{Float64}, Int32, Float64, Int32),
                        mppi.h, xy, size(xy, 2), weight, window))
end

function defineMPPINominalControl!(mppi::MPPISearcher, U::Matrix{Float64})            # setup.jl:70-73
    mppi.s.NominalControl = U
    check(mppi.h, ccall((:mppi_set_nominal, LIB), Int32, (Ptr{Cvoid}, Ptr{Float64}, Int32, Int32),
                        mppi.h, U, size(U, 1), size(U, 2)))                           # Matrix{Float64} is already N x m col-major
end
function defineMPPINominalControl!(mppi::MPPISearcher, U::Matrix{Float64}, Σ)          # setup.jl:74-78
    defineMPPINominalControl!(mppi, U)
    mppi.s.Σ = Float64.(Σ)
    check(mppi.h, ccall((:mppi_set_sigma_lambda, LIB), Int32, (Ptr{Cvoid}, Ptr{Float64}, Float64),
                        mppi.h, mppi.s.Σ, mppi.s.lambda))
end

function ShiftInitialCondition(mppi::MPPISearcher, X0)                                 # MPPIUtils.jl:24-27
    mppi.s.X0 = Float64.(X0)
    check(mppi.h, ccall((:mppi_set_x0, LIB), Int32, (Ptr{Cvoid}, Ptr{Float64}, Int32), mppi.h, mppi.s.X0, length(X0)))
end

function MPPIPlan(mppi::MPPISearcher; noise = nothing)                                 # MPPIUtils.jl:169-203
    r, c = mppi.r, mppi.cres
    GC.@preserve r begin
        c.Control = pointer(r.Control); c.Traj = pointer(r.Traj)                       # filled in place, zero-copy
        np = noise === nothing ? C_NULL : pointer(noise)                               # Array{T,3}(m, K, N)
        check(mppi.h, ccall((:mppi_step, LIB), Int32, (Ptr{Cvoid}, Ptr{Cvoid}, Ref{CResult}), mppi.h, np, c))
    end
    r.Feasibility = c.feasible != 0 ? :Feasible : :InFeasible                          # :195
    r.cost = c.cost; r.RolloutCount = c.rollout_count; r.FeasibleTrajCount = c.feasible_count
    r.time = c.time_s                                                                  # :199-200
    return nothing
end

# B independent plans in one call (BASELINE.json config 5): mppi = defineMPPI(...; B = B); x0s is numStates x B, Us and the
# returned Us_out are N x numControls x B (Julia column-major = the ABI's layout)
function MPPIPlanBatch(mppi::MPPISearcher, x0s::Matrix{Float64}, Us::Array{Float64,3})
    h = mppi.h; B = size(x0s, 2)
    Us_out = similar(Us); costs = zeros(B); feas = zeros(Int32, B)
    check(h, ccall((:mppi_step_batched, LIB), Int32, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Int32}),
                   h, x0s, Us, Us_out, costs, feas))
    return Us_out, costs, feas
end

# ---- the DynamicWindow sibling (OptimalControl/DynamicWindow/src/setup.jl, DWAUtils.jl) -------------------
# Replace include("src/DWAUtils.jl") (DynamicWindow/main.jl:1); lines 26-27 and 49-53 run unchanged; the plant
# step of :61 keeps using the reference's own VehicleDynamics.
function defineDWA(numStates, numControls, X0, goal, SampleNumber, T, XL, XU, CL, CU; precision = 0, device = 0)  # setup.jl:3-52
    K = prod(SampleNumber)
    dwa = defineMPPI(numStates, numControls, 1.0, X0, goal, K, [1.0 0.0; 0.0 1.0], 20, T, XL, XU, CL, CU;   # types.jl:24 N = 20
                     precision = precision, device = device, FeasibilityCount = K, planner = 1,
                     obstacle_penalty = 10000.0 * 712.5)                                                      # DWAUtils.jl:112
    # defineDWAcontrols! (setup.jl:59-85): first control is the slowest index
    g1 = collect(LinRange(CL[1], CU[1], SampleNumber[1])); g2 = collect(LinRange(CL[2], CU[2], SampleNumber[2]))
    cand = Matrix{Float64}(undef, K, 2)
    for i in 1:SampleNumber[1], j in 1:SampleNumber[2]
        cand[(i - 1) * SampleNumber[2] + j, :] = [g1[i], g2[j]]
    end
    check(dwa.h, ccall((:mppi_set_candidates, LIB), Int32, (Ptr{Cvoid}, Ptr{Float64}, Int32, Int32), dwa.h, cand, K, 2))
    return dwa
end

defineDWAobs!(dwa::MPPISearcher, obstacle_list) = defineMPPIobs!(dwa, obstacle_list)    # setup.jl:54-57

function DWAPlan(dwa::MPPISearcher)                                                    # DWAUtils.jl:141-163
    dwa.r.Control = zeros(dwa.s.N, dwa.s.numControls)    # mppi_step writes the N x m (constant) sequence
    MPPIPlan(dwa)
    dwa.r.Control = reshape(dwa.r.Control[1, :], 1, :)   # BestCand.Control is the constant control of the candidate (:156)
    return nothing
end

# mppi_closed_loop_t
struct ClosedLoop
    struct_size::UInt32
    max_iters::Int32; plant_steps::Int32; shift_nominal::Int32
    plant_dt::Float64; update_time::Float64; goal_radius::Float64
end

# The whole driver loop of main.jl:58-83 in one call: returns states_his exactly as main.jl builds it
# ((1 + numStates) x columns, column = [t; states]; CSV.write(..., Tables.table(states_his')) as :87)
# and the per-iteration log (4 x iterations: nominal cost, feasible, RolloutCount, rho).
function MPPIClosedLoop(mppi::MPPISearcher; max_iters = nothing, shift_nominal = 0)
    d = Ref{ClosedLoop}()
    ccall((:mppi_closed_loop_default, LIB), Int32, (Ref{ClosedLoop},), d)
    o = d[]
    iters = max_iters === nothing ? o.max_iters : Int32(max_iters)
    o = ClosedLoop(o.struct_size, iters, o.plant_steps, Int32(shift_nominal), o.plant_dt, o.update_time, o.goal_radius)
    his = zeros(1 + mppi.s.numStates, 1 + iters * o.plant_steps)
    log = zeros(4, iters)
    ncols = Ref{Int32}(0); niters = Ref{Int32}(0)
    check(mppi.h, ccall((:mppi_run_closed_loop, LIB), Int32,
                        (Ptr{Cvoid}, Ref{ClosedLoop}, Ptr{Float64}, Ref{Int32}, Ptr{Float64}, Ref{Int32}),
                        mppi.h, o, his, ncols, log, niters))
    mppi.s.X0 = his[2:end, ncols[]]
    return his[:, 1:ncols[]], log[:, 1:niters[]]
end

# Measurement aid (no counterpart in the reference): n iterations of the three per-iteration calls of main.jl:59-61
# (mppi_set_x0, mppi_set_nominal, mppi_step) from host buffers, each timed inside the library; returns
# (result_us, idle_us): call start -> plan in host memory, -> the handle's stream idle.
function TimeReferenceCalls(mppi::MPPISearcher, X0::Vector{Float64}, U::Matrix{Float64}, n::Integer)
    res = zeros(n); idle = zeros(n)
    check(mppi.h, ccall((:mppi_time_reference_calls, LIB), Int32,
                        (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Int32, Ptr{Cvoid}, Csize_t, Ptr{Float64}, Ptr{Float64}),
                        mppi.h, X0, U, Int32(n), C_NULL, Csize_t(0), res, idle))
    return res, idle
end

end # module
