## ekf-cuda.jl: the filter half of module SLAM on libslamcuda.so (B200, sm_100a) ##
#
# Drop-in for src/ekf.jl and src/data-association.jl of andrewadare/SLAM.jl: same function
# names, argument order and return conventions (see INTEGRATION.md), every operation runs in
# the CUDA library through ccall.  Julia 0.5 / 0.6 syntax, like the rest of the package.
# There is no CPU fallback: a missing library or device raises an error.
#
# Two state types are accepted:
#   EKFSlamState      (src/common.jl:25-28)  host-mirror mode: x / cov are Julia arrays; each
#                     call uploads them, runs the kernels and downloads the result.  This is
#                     what sim! builds (sim/ekfslam-sim.jl:42), so the simulator runs unchanged.
#   CudaEKFSlamState  resident mode: the state lives in HBM behind a handle; x / cov are
#                     fetched on demand (large maps).
#
# NOT EXECUTED in the build container (no Julia there); kept in sync with include/slamcuda.h
# by tests/test_abi.py::test_julia_wrapper_matches_header, which checks every ccall's symbol
# and argument count against the header.

const libslamcuda = get(ENV, "LIBSLAMCUDA", "libslamcuda")

const SLAM_ERR_NAMES = Dict(-1 => "innovation covariance not positive definite",
                            -2 => "landmark id out of range",
                            -3 => "capacity exceeded",
                            -4 => "CUDA failure or no usable device",
                            -5 => "NCCL failure",
                            -6 => "bad argument")

function slam_check(rc::Integer)
    if rc < 0
        msg = unsafe_string(ccall((:slam_last_error, libslamcuda), Cstring, ()))
        error("libslamcuda: ", get(SLAM_ERR_NAMES, Int(rc), "error $rc"), ": ", msg)
    end
    return rc
end

# ---------------------------------------------------------------------------------------
# handles
# ---------------------------------------------------------------------------------------
type CudaHandle
    ptr::Ptr{Void}
    n_max::Int
end

function CudaHandle(n_max::Integer, device::Integer=0)
    out = Ref{Ptr{Void}}(C_NULL)
    slam_check(ccall((:slam_create, libslamcuda), Cint, (Cint, Cint, Ref{Ptr{Void}}),
                     device, n_max, out))
    h = CudaHandle(out[], n_max)
    finalizer(h, x -> ccall((:slam_destroy, libslamcuda), Cint, (Ptr{Void},), x.ptr))
    return h
end

# Resident-mode state: same field names as EKFSlamState (src/common.jl:25-28) so that
# `state.x, state.cov = f(state, ...)` in sim! works unchanged.  x is a host mirror refreshed
# after every call (n doubles; sim! reads state.x[1:3] each step, sim/ekfslam-sim.jl:128);
# cov is a thin AbstractMatrix over the device handle that fetches blocks on getindex.
type DeviceCov <: AbstractMatrix{Float64}
    h::CudaHandle
end

type CudaEKFSlamState <: SlamState
    x::Vector{Float64}
    cov::DeviceCov
    h::CudaHandle
end

function CudaEKFSlamState(x::Vector{Float64}, cov::Matrix{Float64}, n_max::Integer, device::Integer=0)
    h = CudaHandle(n_max, device)
    slam_upload(h, x, cov)
    return CudaEKFSlamState(copy(x), DeviceCov(h), h)
end

slam_length(h::CudaHandle) = Int(slam_check(ccall((:slam_n, libslamcuda), Cint, (Ptr{Void},), h.ptr)))
Base.length(s::CudaEKFSlamState) = slam_length(s.h)
Base.size(c::DeviceCov) = (slam_length(c.h), slam_length(c.h))
Base.getindex(c::DeviceCov, i::Int, j::Int) = cov_block(c.h, i:i, j:j)[1, 1]
Base.getindex(c::DeviceCov, rows::UnitRange{Int}, cols::UnitRange{Int}) = cov_block(c.h, rows, cols)

function slam_upload(h::CudaHandle, x::Vector{Float64}, cov::Matrix{Float64})
    n = length(x)
    slam_check(ccall((:slam_set_state, libslamcuda), Cint,
                     (Ptr{Void}, Ptr{Float64}, Ptr{Float64}, Cint, Cint), h.ptr, x, cov, n, n))
end

function slam_download(h::CudaHandle)
    n = Int(slam_check(ccall((:slam_n, libslamcuda), Cint, (Ptr{Void},), h.ptr)))
    x = Array{Float64}(n)
    P = Array{Float64}(n, n)
    slam_check(ccall((:slam_get_state, libslamcuda), Cint,
                     (Ptr{Void}, Ptr{Float64}, Ptr{Float64}, Cint), h.ptr, x, P, n))
    return x, P
end

"state.x[1:3] without pulling the rest (sim/ekfslam-sim.jl:128)"
function pose(s::CudaEKFSlamState)
    p = Array{Float64}(3)
    slam_check(ccall((:slam_get_pose, libslamcuda), Cint, (Ptr{Void}, Ptr{Float64}), s.h.ptr, p))
    return p
end

"x alone (P stays on the device): slam_get_state with a NULL matrix"
function slam_download_x(h::CudaHandle)
    x = Array{Float64}(slam_length(h))
    slam_check(ccall((:slam_get_state, libslamcuda), Cint,
                     (Ptr{Void}, Ptr{Float64}, Ptr{Float64}, Cint), h.ptr, x, C_NULL, 0))
    return x
end

"cov[i, j] blocks of a resident state (1-based ranges), for front ends"
function cov_block(h::CudaHandle, rows::UnitRange{Int}, cols::UnitRange{Int})
    out = Array{Float64}(length(rows), length(cols))
    slam_check(ccall((:slam_get_cov_block, libslamcuda), Cint,
                     (Ptr{Void}, Cint, Cint, Cint, Cint, Ptr{Float64}, Cint), h.ptr,
                     first(rows) - 1, length(rows), first(cols) - 1, length(cols), out,
                     max(length(rows), 1)))
    return out
end

# one scratch handle per capacity for host-mirror calls
const SCRATCH = Dict{Int,CudaHandle}()

function scratch_handle(n_needed::Integer)
    cap = 128
    while cap < n_needed
        cap *= 2
    end
    if !haskey(SCRATCH, cap)
        SCRATCH[cap] = CudaHandle(cap)
    end
    return SCRATCH[cap]
end

function with_state(state::EKFSlamState, extra::Integer=0)
    h = scratch_handle(length(state.x) + extra)
    slam_upload(h, convert(Vector{Float64}, state.x), convert(Matrix{Float64}, state.cov))
    return h
end
with_state(state::CudaEKFSlamState, extra::Integer=0) = state.h

colmajor4(M::AbstractMatrix) = convert(Vector{Float64}, vec(M))     # 2 x 2, already column-major

# ---------------------------------------------------------------------------------------
# predict -- replaces src/ekf.jl:8-43.  Mutates the state in place and returns (x, P).
# ---------------------------------------------------------------------------------------
function predict(state::SlamState, vehicle::Vehicle, Q::AbstractMatrix, dt::AbstractFloat)
    h = with_state(state)
    slam_check(ccall((:slam_predict, libslamcuda), Cint,
                     (Ptr{Void}, Cdouble, Cdouble, Cdouble, Ptr{Float64}, Cdouble), h.ptr,
                     vehicle.measured_gamma, vehicle.measured_speed, vehicle.wheelbase,
                     colmajor4(Q), dt))
    return finish!(state, h, true)
end

# host-mirror: download; predict writes through the caller's arrays (the reference aliases
# state.x / state.cov, src/ekf.jl:10-11), update / add_features return fresh arrays.
function finish!(state::EKFSlamState, h::CudaHandle, inplace::Bool)
    x, P = slam_download(h)
    if inplace && length(x) == length(state.x)
        state.x[:] = x
        state.cov[:, :] = P
        return state.x, state.cov
    end
    return x, P
end
# resident: P does not move; x (n doubles) is refreshed and `state.x, state.cov = f(...)` in
# sim! re-binds the same DeviceCov
function finish!(state::CudaEKFSlamState, h::CudaHandle, inplace::Bool)
    state.x = slam_download_x(h)
    return state.x, state.cov
end

# ---------------------------------------------------------------------------------------
# associate -- replaces src/data-association.jl:1-51 -> (zf 2 x m, idf 1 x m Int, zn 2 x k)
# ---------------------------------------------------------------------------------------
function associate(state::SlamState, z, R, gate1, gate2)
    nz = size(z, 2)
    zf = Array{Float64}(2, max(nz, 1))
    zn = Array{Float64}(2, max(nz, 1))
    idf = Array{Int64}(max(nz, 1))
    m = Ref{Cint}(0)
    k = Ref{Cint}(0)
    h = with_state(state)
    slam_check(ccall((:slam_associate, libslamcuda), Cint,
                     (Ptr{Void}, Ptr{Float64}, Cint, Ptr{Float64}, Cdouble, Cdouble,
                      Ptr{Float64}, Ptr{Int64}, Ref{Cint}, Ptr{Float64}, Ref{Cint}),
                     h.ptr, convert(Matrix{Float64}, z), nz, colmajor4(R), gate1, gate2,
                     zf, idf, m, zn, k))
    return zf[:, 1:m[]], reshape(convert(Vector{Int}, idf[1:m[]]), 1, Int(m[])), zn[:, 1:k[]]
end

# The spatial pre-gate the reference leaves as a TODO (src/data-association.jl:18-20) is on by default and
# returns exactly what the exhaustive scan returns; associate_exhaustive!(state, true) switches it off.
function associate_exhaustive!(state::SlamState, exhaustive::Bool)
    h = isa(state, CudaEKFSlamState) ? state.h : scratch_handle(length(state.x))   # the flag lives in the handle
    slam_check(ccall((:slam_set_pregate, libslamcuda), Cint, (Ptr{Void}, Cint), h.ptr, exhaustive ? 0 : 1))
    return nothing
end

# compute_association -- replaces src/data-association.jl:53-63 -> (nis, nd)
function compute_association(x, P, z, R, idf)
    h = scratch_handle(length(x))
    slam_upload(h, convert(Vector{Float64}, x), convert(Matrix{Float64}, P))
    nis = Ref{Cdouble}(0)
    nd = Ref{Cdouble}(0)
    slam_check(ccall((:slam_compute_association, libslamcuda), Cint,
                     (Ptr{Void}, Ptr{Float64}, Ptr{Float64}, Int64, Ref{Cdouble}, Ref{Cdouble}),
                     h.ptr, convert(Vector{Float64}, vec(z)), colmajor4(R), idf, nis, nd))
    return nis[], nd[]
end

# ---------------------------------------------------------------------------------------
# update -- replaces src/ekf.jl:46-77 -> new (x, P).  m == 0 is a no-op.
# ---------------------------------------------------------------------------------------
function update(state::SlamState, z, R, idf)
    m = size(z, 2)
    h = with_state(state)
    if m > 0
        slam_check(ccall((:slam_update, libslamcuda), Cint,
                         (Ptr{Void}, Ptr{Float64}, Cint, Ptr{Float64}, Ptr{Int64}), h.ptr,
                         convert(Matrix{Float64}, z), m, colmajor4(R),
                         convert(Vector{Int64}, vec(idf))))
    end
    return finish!(state, h, false)
end

# ---------------------------------------------------------------------------------------
# add_features -- replaces src/ekf.jl:84-122 -> augmented (x, P)
# ---------------------------------------------------------------------------------------
function add_features(state::SlamState, z, R)
    k = size(z, 2)
    h = with_state(state, 2k)
    if k > 0
        # the reference's state grows without bound (hcat / vcat, src/ekf.jl:108-109): double the
        # device capacity when the append would not fit
        need = slam_length(h) + 2k
        if need > h.n_max
            h.n_max = max(need, 2 * h.n_max)
            slam_check(ccall((:slam_reserve, libslamcuda), Cint, (Ptr{Void}, Cint), h.ptr, h.n_max))
        end
        slam_check(ccall((:slam_add_features, libslamcuda), Cint,
                         (Ptr{Void}, Ptr{Float64}, Cint, Ptr{Float64}), h.ptr,
                         convert(Matrix{Float64}, z), k, colmajor4(R)))
    end
    return finish!(state, h, false)
end

# ---------------------------------------------------------------------------------------
# compute_landmark_ellipses for a resident state -- replaces src/common.jl:247-262 without
# pulling P to the host: 2nf x (nsegs+1), rows 2i-1:2i = ellipse i (nsigma = 2, nsegs = 12)
# ---------------------------------------------------------------------------------------
function compute_landmark_ellipses(s::CudaEKFSlamState)
    nsigma = 2
    nsegs = 12
    nf = div(length(s) - 3, 2)
    ellipses = Array{Float64}(2*nf, nsegs + 1)
    slam_check(ccall((:slam_compute_landmark_ellipses, libslamcuda), Cint,
                     (Ptr{Void}, Cdouble, Cint, Ptr{Float64}, Cint), s.h.ptr, nsigma, nsegs,
                     ellipses, max(2*nf, 1)))
    return ellipses
end

function vehicle_ellipse(s::CudaEKFSlamState, nsigma::Real=2, nsegs::Integer=12)
    ell = Array{Float64}(2, nsegs + 1)
    slam_check(ccall((:slam_vehicle_ellipse, libslamcuda), Cint,
                     (Ptr{Void}, Cdouble, Cint, Ptr{Float64}), s.h.ptr, nsigma, nsegs, ell))
    return ell
end
