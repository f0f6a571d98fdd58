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
    # incremental host mirror: the host arrays the device state is known to equal (the ones the previous call
    # returned; holding them here also keeps the garbage collector from recycling their addresses) and an O(n)
    # digest of their contents taken at that moment
    last_x::Vector{Float64}
    last_cov::Matrix{Float64}
    digest::UInt
    mirrored::Bool
end

function CudaHandle(n_max::Integer, device::Integer=0)
    out = Ref{Ptr{Void}}(C_NULL)
    slam_check(ccall((:slam_create, libslamcuda), Cint, (Cint, Cint, Ref{Ptr{Void}}),
                     device, n_max, out))
    h = CudaHandle(out[], n_max, Float64[], Array{Float64}(0, 0), UInt(0), false)
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
    h.mirrored = false
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

# Cheap digest of a host state: all of x, the 3 x 3 vehicle block of cov and up to 512 evenly spaced diagonal entries
# (what a caller that re-seeds a state in place changes); an arbitrary in-place edit of cov needs forget_mirror!().
function state_digest(x, cov)
    n = size(cov, 1)
    step = max(1, div(n, 512))
    d = Float64[cov[i, i] for i in 1:step:n]
    q = min(3, n)
    return hash(x, hash(d, hash(cov[1:q, 1:q])))
end

"forget every device shadow: the next host-mirror call uploads its state in full"
function forget_mirror!()
    for h in values(SCRATCH)
        h.mirrored = false
    end
end

# Host-mirror mode with an incremental mirror: when the arrays handed in ARE the ones the previous call returned
# (`state.x, state.cov = f(state, ...)` in sim! re-binds exactly those, sim/ekfslam-sim.jl:100,117,120) and their
# digest is unchanged, nothing is uploaded.
function with_state(state::EKFSlamState, extra::Integer=0)
    h = scratch_handle(length(state.x) + extra)
    if h.mirrored && h.last_x === state.x && h.last_cov === state.cov &&
       slam_length(h) == length(state.x) && h.digest == state_digest(state.x, state.cov)
        return h
    end
    h.mirrored = false
    slam_upload(h, convert(Vector{Float64}, state.x), convert(Matrix{Float64}, state.cov))
    return h
end

function adopt!(h::CudaHandle, x::Vector{Float64}, cov::Matrix{Float64})
    h.last_x = x
    h.last_cov = cov
    h.digest = state_digest(x, cov)
    h.mirrored = true
    return nothing
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
    n = slam_length(h)
    if inplace && n == length(state.x) && isa(state.x, Vector{Float64}) && isa(state.cov, Matrix{Float64})
        # predict changes x[1:3] and rows / columns 1:3 of cov only (src/ekf.jl:32-41): O(n) doubles come back,
        # written straight into the caller's arrays (column-major, leading dimension n)
        p = Array{Float64}(3)
        slam_check(ccall((:slam_get_pose, libslamcuda), Cint, (Ptr{Void}, Ptr{Float64}), h.ptr, p))
        state.x[1:3] = p
        slam_check(ccall((:slam_get_cov_block, libslamcuda), Cint,
                         (Ptr{Void}, Cint, Cint, Cint, Cint, Ptr{Float64}, Cint), h.ptr, 0, 3, 0, n, state.cov, n))
        slam_check(ccall((:slam_get_cov_block, libslamcuda), Cint,
                         (Ptr{Void}, Cint, Cint, Cint, Cint, Ptr{Float64}, Cint), h.ptr, 0, n, 0, 3, state.cov, n))
        adopt!(h, state.x, state.cov)
        return state.x, state.cov
    end
    x, P = slam_download(h)
    if inplace && length(x) == length(state.x)
        state.x[:] = x
        state.cov[:, :] = P
        return state.x, state.cov
    end
    adopt!(h, x, P)      # update / add_features return fresh arrays (src/ekf.jl:76,121): the next call gets these
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
    if m == 0 && isa(state, EKFSlamState)
        return state.x, state.cov          # no associated observation (first observation step of every run): no-op
    end
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
    if k == 0 && isa(state, EKFSlamState)
        return state.x, state.cov          # the reference's loop body never runs (src/ekf.jl:90)
    end
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

# The browser monitor's payloads for a resident state (sim/browser/wsserver.jl:31-85) without pulling P:
# feature_ellipses -> 5 x nf [cx; cy; sqrt(l1); sqrt(l2); phi] (wsserver.jl:72-85; phi in (-pi/2, pi/2]),
# laser_lines -> 4 x nz [vx; vy; fx; fy] (src/common.jl:265-283).
function feature_ellipses(s::CudaEKFSlamState)
    nf = div(length(s) - 3, 2)
    out = Array{Float64}(5, nf)
    if nf > 0
        slam_check(ccall((:slam_feature_ellipses, libslamcuda), Cint, (Ptr{Void}, Ptr{Float64}), s.h.ptr, out))
    end
    return out
end

function laser_lines(z, s::CudaEKFSlamState)
    nz = size(z, 2)
    lines = Array{Float64}(4, nz)
    if nz > 0
        slam_check(ccall((:slam_laser_lines, libslamcuda), Cint, (Ptr{Void}, Ptr{Float64}, Cint, Ptr{Float64}),
                         s.h.ptr, convert(Matrix{Float64}, z), nz, lines))
    end
    return lines
end

function vehicle_ellipse(s::CudaEKFSlamState, nsigma::Real=2, nsegs::Integer=12)
    ell = Array{Float64}(2, nsegs + 1)
    slam_check(ccall((:slam_vehicle_ellipse, libslamcuda), Cint,
                     (Ptr{Void}, Cdouble, Cint, Ptr{Float64}), s.h.ptr, nsigma, nsegs, ell))
    return ell
end
