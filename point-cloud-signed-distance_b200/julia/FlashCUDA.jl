# FlashCUDA.jl -- the `ccall` shim a maintainer of Flash would add to route the tracking hot path
# through libflashsdf (include/flashsdf.h).  Written against Julia 0.5 syntax to match the reference
# (src/Flash.jl uses `immutable`, `typealias`, `f{T}(x)`); UNTESTED here: no `julia` binary exists in
# the build image (SURVEY 0.7).  The Python package `flash/` is the tested caller of the same ABI.
#
# What changes on the Julia side: nothing in src/tracking.jl, src/models.jl or the notebooks.
# `Flash.GradientDescent.CostFunctor` gains a Float64 fast path that calls the library, and
# `estimate_state` asks for cost AND gradient in one call instead of letting ForwardDiff push
# ceil(nx/10) dual passes through `cost`.
module FlashCUDA

import Flash
import Flash: Manipulator, ManipulatorState, BodyGeometry, ConvexGeometry, RigidInterpolatingSkin,
              DeformableInterpolatingSkin, num_states
using RigidBodyDynamics
import StaticArrays: SVector

const lib = "libflashsdf"   # point-cloud-signed-distance_b200/lib must be on LD_LIBRARY_PATH

# ---- mirror of flash_model_desc (include/flashsdf.h)
immutable ModelDesc
    abi_version::Int32
    n_bodies::Int32
    body_parent::Ptr{Int32}
    joint_type::Ptr{Int32}
    joint_axis::Ptr{Float64}
    joint_to_parent::Ptr{Float64}
    n_surfaces::Int32
    surface_kind::Ptr{Int32}
    surface_body::Ptr{Int32}
    surface_frame::Ptr{Float64}
    point_offset::Ptr{Int32}
    n_surface_points::Ptr{Int32}
    points::Ptr{Float64}
    point_body::Ptr{Int32}
end

type Context
    ptr::Ptr{Void}
    keepalive::Vector{Any}      # the arrays ModelDesc points into
end

function check(ctx::Ptr{Void}, rc::Cint)
    rc == 0 && return
    msg = unsafe_string(ccall((:flash_last_error, lib), Cstring, (Ptr{Void},), ctx))
    error("libflashsdf error $rc: $msg")
end

rowmajor34(R::AbstractMatrix, t::AbstractVector) = Float64[R[1,1], R[1,2], R[1,3], t[1], R[2,1], R[2,2], R[2,3], t[2], R[3,1], R[3,2], R[3,3], t[3]]

"""
Flatten a `Flash.Manipulator` (src/Flash.jl:62-65) into a flash_model_desc and create the context.
Bodies are taken in `toposortedTree` order (skipping the root), so the configuration layout equals
`state.mechanism_state.q` and `x = [q; deformation_data]` (src/gradientdescent.jl:9-17) is passed
through unchanged.
"""
function Context(manip::Manipulator)
    mech = manip.mechanism
    verts = mech.toposortedTree[2:end]
    index = Dict(v.vertexData => Int32(i - 1) for (i, v) in enumerate(verts))
    parent = Int32[get(index, v.parent.vertexData, Int32(-1)) for v in verts]
    jtype = Int32[]; axis = Float64[]; j2p = Float64[]
    for v in verts
        joint = v.edgeToParentData
        jt = joint.jointType
        if isa(jt, Revolute)
            push!(jtype, 1); append!(axis, jt.rotation_axis)
        elseif isa(jt, QuaternionFloating)
            push!(jtype, 2); append!(axis, [0., 0, 1])
        else
            push!(jtype, 0); append!(axis, [0., 0, 1])
        end
        T = mech.jointToJointTransforms[joint]            # frameBefore -> parent body
        append!(j2p, rowmajor34(RigidBodyDynamics.rotationmatrix_normalized_fsa(T.rot), T.trans))
    end
    kind = Int32[]; sbody = Int32[]; sframe = Float64[]; offs = Int32[0]; nsurf = Int32[]
    pts = Float64[]; pbody = Int32[]
    identity12 = Float64[1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0]
    body_of(frame) = index[first(b for b in bodies(mech) if RigidBodyDynamics.is_fixed_to_body(mech, frame, b))]
    for s in manip.surfaces
        if isa(s, ConvexGeometry)
            push!(kind, 0)
            b = first(b for b in bodies(mech) if RigidBodyDynamics.is_fixed_to_body(mech, s.frame, b))
            push!(sbody, get(index, b, Int32(-1)))
            T = RigidBodyDynamics.find_fixed_transform(mech, s.frame, b.frame)   # the visual origin, src/models.jl:154-164
            append!(sframe, rowmajor34(RigidBodyDynamics.rotationmatrix_normalized_fsa(T.rot), T.trans))
            vs = Flash.GeometryTypes.vertices(s.geometry)
            for v in vs; append!(pts, [v[1], v[2], v[3]]); push!(pbody, -1); end
            push!(nsurf, length(vs))
        else
            push!(kind, isa(s, DeformableInterpolatingSkin) ? 2 : 1)
            push!(sbody, -1); append!(sframe, identity12)
            for p in vcat(s.surface_points, s.skeleton_points)      # concatenation order of src/Flash.jl:207-211
                append!(pts, p.v); push!(pbody, body_of(p.frame))
            end
            push!(nsurf, length(s.surface_points))
        end
        push!(offs, div(length(pts), 3))
    end
    keep = Any[parent, jtype, axis, j2p, kind, sbody, sframe, offs, nsurf, pts, pbody]
    desc = ModelDesc(1, length(verts), pointer(parent), pointer(jtype), pointer(axis), pointer(j2p),
                     length(kind), pointer(kind), pointer(sbody), pointer(sframe), pointer(offs),
                     pointer(nsurf), pointer(pts), pointer(pbody))
    out = Ref{Ptr{Void}}(C_NULL)
    rc = ccall((:flash_model_create, lib), Cint, (Ptr{ModelDesc}, Ptr{Void}, Ptr{Ptr{Void}}), &desc, C_NULL, out)
    check(Ptr{Void}(C_NULL), rc)
    ctx = Context(out[], keep)
    finalizer(ctx, c -> ccall((:flash_destroy, lib), Void, (Ptr{Void},), c.ptr))
    ctx
end

"`CostFunctor(manipulator, sensed_points)` (src/gradientdescent.jl:41-47): the points are uploaded once."
function set_points!(ctx::Context, sensed_points::Vector{SVector{3, Float64}})
    check(ctx.ptr, ccall((:flash_points_set, lib), Cint, (Ptr{Void}, Ptr{Float64}, Int64),
                         ctx.ptr, pointer(reinterpret(Float64, sensed_points, (3 * length(sensed_points),))), length(sensed_points)))
end

"""
`(g, x) -> value` of `CostAndGradientFunctor` (src/gradientdescent.jl:59-65): raw cost (not divided by N,
src/tracking.jl:18-20 divides) and its gradient, one kernel pass instead of ForwardDiff's chunk passes.
"""
function cost_and_gradient!(ctx::Context, g::Vector{Float64}, x::Vector{Float64}; deformation_cost_weight=10.0)
    c = Ref{Float64}(0.0)
    check(ctx.ptr, ccall((:flash_cost_grad, lib), Cint,
                         (Ptr{Void}, Ptr{Float64}, Int32, Float64, Ptr{Float64}, Ptr{Float64}, Ptr{Int32}),
                         ctx.ptr, x, 1, deformation_cost_weight, c, g, C_NULL))
    c[]
end

"Batched form: X is nx x C (column-major in Julia == [C][nx] row-major in C)."
function cost_and_gradient!(ctx::Context, G::Matrix{Float64}, X::Matrix{Float64}; deformation_cost_weight=10.0)
    costs = zeros(size(X, 2))
    check(ctx.ptr, ccall((:flash_cost_grad, lib), Cint,
                         (Ptr{Void}, Ptr{Float64}, Int32, Float64, Ptr{Float64}, Ptr{Float64}, Ptr{Int32}),
                         ctx.ptr, X, size(X, 2), deformation_cost_weight, costs, G, C_NULL))
    costs
end

"""
Drop-in for `Flash.Tracking.estimate_state` (src/tracking.jl:8-27) with the same signature.  The loop,
the callback (called with the RAW cost, as in the reference) and the solver object stay in Julia.
"""
function estimate_state(manipulator::Manipulator, sensed_points::Vector{SVector{3, Float64}}, x_estimated::Vector{Float64};
                        callback::Function=(x, c) -> (), solver=nothing, ctx::Context=Context(manipulator))
    set_points!(ctx, sensed_points)
    n = length(sensed_points)
    rate = solver === nothing ? 0.1 : solver.rate
    max_step = solver === nothing ? 0.5 : solver.max_step
    iters = solver === nothing ? 30 : solver.iteration_limit
    x = copy(x_estimated); g = similar(x)
    for i in 1:iters
        c = cost_and_gradient!(ctx, g, x)
        callback(x, c)
        step = -rate .* g ./ n
        s = norm(step)
        s > max_step && (step .*= max_step / s)
        x .+= step
    end
    x
end

"`DepthSensors.raycast_depths` (src/depthsensors.jl:88-97) on the device; NaN marks a miss."
function raycast_depths(ctx::Context, x::Vector{Float64}, sensor::Flash.DepthSensors.DepthSensor, R::Matrix{Float64}, t::Vector{Float64})
    rays = sensor.rays                      # rows x cols of SVector{3}: Julia is column-major, the C side wants
    rm = permutedims(rays, (2, 1))          # row-major ray order (row outer, col inner) like raycast_points (:102-111)
    nr = length(rm)
    depths = Array{Float64}(nr)
    nh = Ref{Int64}(0)
    check(ctx.ptr, ccall((:flash_raycast, lib), Cint,
                         (Ptr{Void}, Ptr{Float64}, Ptr{Float64}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Int64}),
                         ctx.ptr, x, pointer(reinterpret(Float64, vec(rm), (3nr,))), nr, rowmajor34(R, t), depths, C_NULL, nh))
    permutedims(reshape(depths, size(rays, 2), size(rays, 1)), (2, 1))
end

end # module
