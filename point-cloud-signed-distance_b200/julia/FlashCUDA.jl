# FlashCUDA.jl -- the `ccall` shim a maintainer of Flash would add to route the tracking hot path
# through libflashsdf (include/flashsdf.h).  Written against Julia 0.5 syntax to match the reference
# (src/Flash.jl uses `immutable`, `typealias`, `f{T}(x)`); UNTESTED here: no `julia` binary exists in
# the build image (SURVEY 0.7).  The Python package `flash/` is the tested caller of the same ABI.
#
# What changes on the Julia side: nothing in src/tracking.jl, src/models.jl or the notebooks.
# `estimate_state` here has the reference's signature and defaults; it hands the solver the same
# `(g, x) -> value` closure shape as `CostAndGradientFunctor` (src/gradientdescent.jl:59-65), with cost AND gradient
# from one library call instead of ForwardDiff's ceil(nx/10) dual passes through `cost`, and it honours every field of
# the solver object (rate, max_step, iteration_limit, gradient_convergence_tolerance, precondition_divisors).
module FlashCUDA

import Flash
import Flash: Manipulator, ManipulatorState, BodyGeometry, ConvexGeometry, RigidInterpolatingSkin,
              DeformableInterpolatingSkin, num_states
using RigidBodyDynamics
import StaticArrays: SVector
import SimpleGradientDescent
import MathProgBase

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
    desc = ModelDesc(2, length(verts), pointer(parent), pointer(jtype), pointer(axis), pointer(j2p),
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

# ---- contexts are cached per manipulator: building one uploads the model and builds the hulls' BSP trees
const contexts = ObjectIdDict()
context(manip::Manipulator) = get!(() -> Context(manip), contexts, manip)::Context

"""
The closure `src/tracking.jl:16-21` builds, with the gradient: `(g, x) -> value` in the shape of
`CostAndGradientFunctor` (src/gradientdescent.jl:59-65).  As `wrapped_cost` does, it calls `callback(x, c)` with the
RAW cost and returns `c / length(sensed_points)`; `g` receives the gradient of that returned value.
"""
function wrapped_cost_and_gradient(ctx::Context, n_points::Int, callback::Function)
    (g, x) -> begin
        c = cost_and_gradient!(ctx, g, x)
        callback(x, c)
        scale!(g, 1 / n_points)
        c / n_points
    end
end

# mirror of flash_solver_params (include/flashsdf.h)
immutable SolverParams
    rate::Float64
    max_step::Float64
    iteration_limit::Int32
    gradient_convergence_tolerance::Float64
    deformation_cost_weight::Float64
end

field_or(solver, name::Symbol, default) = name in fieldnames(solver) ? getfield(solver, name) : default

"""
`optimize!` of a `SimpleGradientDescent.NaiveSolver` model with an analytic gradient.  SimpleGradientDescent
(@0fcc1f9, REQUIRE.dev:24) is not in the reference tree; the step is the one its keyword names describe
(src/tracking.jl:12-15, examples/irb140.ipynb cell 9, examples/squishable.ipynb cell 9) and the one
flash_estimate_state runs on the device:  g = grad(cost / N);  stop when |g| < gradient_convergence_tolerance;
step = -rate * g ./ precondition_divisors, rescaled so that |step| <= max_step;  x += step.
EVERY field of the solver is honoured; `fg!` is the closure above, so the callback sees every evaluation.
"""
function naive_optimize!(fg!::Function, solver, x0::Vector{Float64})
    x = copy(x0); g = similar(x)
    rate = field_or(solver, :rate, 0.1)
    max_step = field_or(solver, :max_step, 0.5)
    iters = field_or(solver, :iteration_limit, 30)
    tol = field_or(solver, :gradient_convergence_tolerance, 0.0)
    div = field_or(solver, :precondition_divisors, nothing)
    divisors = div === nothing ? ones(length(x)) : Float64[d for d in div]
    for i in 1:iters
        fg!(g, x)
        (tol > 0 && norm(g) < tol) && break
        step = -rate .* g ./ divisors
        s = norm(step)
        s > max_step && scale!(step, max_step / s)
        x .+= step
    end
    x
end

"""
A `MathProgBase.AbstractNLPEvaluator` over the same closure, for solvers that go through MathProgBase
(`IpoptSolver(print_level=1, max_iter=50)`, examples/squishable.ipynb:344): objective and gradient from the library,
no constraints, no Hessian (Ipopt then runs its limited-memory quasi-Newton mode, as it does for the reference's
ForwardDiff gradient).
"""
type FlashEvaluator <: MathProgBase.AbstractNLPEvaluator
    fg!::Function
    scratch::Vector{Float64}
end
MathProgBase.initialize(::FlashEvaluator, requested_features::Vector{Symbol}) =
    all(f -> f in (:Grad,), requested_features) || error("FlashEvaluator provides :Grad only")
MathProgBase.features_available(::FlashEvaluator) = [:Grad]
MathProgBase.eval_f(e::FlashEvaluator, x) = e.fg!(e.scratch, x)
MathProgBase.eval_grad_f(e::FlashEvaluator, g, x) = (e.fg!(g, x); nothing)
MathProgBase.eval_g(::FlashEvaluator, g, x) = nothing
MathProgBase.jac_structure(::FlashEvaluator) = Int[], Int[]
MathProgBase.eval_jac_g(::FlashEvaluator, J, x) = nothing

"""
Drop-in for `Flash.Tracking.estimate_state` (src/tracking.jl:8-27), same signature, same defaults.
 * NaiveSolver with a callback: the loop stays in Julia (`naive_optimize!`), one flash_cost_grad per iteration.
 * NaiveSolver without a callback: the whole loop runs on the device (flash_estimate_state), all solver fields passed.
 * any MathProgBase solver (IpoptSolver): driven through `FlashEvaluator`.
The context of a manipulator is built once and cached (`context`).
"""
function estimate_state(manipulator::Manipulator, sensed_points::Vector{SVector{3, Float64}}, x_estimated::Vector{Float64};
                        callback::Union{Function, Void}=nothing,
                        solver=SimpleGradientDescent.NaiveSolver(num_states(manipulator), rate=0.1, max_step=0.5, iteration_limit=30),
                        ctx::Context=context(manipulator))
    set_points!(ctx, sensed_points)
    n = length(sensed_points)
    if isa(solver, SimpleGradientDescent.NaiveSolver)
        if callback === nothing
            x = copy(x_estimated)
            div = field_or(solver, :precondition_divisors, nothing)
            divisors = div === nothing ? ones(length(x)) : Float64[d for d in div]   # (a named array: ccall roots it for the call)
            sp = SolverParams(solver.rate, solver.max_step, solver.iteration_limit,
                              field_or(solver, :gradient_convergence_tolerance, 0.0), Flash.GradientDescent.default_deformation_cost_weight)
            check(ctx.ptr, ccall((:flash_estimate_state, lib), Cint,
                                 (Ptr{Void}, Ptr{Float64}, Int32, Ptr{SolverParams}, Ptr{Float64}, Ptr{Float64}, Ptr{Int32}),
                                 ctx.ptr, x, 1, &sp, divisors, C_NULL, C_NULL))
            return x
        end
        return naive_optimize!(wrapped_cost_and_gradient(ctx, n, callback), solver, x_estimated)
    end
    fg! = wrapped_cost_and_gradient(ctx, n, callback === nothing ? (x, c) -> () : callback)
    nvars = num_states(manipulator)
    model = MathProgBase.NonlinearModel(solver)
    MathProgBase.loadproblem!(model, nvars, 0, fill(-Inf, nvars), fill(Inf, nvars), Float64[], Float64[], :Min, FlashEvaluator(fg!, zeros(nvars)))
    MathProgBase.setwarmstart!(model, x_estimated)
    MathProgBase.optimize!(model)
    MathProgBase.getsolution(model)
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

"`DepthSensors.raycast_points` (src/depthsensors.jl:99-113): the world points of the rays that hit, in row-major ray order."
function raycast_points(ctx::Context, x::Vector{Float64}, sensor::Flash.DepthSensors.DepthSensor, R::Matrix{Float64}, t::Vector{Float64})
    rm = permutedims(sensor.rays, (2, 1))
    nr = length(rm)
    depths = Array{Float64}(nr)
    hits = Array{Float64}(3nr)
    nh = Ref{Int64}(0)
    check(ctx.ptr, ccall((:flash_raycast, lib), Cint,
                         (Ptr{Void}, Ptr{Float64}, Ptr{Float64}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Int64}),
                         ctx.ptr, x, pointer(reinterpret(Float64, vec(rm), (3nr,))), nr, rowmajor34(R, t), depths, hits, nh))
    [SVector{3, Float64}(hits[3k - 2], hits[3k - 1], hits[3k]) for k in 1:nh[]]
end

"`Flash.skin(state)` (src/Flash.jl:265-268) as a closure over the device field: `skin(ctx, x)(p)` for one point or a vector of points."
function skin(ctx::Context, x::Vector{Float64})
    function field(points::Vector{SVector{3, Float64}})
        n = length(points)
        dist = Array{Float64}(n)
        check(ctx.ptr, ccall((:flash_skin_eval, lib), Cint, (Ptr{Void}, Ptr{Float64}, Ptr{Float64}, Int64, Ptr{Float64}, Ptr{Int32}),
                             ctx.ptr, x, pointer(reinterpret(Float64, points, (3n,))), n, dist, C_NULL))
        dist
    end
    field(p::SVector{3, Float64}) = field([p])[1]
    field
end

# ---- a live stream (examples/irb_and_squishable.ipynb:495-503): frame k+1 uploads and sorts while frame k is tracked
"Start the upload of the NEXT frame; `swap_points!` makes it current.  `sensed_points` must stay alive until then."
function set_points_async!(ctx::Context, sensed_points::Vector{SVector{3, Float64}})
    check(ctx.ptr, ccall((:flash_points_set_async, lib), Cint, (Ptr{Void}, Ptr{Float64}, Int64),
          ctx.ptr, pointer(reinterpret(Float64, sensed_points, (3 * length(sensed_points),))), length(sensed_points)))
end
swap_points!(ctx::Context) = check(ctx.ptr, ccall((:flash_points_swap, lib), Cint, (Ptr{Void},), ctx.ptr))

# ---- `Models.load_urdf` without the Julia mesh packages: URDF + STL / OBJ are read by the library (src/models.jl:154-171)
"A context for the URDF's manipulator (one convex surface per visual mesh); `package_path` as in `Models.load_urdf`."
function Context(urdf::String; package_path::Vector{String}=String[])
    out = Ref{Ptr{Void}}(C_NULL)
    rc = ccall((:flash_model_load_urdf, lib), Cint, (Cstring, Ptr{Cstring}, Int32, Ptr{Void}, Ptr{Ptr{Void}}),
               urdf, package_path, length(package_path), C_NULL, out)
    check(C_NULL, rc)
    ctx = Context(out[], Any[])
    finalizer(ctx, c -> ccall((:flash_destroy, lib), Void, (Ptr{Void},), c.ptr))
    ctx
end

end # module
