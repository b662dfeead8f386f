# KirstineB200.jl -- ccall glue between Kirstine.jl and libkirstine_b200.so.
#
# UNVERIFIED: Julia is not installed in the build container nor on the GPU box, so this file has
# never been executed.  It shows the reference-side binding (INTEGRATION.md) for the C ABI in
# include/kirstine_b200.h; the same ABI is exercised end to end by the Python/ctypes tests.
#
# Usage sketch:
#   using Kirstine, KirstineB200
#   dp = DesignProblem(criterion = DCriterion(), region = DesignInterval(:dose => (0, 1)),
#                      model = KirstineB200.SigEmaxModel(sigma = 1),
#                      covariate_parameterization = JustCopy(:dose), prior_knowledge = prior)
#   objective(d, dp); gateauxderivative(at, dirs, dp); solve(dp, DirectMaximization(...))
# Methods below are more specific than Kirstine's generic ones, so dispatch routes registered model
# types to the GPU and leaves every other model on stock Kirstine.
module KirstineB200

using Kirstine
import Kirstine: objective, gateauxderivative, informationmatrix, solve_with, DesignProblem,
    DesignMeasure, DirectMaximization, DirectMaximizationResult, Pso, OptimizationResult,
    PsoState, SignedMeasure, DCriterion, ACriterion, Identity, DeltaMethod, PriorSample,
    NonlinearRegression, Parameter, trafo_constants, codomain_dimension, parameter_dimension,
    numpoints, region, model, criterion, transformation, prior_knowledge, lowerbound, upperbound

const LIB = get(ENV, "KIRSTINE_B200_LIB", "libkirstine_b200.so")

# ---- registered model families (ids as in kirstine_b200.h) --------------------------------------
abstract type RegisteredModel <: NonlinearRegression end
struct EmaxModel <: RegisteredModel;    sigma_squared::Float64; end
struct SigEmaxModel <: RegisteredModel; sigma::Float64; end
struct TPCModel <: RegisteredModel;     sigma::Float64; end
struct VUMod <: RegisteredModel;        sigma_squared::Float64; m::Int64; end
SigEmaxModel(; sigma) = SigEmaxModel(sigma)
TPCModel(; sigma) = TPCModel(sigma)
model_id(::EmaxModel) = 1;  model_params(m::EmaxModel) = [m.sigma_squared]
model_id(::SigEmaxModel) = 2; model_params(m::SigEmaxModel) = [m.sigma]
model_id(::TPCModel) = 3;   model_params(m::TPCModel) = [m.sigma]
model_id(::VUMod) = 4;      model_params(m::VUMod) = [m.sigma_squared]
Kirstine.unit_length(m::RegisteredModel) = 1
Kirstine.unit_length(m::VUMod) = m.m
cp_id(::RegisteredModel) = 0
cp_id(::VUMod) = 1
crit_id(::DCriterion) = 0
crit_id(::ACriterion) = 1

# ---- the reference's vignette extensions, registered as well ----------------------------------------
# docs/src/extend-criterion.md: det^(1/t)
struct DexpCriterion <: Kirstine.DesignCriterion end
crit_id(::DexpCriterion) = 2
# docs/src/extend-approximation.md: informationmatrix! adds R; sent to the engine once per handle
struct RegularizedFisherMatrix <: Kirstine.NormalApproximation
    R::Matrix{Float64}
end
# docs/src/extend-model.md: logistic regression has its own average_fishermatrix!, so it is a Model, not a
# NonlinearRegression; the handle/evaluator methods below dispatch on the union of both registered supertypes
abstract type RegisteredLogistic <: Kirstine.Model end
struct LinLogReg <: RegisteredLogistic; dim_covariate::Int64; end   # theta = (beta_0, beta_rest...), d = 1..3
struct SinLogReg <: RegisteredLogistic end                          # theta = (a, b, c)
model_id(::LinLogReg) = 6; model_id(::SinLogReg) = 7
model_params(::RegisteredLogistic) = [0.0]                          # ignored by the engine
Kirstine.unit_length(::RegisteredLogistic) = 1
cp_id(::RegisteredLogistic) = 0
const Registered = Union{RegisteredModel,RegisteredLogistic}

struct EngineError <: Exception
    code::Cint
    msg::String
end
function check(rc::Cint)
    rc == 0 || throw(EngineError(rc, unsafe_string(ccall((:kst_last_error, LIB), Cstring, ()))))
    return nothing
end

# ---- problem handle -------------------------------------------------------------------------------
mutable struct Handle
    ptr::Ptr{Cvoid}
end

const HANDLES = IdDict{Any,Handle}()

function handle(dp::DesignProblem{<:Any,<:Any,<:Registered}; device = 0)
    get!(HANDLES, dp) do
        pk = prior_knowledge(dp)
        r = parameter_dimension(pk)
        N = length(pk.p)
        # Vector{P} of isbits parameter structs is already the AoS layout the ABI takes
        theta = reinterpret(Float64, pk.p)
        tc = trafo_constants(transformation(dp), pk)
        t = codomain_dimension(transformation(dp), pk)
        isdelta = transformation(dp) isa DeltaMethod
        dt = isdelta ? reduce(vcat, vec.(tc.jm)) : Float64[]   # column-major t x r per draw
        lb = collect(lowerbound(region(dp)))
        ub = collect(upperbound(region(dp)))
        mp = model_params(model(dp))
        out = Ref{Ptr{Cvoid}}(C_NULL)
        check(ccall((:kst_problem_create, LIB), Cint,
            (Cint, Ptr{Cdouble}, Cint, Cint, Ptr{Cdouble}, Cint, Cint, Cint, Cint, Cint, Cint, Cint, Int64,
             Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Cint, Ref{Ptr{Cvoid}}),
            model_id(model(dp)), mp, length(mp), cp_id(model(dp)), C_NULL, 0, crit_id(criterion(dp)), isdelta ? 1 : 0,
            r, t, Kirstine.unit_length(model(dp)), length(lb), N, theta, pk.weight,
            isdelta ? dt : C_NULL, lb, ub, device, out))
        h = Handle(out[])
        finalizer(x -> ccall((:kst_problem_destroy, LIB), Cint, (Ptr{Cvoid},), x.ptr), h)
        na = Kirstine.normal_approximation(dp)
        if na isa RegularizedFisherMatrix   # Matrix{Float64} is column-major, as the ABI wants
            check(ccall((:kst_problem_set_regularization, LIB), Cint, (Ptr{Cvoid}, Ptr{Cdouble}), h.ptr, na.R))
        end
        h
    end
end

# ---- evaluators -----------------------------------------------------------------------------------
function objective_batch(dp, ds::AbstractVector{DesignMeasure})
    h = handle(dp)
    S, K = length(ds), numpoints(ds[1])
    pts = reduce(hcat, [vec(d.points) for d in ds])        # d x K column-major per particle
    w = reduce(hcat, [d.weights for d in ds])
    fx = Vector{Float64}(undef, S)
    check(ccall((:kst_objective_batch, LIB), Cint,
        (Ptr{Cvoid}, Cint, Cint, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}), h.ptr, S, K, pts, w, fx))
    return fx
end

objective(d::DesignMeasure, dp::DesignProblem{<:Any,<:Any,<:Registered}) = objective_batch(dp, [d])[1]

function gateauxderivative(at::DesignMeasure, directions::AbstractArray{DesignMeasure},
                           dp::DesignProblem{<:Any,<:Any,<:Registered})
    any(d -> numpoints(d) != 1, directions) &&
        throw(ErrorException("derivative only implemented for one-point directions"))
    isinf(objective(at, dp)) && return fill(NaN, size(directions))
    h = handle(dp)
    singular = Ref{Cint}(0)
    check(ccall((:kst_gateaux_prepare, LIB), Cint, (Ptr{Cvoid}, Cint, Ptr{Cdouble}, Ptr{Cdouble}, Ref{Cint}),
        h.ptr, numpoints(at), at.points, at.weights, singular))
    dirs = reduce(hcat, [vec(d.points) for d in directions])
    gd = Vector{Float64}(undef, length(directions))
    check(ccall((:kst_gateaux_eval, LIB), Cint, (Ptr{Cvoid}, Int64, Ptr{Cdouble}, Ptr{Cdouble}),
        h.ptr, length(directions), dirs, gd))
    return reshape(gd, size(directions))
end

# ---- whole PSO on the device ------------------------------------------------------------------------
function solve_with(dp::DesignProblem{<:Any,<:Any,<:Registered}, strategy::DirectMaximization{Pso},
                    trace_state::Bool)
    constraints = Kirstine.DesignConstraints(strategy.prototype, region(dp), strategy.fixedweights,
                                             strategy.fixedpoints)
    h = handle(dp)
    o = strategy.optimizer
    K = numpoints(strategy.prototype)
    d = length(lowerbound(region(dp)))
    status = Ref{Cint}(0)
    t0 = time()
    check(ccall((:kst_pso_init, LIB), Cint,
        (Ptr{Cvoid}, Cint, Cint, Cdouble, Cdouble, Ptr{Cdouble}, Ptr{Cdouble}, Cint, Ptr{Cint}, Ptr{Cint},
         UInt64, Cint, Ref{Cint}),
        h.ptr, o.swarmsize, K, o.c1, o.c2, strategy.prototype.points, strategy.prototype.weights, 1,
        Cint.(constraints.fixw), Cint.(constraints.fixp), rand(UInt64), 0, status))
    status[] == 1 && throw(ErrorException("infinite objective value at best initial point"))
    status[] == 2 && throw(ErrorException("objective evaluates to NaN on all initial points"))
    n = o.iterations - 1
    tfx = Vector{Float64}(undef, n)
    tpts = Array{Float64}(undef, d, K, n)
    tw = Matrix{Float64}(undef, K, n)
    g0 = get_state(h, o.swarmsize, K, d)
    check(ccall((:kst_pso_run, LIB), Cint, (Ptr{Cvoid}, Cint, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}),
        h.ptr, n, tfx, tpts, tw))
    st = get_state(h, o.swarmsize, K, d)
    trace_x = [deepcopy(g0.g); [DesignMeasure(tpts[:, :, i], tw[:, i]) for i in 1:n]]
    trace_fx = [g0.fg; tfx]
    return DirectMaximizationResult(OptimizationResult(st.g, st.fg, trace_x, trace_fx, [g0], st.n_eval, time() - t0))
end

function get_state(h::Handle, S, K, d)
    x = Array{Float64}(undef, d, K, S); xw = Matrix{Float64}(undef, K, S)
    p = similar(x); pw = similar(xw); v = similar(x); vw = similar(xw)
    g = Matrix{Float64}(undef, d, K); gw = Vector{Float64}(undef, K)
    fx = Vector{Float64}(undef, S); fp = similar(fx); fg = Ref{Float64}(0); ne = Ref{Int64}(0)
    check(ccall((:kst_pso_get_state, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble},
         Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ref{Float64}, Ref{Int64}),
        h.ptr, x, xw, p, pw, g, gw, v, vw, fx, fp, fg, ne))
    dm(a, b) = DesignMeasure(copy(a), copy(b))
    zero = SignedMeasure(zeros(d, K), zeros(K))
    return PsoState([dm(x[:, :, i], xw[:, i]) for i in 1:S], [dm(p[:, :, i], pw[:, i]) for i in 1:S],
        dm(g, gw), [SignedMeasure(v[:, :, i], vw[:, i]) for i in 1:S], fx, fp, fg[], zero, deepcopy(zero),
        deepcopy(zero), deepcopy(zero), ne[])
end

end # module
