# KirstineB200.jl -- ccall glue between Kirstine.jl and libkirstine_b200.so.
#
# UNVERIFIED UNDER JULIA: Julia is not installed in the build container nor on the GPU box, so this file has never
# been executed.  What IS checked mechanically (tests/test_julia_glue.py, CPU): every `ccall` in this file names a
# function that include/kirstine_b200.h declares, with the same number of arguments and compatible C types, and the
# struct `ExchangeOptions` mirrors `kst_exchange_options` field by field.  The same ABI is exercised end to end by
# the Python/ctypes tests and by the C example.
#
# Usage sketch:
#   using Kirstine, KirstineB200
#   dp = DesignProblem(criterion = DCriterion(), region = DesignInterval(:dose => (0, 1)),
#                      model = KirstineB200.SigEmaxModel(sigma = 1),
#                      covariate_parameterization = JustCopy(:dose), prior_knowledge = prior)
#   objective(d, dp); gateauxderivative(at, dirs, dp); efficiency(d1, d2, dp); solve(dp, DirectMaximization(...))
# The methods below are more specific than Kirstine's generic ones (they dispatch on the model type parameter of
# DesignProblem), so registered model types go to the GPU and every other model stays on stock Kirstine.
# efficiency (src/designproblem.jl:315-350) and shannon_information (:366-370) need no method of their own: they
# call objective(d, as_doptimality_problem(dp)), which keeps the registered model type and lands here.
module KirstineB200

using Kirstine
import Kirstine: objective, gateauxderivative, informationmatrix, solve_with, DesignProblem,
    DesignMeasure, DirectMaximization, DirectMaximizationResult, Exchange, ExchangeResult, Pso,
    OptimizationResult, PsoState, SignedMeasure, DCriterion, ACriterion, Identity, DeltaMethod, PriorSample,
    NonlinearRegression, Parameter, CovariateParameterization, FisherMatrix, NormalApproximation,
    trafo_constants, codomain_dimension, parameter_dimension, numpoints, region, model, criterion,
    transformation, prior_knowledge, covariate_parameterization, normal_approximation, lowerbound, upperbound,
    check_compatible

const LIB = get(ENV, "KIRSTINE_B200_LIB", "libkirstine_b200.so")

# ---- registered model families (ids as in kirstine_b200.h) ------------------------------------------------------
abstract type RegisteredModel <: NonlinearRegression end
struct EmaxModel <: RegisteredModel;    sigma_squared::Float64; end       # test/example-emax.jl
struct SigEmaxModel <: RegisteredModel; sigma::Float64; end               # docs/src/tutorial.md
struct TPCModel <: RegisteredModel;     sigma::Float64; end               # test/example-compartment.jl
struct VUMod <: RegisteredModel;        sigma_squared::Float64; m::Int64; end                 # test/example-vector.jl
struct VUModAR1 <: RegisteredModel;     sigma_squared::Float64; rho::Float64; m::Int64; end   # dense Sigma(c), kirstine_b200.h
struct DTRMod <: RegisteredModel;       sigma::Float64; m::Int64; end     # test/example-dtr.jl, docs/src/dtr.md
SigEmaxModel(; sigma) = SigEmaxModel(sigma)
TPCModel(; sigma) = TPCModel(sigma)
model_id(::EmaxModel) = 1;    model_params(m::EmaxModel) = [m.sigma_squared]
model_id(::SigEmaxModel) = 2; model_params(m::SigEmaxModel) = [m.sigma]
model_id(::TPCModel) = 3;     model_params(m::TPCModel) = [m.sigma]
model_id(::VUMod) = 4;        model_params(m::VUMod) = [m.sigma_squared]
model_id(::DTRMod) = 5;       model_params(m::DTRMod) = [m.sigma]
model_id(::VUModAR1) = 8;     model_params(m::VUModAR1) = [m.sigma_squared, m.rho]
Kirstine.unit_length(m::RegisteredModel) = 1
Kirstine.unit_length(m::Union{VUMod,VUModAR1,DTRMod}) = m.m

# ---- covariate parameterizations: (cp id, cp parameters) as the ABI wants them -----------------------------------
# Any covariate parameterization that just copies the design point (JustCopy, CopyDose, CopyTime, ...) is id 0.
struct EquiTime <: CovariateParameterization end                                     # test/example-vector.jl:43-50
# docs/src/dtr.md: the five parameterizations of the dose-time-response model
struct CopyBoth <: CovariateParameterization end                                     # :197-203
struct FixedDose <: CovariateParameterization; dose::Float64; end                    # :331-339
struct FixedTimes <: CovariateParameterization; time::Vector{Float64}; end           # :401-409
struct EquidistantTimes <: CovariateParameterization end                             # :484-495
struct LogEquidistantTimes <: CovariateParameterization; b::Float64; end             # :663-681
cp_spec(::RegisteredModel, ::CovariateParameterization) = (0, Float64[])
cp_spec(::Union{VUMod,VUModAR1}, ::CovariateParameterization) = (1, Float64[])
cp_spec(::DTRMod, ::CopyBoth) = (2, Float64[])
cp_spec(::DTRMod, cp::FixedDose) = (3, [cp.dose])
cp_spec(::DTRMod, cp::FixedTimes) = (4, cp.time)
cp_spec(::DTRMod, ::EquidistantTimes) = (5, Float64[])
cp_spec(::DTRMod, cp::LogEquidistantTimes) = (6, [cp.b])
# the two keyword arguments of the registered simplify_unique methods (docs/src/dtr.md:229-247, :498-523)
unique_args(::CovariateParameterization, sargs) = (0.0, 0.0)
unique_args(::CopyBoth, sargs) = (Float64(get(sargs, :minposdose, 0)), Float64(get(sargs, :minpostime, 0)))
unique_args(::Union{EquidistantTimes,LogEquidistantTimes}, sargs) =
    (Float64(get(sargs, :minposdose, 0)), Float64(get(sargs, :minposdelta, 0)))

crit_id(::DCriterion) = 0
crit_id(::ACriterion) = 1

# ---- the reference's vignette extensions, registered as well ----------------------------------------------------
# docs/src/extend-criterion.md: det^(1/t)
struct DexpCriterion <: Kirstine.DesignCriterion end
crit_id(::DexpCriterion) = 2
# docs/src/extend-approximation.md: informationmatrix! adds R; sent to the engine once per handle
struct RegularizedFisherMatrix <: NormalApproximation
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
cp_spec(::RegisteredLogistic, ::CovariateParameterization) = (0, Float64[])
const Registered = Union{RegisteredModel,RegisteredLogistic}

struct EngineError <: Exception
    code::Cint
    msg::String
end
function check(rc::Cint)
    rc == 0 || throw(EngineError(rc, unsafe_string(ccall((:kst_last_error, LIB), Cstring, ()))))
    return nothing
end

# ---- problem handle ---------------------------------------------------------------------------------------------
mutable struct Handle
    ptr::Ptr{Cvoid}
end

const HANDLES = IdDict{Any,Handle}()

function create_handle(m, cp, crit, trafo, pk, na, lb::Vector{Float64}, ub::Vector{Float64}; device = 0)
    r = parameter_dimension(pk)
    N = length(pk.p)
    # Vector{P} of isbits parameter structs is already the AoS layout the ABI takes
    theta = collect(reinterpret(Float64, pk.p))
    tc = trafo_constants(trafo, pk)
    t = codomain_dimension(trafo, pk)
    isdelta = trafo isa DeltaMethod
    dt = isdelta ? reduce(vcat, vec.(tc.jm)) : Float64[]   # column-major t x r per draw
    mp = model_params(m)
    cpid, cpp = cp_spec(m, cp)
    out = Ref{Ptr{Cvoid}}(C_NULL)
    check(ccall((:kst_problem_create, LIB), Cint,
        (Cint, Ptr{Cdouble}, Cint, Cint, Ptr{Cdouble}, Cint, Cint, Cint, Cint, Cint, Cint, Cint, Int64,
         Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Cint, Ref{Ptr{Cvoid}}),
        model_id(m), mp, length(mp), cpid, isempty(cpp) ? C_NULL : cpp, length(cpp), crit_id(crit), isdelta ? 1 : 0,
        r, t, Kirstine.unit_length(m), length(lb), N, theta, pk.weight,
        isdelta ? dt : C_NULL, lb, ub, device, out))
    h = Handle(out[])
    finalizer(x -> ccall((:kst_problem_destroy, LIB), Cint, (Ptr{Cvoid},), x.ptr), h)
    if na isa RegularizedFisherMatrix   # Matrix{Float64} is column-major, as the ABI wants
        check(ccall((:kst_problem_set_regularization, LIB), Cint, (Ptr{Cvoid}, Ptr{Cdouble}), h.ptr, na.R))
    end
    return h
end

function handle(dp::DesignProblem{<:Any,<:Any,<:Registered}; device = 0)
    get!(HANDLES, dp) do
        create_handle(model(dp), covariate_parameterization(dp), criterion(dp), transformation(dp),
                      prior_knowledge(dp), normal_approximation(dp), collect(Float64, lowerbound(region(dp))),
                      collect(Float64, upperbound(region(dp))); device = device)
    end
end

# ---- evaluators -------------------------------------------------------------------------------------------------
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

# objective (src/designproblem.jl:233-237); efficiency and shannon_information arrive here through it
objective(d::DesignMeasure, dp::DesignProblem{<:Any,<:Any,<:Registered}) = objective_batch(dp, [d])[1]

# gateauxderivative (src/designproblem.jl:273-297)
function gateauxderivative(at::DesignMeasure, directions::AbstractArray{DesignMeasure},
                           dp::DesignProblem{<:Any,<:Any,<:Registered})
    any(d -> numpoints(d) != 1, directions) &&
        throw(ErrorException("derivative only implemented for one-point directions"))
    isinf(objective(at, dp)) && return fill(NaN, size(directions))
    h = handle(dp)
    singular = Ref{Cint}(0)
    check(ccall((:kst_gateaux_prepare, LIB), Cint, (Ptr{Cvoid}, Cint, Ptr{Cdouble}, Ptr{Cdouble}, Ref{Cint}),
        h.ptr, numpoints(at), at.points, at.weights, singular))
    singular[] != 0 && throw(Kirstine.LinearAlgebra.SingularException(0))
    dirs = reduce(hcat, [vec(d.points) for d in directions])
    gd = Vector{Float64}(undef, length(directions))
    check(ccall((:kst_gateaux_eval, LIB), Cint, (Ptr{Cvoid}, Int64, Ptr{Cdouble}, Ptr{Cdouble}),
        h.ptr, length(directions), dirs, gd))
    return reshape(gd, size(directions))
end

# informationmatrix (src/design-common.jl:96-112) for one parameter value: a one-draw handle whose region is any
# box around the design's points (the region does not enter the matrix)
function informationmatrix(d::DesignMeasure, m::Registered, cp::CovariateParameterization, p::Parameter,
                           na::NormalApproximation)
    lb = vec(min.(minimum(d.points; dims = 2), 0.0)) .- 1.0
    ub = vec(maximum(d.points; dims = 2)) .+ 1.0
    h = create_handle(m, cp, DCriterion(), Identity(), PriorSample([p]), na, lb, ub)
    r = parameter_dimension(PriorSample([p]))
    M = Matrix{Float64}(undef, r, r)
    check(ccall((:kst_informationmatrix, LIB), Cint,
        (Ptr{Cvoid}, Cint, Ptr{Cdouble}, Ptr{Cdouble}, Int64, Ptr{Cdouble}),
        h.ptr, numpoints(d), d.points, d.weights, 0, M))
    return Kirstine.LinearAlgebra.Symmetric(M)
end

# ---- optimize() on the device -------------------------------------------------------------------------------------
# The packed PsoState of kst_pso_optimize / kst_exchange_run (layout in kirstine_b200.h) as a Kirstine PsoState
function unpack_state(snap::AbstractVector{Float64}, S, K, d)
    o = Ref(0)
    take(n) = (v = snap[(o[] + 1):(o[] + n)]; o[] += n; v)
    x = reshape(take(S * K * d), d, K, S); xw = reshape(take(S * K), K, S)
    p = reshape(take(S * K * d), d, K, S); pw = reshape(take(S * K), K, S)
    v = reshape(take(S * K * d), d, K, S); vw = reshape(take(S * K), K, S)
    g = reshape(take(K * d), d, K); gw = take(K)
    fx = take(S); fp = take(S); fg = take(1)[1]
    dm(a, b) = DesignMeasure(copy(a), copy(b))
    zero = SignedMeasure(zeros(d, K), zeros(K))
    return PsoState([dm(x[:, :, i], xw[:, i]) for i in 1:S], [dm(p[:, :, i], pw[:, i]) for i in 1:S],
        dm(g, gw), [SignedMeasure(v[:, :, i], vw[:, i]) for i in 1:S], fx, fp, fg, zero, deepcopy(zero),
        deepcopy(zero), deepcopy(zero), S)
end

snapshot_length(S, K, d) = 3 * S * K * (d + 1) + K * (d + 1) + 2 * S + 1

# optimize (src/optimize.jl:42-70) with a Pso optimizer in one call; mode 0 = objective, 1 = Gateaux derivative
function optimize_device(h::Handle, o::Pso, prototypes::AbstractVector{DesignMeasure}, fixw, fixp, mode, d)
    K = numpoints(prototypes[1])
    S, it, P = o.swarmsize, o.iterations, length(prototypes)
    pp = reduce(hcat, [vec(q.points) for q in prototypes])
    pw = reduce(hcat, [q.weights for q in prototypes])
    status = Ref{Cint}(0)
    tfx = Vector{Float64}(undef, it)
    tpts = Array{Float64}(undef, d, K, it)
    tw = Matrix{Float64}(undef, K, it)
    snap = Vector{Float64}(undef, snapshot_length(S, K, d))
    t0 = time()
    check(ccall((:kst_pso_optimize, LIB), Cint,
        (Ptr{Cvoid}, Cint, Cint, Cint, Cdouble, Cdouble, Ptr{Cdouble}, Ptr{Cdouble}, Cint, Ptr{Cint}, Ptr{Cint},
         UInt64, Cint, Ref{Cint}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}),
        h.ptr, S, K, it, o.c1, o.c2, pp, pw, P, Cint.(fixw), Cint.(fixp), rand(UInt64), mode, status, tfx, tpts, tw,
        snap))
    status[] == 1 && throw(ErrorException("infinite objective value at best initial point"))
    status[] == 2 && throw(ErrorException("objective evaluates to NaN on all initial points"))
    trace_x = [DesignMeasure(tpts[:, :, i], tw[:, i]) for i in 1:it]
    isfinite(tfx[end]) || @warn "maximum is not finite" trace_x[end] tfx[end]
    return OptimizationResult(trace_x[end], tfx[end], trace_x, tfx, [unpack_state(snap, S, K, d)], S * it, time() - t0)
end

# solve_with for DirectMaximization (src/solve-directmaximization.jl:80-97).  trace_state = true wants every
# iteration's swarm on the host: stock Kirstine's loop does that, calling the routed objective per particle.
function solve_with(dp::DesignProblem{<:Any,<:Any,<:Registered}, strategy::DirectMaximization{Pso},
                    trace_state::Bool)
    trace_state && return invoke(solve_with, Tuple{DesignProblem,DirectMaximization,Bool}, dp, strategy, trace_state)
    constraints = Kirstine.DesignConstraints(strategy.prototype, region(dp), strategy.fixedweights,
                                             strategy.fixedpoints)
    d = length(lowerbound(region(dp)))
    or = optimize_device(handle(dp), strategy.optimizer, [strategy.prototype], constraints.fixw, constraints.fixp, 0, d)
    return DirectMaximizationResult(or)
end

# ---- Exchange (src/solve-exchange.jl:116-179) behind one call -----------------------------------------------------
# kst_exchange_options (kirstine_b200.h), field for field
struct ExchangeOptions
    steps::Cint
    swarm_direction::Cint
    iters_direction::Cint
    c1_direction::Cdouble
    c2_direction::Cdouble
    swarm_weight::Cint
    iters_weight::Cint
    c1_weight::Cdouble
    c2_weight::Cdouble
    maxweight::Cdouble
    maxdist::Cdouble
    unique_a::Cdouble
    unique_b::Cdouble
    seed::UInt64
end

function solve_with(dp::DesignProblem{<:Any,<:Any,<:Registered}, strategy::Exchange{Pso,Pso}, trace_state::Bool)
    trace_state && return invoke(solve_with, Tuple{DesignProblem,Exchange,Bool}, dp, strategy, trace_state)
    cand = strategy.candidate
    check_compatible(cand, region(dp))
    h = handle(dp)
    od, ow, steps = strategy.optimizer_direction, strategy.optimizer_weight, strategy.steps
    sargs = strategy.simplify_args
    ua, ub = unique_args(covariate_parameterization(dp), sargs)
    opt = Ref(ExchangeOptions(steps, od.swarmsize, od.iterations, od.c1, od.c2, ow.swarmsize, ow.iterations, ow.c1, ow.c2,
                              Float64(get(sargs, :maxweight, 0)), Float64(get(sargs, :maxdist, 0)), ua, ub, rand(UInt64)))
    d = length(lowerbound(region(dp)))
    K0 = numpoints(cand)
    Kmax = K0 + steps
    itd, itw, Sd, Sw = od.iterations, ow.iterations, od.swarmsize, ow.swarmsize
    Kout = Ref{Cint}(0); done = Ref{Cint}(0)
    pts = Matrix{Float64}(undef, d, Kmax); w = Vector{Float64}(undef, Kmax)
    dfx = Matrix{Float64}(undef, itd, steps); dx = Array{Float64}(undef, d, itd, steps)
    wfx = Matrix{Float64}(undef, itw, steps)
    wp = Matrix{Float64}(undef, itw * Kmax * d, steps); ww = Matrix{Float64}(undef, itw * Kmax, steps)
    wK = Vector{Cint}(undef, steps)
    nsd, nsw = snapshot_length(Sd, 1, d), snapshot_length(Sw, Kmax, d)
    sd = Matrix{Float64}(undef, nsd, steps); sw = Matrix{Float64}(undef, nsw, steps)
    t0 = time()
    rc = ccall((:kst_exchange_run, LIB), Cint,
        (Ptr{Cvoid}, Ref{ExchangeOptions}, Cint, Ptr{Cdouble}, Ptr{Cdouble}, Cint, Ref{Cint}, Ptr{Cdouble}, Ptr{Cdouble},
         Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cdouble}, Ptr{Cint}, Ptr{Cdouble}, Ptr{Cdouble},
         Ref{Cint}),
        h.ptr, opt, K0, cand.points, cand.weights, Kmax, Kout, pts, w, dfx, dx, wfx, wp, ww, wK, sd, sw, done)
    rc == -7 && throw(Kirstine.LinearAlgebra.SingularException(0))
    rc == -8 && throw(ErrorException(unsafe_string(ccall((:kst_last_error, LIB), Cstring, ()))))
    check(rc)
    secs = (time() - t0) / (2 * steps)
    ORType = OptimizationResult{DesignMeasure,SignedMeasure,PsoState{DesignMeasure,SignedMeasure}}
    ors_d = ORType[]; ors_w = ORType[]
    for s in 1:steps
        tx = [DesignMeasure(reshape(dx[:, i, s], d, 1), [1.0]) for i in 1:itd]
        push!(ors_d, OptimizationResult(tx[end], dfx[end, s], tx, dfx[:, s], [unpack_state(view(sd, :, s), Sd, 1, d)],
                                        Sd * itd, secs))
        k = Int(wK[s])
        P = reshape(wp[1:(itw * k * d), s], d, k, itw); W = reshape(ww[1:(itw * k), s], k, itw)
        tx = [DesignMeasure(P[:, :, i], W[:, i]) for i in 1:itw]
        push!(ors_w, OptimizationResult(tx[end], wfx[end, s], tx, wfx[:, s],
                                        [unpack_state(view(sw, 1:snapshot_length(Sw, k, d), s), Sw, k, d)], Sw * itw, secs))
    end
    return ExchangeResult(ors_d, ors_w)
end

end # module
