# GPT_SGLD.jl — drop-in host module for the hot path of hyunjik11/GPT, backed by libgptb200.so (B200, sm_100a).
#
# Same exported names and positional signatures as the reference module (/root/reference/GPT_SGLD.jl:5):
#   datawhitening, feature, feature2, samplenz, pred, GPTregression, proj, geod
# Every method `ccall`s one entry point of include/gptb200.h; arrays cross the boundary as Julia stores them
# (column-major Float64, I::Matrix{Int32} 1-based), so nothing is copied or transposed on the Julia side.
#
# Differences a caller can see:
#   * feature/feature2 return an RFFFeatures handle (X, Z, b, length_scale, sigma_RBF, phi_scale) instead of the
#     n x D x N array; GPTregression/pred dispatch on it and the kernels generate the features on the fly.
#     Array(f::RFFFeatures) materialises it (gptb_features) for code that indexes phi directly.
#     Methods taking a dense phi::Array{Float64,3} are kept (the reference's calling convention).
#   * Julia 0.3's dSFMT stream cannot be reproduced: w0/U0, permutations and noise come from Philox(param_seed)
#     unless passed explicitly with the keyword arguments w0, U0, perms, noise_w, noise_U (test mode).
#   * On a geodesic NaN the library returns GPTB_GEODESIC_NAN; this module prints the reference's warning and
#     returns zero-filled stores, as GPT_SGLD.jl:24-26,422-424 does.
# Written for Julia >= 1.6 (the reference itself is Julia 0.3 and does not load on 1.x).  Not executed in the
# build container (no julia binary there); the Python mirror gpt_b200/gpt_sgld.py is what the tests drive.
module GPT_SGLD

export datawhitening, feature, feature2, samplenz, pred, GPTregression, proj, geod, RFFFeatures,
       featureNotensor, GPNT_SGLD, NTFeatures, GPT_SGLDERMw, GPT_SGLDERM_RMSprop, GPT_GMC, GPTclassification, GPT_inf, logsumexp, geodboth,
       Model, pred_samples, posterior_mean, gradients, sgld_step!, sgld_advance!, synchronize, get_params, set_params!,
       set_option!, set_domain!, comm_unique_id, comm_init!

using Random, Statistics

const LIB = get(ENV, "GPTB200_LIB", joinpath(@__DIR__, "..", "gpt_b200", "lib", "libgptb200.so"))

const GPTB_OK = Cint(0)
const GPTB_GEODESIC_NAN = Cint(7)
const PREC_FP64, PREC_TC3, PREC_TC1 = Cint(0), Cint(1), Cint(2)
const FEAT_COS, FEAT_SINCOS = Cint(0), Cint(1)
const FLAG_LANGEVIN, FLAG_STIEFEL = Cint(1), Cint(2)

struct Config            # mirrors struct gptb_config
    D::Int32; n::Int32; r::Int32; Q::Int32
    precision::Int32; feature_kind::Int32; device::Int32; reserved::Int32
    seed::UInt64
end

struct RFFFeatures
    X::Matrix{Float64}; length_scale::Vector{Float64}; sigma_RBF::Float64; phi_scale::Float64
    Z::Matrix{Float64}; b::Matrix{Float64}; kind::Cint
end
Base.size(f::RFFFeatures) = (f.kind == FEAT_COS ? size(f.Z, 1) : size(f.Z, 1), size(f.X, 2), size(f.X, 1))

lasterr(h) = unsafe_string(ccall((:gptb_last_error, LIB), Cstring, (Ptr{Cvoid},), h))
function check(st::Cint, h = C_NULL)
    st == GPTB_OK || error("libgptb200: status $st: $(lasterr(h))")
    nothing
end

function create(D, n, r, Q; precision = PREC_TC3, kind = FEAT_COS, device = -1, seed = 0)
    h = Ref{Ptr{Cvoid}}(C_NULL)
    cfg = Ref(Config(D, n, r, Q, precision, kind, device, 0, UInt64(seed)))
    check(ccall((:gptb_create, LIB), Cint, (Ref{Config}, Ref{Ptr{Cvoid}}), cfg, h))
    h[]
end
destroy(h) = ccall((:gptb_destroy, LIB), Cint, (Ptr{Cvoid},), h)

function set_features!(h, f::RFFFeatures)
    ls = f.length_scale
    check(ccall((:gptb_set_features, LIB), Cint,
                (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Int32, Float64, Float64),
                h, f.Z, f.kind == FEAT_COS ? pointer(f.b) : Ptr{Float64}(C_NULL), ls, length(ls), f.sigma_RBF, f.phi_scale), h)
end

# GPT_SGLD.jl:62-67 — host preprocessing, in place like the reference
function datawhitening(X::Array)
    for i = 1:size(X, 2)
        X[:, i] = (X[:, i] .- mean(X[:, i])) ./ std(X[:, i])
    end
    X
end

# GPT_SGLD.jl:71-84 / :88-105 — lazy handles
feature(X::Array, length_scale, sigma_RBF::Real, phi_scale::Real, Z::Array, b::Array) =
    RFFFeatures(Matrix{Float64}(X), Float64[length_scale...], sigma_RBF, phi_scale, Matrix{Float64}(Z), Matrix{Float64}(b), FEAT_COS)
function feature2(X::Array, length_scale, sigma_RBF::Real, phi_scale::Real, Z::Array)
    iseven(size(Z, 1)) || error("n not even")                     # :89-91
    # the library reads the first n/2 rows of Z as the frequencies, exactly as :94 does
    RFFFeatures(Matrix{Float64}(X), Float64[length_scale...], sigma_RBF, phi_scale, Matrix{Float64}(Z), zeros(size(Z)), FEAT_SINCOS)
end

function Base.Array(f::RFFFeatures)
    n, D, N = size(f)
    h = create(D, n, 1, 1; precision = PREC_FP64, kind = f.kind)
    try
        set_features!(h, f)
        phi = Array{Float64}(undef, n, D, N)
        check(ccall((:gptb_features, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Int64, Ptr{Float64}), h, f.X, N, phi), h)
        return phi
    finally
        destroy(h)
    end
end

# GPT_SGLD.jl:181-190 — Q distinct cells of the r^D lattice, little-endian base-r digits + 1
function samplenz(r::Integer, D::Integer, Q::Integer; rng = Random.default_rng())
    L = Set{Int64}()
    while length(L) < Q
        push!(L, rand(rng, 0:Int64(r)^D - 1))
    end
    I = Array{Int32}(undef, Q, D)
    for (q, l) in enumerate(sort!(collect(L)))
        I[q, :] = digits(l, base = r, pad = D) .+ 1
    end
    I
end

# GPT_SGLD.jl:233-243
function pred(w::Array, U::Array, I::Array, phitest::Union{RFFFeatures,Array{Float64,3}}; precision = PREC_TC3)
    n, r, D = size(U); Q = length(w)
    lazy = phitest isa RFFFeatures
    h = create(D, n, r, Q; precision = lazy ? precision : PREC_FP64, kind = lazy ? phitest.kind : FEAT_COS)
    try
        check(ccall((:gptb_set_index, LIB), Cint, (Ptr{Cvoid}, Ptr{Int32}), h, Matrix{Int32}(I)), h)
        if lazy
            set_features!(h, phitest)
            Nt = size(phitest.X, 1); fhat = Vector{Float64}(undef, Nt)
            check(ccall((:gptb_pred, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Int64, Ptr{Float64}),
                        h, w, U, phitest.X, Nt, fhat), h)
        else
            Nt = size(phitest, 3); fhat = Vector{Float64}(undef, Nt)
            check(ccall((:gptb_pred_phi, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Int64, Ptr{Float64}),
                        h, w, U, phitest, Nt, fhat), h)
        end
        return fhat
    finally
        destroy(h)
    end
end

# GPT_SGLD.jl:345-448
function GPTregression(phi::Union{RFFFeatures,Array{Float64,3}}, y::Array, signal_var::Real, I::Array, r::Integer,
                       Q::Integer, m::Integer, epsw::Real, epsU::Real, burnin::Integer, maxepoch::Integer,
                       param_seed::Integer; langevin::Bool = true, stiefel::Bool = true, precision = PREC_TC3,
                       w0 = nothing, U0 = nothing, perms = nothing, noise_w = nothing, noise_U = nothing,
                       devices::Vector{Int} = Int[])
    # devices = [0, 1, ...]: several GPUs driven from this one process (gptb_comm_init_local / gptb_multi_sgld_run); the
    # rows of every minibatch are sharded over them, the chain is the one-GPU chain up to summation order
    n, D, N = size(phi)
    lazy = phi isa RFFFeatures
    devs = isempty(devices) ? [-1] : devices
    hs = Ptr{Cvoid}[]
    try
        yv = Vector{Float64}(vec(y))
        for d in devs
            h = create(D, n, r, Q; precision = lazy ? precision : PREC_FP64, kind = lazy ? phi.kind : FEAT_COS, device = d,
                       seed = param_seed)
            push!(hs, h)
            check(ccall((:gptb_set_index, LIB), Cint, (Ptr{Cvoid}, Ptr{Int32}), h, Matrix{Int32}(I)), h)
            if lazy
                set_features!(h, phi)
                check(ccall((:gptb_set_train, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Int64), h, phi.X, yv, N), h)
            else
                check(ccall((:gptb_set_train_phi, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Int64), h, phi, yv, N), h)
            end
            if w0 === nothing
                check(ccall((:gptb_init_params, LIB), Cint, (Ptr{Cvoid}, UInt64, Int32), h, param_seed, stiefel ? 1 : 0), h)   # :357-369
            else
                check(ccall((:gptb_set_params, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}), h, w0, U0), h)
            end
        end
        h = hs[1]
        numbatches = cld(N, m)                                     # :353
        T = maxepoch * numbatches
        w_store = zeros(Q, T); U_store = zeros(n, r, D, T)         # :358-359
        flags = (langevin ? FLAG_LANGEVIN : Cint(0)) | (stiefel ? FLAG_STIEFEL : Cint(0))
        p(x) = x === nothing ? Ptr{Float64}(C_NULL) : pointer(x)
        pp = perms === nothing ? Ptr{Int64}(C_NULL) : pointer(perms)   # (burnin+maxepoch) x N, 0-based, row-major per epoch
        if length(hs) > 1
            check(ccall((:gptb_comm_init_local, LIB), Cint, (Ptr{Ptr{Cvoid}}, Int32), hs, length(hs)), h)
            st = ccall((:gptb_multi_sgld_run, LIB), Cint,
                       (Ptr{Ptr{Cvoid}}, Int32, Float64, Int64, Float64, Float64, Int32, Int32, Int32, Ptr{Int64}, Ptr{Float64},
                        Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
                       hs, length(hs), signal_var, m, epsw, epsU, burnin, maxepoch, flags, pp, p(noise_w), p(noise_U), w_store, U_store)
        else
            st = ccall((:gptb_sgld_run, LIB), Cint,
                       (Ptr{Cvoid}, Float64, Int64, Float64, Float64, Int32, Int32, Int32, Ptr{Int64}, Ptr{Float64}, Ptr{Float64},
                        Ptr{Float64}, Ptr{Float64}),
                       h, signal_var, m, epsw, epsU, burnin, maxepoch, flags, pp, p(noise_w), p(noise_U), w_store, U_store)
        end
        if st == GPTB_GEODESIC_NAN
            println("Get NaN when moving along Geodesic. Try smaller epsU")    # :25
            return w_store, U_store                                            # zero-filled, :422-424
        end
        check(st, h)
        return w_store, U_store
    finally
        for h in hs; destroy(h); end
    end
end

# GPT_SGLD.jl:14-16
function proj(U::Array, V::Array)
    n, r = size(U); out = similar(U)
    check(ccall((:gptb_proj, LIB), Cint, (Int32, Int32, Int32, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}), -1, n, r, U, V, out))
    out
end

# GPT_SGLD.jl:19-37
function geod(U::Array, mom::Array, t::Real)
    n, r = size(U); out = similar(U)
    st = ccall((:gptb_geod, LIB), Cint, (Int32, Int32, Int32, Ptr{Float64}, Ptr{Float64}, Float64, Ptr{Float64}), -1, n, r, U, mom, t, out)
    if st == GPTB_GEODESIC_NAN
        println("Get NaN when moving along Geodesic. Try smaller epsU")        # :25
        return zeros(n, r)                                                      # :26
    end
    check(st)
    out
end

# GPT_SGLD.jl:1065-1118  GPT_SGLDERMw(phi,y,signal_var,I,r,Q,m,epsw,burnin,maxepoch): SGLD on w only, U fixed at its
# Stiefel draw (GPTB_FLAG_FIX_U).  param_seed names the draws the reference takes from the global RNG.
const FLAG_FIX_U = Cint(4)
function GPT_SGLDERMw(phi::Union{RFFFeatures,Array{Float64,3}}, y::Array, signal_var::Real, I::Array, r::Integer, Q::Integer,
                      m::Integer, epsw::Real, burnin::Integer, maxepoch::Integer; param_seed::Integer = 0, precision = PREC_TC3)
    n, D, N = size(phi)
    lazy = phi isa RFFFeatures
    h = create(D, n, r, Q; precision = lazy ? precision : PREC_FP64, kind = lazy ? phi.kind : FEAT_COS, seed = param_seed)
    try
        check(ccall((:gptb_set_index, LIB), Cint, (Ptr{Cvoid}, Ptr{Int32}), h, Matrix{Int32}(I)), h)
        yv = Vector{Float64}(vec(y))
        if lazy
            set_features!(h, phi)
            check(ccall((:gptb_set_train, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Int64), h, phi.X, yv, N), h)
        else
            check(ccall((:gptb_set_train_phi, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Int64), h, phi, yv, N), h)
        end
        check(ccall((:gptb_init_params, LIB), Cint, (Ptr{Cvoid}, UInt64, Int32), h, param_seed, 1), h)      # :1076-1084
        w_store = zeros(Q, maxepoch * cld(N, m)); U = zeros(n, r, D); w = zeros(Q)
        check(ccall((:gptb_sgld_run, LIB), Cint,
                    (Ptr{Cvoid}, Float64, Int64, Float64, Float64, Int32, Int32, Int32, Ptr{Int64}, Ptr{Float64}, Ptr{Float64},
                     Ptr{Float64}, Ptr{Float64}),
                    h, signal_var, m, epsw, 0.0, burnin, maxepoch, FLAG_LANGEVIN | FLAG_STIEFEL | FLAG_FIX_U, C_NULL, C_NULL, C_NULL,
                    w_store, C_NULL), h)
        check(ccall((:gptb_get_params, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}), h, w, U), h)
        return w_store, U                                                                                 # :1117
    finally
        destroy(h)
    end
end

# ---- full-theta ("no tensor") model: GPT_SGLD.jl:109-120, :809-847 ---------------------------------------------
struct NTFeatures            # lazy result of featureNotensor: its inputs, not the n x N matrix
    X::Matrix{Float64}; length_scale::Vector{Float64}; sigma_RBF::Float64; Z::Matrix{Float64}; b::Vector{Float64}
end
Base.size(f::NTFeatures) = (size(f.Z, 1), size(f.X, 1))

# GPT_SGLD.jl:109  featureNotensor(X,length_scale,sigma_RBF,Z,b), Z=randn(n,D), b=2*pi*rand(n)
featureNotensor(X::Array, length_scale, sigma_RBF::Real, Z::Array, b::Array) =
    NTFeatures(Matrix{Float64}(X), Float64[length_scale...], Float64(sigma_RBF), Matrix{Float64}(Z), vec(Float64.(b)))

function nt_create(f::NTFeatures; precision = PREC_TC3, device = -1, seed = 0)
    n, D = size(f.Z); h = Ref{Ptr{Cvoid}}(C_NULL)
    check(ccall((:gptb_nt_create, LIB), Cint, (Int32, Int32, Int32, Int32, UInt64, Ref{Ptr{Cvoid}}), D, n, precision, device, seed, h))
    st = ccall((:gptb_nt_set_features, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Int32, Float64),
               h[], f.Z, f.b, f.length_scale, length(f.length_scale), f.sigma_RBF)
    st == 0 || error(unsafe_string(ccall((:gptb_nt_last_error, LIB), Cstring, (Ptr{Cvoid},), h[])))
    return h[]
end

function Base.Array(f::NTFeatures)
    h = nt_create(f); n, N = size(f); phi = Matrix{Float64}(undef, n, N)
    try
        st = ccall((:gptb_nt_features, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Int64, Ptr{Float64}), h, f.X, N, phi)
        st == 0 || error(unsafe_string(ccall((:gptb_nt_last_error, LIB), Cstring, (Ptr{Cvoid},), h)))
        return phi
    finally
        ccall((:gptb_nt_destroy, LIB), Cint, (Ptr{Cvoid},), h)
    end
end

# GPT_SGLD.jl:809  GPNT_SGLD(phi,y,signal_var,sigma_theta,m,eps_theta,decay_rate,burnin,maxepoch,param_seed)
function GPNT_SGLD(phi::NTFeatures, y::Array, signal_var::Real, sigma_theta::Real, m::Integer, eps_theta::Real,
                   decay_rate::Real, burnin::Integer, maxepoch::Integer, param_seed::Integer; precision = PREC_TC3)
    n, N = size(phi); T = (maxepoch + burnin) * cld(N, m)
    h = nt_create(phi; precision = precision, seed = param_seed)
    try
        st = ccall((:gptb_nt_set_train, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Int64), h, phi.X, vec(Float64.(y)), N)
        st == 0 || error(unsafe_string(ccall((:gptb_nt_last_error, LIB), Cstring, (Ptr{Cvoid},), h)))
        store = zeros(n, T)
        st = ccall((:gptb_nt_run, LIB), Cint,
                   (Ptr{Cvoid}, Float64, Float64, Int64, Float64, Float64, Int32, Int32, UInt64, Ptr{Float64}, Ptr{Int64}, Ptr{Float64}, Ptr{Float64}),
                   h, signal_var, sigma_theta, m, eps_theta, decay_rate, burnin, maxepoch, param_seed, C_NULL, C_NULL, C_NULL, store)
        if st == GPTB_GEODESIC_NAN
            println("Get NaN in theta. Try smaller epsilon")      # GPT_SGLD.jl:840
            return zeros(n)                                       # :841
        end
        st == 0 || error(unsafe_string(ccall((:gptb_nt_last_error, LIB), Cstring, (Ptr{Cvoid},), h)))
        return store
    finally
        ccall((:gptb_nt_destroy, LIB), Cint, (Ptr{Cvoid},), h)
    end
end

# ---- a persistent handle: what long-running scripts hold instead of re-creating a handle per call ----------------
# (kin40kExperiment.jl:74-87 trains, then predicts with many stored samples; one Model serves both and keeps the
#  training set, the node grid and the sweep's buffers on the device between calls.)
const FLAG_RMSPROP = Cint(8)
const FLAG_CLS_SINGLE_UPDATE = Cint(16)
const OPT_PIPE_GROUPS, OPT_PIPE_RESERVE, OPT_ALGORITHM, OPT_NODE_ORDER, OPT_NODE_THETA, OPT_RMS_ALPHA, OPT_RMS_LAMBDA,
      OPT_PRED_CHUNK_MB, OPT_FRONT_PREFETCH = Cint(0), Cint(1), Cint(2), Cint(3), Cint(4), Cint(5), Cint(6), Cint(7), Cint(8)
const ALG_AUTO, ALG_DIRECT, ALG_NODES = 0.0, 1.0, 2.0

mutable struct Model
    h::Ptr{Cvoid}
    D::Int; n::Int; r::Int; Q::Int; N::Int
    function Model(phi::RFFFeatures, I::Array, r::Integer, Q::Integer; precision = PREC_TC3, device = -1, seed = 0)
        n, D, N = size(phi)
        h = create(D, n, r, Q; precision = precision, kind = phi.kind, device = device, seed = seed)
        m = new(h, D, n, r, Q, 0)
        finalizer(x -> (x.h == C_NULL || (destroy(x.h); x.h = C_NULL)), m)
        set_features!(h, phi)
        check(ccall((:gptb_set_index, LIB), Cint, (Ptr{Cvoid}, Ptr{Int32}), h, Matrix{Int32}(I)), h)
        m
    end
end
Base.close(m::Model) = (m.h == C_NULL || (destroy(m.h); m.h = C_NULL); nothing)

set_option!(m::Model, option::Integer, value::Real) =
    check(ccall((:gptb_set_option, LIB), Cint, (Ptr{Cvoid}, Int32, Float64), m.h, option, value), m.h)
# bounds of every input column for the node grid when only host minibatches are passed (set_train! sets them itself)
set_domain!(m::Model, lo::Vector{Float64}, hi::Vector{Float64}) =
    check(ccall((:gptb_set_domain, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}), m.h, lo, hi), m.h)
function set_train!(m::Model, X::Matrix{Float64}, y::Array)
    m.N = size(X, 1)
    check(ccall((:gptb_set_train, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Int64), m.h, X, Vector{Float64}(vec(y)), m.N), m.h)
end
set_params!(m::Model, w::Vector{Float64}, U::Array{Float64,3}) =
    check(ccall((:gptb_set_params, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}), m.h, w, U), m.h)
init_params!(m::Model, param_seed::Integer; stiefel::Bool = true) =
    check(ccall((:gptb_init_params, LIB), Cint, (Ptr{Cvoid}, UInt64, Int32), m.h, param_seed, stiefel ? 1 : 0), m.h)
function get_params(m::Model)
    w = zeros(m.Q); U = zeros(m.n, m.r, m.D)
    check(ccall((:gptb_get_params, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}), m.h, w, U), m.h)
    w, U
end
function synchronize(m::Model)
    st = ccall((:gptb_synchronize, LIB), Cint, (Ptr{Cvoid},), m.h)
    st == GPTB_GEODESIC_NAN && (println("Get NaN when moving along Geodesic. Try smaller epsU"); return false)
    check(st, m.h); true
end

# The sample loop of kin40kExperiment.jl:75-87 / GPT_SGLD_p.jl:124-132 in one call: pred for every stored sample
# (columns `cols` of the stores), mean over the samples and, on request, the Nt x S matrix of per-sample predictions.
function pred_samples(m::Model, w_store::Matrix{Float64}, U_store::Array{Float64,4}, Xtest::Matrix{Float64};
                      cols = 1:size(w_store, 2), per_sample::Bool = false)
    ws = w_store[:, cols]; Us = U_store[:, :, :, cols]; S = size(ws, 2); Nt = size(Xtest, 1)
    mean_out = Vector{Float64}(undef, Nt)
    per = per_sample ? Matrix{Float64}(undef, Nt, S) : nothing
    check(ccall((:gptb_pred_samples, LIB), Cint,
                (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Int32, Ptr{Float64}, Int64, Ptr{Float64}, Ptr{Float64}),
                m.h, ws, Us, S, Xtest, Nt, mean_out, per_sample ? pointer(per) : Ptr{Float64}(C_NULL)), m.h)
    per_sample ? (mean_out, per) : mean_out
end
# "average the last `last` epochs' samples" exactly as kin40kExperiment.jl:78-86 does: the last sample of each epoch
posterior_mean(m::Model, w_store, U_store, Xtest, numbatches::Integer; last::Integer = 50) =
    pred_samples(m, w_store, U_store, Xtest; cols = [e * numbatches for e in (size(w_store, 2) ÷ numbatches - last + 1):(size(w_store, 2) ÷ numbatches)])

# minibatch gradients of the log posterior (GPT_SGLD.jl:384-408) at the model's current parameters;
# idx: 1-based rows of the training set, or a contiguous range of the current (permuted) order
function gradients(m::Model, signal_var::Real; idx::Union{Nothing,Vector{Int}} = nothing, start::Integer = 1,
                   bs::Integer = m.N, include_prior::Bool = true)
    gw = zeros(m.Q); gU = zeros(m.n, m.r, m.D)
    i0 = idx === nothing ? Ptr{Int64}(C_NULL) : pointer(Int64.(idx .- 1))
    nb = idx === nothing ? bs : length(idx)
    fhat = zeros(nb)
    check(ccall((:gptb_grad, LIB), Cint,
                (Ptr{Cvoid}, Float64, Ptr{Int64}, Int64, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Int32),
                m.h, signal_var, i0, start - 1, nb, gw, gU, fhat, include_prior ? 1 : 0), m.h)
    gw, gU, fhat
end

# one loop body of GPTregression (GPT_SGLD.jl:378-439) on the resident training set; returns false after a geodesic NaN
function sgld_step!(m::Model, signal_var::Real, epsw::Real, epsU::Real; idx::Union{Nothing,Vector{Int}} = nothing,
                    start::Integer = 1, bs::Integer = m.N, langevin::Bool = true, stiefel::Bool = true, rmsprop::Bool = false,
                    noise_w = nothing, noise_U = nothing)
    flags = (langevin ? FLAG_LANGEVIN : Cint(0)) | (stiefel ? FLAG_STIEFEL : Cint(0)) | (rmsprop ? FLAG_RMSPROP : Cint(0))
    i0 = idx === nothing ? Ptr{Int64}(C_NULL) : pointer(Int64.(idx .- 1))
    p(x) = x === nothing ? Ptr{Float64}(C_NULL) : pointer(x)
    st = ccall((:gptb_sgld_step, LIB), Cint,
               (Ptr{Cvoid}, Float64, Ptr{Int64}, Int64, Int64, Float64, Float64, Int32, Ptr{Float64}, Ptr{Float64}),
               m.h, signal_var, i0, start - 1, idx === nothing ? bs : length(idx), epsw, epsU, flags, p(noise_w), p(noise_U))
    st == GPTB_GEODESIC_NAN && (println("Get NaN when moving along Geodesic. Try smaller epsU"); return false)
    check(st, m.h); true
end
# the next nsteps loop bodies, nothing stored (burn-in); asynchronous — call synchronize(m)
sgld_advance!(m::Model, signal_var::Real, mb::Integer, epsw::Real, epsU::Real, nsteps::Integer; langevin = true, stiefel = true) =
    check(ccall((:gptb_sgld_advance, LIB), Cint, (Ptr{Cvoid}, Float64, Int64, Float64, Float64, Int32, Int64), m.h, signal_var, mb,
                epsw, epsU, (langevin ? FLAG_LANGEVIN : Cint(0)) | (stiefel ? FLAG_STIEFEL : Cint(0)), nsteps), m.h)

# ---- several GPUs: one Julia worker process per GPU (`julia -p G`, as the reference's GPT_SGLD_p.jl is run), every
# worker owns a Model on its device; rank 1's id is sent to the others with the usual remotecall machinery.
function comm_unique_id()
    id = Vector{UInt8}(undef, 128)
    check(ccall((:gptb_comm_unique_id, LIB), Cint, (Ptr{UInt8},), id)); id
end
comm_init!(m::Model, nranks::Integer, rank0::Integer, id::Vector{UInt8}) =
    check(ccall((:gptb_comm_init, LIB), Cint, (Ptr{Cvoid}, Int32, Int32, Ptr{UInt8}), m.h, nranks, rank0, id), m.h)

# GPT_SGLD.jl:40-59
function geodboth(U::Array, mom::Array, t::Real)
    n, r = size(U); Un = similar(U); Mn = similar(U)
    st = ccall((:gptb_geodboth, LIB), Cint, (Int32, Int32, Int32, Ptr{Float64}, Ptr{Float64}, Float64, Ptr{Float64}, Ptr{Float64}),
               -1, n, r, U, mom, t, Un, Mn)
    if st == GPTB_GEODESIC_NAN
        println("Get NaN when moving along Geodesic. Try smaller epsU")        # :44
        return zeros(n, r), zeros(n, r)                                         # :45
    end
    check(st); Un, Mn
end

# shared set-up of the samplers below: handle with features, index, training set and initial parameters
function sampler_handle(phi, y, I, r, Q, param_seed, precision)
    n, D, N = size(phi); lazy = phi isa RFFFeatures
    h = create(D, n, r, Q; precision = lazy ? precision : PREC_FP64, kind = lazy ? phi.kind : FEAT_COS, seed = param_seed)
    check(ccall((:gptb_set_index, LIB), Cint, (Ptr{Cvoid}, Ptr{Int32}), h, Matrix{Int32}(I)), h)
    yv = Vector{Float64}(vec(y))
    if lazy
        set_features!(h, phi)
        check(ccall((:gptb_set_train, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Int64), h, phi.X, yv, N), h)
    else
        check(ccall((:gptb_set_train_phi, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Int64), h, phi, yv, N), h)
    end
    check(ccall((:gptb_init_params, LIB), Cint, (Ptr{Cvoid}, UInt64, Int32), h, param_seed, 1), h)
    h, n, D, N
end

# GPT_SGLD.jl:1121-1237  GPT_SGLDERM_RMSprop(phi,y,signal_var,I,r,Q,m,epsilon,alpha,burnin,maxepoch)
function GPT_SGLDERM_RMSprop(phi::Union{RFFFeatures,Array{Float64,3}}, y::Array, signal_var::Real, I::Array, r::Integer, Q::Integer,
                             m::Integer, epsilon::Real, alpha::Real, burnin::Integer, maxepoch::Integer;
                             param_seed::Integer = 0, precision = PREC_TC3)
    h, n, D, N = sampler_handle(phi, y, I, r, Q, param_seed, precision)
    try
        check(ccall((:gptb_set_option, LIB), Cint, (Ptr{Cvoid}, Int32, Float64), h, OPT_RMS_ALPHA, alpha), h)
        T = maxepoch * cld(N, m); w_store = zeros(Q, T); U_store = zeros(n, r, D, T)
        st = ccall((:gptb_sgld_run, LIB), Cint,
                   (Ptr{Cvoid}, Float64, Int64, Float64, Float64, Int32, Int32, Int32, Ptr{Int64}, Ptr{Float64}, Ptr{Float64},
                    Ptr{Float64}, Ptr{Float64}),
                   h, signal_var, m, epsilon, 0.0, burnin, maxepoch, FLAG_LANGEVIN | FLAG_STIEFEL | FLAG_RMSPROP, C_NULL, C_NULL, C_NULL,
                   w_store, U_store)
        st == GPTB_GEODESIC_NAN && (println("Get NaN when moving along Geodesic. Try smaller epsU"); return w_store, U_store)   # :1227-1229
        check(st, h)
        return w_store, U_store
    finally
        destroy(h)
    end
end

# GPT_SGLD.jl:684-805  GPT_GMC(phi,y,signal_var,I,r,Q,epsw,epsU,burnin,maxepoch,L,param_seed)
# restore_U_on_reject=false keeps the reference's behaviour (its `U_old=U` is an alias of the array the leapfrog writes into).
function GPT_GMC(phi::Union{RFFFeatures,Array{Float64,3}}, y::Array, signal_var::Real, I::Array, r::Integer, Q::Integer,
                 epsw::Real, epsU::Real, burnin::Integer, maxepoch::Integer, L::Integer, param_seed::Integer;
                 precision = PREC_TC3, restore_U_on_reject::Bool = false)
    h, n, D, N = sampler_handle(phi, y, I, r, Q, param_seed, precision)
    try
        w_store = zeros(Q, maxepoch); U_store = zeros(n, r, D, maxepoch); accept_prob = zeros(maxepoch + burnin)
        st = ccall((:gptb_gmc_run, LIB), Cint,
                   (Ptr{Cvoid}, Float64, Float64, Float64, Int32, Int32, Int32, Int32, Ptr{Float64}, Ptr{Float64}, Ptr{Float64},
                    Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
                   h, signal_var, epsw, epsU, burnin, maxepoch, L, restore_U_on_reject ? 1 : 0, C_NULL, C_NULL, C_NULL,
                   w_store, U_store, accept_prob)
        st == GPTB_GEODESIC_NAN && (println("Get NaN when moving along Geodesic. Try smaller epsU"); return w_store, U_store, accept_prob)  # :746-748
        check(st, h)
        for a in accept_prob; println("accept_prob=", a); end                                             # :791
        return w_store, U_store, accept_prob
    finally
        destroy(h)
    end
end

# GPT_SGLD.jl export / BloodTransfusionExperiment.jl:53-56
function logsumexp(x::Array)
    a = maximum(x)
    a + log(sum(exp.(x .- a)))
end

# GPT_SGLD.jl:452-680  GPTclassification(phi,y,I,r,Q,m,epsw,epsU,burnin,maxepoch,param_seed;langevin=true,stiefel=true)
# One library handle per class (own parameters, same features / index table / inputs); y holds the labels 1..C (:460).
# double_update=true makes the reference's two moves per minibatch (:624-636 and :639-671).
function GPTclassification(phi::Union{RFFFeatures,Array{Float64,3}}, y::Array, I::Array, r::Integer, Q::Integer, m::Integer,
                           epsw::Real, epsU::Real, burnin::Integer, maxepoch::Integer, param_seed::Integer;
                           langevin = true, stiefel = true, precision = PREC_TC3, double_update::Bool = true)
    labels = Vector{Int32}(vec(round.(Int32, y .- minimum(y))))
    C = Int(maximum(labels)) + 1
    hs = Ptr{Cvoid}[]
    n = D = N = 0
    try
        for c in 1:C
            h, n, D, N = sampler_handle(phi, zeros(length(labels)), I, r, Q, param_seed + 7919 * (c - 1), precision)
            push!(hs, h)
            stiefel || check(ccall((:gptb_init_params, LIB), Cint, (Ptr{Cvoid}, UInt64, Int32), h, param_seed + 7919 * (c - 1), 0), h)
        end
        T = maxepoch * cld(N, m); w_store = zeros(Q, C, T); U_store = zeros(n, r, D, C, T)
        flags = (langevin ? FLAG_LANGEVIN : Cint(0)) | (stiefel ? FLAG_STIEFEL : Cint(0)) |
                (double_update ? Cint(0) : FLAG_CLS_SINGLE_UPDATE)
        st = ccall((:gptb_classify_run, LIB), Cint,
                   (Ptr{Ptr{Cvoid}}, Int32, Ptr{Int32}, Int64, Float64, Float64, Int32, Int32, Int32, Ptr{Int64}, Ptr{Float64},
                    Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
                   hs, C, labels, m, epsw, epsU, burnin, maxepoch, flags, C_NULL, C_NULL, C_NULL, w_store, U_store)
        st == GPTB_GEODESIC_NAN && (println("Get NaN when moving along Geodesic. Try smaller epsU"); return w_store, U_store)   # :631-633
        check(st, hs[1])
        return w_store, U_store
    finally
        for h in hs; destroy(h); end
    end
end

# TGP.jl:37-86  GPT_inf(X,y,sigma,n,r,sigmaRBF,q,generator,num_iterations,burnin): Gibbs sampler of the tensor model with
# Gaussian factors.  Z, b (TGP.feature :6-14: the same draw for every row), I (:57) and the initial factors (:56) come
# from Julia's RNG seeded with `generator`; the row-proportional work runs on the device (gptb_tgp_gibbs), FP64.
function GPT_inf(X::Matrix{Float64}, y::Vector{Float64}, sigma::Real, n::Integer, r::Integer, sigmaRBF::Real, q::Integer,
                 generator::Integer, num_iterations::Integer, burnin::Integer)
    X = datawhitening(copy(X)); yw = vec(datawhitening(reshape(copy(y), :, 1)))            # :39-40
    N, D = size(X)
    Random.seed!(generator)
    Z = randn(n, D); b = rand(n, D)
    f = RFFFeatures(X, [Float64(sigmaRBF)], 1.0, 1.0, Z, 2pi .* b, FEAT_COS)
    U0 = sqrt(1 / r) .* randn(n, r, D)                                                    # :54, :56
    I = Matrix{Int32}(rand(1:r, q, D))                                                    # :57
    h = create(D, n, r, q; precision = PREC_FP64, kind = FEAT_COS, seed = generator)
    try
        check(ccall((:gptb_set_index, LIB), Cint, (Ptr{Cvoid}, Ptr{Int32}), h, I), h)
        set_features!(h, f)
        check(ccall((:gptb_set_train, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Int64), h, X, yw, N), h)
        check(ccall((:gptb_set_params, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}), h, zeros(q), U0), h)
        T = num_iterations - burnin
        W_array = zeros(q, T); V_array = zeros(n, r, D, T)                                # :45-46
        check(ccall((:gptb_tgp_gibbs, LIB), Cint,
                    (Ptr{Cvoid}, Float64, Int32, Int32, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
                    h, sigma, num_iterations, burnin, C_NULL, C_NULL, W_array, V_array), h)
        return W_array, V_array, I
    finally
        destroy(h)
    end
end

end # module
