# shopt_b200.jl -- Julia binding of libshopt_b200.so (C ABI: include/shopt_b200.h) for ShOpt.jl's per-star fitting stage.
#
# include() it next to analyticLBFGS.jl (julia/patch_shopt_jl.py adds the include and replaces the four hot loops of
# shopt/shopt.jl: 130-164 model fit, 194-224 / 281-314 / 316-348 pixel-grid fit, 357-397 learned fit, 534-577 polyfit
# rounds).  It keeps the reference's globals and shopt.yml semantics: every function takes what the loop it replaces
# read from the globals and returns what that loop wrote to them.  Indices are 1-based on this side and 0-based across
# the C ABI.  The library has no CPU fallback: without a CUDA device every call raises with SHOPT_E_NODEVICE (-1).
#
# Julia is not installed in the image this repository is built in, so this file is not executed by the test suite;
# the same symbols with the same argument lists are exercised through Python ctypes (shopt_b200/_lib.py, tests/).
module ShoptB200

export fit_analytic_batch, analytic_cost_grad, s_out_of_bounds, pca_batch, smooth_mode, chisq_mode, polyfit, psf_at,
       field_interp, field_eval, ingest, snr_cut, write_summary, write_summary_full, read_summary, StageParams,
       stage_run, default_params, FLAG_BLACKLISTED, FLAG_PCA_FAILED, FLAG_LEARNED_OOB, FLAG_TRAINING, FLAG_REJECTED

const LIBSHOPT = get(ENV, "LIBSHOPT_B200", joinpath(@__DIR__, "..", "shopt_b200", "libshopt_b200.so"))

const FLAG_BLACKLISTED = Int32(1)    # model s out of [sLowerBound, sUpperBound]        (shopt.jl:166-171)
const FLAG_PCA_FAILED  = Int32(2)    # pixel-grid fit failed -> zeros(r,c), failedStars (shopt.jl:305-309)
const FLAG_LEARNED_OOB = Int32(4)    # learned s out of bounds -> failedStars           (shopt.jl:384-386)
const FLAG_TRAINING    = Int32(8)
const FLAG_REJECTED    = Int32(16)
const FIT_THREW        = Int32(-1)   # the reference's catch branch: s = g1 = g2 = 0     (shopt.jl:155-162, 387-393)

last_error() = unsafe_string(ccall((:shopt_last_error, LIBSHOPT), Cstring, ()))
function check(rc)
    rc == 0 && return nothing
    error("libshopt_b200: code $rc: " * last_error())
end

"(r, c, N) Float64 cube from the reference's Vector{Matrix{Float64}}: column-major, exactly the C layout"
cube(stars)::Array{Float64,3} = isempty(stars) ? zeros(0, 0, 0) : Array{Float64,3}(cat(stars...; dims = 3))
uncube(X) = [X[:, :, i] for i in 1:size(X, 3)]

# ---- shopt.jl:130-164 (cost / g!, model stars) and :357-397 (costD / gD!, learned stars) ------------------------------
"""
    fit_analytic_batch(stars, n_a, g_tol; x0 = rand(3, N), max_iter = 1000) -> s, g1, g2, status, iters

One `Optim.optimize(cost, g!, x0[:, i], LBFGS(), Optim.Options(g_tol = g_tol))` per star, then s = x[1]^2 and
(g1, g2) from (e1, e2) as at shopt.jl:145-153.  `status[i] == FIT_THREW` is the reference's catch branch (zeros).
"""
function fit_analytic_batch(stars, n_a::Integer, g_tol::Real; x0::Matrix{Float64} = rand(3, length(stars)), max_iter::Integer = 1000)
    X = cube(stars); r, c, N = size(X)
    s, g1, g2 = zeros(N), zeros(N), zeros(N); status, iters = zeros(Int32, N), zeros(Int32, N)
    GC.@preserve X x0 s g1 g2 status iters check(ccall((:shopt_fit_analytic_batch, LIBSHOPT), Cint,
        (Ptr{Float64}, Cint, Cint, Int64, Cint, Ptr{Float64}, Float64, Cint,
         Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Int32}, Ptr{Int32}),
        X, r, c, N, n_a, x0, g_tol, max_iter, s, g1, g2, status, iters))
    return s, g1, g2, status, iters
end

"cost(params) and g!(storage, params) of analyticLBFGS.jl:8-46, 89-94 for one [σ, e1, e2] per star: f (N), g (3, N)"
function analytic_cost_grad(stars, n_a::Integer, x::Matrix{Float64})
    X = cube(stars); r, c, N = size(X); f = zeros(N); g = zeros(3, N)
    GC.@preserve X x f g check(ccall((:shopt_analytic_cost_grad_batch, LIBSHOPT), Cint,
        (Ptr{Float64}, Cint, Cint, Int64, Cint, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}), X, r, c, N, n_a, x, f, g))
    return f, g
end

"s-blacklist of shopt.jl:166-171 / failed list of :384-386: 1-based indices with s < lo or s > hi, ascending"
function s_out_of_bounds(s::Vector{Float64}, lo::Real, hi::Real)
    idx = zeros(Int64, length(s))
    n = GC.@preserve s idx ccall((:shopt_s_out_of_bounds, LIBSHOPT), Int64, (Ptr{Float64}, Int64, Float64, Float64, Ptr{Int64}),
                                 s, length(s), lo, hi, idx)
    n < 0 && check(n)
    return idx[1:n] .+ 1
end

# ---- shopt.jl:281-314 (mode == "PCA") --------------------------------------------------------------------------------
"nanToZero -> pca_image(., PCAterms) -> smooth(., lanczos) (or ./ sum when unity_sum) per star; returns pixelGridFits, failedStars"
function pca_batch(stars, PCAterms::Integer, lanczos::Integer, unity_sum::Bool; pratio::Real = 0.99)
    X = cube(stars); r, c, N = size(X)
    pgf = similar(X); k, status = zeros(Int32, N), zeros(Int32, N)
    GC.@preserve X pgf k status check(ccall((:shopt_pca_batch, LIBSHOPT), Cint,
        (Ptr{Float64}, Cint, Cint, Int64, Cint, Float64, Cint, Cint, Ptr{Float64}, Ptr{Int32}, Ptr{Int32}),
        X, r, c, N, PCAterms, pratio, lanczos, unity_sum, pgf, k, status))
    return uncube(pgf), findall(status .== 1), k          # status 1: zeros(r, c) + failedStars (shopt.jl:305-309)
end

# ---- shopt.jl:316-348 (mode == "smooth") and :194-224 (mode == "chisq") ------------------------------------------------
function smooth_mode(stars, lanczos::Integer, unity_sum::Bool)
    X = cube(stars); out = similar(X); r, c, N = size(X)
    GC.@preserve X out check(ccall((:shopt_smooth_mode_batch, LIBSHOPT), Cint,
        (Ptr{Float64}, Cint, Cint, Int64, Cint, Cint, Ptr{Float64}), X, r, c, N, lanczos, unity_sum, out))
    return uncube(out)
end
"x0 = rand(r, c, N): the reference's start values, kept by the pixels that have no finite chi-square term"
function chisq_mode(stars, errVignets, unity_sum::Bool; x0::Array{Float64,3} = rand(size(stars[1])..., length(stars)))
    X = cube(stars); W = cube(errVignets); out = similar(X); r, c, N = size(X)
    GC.@preserve X W x0 out check(ccall((:shopt_chisq_batch, LIBSHOPT), Cint,
        (Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Float64, Cint, Cint, Int64, Cint, Ptr{Float64}),
        X, W, x0, 0.0, r, c, N, unity_sum, out))
    return uncube(out)
end

# ---- shopt.jl:492-577 (polynomial_optimizer per pixel, MSE, worst_10_percent, the drop loop) ---------------------------
"""
    polyfit(pixelGridFits, u, v, training_indices, degree, iterations; median_constraint, outlier_sigma_clipping)
        -> PolynomialMatrix (r, c, K), kept (1-based training stars after the last applied rejection), n_per_round
"""
function polyfit(pixelGridFits, u, v, training_indices::Vector{Int}, degree::Integer, iterations::Integer;
                 median_constraint::Bool = false, outlier_sigma_clipping::Bool = false)
    P = cube(pixelGridFits); r, c, N = size(P); K = (degree + 1) * (degree + 2) ÷ 2
    uu, vv = Vector{Float64}(u), Vector{Float64}(v)
    poly = zeros(r, c, K); idx = Int64.(training_indices .- 1)
    mse = zeros(length(idx)); kept = zeros(Int64, length(idx)); nkept = Ref{Int64}(0); npr = zeros(Int64, max(iterations, 1))
    GC.@preserve P uu vv idx poly mse kept npr check(ccall((:shopt_polyfit, LIBSHOPT), Cint,
        (Ptr{Float64}, Cint, Cint, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Int64}, Int64, Cint, Cint, Cint, Cint,
         Ptr{Float64}, Ptr{Float64}, Ptr{Int64}, Ref{Int64}, Ptr{Int64}),
        P, r, c, N, uu, vv, idx, length(idx), degree, iterations, median_constraint, outlier_sigma_clipping,
        poly, mse, kept, nkept, npr))
    return poly, Int.(kept[1:nkept[]] .+ 1), npr[1:iterations]
end

# ---- reader.jl:40-68 --------------------------------------------------------------------------------------------------
"p(u, v, polMatrix, degree) of reader.jl batched over M positions: (r, c, M), each stamp divided by its sum"
function psf_at(u::Vector{Float64}, v::Vector{Float64}, polMatrix::Array{Float64,3}, degree::Integer)
    r, c, _ = size(polMatrix); M = length(u); out = zeros(r, c, M)
    GC.@preserve polMatrix u v out check(ccall((:shopt_psf_at_batch, LIBSHOPT), Cint,
        (Ptr{Float64}, Cint, Cint, Cint, Ptr{Float64}, Ptr{Float64}, Int64, Ptr{Float64}),
        polMatrix, r, c, degree, u, v, M, out))
    return out
end

# ---- shopt.jl:422-466 (interpolate.jl:7-87) ------------------------------------------------------------------------------
"least-squares cubic s(u,v), g1(u,v), g2(u,v): the limit of the reference's three LBFGS runs from rand(10)"
function field_interp(u, v, s_data, g1_data, g2_data)
    uu, vv = Vector{Float64}(u), Vector{Float64}(v)
    a, b, d = Vector{Float64}(s_data), Vector{Float64}(g1_data), Vector{Float64}(g2_data)
    coef = zeros(10, 3)
    GC.@preserve uu vv a b d coef begin
        vals = [pointer(a), pointer(b), pointer(d)]
        rank = GC.@preserve vals ccall((:shopt_field_interp, LIBSHOPT), Cint,
            (Ptr{Float64}, Ptr{Float64}, Ptr{Ptr{Float64}}, Cint, Int64, Ptr{Float64}), uu, vv, vals, 3, length(uu), coef)
        rank < 0 && check(rank)
    end
    return coef[:, 1], coef[:, 2], coef[:, 3]              # sC, g1C, g2C in the reference's term order
end
function field_eval(coef::Vector{Float64}, u::Vector{Float64}, v::Vector{Float64})
    out = zeros(length(u))
    GC.@preserve coef u v out check(ccall((:shopt_field_eval, LIBSHOPT), Cint,
        (Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Int64, Ptr{Float64}), coef, u, v, length(u), out))
    return out
end

# ---- cataloging() without the FITS decoding (dataPreprocessing.jl:342-397) ---------------------------------------------
function ingest(vignets::Array{Float32,3}, unity_sum::Bool, new_img_dim::Integer, lanczos::Integer)
    r0, c0, N = size(vignets); nd = new_img_dim
    out = nd == r0 ? zeros(r0, c0, N) : zeros(nd, nd, N)
    GC.@preserve vignets out check(ccall((:shopt_ingest_batch, LIBSHOPT), Cint,
        (Ptr{Cvoid}, Cint, Cint, Cint, Int64, Cint, Cint, Cint, Ptr{Float64}), vignets, 1, r0, c0, N, unity_sum, nd, lanczos, out))
    return uncube(out)
end
"outliers_filter (outliers.jl:16-34): 1-based indices of the stars with snr < quantile(snr, k), and the cutoff"
function snr_cut(snr::Vector{Float32}, k::Real)
    idx = zeros(Int64, length(snr)); q1 = Ref{Float64}(0.0)
    n = GC.@preserve snr idx ccall((:shopt_snr_cut, LIBSHOPT), Int64, (Ptr{Float32}, Int64, Float64, Ptr{Int64}, Ref{Float64}),
                                   snr, length(snr), k, idx, q1)
    n < 0 && check(n)
    return idx[1:n] .+ 1, q1[]
end

# ---- summary.shopt (dataOutprocessing.jl:52-84 truncated, :116-168 full; reader.jl:13-38) -------------------------------
write_summary(path, PolynomialMatrix::Array{Float64,3}, degree, sC, g1C, g2C) =
    GC.@preserve PolynomialMatrix sC g1C g2C check(ccall((:shopt_summary_write, LIBSHOPT), Cint,
        (Cstring, Ptr{Float64}, Cint, Cint, Cint, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
        path, PolynomialMatrix, size(PolynomialMatrix, 1), size(PolynomialMatrix, 2), degree, sC, g1C, g2C))

# mirror of `struct shopt_summary_full` (include/shopt_b200.h)
struct SummaryFull
    poly::Ptr{Float64}; r::Cint; c::Cint; degree::Cint
    sC::Ptr{Float64}; g1C::Ptr{Float64}; g2C::Ptr{Float64}
    n_stars::Int64
    u::Ptr{Float64}; v::Ptr{Float64}; s_model::Ptr{Float64}; g1_model::Ptr{Float64}; g2_model::Ptr{Float64}
    s_data::Ptr{Float64}; g1_data::Ptr{Float64}; g2_data::Ptr{Float64}
    stars::Ptr{Float64}; pgf::Ptr{Float64}
    n_train::Int64; train_u::Ptr{Float64}; train_v::Ptr{Float64}
    n_valid::Int64; valid_u::Ptr{Float64}; valid_v::Ptr{Float64}; valid_stars::Ptr{Float64}
    n_flag_snr::Int64; n_flag_blacklist::Int64; n_flag_failed::Int64
    flag_snr::Ptr{Float64}; flag_blacklist::Ptr{Float64}; flag_failed::Ptr{Float64}
end
ptr_or_null(a) = isempty(a) ? Ptr{Float64}(C_NULL) : pointer(a)

"truncate_summary_file: false -- the seven-HDU layout with the star cubes (dataOutprocessing.jl:86-168)"
function write_summary_full(path, PolynomialMatrix::Array{Float64,3}, degree, sC, g1C, g2C, s_model, g1_model, g2_model,
                            s_data, g1_data, g2_data, u, v, starCatalog, pixelGridFits, training_u, training_v,
                            validation_stars, validation_u, validation_v, outlier_indices, s_blacklist, failedStars)
    V = cube(starCatalog); P = cube(pixelGridFits); VS = cube(validation_stars)
    f64(x) = Vector{Float64}(x)
    cols = map(f64, (u, v, s_model, g1_model, g2_model, s_data, g1_data, g2_data))
    tu, tv, vu, vv = f64(training_u), f64(training_v), f64(validation_u), f64(validation_v)
    fl = map(f64, (outlier_indices, s_blacklist, failedStars))
    sc, g1c, g2c = f64(sC), f64(g1C), f64(g2C)
    GC.@preserve PolynomialMatrix sc g1c g2c cols tu tv vu vv fl V P VS begin
        f = SummaryFull(pointer(PolynomialMatrix), size(PolynomialMatrix, 1), size(PolynomialMatrix, 2), degree,
                        pointer(sc), pointer(g1c), pointer(g2c), size(V, 3),
                        ptr_or_null(cols[1]), ptr_or_null(cols[2]), ptr_or_null(cols[3]), ptr_or_null(cols[4]),
                        ptr_or_null(cols[5]), ptr_or_null(cols[6]), ptr_or_null(cols[7]), ptr_or_null(cols[8]),
                        ptr_or_null(V), ptr_or_null(P), length(tu), ptr_or_null(tu), ptr_or_null(tv),
                        size(VS, 3), ptr_or_null(vu), ptr_or_null(vv), ptr_or_null(VS),
                        length(fl[1]), length(fl[2]), length(fl[3]), ptr_or_null(fl[1]), ptr_or_null(fl[2]), ptr_or_null(fl[3]))
        check(ccall((:shopt_summary_write_full, LIBSHOPT), Cint, (Cstring, Ref{SummaryFull}), path, f))
    end
end

"read_shopt of reader.jl:13-38 without astropy: PolynomialMatrix, degree, sC, g1C, g2C"
function read_summary(path)
    r, c, deg = Ref{Cint}(0), Ref{Cint}(0), Ref{Cint}(0); sC, g1C, g2C = zeros(10), zeros(10), zeros(10)
    check(ccall((:shopt_summary_read, LIBSHOPT), Cint,
        (Cstring, Ptr{Float64}, Int64, Ref{Cint}, Ref{Cint}, Ref{Cint}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
        path, C_NULL, 0, r, c, deg, sC, g1C, g2C))
    K = (deg[] + 1) * (deg[] + 2) ÷ 2; poly = zeros(r[], c[], K)
    GC.@preserve poly check(ccall((:shopt_summary_read, LIBSHOPT), Cint,
        (Cstring, Ptr{Float64}, Int64, Ref{Cint}, Ref{Cint}, Ref{Cint}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
        path, poly, length(poly), r, c, deg, sC, g1C, g2C))
    return poly, Int(deg[]), sC, g1C, g2C
end

# ---- the whole stage in one call (shopt.jl:130-577): everything stays on the device between the loops ----------------------
# Field order and types mirror `struct shopt_params` / `struct shopt_stage_io` of include/shopt_b200.h (isbits, C layout).
Base.@kwdef mutable struct StageParams
    n_a::Cint = 64                       # AnalyticFitParams.analyticFitStampSize
    g_tol_model::Cdouble = 1e-6          # minGradientAnalyticModel
    g_tol_learned::Cdouble = 1e-6        # minGradientAnalyticLearned
    max_iter::Cint = 1000
    s_lo::Cdouble = 0.075                # dataProcessing.sLowerBound
    s_hi::Cdouble = 1.0                  # sUpperBound
    pca_terms::Cint = 50                 # NNparams.PCAterms
    pratio::Cdouble = 0.99
    lanczos::Cint = 3
    unity_sum::Cint = 0                  # sum_pixel_grid_and_inputs_to_unity
    degree::Cint = 3                     # polynomialDegree
    iterations::Cint = 3                 # iterations
    training_ratio::Cdouble = 0.8
    median_constraint::Cint = 0
    sigma_clip::Cint = 0                 # outlier_sigma_clipping
    mode::Cint = 0                       # 0 "PCA", 1 "smooth", 2 "chisq"
    chisq_fill::Cdouble = 0.0
    profile::Cint = 0                    # 0 fGaussian (what the reference calls), 1 fkolmogorov (radialProfiles.jl:16-24)
end
const MODES = Dict("PCA" => 0, "smooth" => 1, "chisq" => 2)
"StageParams from the shopt.yml globals of dataPreprocessing.jl:4-38"
default_params(; kw...) = StageParams(; kw...)

mutable struct StageIO
    stamps::Ptr{Float64}; r::Cint; c::Cint; n_local::Int64
    u::Ptr{Float64}; v::Ptr{Float64}; x0_model::Ptr{Float64}; x0_learned::Ptr{Float64}
    n_global::Int64; offset::Int64; perm::Ptr{Int64}
    model::Ptr{Float64}; learned::Ptr{Float64}; pgf::Ptr{Float64}; flags::Ptr{Int32}; pca_k::Ptr{Int32}
    poly::Ptr{Float64}; n_per_round::Ptr{Int64}; counts::NTuple{4,Int64}
    weights::Ptr{Float64}; x0_pixels::Ptr{Float64}
    field::NTuple{30,Float64}; field_rank::Int32
    polyfit_rank::Int32; polyfit_info::Int32; stamps_f32::Int32
    perm_dev::Ptr{Int64}
    model_status::Ptr{Int32}; model_iters::Ptr{Int32}; learned_status::Ptr{Int32}; learned_iters::Ptr{Int32}
end

"""
    stage_run(p, stars, u, v; x0_model, x0_learned, perm, errVignets = nothing, want_pgf = true) -> NamedTuple

Model fit -> s-blacklist -> pixel-grid fit -> learned fit -> failed list -> s/g1/g2 field fit -> training split ->
polyfit rounds, on ONE GPU with host arrays (`shopt_stage_run`).  `stars` is the Vector{Matrix{Float64}} catalogue or a
Float32 (r, c, N) VIGNET cube (uploaded as such and ingested on the device).  Stars keep their ORIGINAL indices: the
splices of shopt.jl:177-184, 405-417 become `findall(flags .& 0x7 .== 0)`.
"""
function stage_run(p::StageParams, stars, u, v; x0_model = rand(3, nstars(stars)), x0_learned = rand(3, nstars(stars)),
                   perm = Random_randperm(nstars(stars)), errVignets = nothing, x0_pixels = nothing, want_pgf::Bool = true)
    f32 = stars isa Array{Float32,3}
    X = f32 ? stars : cube(stars); r, c, N = size(X); K = (p.degree + 1) * (p.degree + 2) ÷ 2
    uu, vv = Vector{Float64}(u), Vector{Float64}(v); perm0 = Int64.(perm .- 1)
    model, learned = zeros(N, 3), zeros(N, 3); flags, pca_k = zeros(Int32, N), zeros(Int32, N)
    pgf = want_pgf ? zeros(r, c, N) : zeros(0, 0, 0); poly = zeros(r, c, K); npr = zeros(Int64, max(p.iterations, 1))
    W = errVignets === nothing ? zeros(0, 0, 0) : cube(errVignets)
    X0 = x0_pixels === nothing ? zeros(0, 0, 0) : x0_pixels
    mst, mit, lst, lit = zeros(Int32, N), zeros(Int32, N), zeros(Int32, N), zeros(Int32, N)
    GC.@preserve X uu vv x0_model x0_learned perm0 model learned flags pca_k pgf poly npr W X0 mst mit lst lit begin
        io = StageIO(Ptr{Float64}(pointer(X)), r, c, N, pointer(uu), pointer(vv), pointer(x0_model), pointer(x0_learned),
                     N, 0, pointer(perm0), pointer(model), pointer(learned), want_pgf ? pointer(pgf) : C_NULL,
                     pointer(flags), pointer(pca_k), pointer(poly), pointer(npr), (0, 0, 0, 0),
                     isempty(W) ? C_NULL : pointer(W), isempty(X0) ? C_NULL : pointer(X0),
                     ntuple(_ -> 0.0, 30), 0, 0, 0, f32 ? 1 : 0, C_NULL, pointer(mst), pointer(mit), pointer(lst), pointer(lit))
        check(ccall((:shopt_stage_run, LIBSHOPT), Cint, (Ref{StageParams}, Ref{StageIO}), p, io))
        fld = collect(io.field)
        return (s_model = model[:, 1], g1_model = model[:, 2], g2_model = model[:, 3],
                s_data = learned[:, 1], g1_data = learned[:, 2], g2_data = learned[:, 3],
                flags = flags, pca_k = pca_k, pixelGridFits = want_pgf ? uncube(pgf) : nothing, PolynomialMatrix = poly,
                sC = fld[1:10], g1C = fld[11:20], g2C = fld[21:30], n_per_round = npr[1:p.iterations],
                counts = (blacklisted = io.counts[1], failed = io.counts[2], training = io.counts[3], kept = io.counts[4]),
                model_status = mst, learned_status = lst, polyfit_rank = io.polyfit_rank, polyfit_info = io.polyfit_info)
    end
end
nstars(stars) = stars isa AbstractArray{<:Real,3} ? size(stars, 3) : length(stars)
Random_randperm(n) = Base.invokelatest(getfield(Base.require(Base.PkgId(Base.UUID("9a3f8284-a2c9-5f02-9a11-845980a1fd5c"), "Random")), :randperm), n)

end # module
