# MDB200.jl -- Julia glue for libmdb200.so: `ccall` wrappers + `ChainRulesCore.rrule`s that put the
# B200 kernels behind the reference's own function signatures (jumerckx/Monodepth2.jl).
#
# NOT EXECUTED on the build / GPU boxes of this repository: neither has a Julia runtime (SURVEY F2).
# The same C ABI is exercised end to end from Python (monodepth2.jl_b200/mdb200, tests/).  This file
# is the binding a maintainer of the reference adds -- `include("MDB200.jl")` as the LAST include of
# src/Monodepth.jl:62-74 (after render.jl / simple_depth.jl, so every function named below exists).
#
# Dispatch: the functions are IMPORTED from the enclosing module (`import ..Monodepth: ...`), so the definitions
# below are new METHODS of the reference's own functions, specialised on `CuArray{Float32}`; the reference's call
# sites (`train_loss`, `slow_depth`, `find_static`, the test suite) reach them through ordinary dispatch, and every
# other argument type (CPU `Array`s, Float64) keeps running the reference's pure-Julia methods.  Three names have no
# generic function to extend in the fork and are DEFINED here for it: `warp` (called at src/simple_depth.jl:30-32
# but missing, SURVEY F4), and the in-place-mutating `sample` / non-differentiable `render_tgt_rgb_depth`, whose
# CuArray methods are replaced outright (same signature).
#
# Layout note: Julia's column-major (W,H,C,L,N) CuArray is passed as-is; the library reads the same
# bytes as C order [N][L][C][H][W].  `K`/`invK` are passed as 9 host floats in column-major order.
module MDB200

using CUDA
using ChainRulesCore
import ..Monodepth: Pose, Params, TrainCache, SSIM, Backproject, Project
import ..Monodepth: hat, so3_exp_map, composeT, disparity_to_depth, smooth_loss,
                    photometric_loss, prediction_loss, automasking_loss, _apply_mask, train_loss,
                    inv, create_meshgrid, get_src_xyz_from_plane_disparity, get_tgt_xyz_from_plane_disparity,
                    sample, plane_volume_rendering, render_tgt_rgb_depth, slow_depth, save_disparity, cpu
using Statistics: mean

const libmdb200 = get(ENV, "MDB200_LIB", joinpath(@__DIR__, "..", "mdb200", "libmdb200.so"))

const MAX_SCALES = 8
const MAX_SOURCES = 2

# ---- mirrors of the C structs in include/mdb200.h -----------------------------------------------
struct LossParams
    min_depth::Cdouble
    max_depth::Cdouble
    disparity_smoothness::Cdouble
    alpha::Cfloat
    ssim_c1::Cfloat
    ssim_c2::Cfloat
    automasking::Int32
    normalize_disparity::Int32
    smooth_weight_mode::Int32
    upsample_align_corners::Int32
    n_global::Int32
    abs_subgradient::Int32       # 0: sign(0) = 0 (ChainRules); 1: ForwardDiff's Dual rule, what Zygote's CuArray broadcast uses
end

struct LossIO
    x::CuPtr{Cfloat}
    W::Int32; H::Int32; C::Int32; L::Int32; N::Int32
    target_id::Int32
    S::Int32
    source_ids::NTuple{MAX_SOURCES,Int32}
    invert::NTuple{MAX_SOURCES,Int32}
    rvec::NTuple{MAX_SOURCES,CuPtr{Cfloat}}
    tvec::NTuple{MAX_SOURCES,CuPtr{Cfloat}}
    Ls::Int32
    disparity::NTuple{MAX_SCALES,CuPtr{Cfloat}}
    dw::NTuple{MAX_SCALES,Int32}
    dh::NTuple{MAX_SCALES,Int32}
    scale::NTuple{MAX_SCALES,Cfloat}
    auto_loss::CuPtr{Cfloat}
    K::NTuple{9,Cfloat}
    invK::NTuple{9,Cfloat}
    loss::CuPtr{Cfloat}
    d_disparity::NTuple{MAX_SCALES,CuPtr{Cfloat}}
    d_rvec::NTuple{MAX_SOURCES,CuPtr{Cfloat}}
    d_tvec::NTuple{MAX_SOURCES,CuPtr{Cfloat}}
    vis_warp_loss::CuPtr{Cfloat}
    x_format::Int32          # 0: Float32 frames; 1: UInt8 (N0f8) frames, same layout
end

# mdb_host_io: every pointer is a HOST pointer (mdb_train_loss_host)
struct HostIO
    x::Ptr{Cvoid}
    flip::Ptr{UInt8}
    disparity::NTuple{MAX_SCALES,Ptr{Cfloat}}
    rvec::NTuple{MAX_SOURCES,Ptr{Cfloat}}
    tvec::NTuple{MAX_SOURCES,Ptr{Cfloat}}
    auto_loss::Ptr{Cfloat}
    loss::Ptr{Cfloat}
    d_disparity::NTuple{MAX_SCALES,Ptr{Cfloat}}
    d_rvec::NTuple{MAX_SOURCES,Ptr{Cfloat}}
    d_tvec::NTuple{MAX_SOURCES,Ptr{Cfloat}}
end

# ---- handle + error convention (status codes -> Julia errors) ------------------------------------
const HANDLE = Ref{Ptr{Cvoid}}(C_NULL)

function handle()
    if HANDLE[] == C_NULL
        rc = ccall((:mdb_create, libmdb200), Int32, (Ref{Ptr{Cvoid}}, Int32), HANDLE, CUDA.deviceid(CUDA.device()))
        rc == 0 || error("mdb_create failed with status $rc (needs an sm_100 device; there is no CPU fallback)")
    end
    HANDLE[]
end

function check(rc::Int32)
    rc == 0 && return nothing
    msg = unsafe_string(ccall((:mdb_last_error, libmdb200), Cstring, (Ptr{Cvoid},), handle()))
    error("mdb200 error $rc: $msg")
end

stream_ptr() = Base.unsafe_convert(Ptr{Cvoid}, CUDA.stream().handle)
pad(t, n, fill) = ntuple(i -> i <= length(t) ? t[i] : fill, n)
k9(K) = NTuple{9,Cfloat}(Cfloat.(vec(Array(K))))           # column-major, as Julia stores it
nullp() = CuPtr{Cfloat}(0)

# ---- fine-grained entry points (same names as the reference) -------------------------------------
function so3_exp_map(rvec::CuArray{Float32,2})              # src/utils.jl:106-121
    N = size(rvec, 2)
    R = CUDA.zeros(Float32, 3, 3, N)
    GC.@preserve rvec R check(ccall((:mdb_so3_exp_map, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, Int32, CuPtr{Cfloat}, Ptr{Cvoid}), handle(), rvec, N, R, stream_ptr()))
    R
end

function composeT(rvec::CuArray{Float32,2}, t::CuArray{Float32,3}, invert::Bool)   # src/utils.jl:185-192
    N = size(rvec, 2)
    R = CUDA.zeros(Float32, 3, 3, N); to = CUDA.zeros(Float32, 3, 1, N)
    GC.@preserve rvec t R to check(ccall((:mdb_compose_T, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), rvec, t, Int32(invert), N, R, to, stream_ptr()))
    R, to
end

function ChainRulesCore.rrule(::typeof(composeT), rvec::CuArray{Float32,2}, t::CuArray{Float32,3}, invert::Bool)
    R, to = composeT(rvec, t, invert)
    function composeT_pullback(Δ)
        dR, dt = unthunk(Δ[1]), unthunk(Δ[2])
        N = size(rvec, 2)
        dr = CUDA.zeros(Float32, 3, N); dtv = CUDA.zeros(Float32, 3, 1, N)
        dRc = dR isa AbstractZero ? CUDA.zeros(Float32, 3, 3, N) : CuArray{Float32}(dR)
        dtc = dt isa AbstractZero ? CUDA.zeros(Float32, 3, 1, N) : CuArray{Float32}(dt)
        GC.@preserve rvec t dRc dtc dr dtv check(ccall((:mdb_compose_T_bwd, libmdb200), Int32,
            (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cvoid}),
            handle(), rvec, t, Int32(invert), N, dRc, dtc, dr, dtv, stream_ptr()))
        NoTangent(), dr, dtv, NoTangent()
    end
    (R, to), composeT_pullback
end

function (ssim::SSIM)(x::CuArray{Float32,4}, y::CuArray{Float32,4})               # src/utils.jl:29-43
    W, H, C, N = size(x)
    out = similar(x)
    GC.@preserve x y out check(ccall((:mdb_ssim, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, Int32, Int32, Cfloat, Cfloat, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), x, y, W, H, C, N, Cfloat(ssim.c1), Cfloat(ssim.c2), out, stream_ptr()))
    out
end

# automasking_loss(ssim, inputs, target; source_ids) -- src/training.jl:9-11, the reference's signature: `target` is the
# target ARRAY (src/dtk.jl:59-61 passes x[:, :, :, target_id, :], a dense copy)
function automasking_loss(ssim::SSIM, inputs::CuArray{Float32,5}, target::CuArray{Float32,4}; source_ids)
    W, H, C, L, N = size(inputs)
    out = CUDA.zeros(Float32, W, H, 1, N)
    ids = Int32.(collect(source_ids))
    GC.@preserve inputs target out ids check(ccall((:mdb_automasking_loss, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int64, Ptr{Int32}, Int32, Int32, Int32, Int32, Int32, Int32, Cfloat, Cfloat, Cfloat,
         CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), inputs, target, Int64(0), ids, length(ids), W, H, C, L, N, 0.85f0, Cfloat(ssim.c1), Cfloat(ssim.c2), out, stream_ptr()))
    out
end

include("MDB200_ops.jl")      # the other fine-grained operators and their rrules

# ---- the fused hot path ---------------------------------------------------------------------------
"""
    fused_loss(x, disparities, poses, auto_loss, cache, parameters; kw...) -> (loss, ∇disparities, ∇poses)

Value and gradient of `train_loss`'s arithmetic (src/training.jl:27-77) in one pass of the fused
kernels.  `n_global` is the batch size the means divide by (multi-GPU shards pass the global N).
"""
# `x` may also be the UInt8 (N0f8) stack of the same shape (`depth10k_batch_u8`): the kernel converts on the fly, bit-identically
function fused_loss(x::Union{CuArray{Float32,5},CuArray{UInt8,5}}, disparities::Vector{<:CuArray{Float32,4}}, poses::Vector{<:Pose},
                    auto_loss, cache::TrainCache, parameters::Params;
                    normalize_disparity::Bool=true, smooth_weight_mode::Integer=0, n_global::Integer=0,
                    abs_subgradient::Integer=0, packed_send::Union{Nothing,CuArray{Float32,1}}=nothing, first_item::Integer=0)
    W, H, C, L, N = size(x)
    S, Ls = length(poses), length(disparities)
    loss = CUDA.zeros(Float32, 1)
    dd = [similar(d) for d in disparities]
    dr = [similar(p.rvec) for p in poses]
    dt = [similar(p.tvec) for p in poses]
    io = LossIO(reinterpret(CuPtr{Cfloat}, pointer(x)), W, H, C, L, N, cache.target_id, S,
        pad(Int32.(cache.source_ids), MAX_SOURCES, Int32(0)),
        pad(Int32.(cache.source_ids .< cache.target_id), MAX_SOURCES, Int32(0)),   # src/training.jl:28
        pad(map(p -> pointer(p.rvec), poses), MAX_SOURCES, nullp()),
        pad(map(p -> pointer(p.tvec), poses), MAX_SOURCES, nullp()),
        Ls, pad(map(pointer, disparities), MAX_SCALES, nullp()),
        pad(Int32.(map(d -> size(d, 1), disparities)), MAX_SCALES, Int32(0)),
        pad(Int32.(map(d -> size(d, 2), disparities)), MAX_SCALES, Int32(0)),
        pad(Cfloat.(cache.scales), MAX_SCALES, 0f0),
        parameters.automasking ? pointer(auto_loss) : nullp(),
        k9(cache.K), k9(cache.invK), pointer(loss),
        pad(map(pointer, dd), MAX_SCALES, nullp()), pad(map(pointer, dr), MAX_SOURCES, nullp()),
        pad(map(pointer, dt), MAX_SOURCES, nullp()), nullp(), Int32(eltype(x) === UInt8 ? 1 : 0))
    p = LossParams(parameters.min_depth, parameters.max_depth, parameters.disparity_smoothness, 0.85f0,
        Cfloat(cache.ssim.c1), Cfloat(cache.ssim.c2), Int32(parameters.automasking), Int32(normalize_disparity),
        Int32(smooth_weight_mode), Int32(1), Int32(n_global), Int32(abs_subgradient))
    nbytes = ccall((:mdb_train_loss_workspace_bytes, libmdb200), Csize_t, (Ref{LossIO},), io)
    nbytes > 0 || error("mdb200: unsupported shape for the fused loss")
    ws = CuArray{UInt8}(undef, nbytes)
    ioref = Ref(io)
    if packed_send !== nothing
        # a shard of a batch of n_global items: loss and pose gradients are written straight into this rank's slots of the
        # packed all-reduce buffer (packed_size(n_global, S) floats, the other slots zero); `allreduce_loss_posegrad!` is next
        n_global > 0 || error("fused_loss: packed_send needs n_global")
        GC.@preserve packed_send check(ccall((:mdb_packed_bind, libmdb200), Int32, (Ref{LossIO}, CuPtr{Cfloat}, Int32, Int32),
            ioref, packed_send, n_global, first_item))
    end
    GC.@preserve x disparities poses auto_loss loss dd dr dt ws packed_send begin
        check(ccall((:mdb_train_loss_fwdbwd, libmdb200), Int32,
            (Ptr{Cvoid}, Ref{LossIO}, Ref{LossParams}, CuPtr{Cvoid}, Csize_t, Ptr{Cvoid}),
            handle(), ioref, p, ws, nbytes, stream_ptr()))
    end
    loss, dd, [(rvec=dr[i], tvec=dt[i]) for i in 1:S]
end

# ---- the same step for a batch that is still in HOST memory ---------------------------------------------------------------
# The reference moves a batch with `transfer(x)` = `gpu ∘ f32` (src/utils.jl:1-3) and then calls train_loss; here the copy-in,
# the fused loss and the copy-out of the gradients are one C call that pipelines chunks of items over three streams.
"""
    HostPlan(W, H, C, N, disp_sizes, cache, parameters; x_u8_strips = false, chunk_items = 0, first_item = -1, kw...)

Owns the device buffers of `fused_loss_host` for one batch shape (`cudaMalloc` once, at construction).  `x_u8_strips`:
`x` is the dataset's UInt8 triplet strips `(3, 3W, H, N)` instead of the Float32 stack `(W,H,C,3,N)`.  `first_item >= 0`:
the batch is a shard of `n_global` items starting there (see `comm_init` / `peers_attach` in MDB200_ops.jl).
"""
mutable struct HostPlan
    ptr::Ptr{Cvoid}
    W::Int; H::Int; C::Int; N::Int; S::Int
    disp_sizes::Vector{Tuple{Int,Int}}
    n_pose::Int
end

function HostPlan(W::Integer, H::Integer, C::Integer, N::Integer, disp_sizes, cache::TrainCache, parameters::Params;
                  x_u8_strips::Bool = false, chunk_items::Integer = 0, first_item::Integer = -1,
                  normalize_disparity::Bool = true, smooth_weight_mode::Integer = 0, n_global::Integer = 0,
                  abs_subgradient::Integer = 0)
    S, Ls = length(cache.source_ids), length(disp_sizes)
    shape = LossIO(nullp(), W, H, C, 3, N, cache.target_id, S,
        pad(Int32.(cache.source_ids), MAX_SOURCES, Int32(0)), pad(Int32.(cache.source_ids .< cache.target_id), MAX_SOURCES, Int32(0)),
        pad((), MAX_SOURCES, nullp()), pad((), MAX_SOURCES, nullp()), Ls, pad((), MAX_SCALES, nullp()),
        pad(Int32.(first.(disp_sizes)), MAX_SCALES, Int32(0)), pad(Int32.(last.(disp_sizes)), MAX_SCALES, Int32(0)),
        pad(Cfloat.(cache.scales), MAX_SCALES, 0f0), nullp(), k9(cache.K), k9(cache.invK), nullp(),
        pad((), MAX_SCALES, nullp()), pad((), MAX_SOURCES, nullp()), pad((), MAX_SOURCES, nullp()), nullp(), Int32(0))
    p = LossParams(parameters.min_depth, parameters.max_depth, parameters.disparity_smoothness, 0.85f0,
        Cfloat(cache.ssim.c1), Cfloat(cache.ssim.c2), Int32(parameters.automasking), Int32(normalize_disparity),
        Int32(smooth_weight_mode), Int32(1), Int32(n_global), Int32(abs_subgradient))
    out = Ref{Ptr{Cvoid}}(C_NULL)
    check(ccall((:mdb_host_plan_create, libmdb200), Int32,
        (Ptr{Cvoid}, Ref{LossIO}, Ref{LossParams}, Int32, Int32, Int32, Ref{Ptr{Cvoid}}),
        handle(), shape, p, chunk_items, x_u8_strips ? 1 : 0, first_item, out))
    info = zeros(Int64, 4)
    check(ccall((:mdb_host_plan_info, libmdb200), Int32, (Ptr{Cvoid}, Ptr{Int64}), out[], info))
    plan = HostPlan(out[], W, H, C, N, S, [(Int(a), Int(b)) for (a, b) in disp_sizes], Int(info[4]))
    finalizer(pl -> (pl.ptr == C_NULL || ccall((:mdb_host_plan_destroy, libmdb200), Int32, (Ptr{Cvoid},), pl.ptr); pl.ptr = C_NULL), plan)
    plan
end

"""
    fused_loss_host(plan, x, disparities, poses, auto_loss; flips = nothing) -> (loss, ∇disparities, ∇poses)   (host Arrays)

`x::Array{Float32,5}` (or the UInt8 strips the plan was built for), `disparities::Vector{Array{Float32,4}}`,
`poses::Vector` of `(rvec (3,N), tvec (3,1,N))` host arrays.  Pinned arrays (`CUDA.pin`) reach the PCIe rate.
"""
function fused_loss_host(plan::HostPlan, x::Array, disparities::Vector{<:Array{Float32}}, poses::Vector, auto_loss;
                         flips::Union{Nothing,Vector{UInt8}} = nothing)
    loss = zeros(Float32, 1)
    dd = [similar(d) for d in disparities]
    dr = [zeros(Float32, 3, plan.n_pose) for _ in 1:plan.S]; dt = [zeros(Float32, 3, 1, plan.n_pose) for _ in 1:plan.S]
    hp(a) = Ptr{Cfloat}(pointer(a)); hnull = Ptr{Cfloat}(C_NULL)
    io = HostIO(Ptr{Cvoid}(pointer(x)), flips === nothing ? Ptr{UInt8}(C_NULL) : pointer(flips),
        pad(map(hp, disparities), MAX_SCALES, hnull), pad(map(p -> hp(p.rvec), poses), MAX_SOURCES, hnull),
        pad(map(p -> hp(p.tvec), poses), MAX_SOURCES, hnull), auto_loss === nothing ? hnull : hp(auto_loss), hp(loss),
        pad(map(hp, dd), MAX_SCALES, hnull), pad(map(hp, dr), MAX_SOURCES, hnull), pad(map(hp, dt), MAX_SOURCES, hnull))
    GC.@preserve x disparities poses auto_loss flips loss dd dr dt check(ccall((:mdb_train_loss_host, libmdb200), Int32,
        (Ptr{Cvoid}, Ref{HostIO}, Ptr{Cvoid}), plan.ptr, io, stream_ptr()))
    loss[1], dd, [(rvec = dr[i], tvec = dt[i]) for i in 1:plan.S]
end

# The scalar loss is the only output, so the rrule's forward runs forward+backward with unit seed and
# the pullback scales the cached gradients by the incoming cotangent (SURVEY 8b).
# (Zygote needs the scalar on the host: this read is the step's one synchronisation, as `loss` is in the reference)
loss_value(x, disparities, poses, auto_loss, cache, parameters) =
    Array(fused_loss(x, disparities, poses, auto_loss, cache, parameters)[1])[1]

function ChainRulesCore.rrule(::typeof(loss_value), x, disparities, poses, auto_loss, cache, parameters)
    loss, dd, dp = fused_loss(x, disparities, poses, auto_loss, cache, parameters)
    function loss_value_pullback(Δ)
        δ = Float32(unthunk(Δ))
        NoTangent(), NoTangent(), [δ .* g for g in dd],
        [Tangent{Pose}(rvec=δ .* g.rvec, tvec=δ .* g.tvec) for g in dp], NoTangent(), NoTangent(), NoTangent()
    end
    Array(loss)[1], loss_value_pullback
end

"""
Drop-in `train_loss` (src/training.jl:21-78): same arguments and return tuple; only lines 27-69 of
the reference body are replaced by `loss_value`.
"""
function train_loss(model, x::CuArray{Float32,5}, auto_loss, cache::TrainCache, parameters::Params, do_visualization)
    disparities, poses = model(x, cache.source_ids, cache.target_id)           # src/training.jl:26
    loss = loss_value(x, disparities, poses, auto_loss, cache, parameters)
    vis_disparity = do_visualization ? Array(disparities[end]) : nothing
    loss, vis_disparity, nothing, nothing
end

# ---- render path ------------------------------------------------------------------------------------
function render_tgt_rgb_depth(rgb::CuArray{Float32,5}, sigma::CuArray{Float32,5}, disparity_src::CuArray{Float32,2},
                              xyz_tgt::CuArray{Float32,5}, pose::Pose, K_inv, K)            # src/render.jl:92-114
    W, H, _, Np, B = size(rgb)
    rgb_out = CUDA.zeros(Float32, W, H, 3, B); depth = CUDA.zeros(Float32, W, H, 1, Np, B)
    mask = CUDA.zeros(Int32, W, H, 1, 1, B)
    nbytes = ccall((:mdb_render_workspace_bytes, libmdb200), Csize_t, (Int32, Int32, Int32, Int32), W, H, Np, B)
    ws = CuArray{UInt8}(undef, nbytes)
    Ki, Kk = collect(k9(K_inv)), collect(k9(K))
    GC.@preserve rgb sigma disparity_src xyz_tgt pose rgb_out depth mask ws check(ccall((:mdb_render_tgt_rgb_depth_fwd, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cfloat}, Ptr{Cfloat},
         Int32, Int32, Int32, Int32, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Int32}, CuPtr{Cvoid}, Csize_t, Ptr{Cvoid}),
        handle(), rgb, sigma, disparity_src, xyz_tgt, pose.rvec, pose.tvec, Ki, Kk, W, H, Np, B, rgb_out, depth, mask, ws, nbytes, stream_ptr()))
    rgb_out, depth, mask
end

function ChainRulesCore.rrule(::typeof(render_tgt_rgb_depth), rgb, sigma, disparity_src, xyz_tgt, pose::Pose, K_inv, K)
    rgb_out, depth, mask = render_tgt_rgb_depth(rgb, sigma, disparity_src, xyz_tgt, pose, K_inv, K)
    function render_pullback(Δ)
        W, H, _, Np, B = size(rgb)
        g_rgb = Δ[1] isa AbstractZero ? CUDA.zeros(Float32, size(rgb_out)) : CuArray{Float32}(unthunk(Δ[1]))
        g_depth = Δ[2] isa AbstractZero ? CUDA.zeros(Float32, size(depth)) : CuArray{Float32}(unthunk(Δ[2]))
        d_rgb, d_sigma, d_xyz = similar(rgb), similar(sigma), similar(xyz_tgt)
        d_disp, d_r, d_t = similar(disparity_src), similar(pose.rvec), similar(pose.tvec)
        nbytes = ccall((:mdb_render_workspace_bytes, libmdb200), Csize_t, (Int32, Int32, Int32, Int32), W, H, Np, B)
        ws = CuArray{UInt8}(undef, nbytes)
        Ki, Kk = collect(k9(K_inv)), collect(k9(K))
        GC.@preserve rgb sigma disparity_src xyz_tgt pose depth g_rgb g_depth d_rgb d_sigma d_xyz d_disp d_r d_t ws check(ccall(
            (:mdb_render_tgt_rgb_depth_bwd, libmdb200), Int32,
            (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cfloat}, Ptr{Cfloat},
             Int32, Int32, Int32, Int32, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat},
             CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cvoid}, Csize_t, Ptr{Cvoid}),
            handle(), rgb, sigma, disparity_src, xyz_tgt, pose.rvec, pose.tvec, Ki, Kk, W, H, Np, B, depth, g_rgb, g_depth,
            d_rgb, d_sigma, d_xyz, d_disp, d_r, d_t, ws, nbytes, stream_ptr()))
        NoTangent(), d_rgb, d_sigma, d_disp, d_xyz, Tangent{Pose}(rvec=d_r, tvec=d_t), NoTangent(), NoTangent()
    end
    (rgb_out, depth, mask), render_pullback
end

# ---- SURVEY 8(f) rank 2: render variants (mdb_render_opts) ------------------------------------------
struct RenderOpts          # mdb_render_opts
    mine_grid::Int32       # grid = (x + 0.5) / (W/2) - 1, align_corners=false   (test/test_sample.jl:131-137)
    mine_mask::Int32       # valid = -1 < x < W                                    (test/test_sample.jl:123-125)
    depth_eps::Cfloat      # expected depth = sum w z / (sum w + depth_eps)        (src/render.jl:47 TODO)
    reserved::Int32
end

"""
    render_tgt_rgb_depth_ex(rgb, sigma, disparity_src, xyz_tgt_or_nothing, pose, K_inv, K;
                            mine_grid=false, mine_mask=false, depth_eps=1f-5)

-> (rgb, transparency_acc, mask, expected_depth, weights_sum).  `xyz_tgt === nothing` folds
`get_src_xyz_from_plane_disparity` / `get_tgt_xyz_from_plane_disparity` (src/render.jl:25-30, 51-62) into the
kernels.  The rrule below calls `mdb_render_tgt_rgb_depth_bwd_ex` with the saved outputs.
"""
function render_tgt_rgb_depth_ex(rgb::CuArray{Float32,5}, sigma::CuArray{Float32,5}, disparity_src::CuArray{Float32,2},
                                 xyz_tgt::Union{Nothing,CuArray{Float32,5}}, pose::Pose, K_inv, K;
                                 mine_grid::Bool=false, mine_mask::Bool=false, depth_eps::Real=1f-5)
    W, H, _, Np, B = size(rgb)
    rgb_out = CUDA.zeros(Float32, W, H, 3, B); depth = CUDA.zeros(Float32, W, H, 1, Np, B)
    mask = CUDA.zeros(Int32, W, H, 1, 1, B)
    zexp = CUDA.zeros(Float32, W, H, 1, B); wsum = CUDA.zeros(Float32, W, H, 1, B)
    nbytes = ccall((:mdb_render_workspace_bytes, libmdb200), Csize_t, (Int32, Int32, Int32, Int32), W, H, Np, B)
    ws = CuArray{UInt8}(undef, nbytes)
    Ki, Kk = collect(k9(K_inv)), collect(k9(K))
    opts = Ref(RenderOpts(mine_grid, mine_mask, depth_eps, 0))
    xp = xyz_tgt === nothing ? CuPtr{Cfloat}(0) : pointer(xyz_tgt)
    GC.@preserve rgb sigma disparity_src xyz_tgt pose rgb_out depth mask zexp wsum ws opts check(ccall(
        (:mdb_render_tgt_rgb_depth_fwd_ex, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cfloat}, Ptr{Cfloat},
         Int32, Int32, Int32, Int32, Ptr{RenderOpts}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Int32}, CuPtr{Cfloat}, CuPtr{Cfloat},
         CuPtr{Cvoid}, Csize_t, Ptr{Cvoid}),
        handle(), rgb, sigma, disparity_src, xp, pose.rvec, pose.tvec, Ki, Kk, W, H, Np, B, opts, rgb_out, depth, mask, zexp, wsum,
        ws, nbytes, stream_ptr()))
    rgb_out, depth, mask, zexp, wsum
end

function ChainRulesCore.rrule(::typeof(render_tgt_rgb_depth_ex), rgb, sigma, disparity_src, xyz_tgt, pose::Pose, K_inv, K;
                              mine_grid::Bool=false, mine_mask::Bool=false, depth_eps::Real=1f-5)
    outs = render_tgt_rgb_depth_ex(rgb, sigma, disparity_src, xyz_tgt, pose, K_inv, K; mine_grid, mine_mask, depth_eps)
    rgb_out, depth, mask, zexp, wsum = outs
    function render_ex_pullback(Δ)
        W, H, _, Np, B = size(rgb)
        cot(i, like) = Δ[i] isa AbstractZero ? CUDA.zeros(Float32, size(like)) : CuArray{Float32}(unthunk(Δ[i]))
        g_rgb, g_depth, g_z = cot(1, rgb_out), cot(2, depth), cot(4, zexp)
        d_rgb, d_sigma = similar(rgb), similar(sigma)
        d_xyz = xyz_tgt === nothing ? nothing : similar(xyz_tgt)
        d_disp, d_r, d_t = similar(disparity_src), similar(pose.rvec), similar(pose.tvec)
        nbytes = ccall((:mdb_render_workspace_bytes, libmdb200), Csize_t, (Int32, Int32, Int32, Int32), W, H, Np, B)
        ws = CuArray{UInt8}(undef, nbytes)
        Ki, Kk = collect(k9(K_inv)), collect(k9(K))
        opts = Ref(RenderOpts(mine_grid, mine_mask, depth_eps, 0))
        null = CuPtr{Cfloat}(0)
        xp, dxp = xyz_tgt === nothing ? (null, null) : (pointer(xyz_tgt), pointer(d_xyz))
        GC.@preserve rgb sigma disparity_src xyz_tgt pose depth zexp wsum g_rgb g_depth g_z d_rgb d_sigma d_xyz d_disp d_r d_t ws opts check(ccall(
            (:mdb_render_tgt_rgb_depth_bwd_ex, libmdb200), Int32,
            (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cfloat}, Ptr{Cfloat},
             Int32, Int32, Int32, Int32, Ptr{RenderOpts}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat},
             CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cvoid},
             Csize_t, Ptr{Cvoid}),
            handle(), rgb, sigma, disparity_src, xp, pose.rvec, pose.tvec, Ki, Kk, W, H, Np, B, opts, depth, zexp, wsum, g_rgb, g_depth,
            g_z, d_rgb, d_sigma, dxp, d_disp, d_r, d_t, ws, nbytes, stream_ptr()))
        NoTangent(), d_rgb, d_sigma, d_disp, (d_xyz === nothing ? NoTangent() : d_xyz), Tangent{Pose}(rvec=d_r, tvec=d_t),
        NoTangent(), NoTangent()
    end
    outs, render_ex_pullback
end

# ---- SURVEY 8(f) rank 1: slow_depth's optimiser loop on the device (src/simple_depth.jl:1-62) ---------
struct AdamParams          # mdb_adam_params
    eta::Cdouble
    beta1::Cdouble
    beta2::Cdouble
    eps::Cdouble
end

"""
    slow_depth(x, ssim, backprojections, projections, invKs, Ks, transfer; target_id, source_ids,
               min_depth, max_depth, log_dir)

Same signature as the reference.  `disp`, the two `Pose`s and the ADAM moments live on the device; one
iteration (loss forward+backward + `update!`) is a CUDA graph that `mdb_slow_depth_run` replays.  The
reference's logging at iteration 1, 5, 10, ... is kept by running the graph in chunks.
"""
function slow_depth(x::CuArray{Float32,5}, ssim, backprojections, projections, invKs, Ks, transfer;
                    target_id, source_ids, min_depth, max_depth, log_dir, iters::Integer=500, log_step::Integer=5)
    W, H, C, L, N = size(x)
    S = length(source_ids)
    disp = CUDA.fill(0.5f0, (W, H, 1, N))                                   # (1,W,H,1) in the reference: same bytes
    poses = [Pose(CuArray(repeat(Float32[0, 0, 0.01], 1, N)), CUDA.zeros(Float32, 3, 1, N)) for _ in 1:S]   # :9-13
    loss = CUDA.zeros(Float32, 1)
    dd, dr, dt = similar(disp), [similar(p.rvec) for p in poses], [similar(p.tvec) for p in poses]
    history = CUDA.zeros(Float32, iters)
    io = LossIO(pointer(x), W, H, C, L, N, target_id, S,
        pad(Int32.(source_ids), MAX_SOURCES, Int32(0)), pad(Int32.(source_ids .< target_id), MAX_SOURCES, Int32(0)),
        pad(map(p -> pointer(p.rvec), poses), MAX_SOURCES, nullp()), pad(map(p -> pointer(p.tvec), poses), MAX_SOURCES, nullp()),
        1, pad([pointer(disp)], MAX_SCALES, nullp()), pad(Int32[W], MAX_SCALES, Int32(0)), pad(Int32[H], MAX_SCALES, Int32(0)),
        pad(Cfloat[1], MAX_SCALES, 0f0), nullp(), k9(Ks), k9(invKs), pointer(loss),
        pad([pointer(dd)], MAX_SCALES, nullp()), pad(map(pointer, dr), MAX_SOURCES, nullp()),
        pad(map(pointer, dt), MAX_SOURCES, nullp()), nullp(), Int32(0))
    p = LossParams(min_depth, max_depth, 1e-3, 0.85f0, Cfloat(ssim.c1), Cfloat(ssim.c2), Int32(0),
        Int32(0), Int32(1), Int32(1), Int32(0), Int32(0))                   # un-normalised, weight 1  (:39)
    adam = AdamParams(3e-4, 0.9, 0.999, 1e-8)                               # ADAM(3e-4)              (:16)
    ws = CuArray{UInt8}(undef, ccall((:mdb_train_loss_workspace_bytes, libmdb200), Csize_t, (Ref{LossIO},), io))
    state = CUDA.zeros(UInt8, ccall((:mdb_slow_depth_state_bytes, libmdb200), Csize_t, (Ref{LossIO},), io))
    plan = Ref{Ptr{Cvoid}}(C_NULL)
    GC.@preserve x disp poses loss dd dr dt history ws state begin
        check(ccall((:mdb_slow_depth_create, libmdb200), Int32,
            (Ptr{Cvoid}, Ref{LossIO}, Ref{LossParams}, Ref{AdamParams}, CuPtr{Cvoid}, Csize_t, CuPtr{Cvoid}, Csize_t,
             CuPtr{Cfloat}, Int32, Ptr{Ptr{Cvoid}}),
            handle(), io, p, adam, state, sizeof(state), ws, sizeof(ws), history, Int32(iters), plan))
        done = 0
        while done < iters
            nxt = done == 0 ? 1 : min(iters, (done ÷ log_step + 1) * log_step)          # :23
            check(ccall((:mdb_slow_depth_run, libmdb200), Int32, (Ptr{Cvoid}, Int32, Ptr{Cvoid}),
                plan[], Int32(nxt - done), stream_ptr()))
            done = nxt
            save_disparity(reshape(cpu(disp), (W, H)), joinpath(log_dir, "d-$done.png"))   # :45-47
            println(done, " ", mean(disp))                                                  # :55
        end
        ccall((:mdb_slow_depth_destroy, libmdb200), Int32, (Ptr{Cvoid},), plan[])
    end
    disp, poses, Array(history)
end

# ---- datasets: ship the decoder's 8-bit raster, do the rest of `getindex` on the device -------------------------------
# Both take ROW-MAJOR rasters as the PNG decoder produces them: a Julia `Matrix` from `load` is (rows, cols) column-major,
# so pass `permutedims(rawview(channelview(img)), ...)` with the width axis first -- or decode straight into such a buffer.

"""
    depth10k_batch(strips::CuArray{UInt8,4}; flips = nothing, grayscale = false) -> CuArray{Float32,5} (W,H,C,3,N)

`Depth10k.getindex` (src/dtk.jl:29-46) for a batch: `strips` is (3, 3W, H, N) = RGB-interleaved rows of the triplet image.
"""
function depth10k_batch(strips::CuArray{UInt8,4}; flips::Union{Nothing,CuArray{UInt8,1}} = nothing, grayscale::Bool = false)
    _, W3, H, N = size(strips)
    W = W3 ÷ 3
    x = CuArray{Float32}(undef, W, H, grayscale ? 1 : 3, 3, N)
    GC.@preserve strips flips x check(ccall((:mdb_ingest_triplet_u8, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{UInt8}, CuPtr{UInt8}, Int32, Int32, Int32, Int32, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), strips, flips === nothing ? CuPtr{UInt8}(0) : pointer(flips), W, H, N, Int32(grayscale), x, stream_ptr()))
    x
end

"""
    depth10k_batch_u8(strips::CuArray{UInt8,4}; flips = nothing) -> CuArray{UInt8,5} (W,H,3,3,N)

The same split / FlipX / transpose without the conversion: the N0f8 stack `fused_loss` reads directly (a quarter of the bytes).
"""
function depth10k_batch_u8(strips::CuArray{UInt8,4}; flips::Union{Nothing,CuArray{UInt8,1}} = nothing)
    _, W3, H, N = size(strips)
    W = W3 ÷ 3
    x = CuArray{UInt8}(undef, W, H, 3, 3, N)
    GC.@preserve strips flips x check(ccall((:mdb_ingest_triplet_u8_planar, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{UInt8}, CuPtr{UInt8}, Int32, Int32, Int32, CuPtr{UInt8}, Ptr{Cvoid}),
        handle(), strips, flips === nothing ? CuPtr{UInt8}(0) : pointer(flips), W, H, N, x, stream_ptr()))
    x
end

"""
    kitti_batch(frames::CuArray{UInt8,4}, width, height; flips = nothing) -> CuArray{Float32,5} (W,H,1,3,N)

`KittyDataset.getindex` (src/kitty.jl:45-61) for a batch: `frames` is (W_orig, H_orig, 3, N), the three consecutive gray
`image_0` rasters of every sample; `imresize` to `d.resolution`, `Float32.`, transpose and stack happen on the device.
"""
function kitti_batch(frames::CuArray{UInt8,4}, width::Integer, height::Integer; flips::Union{Nothing,CuArray{UInt8,1}} = nothing)
    Wo, Ho, L, N = size(frames)
    L == 3 || error("kitti_batch: need three frames per sample")
    x = CuArray{Float32}(undef, width, height, 1, 3, N)
    GC.@preserve frames flips x check(ccall((:mdb_ingest_kitti_u8, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{UInt8}, CuPtr{UInt8}, Int32, Int32, Int32, Int32, Int32, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), frames, flips === nothing ? CuPtr{UInt8}(0) : pointer(flips), Wo, Ho, width, height, N, x, stream_ptr()))
    x
end

end # module
