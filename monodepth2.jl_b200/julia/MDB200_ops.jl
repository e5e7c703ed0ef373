# MDB200_ops.jl -- included by MDB200.jl: CuArray{Float32} methods of the fine-grained functions the reference's
# tests and drivers call by name (test/runtests.jl:19-21,36,74,87,110-111; test/test_sample.jl:66-72;
# test/test_plane_volume_rendering.jl:1; src/simple_depth.jl:30-39), each a `ccall` into libmdb200.so, with a
# `ChainRulesCore.rrule` wherever the C ABI exports a pullback.  NOT EXECUTED here (no Julia runtime, SURVEY F2).

const F32 = CuArray{Float32}
fptr(a) = a === nothing ? CuPtr{Cfloat}(0) : pointer(a)
zeros_or(Δ, like) = Δ isa AbstractZero ? CUDA.zeros(Float32, size(like)) : CuArray{Float32}(unthunk(Δ))

# ---- so3 ------------------------------------------------------------------------------------------------------------
function hat(rvec::CuArray{Float32,2})                                    # src/utils.jl:123-132
    N = size(rvec, 2)
    S = CUDA.zeros(Float32, 3, 3, N)
    GC.@preserve rvec S check(ccall((:mdb_hat, libmdb200), Int32, (Ptr{Cvoid}, CuPtr{Cfloat}, Int32, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), rvec, N, S, stream_ptr()))
    S
end

function ChainRulesCore.rrule(::typeof(hat), rvec::CuArray{Float32,2})   # src/utils.jl:134-145
    function hat_pullback(Δ)
        d = zeros_or(Δ, CUDA.zeros(Float32, 3, 3, size(rvec, 2)))
        g = similar(rvec)
        GC.@preserve d g check(ccall((:mdb_hat_pullback, libmdb200), Int32, (Ptr{Cvoid}, CuPtr{Cfloat}, Int32, CuPtr{Cfloat}, Ptr{Cvoid}),
            handle(), d, size(rvec, 2), g, stream_ptr()))
        NoTangent(), g
    end
    hat(rvec), hat_pullback
end

# ---- elementwise / geometry -------------------------------------------------------------------------------------------
function disparity_to_depth(disparity::F32, min_depth, max_depth)        # src/utils.jl:179-183
    out = similar(disparity)
    GC.@preserve disparity out check(ccall((:mdb_disparity_to_depth, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, Int64, Cdouble, Cdouble, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), disparity, length(disparity), Cdouble(min_depth), Cdouble(max_depth), out, stream_ptr()))
    out
end

function (b::Backproject)(depth::CuArray{Float32,3}, invK)                # src/utils.jl:67-69; depth (1,WH,N)
    WH, N = size(depth, 2), size(depth, 3)
    W = Int(maximum(Array(b.coordinates)[1, :])); H = WH ÷ W
    pts = CUDA.zeros(Float32, 3, WH, N)
    Ki = collect(k9(invK))
    GC.@preserve depth pts check(ccall((:mdb_backproject, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, Ptr{Cfloat}, Int32, Int32, Int32, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), depth, Ki, W, H, N, pts, stream_ptr()))
    pts
end

function (p::Project)(points::CuArray{Float32,3}, K, R::CuArray{Float32,3}, t::CuArray{Float32,3})   # src/utils.jl:99-103
    WH, N = size(points, 2), size(points, 3)
    nz = Array(p.normalizer); W = Int(nz[1]) + 1; H = Int(nz[2]) + 1      # normalizer = [W-1, H-1] (src/utils.jl:75-77)
    uv = CUDA.zeros(Float32, 2, WH, N)
    Kk = collect(k9(K))
    GC.@preserve points R t uv check(ccall((:mdb_project, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, Ptr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, Int32, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), points, Kk, R, t, W, H, N, uv, stream_ptr()))
    uv
end

# ---- warp (P13): missing in the fork, called at src/simple_depth.jl:30-32 ---------------------------------------------
"""
    warp(disp, x, Ps, backprojections, projections, invKs, Ks; min_depth, max_depth, source_ids)

`disp (1,W,H,N)`, `x (W,H,C,L,N)`, `Ps::Vector{(R (3,3,N), t (3,1,N))}` -> `Vector{(W,H,C,N)}`: the body the fork
inlined at src/training.jl:48-57.  `backprojections` / `projections` are accepted for signature compatibility.
"""
function warp(disp::CuArray{Float32,4}, x::CuArray{Float32,5}, Ps, backprojections, projections, invKs, Ks;
              min_depth, max_depth, source_ids)
    W, H, C, L, N = size(x)
    S = length(Ps)
    outs = [CUDA.zeros(Float32, W, H, C, N) for _ in 1:S]
    Rp = [pointer(P[1]) for P in Ps]; tp = [pointer(P[2]) for P in Ps]; op = map(pointer, outs)
    ids = Int32.(collect(source_ids))
    Kk, Ki = collect(k9(Ks)), collect(k9(invKs))
    GC.@preserve disp x Ps outs Rp tp op ids check(ccall((:mdb_warp, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{CuPtr{Cfloat}}, Ptr{CuPtr{Cfloat}}, Ptr{Int32}, Int32, Ptr{Cfloat}, Ptr{Cfloat},
         Int32, Int32, Int32, Int32, Int32, Cdouble, Cdouble, Ptr{CuPtr{Cfloat}}, Ptr{Cvoid}),
        handle(), disp, x, Rp, tp, ids, S, Kk, Ki, W, H, C, L, N, Cdouble(min_depth), Cdouble(max_depth), op, stream_ptr()))
    outs
end

function ChainRulesCore.rrule(::typeof(warp), disp::CuArray{Float32,4}, x::CuArray{Float32,5}, Ps, backprojections, projections,
                              invKs, Ks; min_depth, max_depth, source_ids)
    outs = warp(disp, x, Ps, backprojections, projections, invKs, Ks; min_depth, max_depth, source_ids)
    function warp_pullback(Δ)
        W, H, C, L, N = size(x)
        S = length(Ps)
        gs = [zeros_or(Δ[s], outs[s]) for s in 1:S]
        d_disp = similar(disp)
        dR = [similar(P[1]) for P in Ps]; dt = [similar(P[2]) for P in Ps]
        Rp = [pointer(P[1]) for P in Ps]; tp = [pointer(P[2]) for P in Ps]
        gp, dRp, dtp = map(pointer, gs), map(pointer, dR), map(pointer, dt)
        ids = Int32.(collect(source_ids))
        Kk, Ki = collect(k9(Ks)), collect(k9(invKs))
        GC.@preserve disp x Ps gs d_disp dR dt Rp tp gp dRp dtp ids check(ccall((:mdb_warp_bwd, libmdb200), Int32,
            (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{CuPtr{Cfloat}}, Ptr{CuPtr{Cfloat}}, Ptr{Int32}, Int32, Ptr{Cfloat}, Ptr{Cfloat},
             Int32, Int32, Int32, Int32, Int32, Cdouble, Cdouble, Ptr{CuPtr{Cfloat}}, CuPtr{Cfloat}, Ptr{CuPtr{Cfloat}},
             Ptr{CuPtr{Cfloat}}, Ptr{Cvoid}),
            handle(), disp, x, Rp, tp, ids, S, Kk, Ki, W, H, C, L, N, Cdouble(min_depth), Cdouble(max_depth), gp, d_disp, dRp, dtp,
            stream_ptr()))
        NoTangent(), d_disp, NoTangent(), [(dR[s], dt[s]) for s in 1:S], NoTangent(), NoTangent(), NoTangent(), NoTangent()
    end
    outs, warp_pullback
end

# ---- SSIM / photometric / min-reprojection / mask ---------------------------------------------------------------------
ssim_ws(W, H, C, N) = CuArray{UInt8}(undef, ccall((:mdb_ssim_bwd_workspace_bytes, libmdb200), Csize_t, (Int32, Int32, Int32, Int32), W, H, C, N))

function ChainRulesCore.rrule(ssim::SSIM, x::CuArray{Float32,4}, y::CuArray{Float32,4})     # pullback w.r.t. x
    out = ssim(x, y)
    function ssim_pullback(Δ)
        W, H, C, N = size(x)
        g = zeros_or(Δ, out); dx = similar(x); ws = ssim_ws(W, H, C, N)
        GC.@preserve x y g dx ws check(ccall((:mdb_ssim_bwd, libmdb200), Int32,
            (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, Int32, Int32, Cfloat, Cfloat, CuPtr{Cfloat},
             CuPtr{Cvoid}, Csize_t, Ptr{Cvoid}),
            handle(), x, y, g, W, H, C, N, Cfloat(ssim.c1), Cfloat(ssim.c2), dx, ws, sizeof(ws), stream_ptr()))
        NoTangent(), dx, NoTangent()
    end
    out, ssim_pullback
end

function photometric_loss(ssim::SSIM, predicted::CuArray{Float32,4}, target::CuArray{Float32,4}; α=0.85)   # src/training.jl:1-7
    W, H, C, N = size(predicted)
    out = CUDA.zeros(Float32, W, H, 1, N)
    GC.@preserve predicted target out check(ccall((:mdb_photometric_loss, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, Int32, Int32, Cfloat, Cfloat, Cfloat, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), predicted, target, W, H, C, N, Cfloat(α), Cfloat(ssim.c1), Cfloat(ssim.c2), out, stream_ptr()))
    out
end

function prediction_loss(ssim::SSIM, predictions::Vector{<:CuArray{Float32,4}}, target::CuArray{Float32,4})   # src/training.jl:13-15
    W, H, C, N = size(target)
    S = length(predictions)
    out = CUDA.zeros(Float32, W, H, 1, N)
    pp = map(pointer, predictions)
    GC.@preserve predictions target out pp check(ccall((:mdb_prediction_loss, libmdb200), Int32,
        (Ptr{Cvoid}, Ptr{CuPtr{Cfloat}}, Int32, CuPtr{Cfloat}, Int32, Int32, Int32, Int32, Cfloat, Cfloat, Cfloat, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), pp, S, target, W, H, C, N, 0.85f0, Cfloat(ssim.c1), Cfloat(ssim.c2), out, stream_ptr()))
    out
end

function ChainRulesCore.rrule(::typeof(prediction_loss), ssim::SSIM, predictions::Vector{<:CuArray{Float32,4}}, target::CuArray{Float32,4})
    out = prediction_loss(ssim, predictions, target)
    function prediction_loss_pullback(Δ)
        W, H, C, N = size(target)
        S = length(predictions)
        g = zeros_or(Δ, out)
        ds = [similar(p) for p in predictions]
        ws = ssim_ws(W, H, C, N)
        pp, dp = map(pointer, predictions), map(pointer, ds)
        GC.@preserve predictions target g ds ws pp dp check(ccall((:mdb_prediction_loss_bwd, libmdb200), Int32,
            (Ptr{Cvoid}, Ptr{CuPtr{Cfloat}}, Int32, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, Int32, Int32, Cfloat, Cfloat, Cfloat, Int32,
             Ptr{CuPtr{Cfloat}}, CuPtr{Cvoid}, Csize_t, Ptr{Cvoid}),
            handle(), pp, S, target, g, W, H, C, N, 0.85f0, Cfloat(ssim.c1), Cfloat(ssim.c2), Int32(0), dp, ws, sizeof(ws), stream_ptr()))
        NoTangent(), NoTangent(), ds, NoTangent()
    end
    out, prediction_loss_pullback
end

function _apply_mask(mask::CuArray{Float32,4}, warp_loss::CuArray{Float32,4})             # src/training.jl:17-19
    out = similar(warp_loss)
    GC.@preserve mask warp_loss out check(ccall((:mdb_apply_mask, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int64, CuPtr{Cfloat}, Ptr{Cvoid}), handle(), mask, warp_loss, length(out), out, stream_ptr()))
    out
end

function ChainRulesCore.rrule(::typeof(_apply_mask), mask::CuArray{Float32,4}, warp_loss::CuArray{Float32,4})
    out = _apply_mask(mask, warp_loss)
    function apply_mask_pullback(Δ)
        g = zeros_or(Δ, out); dm = similar(mask); dw = similar(warp_loss)
        GC.@preserve mask warp_loss g dm dw check(ccall((:mdb_apply_mask_bwd, libmdb200), Int32,
            (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int64, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cvoid}),
            handle(), mask, warp_loss, g, length(out), dm, dw, stream_ptr()))
        NoTangent(), dm, dw
    end
    out, apply_mask_pullback
end

# ---- smooth_loss (value and gradient come from one launch) -------------------------------------------------------------
function smooth_loss_fwdbwd(disparity::CuArray{Float32,3}, image::CuArray{Float32,4}; abs_subgradient::Integer=0)
    W, H, C, N = size(image)
    loss = CUDA.zeros(Float32, 1); grad = similar(disparity)
    ws = CuArray{UInt8}(undef, ccall((:mdb_smooth_loss_workspace_bytes, libmdb200), Csize_t, (Int32, Int32, Int32), W, H, N))
    GC.@preserve disparity image loss grad ws check(ccall((:mdb_smooth_loss_ex, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, Int32, Int32, Int32, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cvoid}, Csize_t, Ptr{Cvoid}),
        handle(), disparity, image, W, H, C, N, Int32(abs_subgradient), loss, grad, ws, sizeof(ws), stream_ptr()))
    Array(loss)[1], grad
end
smooth_loss(disparity::CuArray{Float32,3}, image::CuArray{Float32,4}) = smooth_loss_fwdbwd(disparity, image)[1]   # src/utils.jl:163-177
function ChainRulesCore.rrule(::typeof(smooth_loss), disparity::CuArray{Float32,3}, image::CuArray{Float32,4})
    l, grad = smooth_loss_fwdbwd(disparity, image)
    l, Δ -> (NoTangent(), Float32(unthunk(Δ)) .* grad, NoTangent())
end

# ---- render-side operators (src/render.jl) --------------------------------------------------------------------------------
function inv(A::CuArray{Float32})                                         # Monodepth.inv, src/render.jl:3-17: (3,3) or (3,3,M)
    out = similar(A)
    GC.@preserve A out check(ccall((:mdb_inv3x3_batched, libmdb200), Int32, (Ptr{Cvoid}, CuPtr{Cfloat}, Int32, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), A, length(A) ÷ 9, out, stream_ptr()))
    out
end

# the reference builds the meshgrid on the CPU (src/render.jl:21-23) and `transfer`s it; this variant builds it on the device
function create_meshgrid_gpu(H::Integer, W::Integer)
    out = CUDA.zeros(Float32, 3, W, H)
    GC.@preserve out check(ccall((:mdb_create_meshgrid, libmdb200), Int32, (Ptr{Cvoid}, Int32, Int32, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), H, W, out, stream_ptr()))
    out
end

function get_src_xyz_from_plane_disparity(meshgrid::CuArray{Float32,3}, disparity::CuArray{Float32,2}, K_inv)   # src/render.jl:25-30
    _, W, H = size(meshgrid); Np, B = size(disparity)
    out = CUDA.zeros(Float32, 3, W, H, Np, B)
    Ki = collect(k9(K_inv))
    GC.@preserve meshgrid disparity out check(ccall((:mdb_plane_xyz_src, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cfloat}, Int32, Int32, Int32, Int32, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), meshgrid, disparity, Ki, W, H, Np, B, out, stream_ptr()))
    out
end

function get_tgt_xyz_from_plane_disparity(xyz_src::CuArray{Float32,5}, pose::Pose)                             # src/render.jl:51-62
    _, W, H, Np, B = size(xyz_src)
    out = similar(xyz_src)
    GC.@preserve xyz_src pose out check(ccall((:mdb_plane_xyz_tgt, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, Int32, Int32, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), xyz_src, pose.rvec, pose.tvec, W, H, Np, B, out, stream_ptr()))
    out
end

render_ws(W, H, Np, B) = CuArray{UInt8}(undef, ccall((:mdb_render_workspace_bytes, libmdb200), Csize_t, (Int32, Int32, Int32, Int32), W, H, Np, B))

# sample(src (W,H,C,Np*B), depth_src (Np,B), pose, K, K_inv) -> tgt, valid_mask (W*H, Np*B)       src/render.jl:64-90
function sample(src::CuArray{Float32,4}, depth_src::CuArray{Float32,2}, pose::Pose, K, K_inv)
    W, H, Cs, _ = size(src); Np, B = size(depth_src)
    tgt = similar(src); valid = CUDA.zeros(UInt8, W * H, Np * B)
    ws = render_ws(W, H, Np, B)
    Kk, Ki = collect(k9(K)), collect(k9(K_inv))
    GC.@preserve src depth_src pose tgt valid ws check(ccall((:mdb_sample, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cfloat}, Ptr{Cfloat}, Int32, Int32, Int32, Int32, Int32,
         CuPtr{Cfloat}, CuPtr{UInt8}, CuPtr{Cvoid}, Csize_t, Ptr{Cvoid}),
        handle(), src, depth_src, pose.rvec, pose.tvec, Kk, Ki, W, H, Cs, Np, B, tgt, valid, ws, sizeof(ws), stream_ptr()))
    tgt, valid
end

function ChainRulesCore.rrule(::typeof(sample), src::CuArray{Float32,4}, depth_src::CuArray{Float32,2}, pose::Pose, K, K_inv)
    tgt, valid = sample(src, depth_src, pose, K, K_inv)
    function sample_pullback(Δ)
        W, H, Cs, _ = size(src); Np, B = size(depth_src)
        g = zeros_or(Δ[1], tgt); d_src = similar(src); ws = render_ws(W, H, Np, B)
        Kk, Ki = collect(k9(K)), collect(k9(K_inv))
        GC.@preserve g depth_src pose d_src ws check(ccall((:mdb_sample_bwd, libmdb200), Int32,
            (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cfloat}, Ptr{Cfloat}, Int32, Int32, Int32, Int32, Int32,
             CuPtr{Cfloat}, CuPtr{Cvoid}, Csize_t, Ptr{Cvoid}),
            handle(), g, depth_src, pose.rvec, pose.tvec, Kk, Ki, W, H, Cs, Np, B, d_src, ws, sizeof(ws), stream_ptr()))
        NoTangent(), d_src, NoTangent(), NoTangent(), NoTangent(), NoTangent()
    end
    (tgt, valid), sample_pullback
end

# plane_volume_rendering(rgb (W,H,3,Np,B), sigma (W,H,1,Np,B), xyz (3,W,H,Np,B)) -> rgb_out, transparency_acc, weights   :32-49
function plane_volume_rendering(rgb::CuArray{Float32,5}, sigma::CuArray{Float32,5}, xyz::CuArray{Float32,5})
    W, H, _, Np, B = size(rgb)
    rgb_out = CUDA.zeros(Float32, W, H, 3, B); tacc = similar(sigma); wts = similar(sigma)
    GC.@preserve rgb sigma xyz rgb_out tacc wts check(ccall((:mdb_plane_volume_rendering, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, Int32, Int32, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), rgb, sigma, xyz, W, H, Np, B, rgb_out, tacc, wts, stream_ptr()))
    rgb_out, tacc, wts
end

function ChainRulesCore.rrule(::typeof(plane_volume_rendering), rgb::CuArray{Float32,5}, sigma::CuArray{Float32,5}, xyz::CuArray{Float32,5})
    outs = plane_volume_rendering(rgb, sigma, xyz)
    function pvr_pullback(Δ)
        W, H, _, Np, B = size(rgb)
        g = [Δ[i] isa AbstractZero ? nothing : CuArray{Float32}(unthunk(Δ[i])) for i in 1:3]
        d_rgb, d_sigma, d_xyz = similar(rgb), similar(sigma), similar(xyz)
        GC.@preserve rgb sigma xyz g d_rgb d_sigma d_xyz check(ccall((:mdb_plane_volume_rendering_bwd, libmdb200), Int32,
            (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, Int32, Int32, CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat},
             CuPtr{Cfloat}, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cvoid}),
            handle(), rgb, sigma, xyz, W, H, Np, B, fptr(g[1]), fptr(g[2]), fptr(g[3]), d_rgb, d_sigma, d_xyz, stream_ptr()))
        NoTangent(), d_rgb, d_sigma, d_xyz
    end
    outs, pvr_pullback
end

# ---- SURVEY 8(e): the exchange step of a batch sharded over the GPUs of one box -------------------------------------------
# One Julia process (or task) per GPU; the launcher (MPI.jl, Distributed, ...) carries the 128-byte NCCL id or the 64-byte
# CUDA IPC handles between the ranks, the library does the rest.
packed_size(n_global::Integer, S::Integer) = ccall((:mdb_packed_size, libmdb200), Int64, (Int32, Int32), n_global, S)
nccl_unique_id() = (id = zeros(UInt8, 128); check(ccall((:mdb_comm_get_unique_id, libmdb200), Int32, (Ptr{UInt8}, Csize_t), id, 128)); id)
comm_init(id::Vector{UInt8}, rank::Integer, world::Integer) =
    check(ccall((:mdb_comm_init, libmdb200), Int32, (Ptr{Cvoid}, Ptr{UInt8}, Int32, Int32), handle(), id, rank, world))
peers_export(world::Integer, max_count::Integer) = (hd = zeros(UInt8, 64);
    check(ccall((:mdb_comm_peers_export, libmdb200), Int32, (Ptr{Cvoid}, Int32, Int64, Ptr{UInt8}, Csize_t), handle(), world, max_count, hd, 64)); hd)
peers_attach(handles::Vector{UInt8}, rank::Integer) =
    check(ccall((:mdb_comm_peers_attach, libmdb200), Int32, (Ptr{Cvoid}, Ptr{UInt8}, Int32), handle(), handles, rank))
allreduce_loss_posegrad!(recv::CuArray{Float32,1}, send::CuArray{Float32,1}) = (GC.@preserve send recv check(ccall(
    (:mdb_allreduce_loss_posegrad, libmdb200), Int32, (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int64, Ptr{Cvoid}),
    handle(), send, recv, length(send), stream_ptr())); recv)

# ---- SURVEY 8(f) rank 4: the operators either side of the loss, on the network's side ----------------------------------------
"""
    pose_head(conv_out; scale = 1f-2) -> Pose

The tail of `(decoder::PoseDecoder)(features)` (src/pose_decoder.jl:29-31): `conv_out (w,h,6,N)` is `decoder.pose(squeezed)`;
spatial mean, `× 1e-2`, split into `Pose(rvec (3,N), tvec (3,1,N))`.  A maintainer replaces those three lines with
`pose_head(decoder.pose(squeezed))`; the rrule below carries the loss path's pose gradients back to the conv output.
"""
function pose_head(conv_out::CuArray{Float32,4}; scale::Real = 1f-2)
    w, h, c, N = size(conv_out)
    c == 6 || error("pose_head: expected 6 channels, got $c")
    rvec = CUDA.zeros(Float32, 3, N); tvec = CUDA.zeros(Float32, 3, 1, N)
    GC.@preserve conv_out rvec tvec check(ccall((:mdb_pose_head, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, Int32, Int32, Int32, Cfloat, CuPtr{Cfloat}, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), conv_out, w, h, N, Cfloat(scale), rvec, tvec, stream_ptr()))
    Pose(rvec, tvec)
end

function ChainRulesCore.rrule(::typeof(pose_head), conv_out::CuArray{Float32,4}; scale::Real = 1f-2)
    w, h, _, N = size(conv_out)
    function pose_head_pullback(Δ)
        Δ = unthunk(Δ)
        d_r = zeros_or(Δ.rvec, CUDA.zeros(Float32, 3, N)); d_t = zeros_or(Δ.tvec, CUDA.zeros(Float32, 3, 1, N))
        d_conv = similar(conv_out)
        GC.@preserve d_r d_t d_conv check(ccall((:mdb_pose_head_bwd, libmdb200), Int32,
            (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, Int32, Cfloat, CuPtr{Cfloat}, Ptr{Cvoid}),
            handle(), d_r, d_t, w, h, N, Cfloat(scale), d_conv, stream_ptr()))
        NoTangent(), d_conv
    end
    pose_head(conv_out; scale), pose_head_pullback
end

"""
    disparity_heads(logits::Vector) -> Vector

The `σ` that ends every scale level of the DepthDecoder (`Conv((3,3), ch => 1, σ)`, src/depth_decoder.jl:46), for all levels
in one launch: `logits[k] (w_k,h_k,1,N)` are the convolutions' pre-activations (build the decoder's last `Conv` with
`identity`), the result is the `disparities` vector `train_loss` consumes (src/training.jl:27,41-46).
"""
function disparity_heads(logits::Vector{<:CuArray{Float32,4}})
    Ls = length(logits); N = size(logits[1], 4)
    outs = [similar(l) for l in logits]
    dw = Int32[size(l, 1) for l in logits]; dh = Int32[size(l, 2) for l in logits]
    lp = map(pointer, logits); op = map(pointer, outs)
    GC.@preserve logits outs lp op dw dh check(ccall((:mdb_disparity_heads, libmdb200), Int32,
        (Ptr{Cvoid}, Ptr{CuPtr{Cfloat}}, Ptr{Int32}, Ptr{Int32}, Int32, Int32, Ptr{CuPtr{Cfloat}}, Ptr{Cvoid}),
        handle(), lp, dw, dh, Ls, N, op, stream_ptr()))
    outs
end

function ChainRulesCore.rrule(::typeof(disparity_heads), logits::Vector{<:CuArray{Float32,4}})
    disp = disparity_heads(logits)
    function disparity_heads_pullback(Δ)
        Δ = unthunk(Δ)
        Ls = length(disp); N = size(disp[1], 4)
        g = [zeros_or(Δ[k], disp[k]) for k in 1:Ls]
        d_logits = [similar(d) for d in disp]
        dw = Int32[size(d, 1) for d in disp]; dh = Int32[size(d, 2) for d in disp]
        dp = map(pointer, disp); gp = map(pointer, g); lp = map(pointer, d_logits)
        GC.@preserve disp g d_logits dp gp lp dw dh check(ccall((:mdb_disparity_heads_bwd, libmdb200), Int32,
            (Ptr{Cvoid}, Ptr{CuPtr{Cfloat}}, Ptr{CuPtr{Cfloat}}, Ptr{Int32}, Ptr{Int32}, Int32, Int32, Ptr{CuPtr{Cfloat}}, Ptr{Cvoid}),
            handle(), dp, gp, dw, dh, Ls, N, lp, stream_ptr()))
        NoTangent(), d_logits
    end
    disp, disparity_heads_pullback
end

"""
    item_mean(map) -> Vector{Float32} (host), one mean per item (the last dimension), summed in a fixed order.
"""
function item_mean(map::CuArray{Float32})
    N = size(map, ndims(map)); per = length(map) ÷ N
    out = CUDA.zeros(Float32, N)
    GC.@preserve map out check(ccall((:mdb_item_mean, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, Int64, Int32, CuPtr{Cfloat}, Ptr{Cvoid}), handle(), map, Int64(per), N, out, stream_ptr()))
    Array(out)
end

"""
    static_scores(ssim, x, target_id; source_ids) -> Vector{Float32}

`find_static`'s statistic (src/dtk.jl:51-69) for a whole batch at once: `mean(automasking_loss(ssim, x, target; source_ids))`
per item of `x (W,H,C,L,N)`; the dataset filter keeps the files whose score exceeds `α`.
"""
function static_scores(ssim::SSIM, x::CuArray{Float32,5}, target_id::Integer; source_ids)
    item_mean(automasking_loss(ssim, x, x[:, :, :, target_id, :]; source_ids))
end

"""
    grid_sample_border(input (W,H,C,N), grid (2,Wo,Ho,N)) -> (Wo,Ho,C,N)

`NNlib.grid_sample(input, grid; padding_mode=:border)` with NNlib's align-corners convention, as both call sites use it
(src/training.jl:56, src/render.jl:88).  Not a method of `NNlib.grid_sample`: the glue does not pirate NNlib's function.
"""
function grid_sample_border(input::CuArray{Float32,4}, grid::CuArray{Float32,4})
    W, H, C, N = size(input); _, Wo, Ho, _ = size(grid)
    out = CUDA.zeros(Float32, Wo, Ho, C, N)
    GC.@preserve input grid out check(ccall((:mdb_grid_sample, libmdb200), Int32,
        (Ptr{Cvoid}, CuPtr{Cfloat}, CuPtr{Cfloat}, Int32, Int32, Int32, Int32, Int32, Int32, CuPtr{Cfloat}, Ptr{Cvoid}),
        handle(), input, grid, W, H, C, N, Wo, Ho, out, stream_ptr()))
    out
end

"release the library's per-process state (events, workspaces, the communicator); the next call creates it again"
function destroy()
    HANDLE[] == C_NULL && return nothing
    ccall((:mdb_destroy, libmdb200), Int32, (Ptr{Cvoid},), HANDLE[])
    HANDLE[] = C_NULL
    nothing
end

"1 if a one-shot peer reduce gave up waiting for a peer (synchronises the device)"
comm_status() = (f = Ref{Int32}(0); check(ccall((:mdb_comm_status, libmdb200), Int32, (Ptr{Cvoid}, Ref{Int32}), handle(), f)); f[])
comm_destroy() = check(ccall((:mdb_comm_destroy, libmdb200), Int32, (Ptr{Cvoid},), handle()))
version() = ccall((:mdb_version, libmdb200), Int32, ())
"kernels launched through the library's handle since it was created"
launch_count() = ccall((:mdb_launch_count, libmdb200), Int64, (Ptr{Cvoid},), handle())
"test hooks of include/mdb200.h (`strip_segs`, `strip_taper`, `render_precise`, ...); 0 restores the library's choice"
debug_option(name::AbstractString, value::Integer) = check(ccall((:mdb_debug_option, libmdb200), Int32, (Cstring, Int64), name, value))
